"""
The resize oracle (oracle/resize_ref.py) against cv2 itself - the reference's own implementation of the step
(gance/image_sources/video_common.py:399-429 calls cv2.resize(..., INTER_CUBIC)) - and the host-side coefficient
tables of the CUDA path against the oracle's. No GPU needed.
"""

import numpy as np
import pytest

from gance_b200.image_sources import video_common
from oracle import resize_ref

cv2 = pytest.importorskip("cv2")

CASES = [(64, 64, 135, 135), (32, 48, 37, 37), (100, 100, 37, 41), (256, 256, 540, 540), (17, 23, 100, 61),
         (64, 64, 64, 64), (50, 50, 51, 49), (256, 256, 128, 128)]


@pytest.fixture()
def opencv_own_path():
    """cv2 with Intel IPP dispatch off: OpenCV's own resize.cpp code, the algorithm the oracle restates."""
    before = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)
    yield
    cv2.ipp.setUseIPP(before)


@pytest.mark.parametrize("sh,sw,dh,dw", CASES)
def test_restatement_is_bit_exact_with_opencv(opencv_own_path, sh, sw, dh, dw):
    image = np.random.RandomState(sh * 1000 + dw).randint(0, 256, (sh, sw, 3), dtype=np.uint8)
    want = cv2.resize(image, (dw, dh), interpolation=cv2.INTER_CUBIC)
    got = resize_ref.resize_cubic_u8(image, dh, dw)
    assert got.shape == want.shape == (dh, dw, 3)
    assert np.array_equal(got, want)


def test_full_size_frame_is_bit_exact_with_opencv(opencv_own_path):
    """BASELINE config 5: 1024^2 -> 2160^2 (6480 elements per row = 810 vector groups, no integer tail)."""
    yy, xx = np.mgrid[0:1024, 0:1024]
    image = np.stack([127.5 + 127.5 * np.sin(xx / 37.0 + yy / 91.0), 127.5 + 127.5 * np.cos(xx / 13.0 - yy / 57.0),
                      ((xx + yy) / 8) % 256], axis=-1).astype(np.uint8)
    image[::7, ::5] = np.random.RandomState(3).randint(0, 256, image[::7, ::5].shape, dtype=np.uint8)
    want = cv2.resize(image, (2160, 2160), interpolation=cv2.INTER_CUBIC)
    assert np.array_equal(resize_ref.resize_cubic_u8(image, 2160, 2160), want)


def test_ipp_dispatch_differs_by_at_most_one_lsb():
    """With IPP on (this wheel's default) cv2 evaluates the same kernel in floating point: +-1 LSB on a few % of pixels."""
    if not cv2.ipp.useIPP():
        pytest.skip("this cv2 build has no IPP dispatch")
    image = np.random.RandomState(5).randint(0, 256, (128, 128, 3), dtype=np.uint8)
    ipp = cv2.resize(image, (270, 270), interpolation=cv2.INTER_CUBIC).astype(np.int32)
    own = resize_ref.resize_cubic_u8(image, 270, 270).astype(np.int32)
    assert np.abs(ipp - own).max() <= 1
    assert (ipp != own).mean() < 0.08


def test_saturation_and_constant_images():
    """Overshoot of the negative lobes saturates to [0,255]; a constant image stays constant (the rounded coefficients of
    a tap set sum to 2048 +- 1, as in OpenCV, which never moves a uint8 constant after the >> 22 rounding)."""
    checker = (np.indices((32, 32)).sum(axis=0) % 2 * 255).astype(np.uint8)[..., None].repeat(3, axis=2)
    out = resize_ref.resize_cubic_u8(checker, 67, 67)
    assert out.min() == 0 and out.max() == 255
    flat = np.full((16, 16, 3), 201, dtype=np.uint8)
    assert (resize_ref.resize_cubic_u8(flat, 40, 40) == 201).all()
    _, coef = resize_ref.cubic_tables(1024, 2160)
    assert (np.abs(coef.astype(np.int32).sum(axis=1) - 2048) <= 1).all()


@pytest.mark.parametrize("src,dst", [(1024, 2160), (512, 2160), (64, 135), (100, 37)])
def test_host_tables_equal_the_oracle_tables(src, dst):
    ofs, coef = video_common.cubic_tables(src, dst)
    ref_ofs, ref_coef = resize_ref.cubic_tables(src, dst)
    assert np.array_equal(ofs, ref_ofs) and np.array_equal(coef, ref_coef)
    assert ofs.dtype == np.int32 and coef.dtype == np.int16
    window = video_common._window(ofs, video_common.TILE_W)
    assert 4 <= window <= int(np.ceil(video_common.TILE_W * src / dst)) + 5


def test_rejects_non_uint8():
    with pytest.raises(ValueError):
        resize_ref.resize_cubic_u8(np.zeros((4, 4, 3), dtype=np.float32), 8, 8)
