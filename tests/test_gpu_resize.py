"""
gance_resize_cubic_u8 (csrc/resize_kernels.cu) through the C-ABI against the oracle / cv2: bit-exact uint8 frames.
Reference step: gance/image_sources/video_common.py:399-455 (resize_source, scale_square_source_duplicate).
"""

import numpy as np
import pytest
import torch

from oracle import resize_ref

pytestmark = pytest.mark.gpu

CASES = [(64, 64, 135, 135, 2), (32, 48, 37, 37, 1), (100, 100, 37, 41, 3), (256, 256, 540, 540, 2), (17, 23, 100, 61, 1),
         (64, 64, 64, 64, 1), (256, 256, 128, 128, 2), (512, 512, 2160, 2160, 1)]


@pytest.fixture(scope="module")
def resizer():
    from gance_b200.image_sources.video_common import FrameResizer

    return FrameResizer(torch.device("cuda", 0))


@pytest.mark.parametrize("sh,sw,dh,dw,batch", CASES)
def test_resize_is_bit_exact_with_the_oracle(resizer, sh, sw, dh, dw, batch):
    images = np.random.RandomState(sh + dw).randint(0, 256, (batch, sh, sw, 3), dtype=np.uint8)
    got = resizer.resize(torch.as_tensor(images).cuda(), dh, dw).cpu().numpy()
    for b in range(batch):
        assert np.array_equal(got[b], resize_ref.resize_cubic_u8(images[b], dh, dw)), f"frame {b}"


def test_full_size_batch_against_opencv_own_path(resizer):
    """BASELINE config 5 geometry, 1024^2 -> 2160^2, checked against cv2 itself (IPP dispatch off) on the box."""
    cv2 = pytest.importorskip("cv2")
    before = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)
    try:
        images = np.random.RandomState(11).randint(0, 256, (3, 1024, 1024, 3), dtype=np.uint8)
        images[1] = np.linspace(0, 255, 1024, dtype=np.uint8)[None, :, None]  # smooth ramp
        got = resizer.resize(torch.as_tensor(images).cuda(), 2160, 2160).cpu().numpy()
        for b in range(3):
            assert np.array_equal(got[b], cv2.resize(images[b], (2160, 2160), interpolation=cv2.INTER_CUBIC))
    finally:
        cv2.ipp.setUseIPP(before)


def test_constant_frames_stay_constant_and_checksum_is_batch_invariant(resizer):
    """Size-independent properties at full size: constants are preserved; a frame's result does not depend on its batch slot."""
    flat = torch.full((2, 1024, 1024, 3), 77, dtype=torch.uint8, device="cuda")
    assert bool((resizer.resize(flat, 2160, 2160) == 77).all())
    images = torch.randint(0, 256, (4, 1024, 1024, 3), dtype=torch.uint8, device="cuda")
    together = resizer.resize(images, 2160, 2160)
    alone = resizer.resize(images[2:3], 2160, 2160)
    assert torch.equal(together[2], alone[0])


def test_resize_source_keeps_order_and_passes_through_matching_frames():
    from gance_b200.image_sources import video_common
    from gance_b200.image_sources.image_sources_common import ImageResolution

    rng = np.random.RandomState(2)
    frames = [rng.randint(0, 256, (48, 48, 3), dtype=np.uint8) for _ in range(5)]
    frames.insert(2, rng.randint(0, 256, (96, 96, 3), dtype=np.uint8))  # already at the target resolution
    out = list(video_common.resize_source(iter(frames), ImageResolution(96, 96), batch=4))
    assert len(out) == 6 and out[2] is frames[2]
    for got, image in zip(out, frames):
        if image.shape[0] != 96:
            assert np.array_equal(got, resize_ref.resize_cubic_u8(image, 96, 96))
    doubled = list(video_common.scale_square_source_duplicate(iter(frames[:2]), 96, frame_multiplier=3))
    assert len(doubled) == 6 and all(np.array_equal(doubled[i], doubled[0]) for i in range(3))


def test_argument_validation():
    from gance_b200 import _lib

    lib = _lib.init(0)
    x = torch.zeros(16, dtype=torch.uint8, device="cuda")
    assert lib.gance_resize_cubic_u8(None, 1, 4, 4, x.data_ptr(), 4, 4, x.data_ptr(), x.data_ptr(), x.data_ptr(),
                                     x.data_ptr(), 4, 4, None) == 1
    assert lib.gance_resize_cubic_u8(x.data_ptr(), 1, 4, 4, x.data_ptr(), 0, 4, x.data_ptr(), x.data_ptr(), x.data_ptr(),
                                     x.data_ptr(), 4, 4, None) == 2


def test_renderer_scales_on_the_device(tmp_path):
    """MultiNetwork.indexed_render_batches(output_side_length=S): synthesis -> bicubic resize on the device -> host,
    identical to resizing the un-scaled frames of the same call with the oracle."""
    from gance_b200.network_interface import network_functions
    from gance_b200.networks import container

    path = tmp_path / "net64.pkl"
    container.write_synthetic_network(path, 64, seed=3, perturbed=True)
    wplus = np.random.RandomState(4).randn(2, 3, 12, 512).astype(np.float32)  # 2 batches of 3 matrices (const noise)
    with network_functions.MultiNetwork([path]) as networks:
        plain = [frames.copy() for frames in networks.indexed_render_batches(0, iter(wplus), max_batch=4)]
        scaled = [frames.copy() for frames in networks.indexed_render_batches(0, iter(wplus), max_batch=4, output_side_length=135)]
    assert len(plain) == len(scaled) == 2
    for small, big in zip(plain, scaled):
        assert small.shape == (3, 64, 64, 3) and big.shape == (3, 135, 135, 3)
        for b in range(3):
            assert np.array_equal(big[b], resize_ref.resize_cubic_u8(small[b], 135, 135))
