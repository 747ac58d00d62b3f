"""
TEST INFRASTRUCTURE (oracle): CPU restatement of the frame resize on the path's output side.

The reference resizes every synthesized frame with `cv2.resize(image, (side, side), interpolation=cv2.INTER_CUBIC)`
(/root/reference/gance/image_sources/video_common.py:399-429, called from `scale_square_source_duplicate`, 432-455, by
gance/projection_file_blend.py:184-199). The arithmetic lives in a third-party dependency that is not under
/root/reference: **opencv-python (unpinned, requirements/prod.txt:17; 4.13.0 in this image)**, file
`modules/imgproc/src/resize.cpp`. Its published algorithm for 8-bit images, restated here:

* `resizeGeneric_` set-up: `scale = 1 / (dst / src)` (double); for each destination index d:
  `f = float((d + 0.5) * scale - 0.5)`, `s = floor(f)`, `f -= s`; `interpolateCubic(f)` with A = -0.75 in float:
  `c0 = ((A(f+1) - 5A)(f+1) + 8A)(f+1) - 4A`, `c1 = ((A+2)f - (A+3))f^2 + 1`, `c2 = ((A+2)(1-f) - (A+3))(1-f)^2 + 1`,
  `c3 = 1 - c0 - c1 - c2`; fixed point: `saturate_cast<short>(c * 2048)` (round half to even). Taps are source
  indices s-1 .. s+2, clamped to the image (replicated border).
* `HResizeCubic<uchar,int,short>`: integer row pass, `row[d] = sum_k src[clamp(s-1+k)] * alpha[d][k]` (int32).
* `VResizeCubic<uchar,int,short,FixedPtCast<int,uchar,22>, VResizeCubicVec_32s8u>`: the vector body handles the first
  `floor(n / 8) * 8` elements of a destination row (n = width * channels; 8 = lanes of the baseline-SSE `v_uint16x8`)
  in FLOAT: `b_k = beta_k * (1 / 2^22)`, `x = S3*b3; x = S2*b2 + x; x = S1*b1 + x; x = S0*b0 + x` (separately rounded
  multiply and add: the baseline build has no FMA), `v_round` (half to even), saturating packs to uint8. The scalar
  tail uses integers: `saturate_cast<uchar>((S0*b0 + S1*b1 + S2*b2 + S3*b3 + 2^21) >> 22)`.

Pinned: `tests/test_oracle_resize.py` checks this file against cv2 itself (run here and on the GPU box: cv2 ships in
the image) with Intel IPP dispatch switched off (`cv2.ipp.setUseIPP(False)`), i.e. against OpenCV's own code above,
bit for bit. With IPP on (this wheel's default) cv2 calls the closed-source `ippiResizeCubic` instead, which evaluates
the same kernel in floating point without the 11-bit coefficient quantisation: it differs from OpenCV's own result by
exactly 1 LSB on ~4.5 % of the pixels of a random image (measured, same test) and is not reproducible bit for bit.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may import this module.
"""

from typing import Tuple

import numpy as np

INTER_RESIZE_COEF_BITS = 11
INTER_RESIZE_COEF_SCALE = 1 << INTER_RESIZE_COEF_BITS
VECTOR_LANES = 8  # v_uint16x8 of the baseline (SSE) universal-intrinsics build


def cubic_tables(src: int, dst: int) -> Tuple[np.ndarray, np.ndarray]:
    """(offsets [dst] int32 = index of tap 1, coefficients [dst,4] int16) of one axis (resize.cpp, resizeGeneric_ set-up)."""
    scale = 1.0 / (dst / src)
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f)
    f = (f - s).astype(np.float32)
    a = np.float32(-0.75)
    one, two, three = np.float32(1), np.float32(2), np.float32(3)
    c0 = ((a * (f + one) - np.float32(5) * a) * (f + one) + np.float32(8) * a) * (f + one) - np.float32(4) * a
    c1 = ((a + two) * f - (a + three)) * f * f + one
    c2 = ((a + two) * (one - f) - (a + three)) * (one - f) * (one - f) + one
    c3 = one - c0 - c1 - c2
    coef = np.stack([c0, c1, c2, c3], axis=1).astype(np.float32) * np.float32(INTER_RESIZE_COEF_SCALE)
    return s.astype(np.int32), np.clip(np.rint(coef), -32768, 32767).astype(np.int16)


def resize_cubic_u8(image: np.ndarray, dst_height: int, dst_width: int) -> np.ndarray:
    """cv2.resize(image, (dst_width, dst_height), interpolation=cv2.INTER_CUBIC) of an (H,W,C) uint8 image, OpenCV's own path."""
    if image.dtype != np.uint8 or image.ndim != 3:
        raise ValueError("expected an (H,W,C) uint8 image")
    sh, sw, cn = image.shape
    xofs, alpha = cubic_tables(sw, dst_width)
    yofs, beta = cubic_tables(sh, dst_height)
    xi = np.clip(xofs[:, None] + np.arange(-1, 3)[None, :], 0, sw - 1)
    rows = (image.astype(np.int32)[:, xi, :] * alpha.astype(np.int32)[None, :, :, None]).sum(axis=2, dtype=np.int32)
    rows = rows.reshape(sh, dst_width * cn)  # one destination row of the horizontal pass = dst_width * cn int32
    yi = np.clip(yofs[:, None] + np.arange(-1, 3)[None, :], 0, sh - 1)
    taps = [rows[yi[:, k]] for k in range(4)]  # [dst_height, n] int32 each
    n = dst_width * cn
    body = (n // VECTOR_LANES) * VECTOR_LANES
    out = np.empty((dst_height, n), dtype=np.uint8)
    b = beta.astype(np.float32) * np.float32(1.0 / (INTER_RESIZE_COEF_SCALE * INTER_RESIZE_COEF_SCALE))
    x = taps[3][:, :body].astype(np.float32) * b[:, 3, None]
    for k in (2, 1, 0):
        x = taps[k][:, :body].astype(np.float32) * b[:, k, None] + x
    out[:, :body] = np.clip(np.rint(x), 0, 255).astype(np.uint8)
    if body < n:
        acc = np.zeros((dst_height, n - body), dtype=np.int32)
        for k in range(4):
            acc = acc + taps[k][:, body:] * beta.astype(np.int32)[:, k, None]
        out[:, body:] = np.clip((acc + (1 << 21)) >> 22, 0, 255).astype(np.uint8)
    return out.reshape(dst_height, dst_width, cn)
