"""
Output-side frame scaling of the hot path: `resize_source` / `scale_square_source_duplicate` with the reference's
signatures (gance/image_sources/video_common.py:399-455), backed by `gance_resize_cubic_u8` (csrc/resize_kernels.cu),
plus the batched device twin the renderers use (`FrameResizer`). Video decode / encode / audio mux of the same
reference file are out of scope (ffmpeg / vidgear I/O).

The kernel reproduces OpenCV's own 8-bit INTER_CUBIC arithmetic bit for bit (oracle/resize_ref.py states the algorithm
and what it is pinned against); the fixed-point coefficient tables are built here, on the host, once per geometry.
"""

import itertools
from typing import Dict, Iterator, Tuple

import numpy as np
import torch

from gance_b200 import _lib
from gance_b200.image_sources.image_sources_common import ImageResolution, image_resolution

INTER_RESIZE_COEF_SCALE = 2048  # OpenCV INTER_RESIZE_COEF_BITS = 11
TILE_W, TILE_H = 128, 16  # destination tile of resize_cubic_u8_kernel


def cubic_tables(src: int, dst: int) -> Tuple[np.ndarray, np.ndarray]:
    """
    Per-axis tables of OpenCV's 8-bit bicubic resize (resize.cpp, resizeGeneric_ set-up + interpolateCubic, A = -0.75):
    offsets [dst] int32 (source index of tap 1; taps are offset-1 .. offset+2, clamped by the kernel) and coefficients
    [dst,4] int16 = round-half-even(c * 2048), everything evaluated in float32 from a double-precision coordinate.
    """
    scale = 1.0 / (dst / src)
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f)
    f = (f - s).astype(np.float32)
    a, one = np.float32(-0.75), np.float32(1)
    c0 = ((a * (f + one) - np.float32(5) * a) * (f + one) + np.float32(8) * a) * (f + one) - np.float32(4) * a
    c1 = ((a + np.float32(2)) * f - (a + np.float32(3))) * f * f + one
    c2 = ((a + np.float32(2)) * (one - f) - (a + np.float32(3))) * (one - f) * (one - f) + one
    c3 = one - c0 - c1 - c2
    coef = np.stack([c0, c1, c2, c3], axis=1).astype(np.float32) * np.float32(INTER_RESIZE_COEF_SCALE)
    return s.astype(np.int32), np.clip(np.rint(coef), -32768, 32767).astype(np.int16)


def _window(offsets: np.ndarray, tile: int) -> int:
    """Largest source extent (taps included) any destination tile of `tile` indices needs along one axis."""
    starts = np.arange(0, len(offsets), tile)
    stops = np.minimum(starts + tile, len(offsets)) - 1
    return int((offsets[stops] + 2 - (offsets[starts] - 1) + 1).max())


class FrameResizer:
    """uint8 NHWC frames on the device -> bicubic-resized uint8 NHWC frames on the device (tables cached per geometry)."""

    def __init__(self, device: torch.device) -> None:
        if device.type != "cuda":
            raise RuntimeError("FrameResizer needs a CUDA device: the hot path has no CPU fallback")
        self.device = device
        self.lib = _lib.init(device.index or 0)
        self._tables: Dict[Tuple[int, int, int, int], tuple] = {}

    def _geometry(self, sh: int, sw: int, dh: int, dw: int):
        key = (sh, sw, dh, dw)
        if key not in self._tables:
            xofs, xcoef = cubic_tables(sw, dw)
            yofs, ycoef = cubic_tables(sh, dh)
            on_device = [torch.as_tensor(t).to(self.device).contiguous() for t in (xofs, xcoef, yofs, ycoef)]
            self._tables[key] = (*on_device, _window(xofs, TILE_W), _window(yofs, TILE_H))
        return self._tables[key]

    def resize(self, frames: torch.Tensor, height: int, width: int, out: torch.Tensor = None) -> torch.Tensor:
        """frames [B,H,W,3] uint8 (device, contiguous) -> [B,height,width,3] uint8."""
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[3] != 3 or not frames.is_cuda:
            raise ValueError(f"expected a [B,H,W,3] uint8 CUDA tensor, got {tuple(frames.shape)} {frames.dtype} {frames.device}")
        frames = frames.contiguous()
        batch, sh, sw, _ = frames.shape
        if out is None:
            out = torch.empty((batch, height, width, 3), dtype=torch.uint8, device=frames.device)
        xofs, xcoef, yofs, ycoef, win_w, win_h = self._geometry(sh, sw, height, width)
        stream = torch.cuda.current_stream(frames.device).cuda_stream
        _lib.check(
            self.lib.gance_resize_cubic_u8(frames.data_ptr(), batch, sh, sw, out.data_ptr(), height, width, xofs.data_ptr(),
                                           xcoef.data_ptr(), yofs.data_ptr(), ycoef.data_ptr(), win_w, win_h, stream),
            "gance_resize_cubic_u8",
        )
        return out


_RESIZERS: Dict[int, FrameResizer] = {}


def _resizer(device: torch.device) -> FrameResizer:
    index = device.index or 0
    if index not in _RESIZERS:
        _RESIZERS[index] = FrameResizer(torch.device("cuda", index))
    return _RESIZERS[index]


def resize_source(source: Iterator[np.ndarray], resolution: ImageResolution, batch: int = 8) -> Iterator[np.ndarray]:
    """
    For each (H,W,3) uint8 image of `source`, scale it to `resolution` (reference: video_common.py:399-429; images that
    already have the resolution pass through untouched, as there). Frames are staged `batch` at a time so one launch
    serves several frames; order is preserved. Like the reference's call `cv2.resize(image, (resolution.height,
    resolution.width))`, whose dsize argument is (width, height), the output has `resolution.height` COLUMNS and
    `resolution.width` ROWS - the two only differ for non-square targets, which GANCE never requests.
    """
    if not torch.cuda.is_available():
        raise RuntimeError("resize_source needs a CUDA device: the hot path has no CPU fallback")
    device = torch.device("cuda", torch.cuda.current_device())
    resizer = _resizer(device)
    out_rows, out_cols = resolution.width, resolution.height
    iterator = iter(source)
    while True:
        chunk = list(itertools.islice(iterator, batch))
        if not chunk:
            return
        index = 0
        while index < len(chunk):
            image = chunk[index]
            if image_resolution(image) == resolution:
                yield image
                index += 1
                continue
            # a run of equally shaped frames that need scaling goes to the device together
            stop = index + 1
            while stop < len(chunk) and chunk[stop].shape == image.shape and image_resolution(chunk[stop]) != resolution:
                stop += 1
            staged = torch.as_tensor(np.stack(chunk[index:stop])).to(device)
            resized = resizer.resize(staged, out_rows, out_cols).cpu().numpy()
            for frame in resized:
                yield frame
            index = stop


def scale_square_source_duplicate(source: Iterator[np.ndarray], output_side_length: int,
                                  frame_multiplier: int = 1) -> Iterator[np.ndarray]:
    """
    Scale the resolution and the number of frames of a source (reference: video_common.py:432-455): every frame is
    resized to output_side_length^2 and then repeated `frame_multiplier` times.
    """
    resized = resize_source(source, ImageResolution(output_side_length, output_side_length))
    if frame_multiplier == 1:
        return resized
    return (frame for image in resized for frame in itertools.repeat(image, frame_multiplier))
