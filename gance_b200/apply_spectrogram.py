"""
Spectrogram vectors on the GPU (mirrors /root/reference/gance/apply_spectrogram.py: same function
names, arguments and return shapes). The reference computes these with numpy / scipy / sklearn on
the host - a 510-point FFT per frame (49-82), a Python loop of `scipy.signal.resample` (20-46) and
Python loops of `savgol_filter` (vector_sources_common.py:136-188); here each step is a float64
CUDA kernel (gance_b200/audio_device.py) and the arrays stay in HBM between steps.
"""
from typing import Optional, Tuple

import numpy as np
import torch

from gance_b200 import audio_device
from gance_b200.vector_sources.vector_types import ConcatenatedVectors


def _as_mono_device(data) -> torch.Tensor:
    if isinstance(data, torch.Tensor):
        return data if data.dim() == 1 else data.to(torch.float64).mean(dim=1)
    array = np.asarray(data)
    if array.ndim > 1:  # apply_spectrogram.py:62-65: "stereo" data is averaged to mono
        array = np.mean(array, axis=1)
    return torch.as_tensor(np.ascontiguousarray(array, dtype=np.float32)).to(audio_device._device())  # pylint: disable=protected-access


def compute_spectrogram_smooth_scale_device(
    data, vector_length: int, amplitude_range: Optional[Tuple[int, int]] = None,
    roll_shift: Optional[np.ndarray] = None,
) -> torch.Tensor:
    """As `compute_spectrogram_smooth_scale`, but returns the float64 [T, vector_length] device tensor."""
    return audio_device.spectrogram_vectors(_as_mono_device(data), vector_length, amplitude_range, roll_shift)


def compute_spectrogram_smooth_scale(
    data: ConcatenatedVectors, vector_length: int, amplitude_range: Optional[Tuple[int, int]] = None
) -> ConcatenatedVectors:
    """
    apply_spectrogram.py:85-118: windowed FFT -> dB -> resample to `vector_length` -> global min-max
    scale to `amplitude_range` -> Savitzky-Golay across vectors (7,3) and within each vector (5,3).
    Returns float64 (T*vector_length,) on the host, like the reference.
    """
    vectors = compute_spectrogram_smooth_scale_device(data, vector_length, amplitude_range)
    return ConcatenatedVectors(vectors.reshape(-1).cpu().numpy())
