"""
Functions that produce `VisualizationInput`s (mirrors
/root/reference/gance/data_into_network_visualization/visualization_inputs.py: same names,
arguments and return type). The spectrogram, roll, smoothing and alpha blends run on the GPU
(gance_b200/audio_device.py); the T per-frame scalars - FFT roll amounts and network indices -
are computed on the host with the reference's own numpy / pandas / scipy operations because they
end in `np.rint` (gance_b200/vector_sources/vector_reduction.py).

Each function has a `*_device` twin that returns the latents as float32 device tensors in the
generator's layout ([T,512] / [T,18,512]), so a whole track can go from WAV to frames without the
latents visiting the host.
"""
from typing import List, NamedTuple, Optional, Tuple

import numpy as np
import torch
from sklearn.preprocessing import minmax_scale
from typing_extensions import Protocol

from gance_b200 import audio_device
from gance_b200.apply_spectrogram import compute_spectrogram_smooth_scale_device
from gance_b200.data_into_network_visualization.visualization_common import ResultLayers, VisualizationInput
from gance_b200.vector_sources import vector_sources_common
from gance_b200.vector_sources.primatives import NoisePrefetch, Sigmas, gaussian_data
from gance_b200.vector_sources.vector_reduction import quantize_results_layers, reduce_vector_rms_rolling_average
from gance_b200.vector_sources.vector_types import (
    ConcatenatedMatrices,
    ConcatenatedVectors,
    MatricesLabel,
    VectorsLabel,
)

W_PLUS_DEPTH = 18  # the literal in visualization_inputs.py:242


class CreateVisualizationInput(Protocol):  # pylint: disable=too-few-public-methods
    """Standard shape of a visualization input function (visualization_inputs.py:32-50)."""

    def __call__(
        self: "CreateVisualizationInput",
        time_series_audio_vectors: np.ndarray,
        vector_length: int,
        network_indices: List[int],
    ) -> VisualizationInput:
        ...


class DeviceLatents(NamedTuple):
    """What the synthesis loop needs, with the latents still in HBM."""

    latents: torch.Tensor  # float32 [T,512] (vector path) or [T,depth,512] (matrix path)
    network_indices: np.ndarray  # int (T,)
    spectrogram: torch.Tensor  # float64 [T,512]


def _roll_values(time_series_audio_vectors: np.ndarray, vector_length: int) -> ResultLayers:
    """visualization_inputs.py:74-80: RMS -> rolling mean -> savgol(7,3) -> quantised onto 0..2."""
    return quantize_results_layers(
        results_layers=reduce_vector_rms_rolling_average(
            time_series_audio_vectors=time_series_audio_vectors, vector_length=vector_length
        ),
        network_indices=list(np.arange(0, 3)),
    )


def _create_spectrogram_device(
    time_series_audio_vectors: np.ndarray, vector_length: int, fft_amplitude_range: Tuple[int, int], fft_roll_enabled: bool
) -> torch.Tensor:
    """visualization_inputs.py:53-91 -> float64 [T, vector_length] on the device."""
    roll_shift = None
    if fft_roll_enabled:
        # rotate_vectors_over_time (vector_sources_common.py:408-428): frame t is rolled by -cumsum(roll)[t]
        roll_shift = np.cumsum(_roll_values(time_series_audio_vectors, vector_length).result.data)
    return compute_spectrogram_smooth_scale_device(
        time_series_audio_vectors, vector_length, fft_amplitude_range, roll_shift=roll_shift
    )


def _create_spectrogram(
    time_series_audio_vectors: ConcatenatedVectors, vector_length: int, fft_amplitude_range: Tuple[int, int], fft_roll_enabled: bool
) -> ConcatenatedVectors:
    device = _create_spectrogram_device(time_series_audio_vectors, vector_length, fft_amplitude_range, fft_roll_enabled)
    return ConcatenatedVectors(device.reshape(-1).cpu().numpy())


def _noise(vector_length: int, num_vectors: int, prefetched: Optional[NoisePrefetch] = None) -> np.ndarray:
    """visualization_inputs.py:135-142: Gaussian-smoothed noise, min-max scaled to (-4, 4); float32, host."""
    return minmax_scale(
        gaussian_data(vector_length=vector_length, num_vectors=num_vectors, sigmas=Sigmas(across_vectors=50, within_vectors=0),
                      prefetched=prefetched),
        feature_range=(-4, 4),
    )


def _indices(time_series_audio_vectors, vector_length, network_indices, **savgol) -> ResultLayers:
    return quantize_results_layers(
        results_layers=reduce_vector_rms_rolling_average(
            time_series_audio_vectors=time_series_audio_vectors, vector_length=vector_length, **savgol
        ),
        network_indices=network_indices,
    )


def alpha_blend_vectors_max_rms_power_audio_device(
    alpha: float, fft_roll_enabled: bool, fft_amplitude_range: Tuple[int, int], time_series_audio_vectors: np.ndarray,
    vector_length: int, network_indices: List[int], noise_prefetch: Optional[NoisePrefetch] = None,
) -> DeviceLatents:
    """noise-blend latents as float32 [T,512] on the device (the vector path's z)."""
    spectrogram = _create_spectrogram_device(time_series_audio_vectors, vector_length, fft_amplitude_range, fft_roll_enabled)
    noise = _noise(vector_length, spectrogram.shape[0], noise_prefetch)
    _, latents = audio_device.alpha_blend_noise(spectrogram, noise, alpha, want_f64=False)
    indices = _indices(time_series_audio_vectors, vector_length, network_indices)
    return DeviceLatents(latents, np.asarray(indices.result.data), spectrogram)


def alpha_blend_vectors_max_rms_power_audio(
    alpha: float,
    fft_roll_enabled: bool,
    fft_amplitude_range: Tuple[int, int],
    time_series_audio_vectors: np.ndarray,
    vector_length: int,
    network_indices: List[int],
) -> VisualizationInput:
    """visualization_inputs.py:94-166: combined = noise*(1-alpha) + spectrogram*alpha, (T*512,) float64."""
    spectrogram = _create_spectrogram_device(time_series_audio_vectors, vector_length, fft_amplitude_range, fft_roll_enabled)
    noise = _noise(vector_length, spectrogram.shape[0])
    combined, _ = audio_device.alpha_blend_noise(spectrogram, noise, alpha, want_f32=False)
    return VisualizationInput(
        a_vectors=VectorsLabel(data=spectrogram.reshape(-1).cpu().numpy(), vector_length=vector_length, label="Audio Spectrogram"),
        b_vectors=VectorsLabel(data=noise, vector_length=vector_length, label="Gaussian Smoothed Noise"),
        combined=VectorsLabel(data=combined.cpu().numpy(), vector_length=vector_length, label=f"Combined w/ Alpha Blending, a={alpha}"),
        network_indices=_indices(time_series_audio_vectors, vector_length, network_indices),
    )


def _projected_rows(final_latents_matrices_label: MatricesLabel, vector_length: int, num_vectors: int):
    """Row 0 of the projection file as [F, L] float32 and the exact repeat count (visualization_inputs.py:220-231)."""
    row = vector_sources_common.demote_to_vector_select(final_latents_matrices_label.data, index_to_take=0)
    frames = vector_sources_common.sub_vectors(np.asarray(row), vector_length)
    return np.ascontiguousarray(frames, dtype=np.float32), vector_sources_common.duplication_factor(len(frames), num_vectors)


def alpha_blend_projection_file_device(
    final_latents_matrices_label: MatricesLabel, alpha: float, fft_roll_enabled: bool, fft_amplitude_range: Tuple[int, int],
    blend_depth: int, time_series_audio_vectors: np.ndarray, vector_length: int, network_indices: List[int],
    depth: Optional[int] = None,
) -> DeviceLatents:
    """projection-file-blend latents as float32 [T,depth,512] on the device (the matrix path's w+)."""
    spectrogram = _create_spectrogram_device(time_series_audio_vectors, vector_length, fft_amplitude_range, fft_roll_enabled)
    projected, duplication = _projected_rows(final_latents_matrices_label, vector_length, spectrogram.shape[0])
    rows = min(final_latents_matrices_label.data.shape[0], W_PLUS_DEPTH) if depth is None else depth
    _, latents = audio_device.alpha_blend_wplus(spectrogram, projected, duplication, alpha, blend_depth, rows, want_f64=False)
    indices = _indices(time_series_audio_vectors, vector_length, network_indices, savgol_window_length=3, savgol_polyorder=2)
    return DeviceLatents(latents, np.asarray(indices.result.data), spectrogram)


def alpha_blend_projection_file(  # pylint: disable=too-many-locals
    final_latents_matrices_label: MatricesLabel,
    alpha: float,
    fft_roll_enabled: bool,
    fft_amplitude_range: Tuple[int, int],
    blend_depth: int,
    time_series_audio_vectors: ConcatenatedVectors,
    vector_length: int,
    network_indices: List[int],
) -> VisualizationInput:
    """
    visualization_inputs.py:169-270: rows [0, blend_depth) = w*(1-alpha) + spectrogram*alpha, rows
    [blend_depth, 18) = the projected w; combined is (rows, T*512) float64.
    """
    spectrogram = _create_spectrogram_device(time_series_audio_vectors, vector_length, fft_amplitude_range, fft_roll_enabled)
    num_vectors = spectrogram.shape[0]
    projected, duplication = _projected_rows(final_latents_matrices_label, vector_length, num_vectors)
    file_depth = final_latents_matrices_label.data.shape[0]
    # np.concatenate((blended[:blend_depth], projected[blend_depth:18])): the result has
    # blend_depth + max(0, min(file_depth, 18) - blend_depth) rows
    depth = blend_depth + max(0, min(file_depth, W_PLUS_DEPTH) - blend_depth)
    combined, _ = audio_device.alpha_blend_wplus(spectrogram, projected, duplication, alpha, blend_depth, depth, want_f32=False)
    projected_vectors = ConcatenatedMatrices(
        np.tile(np.repeat(projected, duplication, axis=0).reshape(-1).astype(final_latents_matrices_label.data.dtype), (file_depth, 1))
    )
    return VisualizationInput(
        a_vectors=VectorsLabel(data=spectrogram.reshape(-1).cpu().numpy(), vector_length=vector_length, label="Rolled Audio Spectrogram"),
        b_vectors=MatricesLabel(data=projected_vectors, vector_length=vector_length, label=final_latents_matrices_label.label),
        combined=MatricesLabel(data=combined.cpu().numpy(), vector_length=vector_length, label=f"Combined w/ Alpha Blending, a={alpha}"),
        network_indices=_indices(time_series_audio_vectors, vector_length, network_indices, savgol_window_length=3, savgol_polyorder=2),
    )
