"""
Per-frame RMS -> network index / FFT roll amount (mirrors the hot-path part of
/root/reference/gance/vector_sources/vector_reduction.py:22-124,161-194).

These T scalars are computed on the HOST with the reference's own numpy / pandas / scipy
operations: the result is `np.rint` of a rescaled float, ties flip on a 1-ulp difference
(SURVEY.md §7 "Index/roll determinism"), and frame indexing has to be bit-exact.
"""
from typing import List

import numpy as np
import pandas as pd
from scipy.signal import savgol_filter

from gance_b200.data_into_network_visualization.visualization_common import DataLabel, ResultLayers
from gance_b200.vector_sources.vector_sources_common import remap_values_into_range
from gance_b200.vector_sources.vector_types import ConcatenatedVectors

RMS_HOP_LENGTH = 512  # librosa.feature.rms default hop, whatever frame_length is


def _compute_raw_rms(time_series_audio_vectors: ConcatenatedVectors, vector_length: int) -> np.ndarray:
    """
    librosa.feature.rms(y, frame_length=vector_length, center=False)[0] (vector_reduction.py:22-35):
    frames y[i*512 : i*512+vector_length], sqrt(mean(|frame|^2)) in the input's dtype.
    """
    y = np.ascontiguousarray(time_series_audio_vectors)
    n_frames = 1 + (y.shape[-1] - vector_length) // RMS_HOP_LENGTH
    if n_frames < 1:
        raise ValueError("Audio is shorter than one frame")
    frames = np.lib.stride_tricks.as_strided(
        y, shape=(vector_length, n_frames), strides=(y.itemsize, RMS_HOP_LENGTH * y.itemsize), writeable=False
    )
    return np.sqrt(np.mean(np.abs(frames) ** 2, axis=0, keepdims=True))[0]


def _smoothed_rolling_average(
    input_values: DataLabel, rolling_average_window: int = 3, savgol_window_length: int = 7, savgol_polyorder: int = 3
) -> ResultLayers:
    """Rolling mean (NaN -> series mean) then Savitzky-Golay (vector_reduction.py:61-99)."""
    as_series = pd.Series(input_values.data)
    rolling_average = as_series.rolling(rolling_average_window).mean().fillna(as_series.mean()).to_numpy()
    smoothed = savgol_filter(x=rolling_average, window_length=savgol_window_length, polyorder=savgol_polyorder)
    return ResultLayers(
        result=DataLabel(smoothed, f"Savgol Smoothing Filter (window={savgol_window_length}, polyorder={savgol_polyorder})"),
        layers=[DataLabel(rolling_average, f"Rolling Average (window={rolling_average_window})"), input_values],
    )


def reduce_vector_rms_rolling_average(
    time_series_audio_vectors: ConcatenatedVectors,
    vector_length: int,
    rolling_average_window: int = 3,
    savgol_window_length: int = 7,
    savgol_polyorder: int = 3,
) -> ResultLayers:
    """vector_reduction.py:102-124."""
    return _smoothed_rolling_average(
        DataLabel(_compute_raw_rms(time_series_audio_vectors, vector_length), "Raw RMS Power"),
        rolling_average_window=rolling_average_window,
        savgol_window_length=savgol_window_length,
        savgol_polyorder=savgol_polyorder,
    )


def quantize_results_layers(results_layers: ResultLayers, network_indices: List[int]) -> ResultLayers:
    """Scale min..max onto 0..n-1 and round half to even (vector_reduction.py:161-194)."""
    values = results_layers.result.data
    scaled = remap_values_into_range(
        data=values, input_range=(min(values), max(values)), output_range=(0, len(network_indices) - 1)
    )
    quantized = np.rint(scaled).astype(int)
    return ResultLayers(
        result=DataLabel(quantized, f"{results_layers.result.label} Scaled, Quantized"),
        layers=[results_layers.result] + results_layers.layers,
    )
