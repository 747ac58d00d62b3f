"""
ORACLE (test infrastructure, not product): CPU restatement (numpy / scipy / pandas, float64) of
the reference's audio -> per-frame latents -> blend path, SURVEY.md §8 rows a1-a12. It is
vectorised where the reference loops in Python, but keeps the reference's arithmetic and
operation order; every function cites the reference file:line it follows.

PINNED: `tests/test_oracle_audio.py` checks every function here against the reference's own
Python (imported from /root/reference through oracle/reference_shim.py, when present) and against
the golden vectors in tests/golden/ generated from it by tests/golden/make_golden.py; the
resampler / RMS restatements reproduce the reference's known answer 0.00298562
(test/test_dynamic_model_switching.py:39) and length law (test/test_vector_source_music.py:24).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""

from typing import List, NamedTuple, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import scipy.ndimage
from scipy.signal import resample as fourier_resample
from scipy.signal import savgol_filter

from oracle import third_party_ref

DEFAULT_RANDOM_SEED = 1234  # gance/vector_sources/primatives.py


class OracleInput(NamedTuple):
    """What the two alpha_blend_* functions return that the synthesis path consumes."""

    spectrogram: np.ndarray  # (T*512,) float64  -> VisualizationInput.a_vectors.data
    combined: np.ndarray  # (T*512,) or (18, T*512) float64 -> VisualizationInput.combined.data
    network_indices: np.ndarray  # (T,) int -> VisualizationInput.network_indices.result.data
    roll_values: Optional[np.ndarray]  # (T,) int, None when fft_roll_enabled is False


# -- a1: music.read_wav_file + vector_sources_common.remap_values_into_range ------------------
def remap_values_into_range(data: np.ndarray, input_range, output_range) -> np.ndarray:
    """
    vector_sources_common.py:44-61: scipy interp1d(list(input_range), list(output_range)) applied
    per value: slope = (y_hi-y_lo)/(x_hi-x_lo); y = slope*(x-x_lo) + y_lo, in float64.
    """
    x_lo, x_hi = float(input_range[0]), float(input_range[1])
    y_lo, y_hi = float(output_range[0]), float(output_range[1])
    slope = (y_hi - y_lo) / (x_hi - x_lo)
    return slope * (np.asarray(data, dtype=np.float64) - x_lo) + y_lo


def wav_to_float32(wav_data: np.ndarray) -> np.ndarray:
    """music.py:190-205: int16 -> [-1,1] through interp1d, then .astype(float32)."""
    if wav_data.dtype == np.float32:
        return wav_data
    ranges = {np.dtype(np.int32): (-2147483648, 2147483647), np.dtype(np.int16): (-32768, 32767), np.dtype(np.int8): (0, 255)}
    if wav_data.dtype not in ranges:
        raise ValueError(f"Cannot safely convert wav data to np.float32, unknown input format: {wav_data.dtype}")
    low, high = ranges[wav_data.dtype]
    if wav_data.size and (wav_data.min() < low or wav_data.max() > high):  # interp1d(bounds_error=True) in the reference
        raise ValueError("A value in x_new is below the interpolation range.")
    return remap_values_into_range(wav_data, ranges[wav_data.dtype], (-1, 1)).astype(np.float32)


# -- a2: music.read_wavs_scale_for_video -------------------------------------------------------
def scale_wav_for_video(
    wav_data: np.ndarray,
    sample_rate: int,
    vector_length: int,
    frames_per_second: Optional[float] = None,
    target_num_vectors: Optional[int] = None,
    pad_to_length: bool = True,
) -> np.ndarray:
    """
    music.py:60-169 for one already-decoded wav: mono mean (108-119), new sample rate (127-146),
    resampy kaiser_best (212-230), zero pad to a multiple of vector_length (148-155).
    """
    if frames_per_second is not None and target_num_vectors is not None:
        raise ValueError("Can't use both FPS mode and target vector count mode.")
    if frames_per_second is None and target_num_vectors is None:
        raise ValueError("Need to use FPS mode or target vector count mode.")
    data = wav_data.mean(axis=1) if wav_data.ndim > 1 else wav_data
    n = data.shape[0]
    if frames_per_second is not None:
        new_rate = int(sample_rate * (vector_length * (frames_per_second * (n / sample_rate))) / n)
    else:
        new_rate = float(sample_rate) * (target_num_vectors / (n / vector_length))
    scaled = third_party_ref.resample(data, sr_orig=sample_rate, sr_new=new_rate)
    if not pad_to_length:
        return scaled
    size = int(np.ceil(scaled.shape[0] / vector_length) * vector_length)
    return np.pad(scaled, pad_width=(0, size - len(scaled)), mode="constant", constant_values=0)


# -- a3: apply_spectrogram.compute_spectrogram -----------------------------------------------
def compute_spectrogram(data: np.ndarray, num_frequency_bins: int) -> np.ndarray:
    """
    apply_spectrogram.py:49-82: windows of m = bins-2 samples at hop = bins (68-69), periodic
    Hann np.hanning(m+1)[:-1] (70), complex FFT along the window (73), first m//2 bins (76),
    20*log10(|X| / max|X|) (80-81). Returns (m//2, T) float64.
    """
    m = num_frequency_bins - 1 * 2
    slices = third_party_ref.view_as_windows(data, (m,), step=num_frequency_bins)
    win = np.hanning(m + 1)[:-1]
    slices = (slices * win).T
    spectrum = np.fft.fft(slices, axis=0)[: (m // 2)]
    s = np.abs(spectrum)
    return 20 * np.log10(s / np.max(s))


def minmax_scale_1d(flat: np.ndarray, feature_range) -> np.ndarray:
    """
    sklearn.preprocessing.minmax_scale on a 1-D array = one feature: scale = (r1-r0)/(max-min);
    min_ = r0 - min*scale; X*scale + min_ (MinMaxScaler.fit/transform), dtype preserved.
    """
    lo, hi = np.min(flat), np.max(flat)
    data_range = hi - lo
    if data_range == 0:  # sklearn _handle_zeros_in_scale
        data_range = flat.dtype.type(1)
    scale = (feature_range[1] - feature_range[0]) / data_range
    min_ = feature_range[0] - lo * scale
    return (flat * scale + min_).astype(flat.dtype)


# -- a4: apply_spectrogram.reshape_spectrogram_to_vectors ------------------------------------
def reshape_spectrogram_to_vectors(spectrogram: np.ndarray, vector_length: int, amplitude_range) -> np.ndarray:
    """
    apply_spectrogram.py:20-46 + vector_sources_common.py:211-230: each transposed row (255) is
    Fourier-resampled to vector_length with scipy.signal.resample, rows are concatenated, then the
    WHOLE flat array is min-max scaled to amplitude_range (43).
    """
    resampled = fourier_resample(np.transpose(spectrogram), vector_length, axis=1)
    flat = resampled.reshape(-1)
    return minmax_scale_1d(flat, amplitude_range) if amplitude_range is not None else flat


# -- a5 / a6: savgol smoothing ----------------------------------------------------------------
def smooth_across_vectors(data: np.ndarray, vector_length: int, window_length: int = 7, polyorder: int = 3) -> np.ndarray:
    """vector_sources_common.py:136-166: savgol along TIME for each of the vector positions."""
    vectors = data.reshape(-1, vector_length)
    return savgol_filter(vectors, window_length, polyorder, axis=0).reshape(-1)


def smooth_each_vector(data: np.ndarray, vector_length: int, window_length: int = 51, polyorder: int = 2) -> np.ndarray:
    """vector_sources_common.py:169-188: savgol WITHIN each vector."""
    vectors = data.reshape(-1, vector_length)
    return savgol_filter(vectors, window_length, polyorder, axis=1).reshape(-1)


def compute_spectrogram_smooth_scale(data: np.ndarray, vector_length: int, amplitude_range) -> np.ndarray:
    """apply_spectrogram.py:85-118: spectrogram -> vectors -> savgol time (7,3) -> savgol vec (5,3)."""
    spectrogram = compute_spectrogram(data, vector_length)
    vectors = reshape_spectrogram_to_vectors(spectrogram, vector_length, amplitude_range)
    vectors = smooth_across_vectors(vectors, vector_length, window_length=7, polyorder=3)
    return smooth_each_vector(vectors, vector_length, window_length=5, polyorder=3)


# -- a7: vector_reduction ---------------------------------------------------------------------
def compute_raw_rms(audio: np.ndarray, vector_length: int) -> np.ndarray:
    """vector_reduction.py:22-35: librosa.feature.rms(frame_length=vector_length, center=False)[0]."""
    return third_party_ref.librosa_rms(y=audio, frame_length=vector_length, center=False)[0]


def smoothed_rolling_average(values: np.ndarray, rolling_average_window=3, savgol_window_length=7, savgol_polyorder=3) -> np.ndarray:
    """vector_reduction.py:61-99: pandas rolling mean, NaN -> series mean, then savgol."""
    series = pd.Series(values)
    rolling = series.rolling(rolling_average_window).mean().fillna(series.mean()).to_numpy()
    return savgol_filter(x=rolling, window_length=savgol_window_length, polyorder=savgol_polyorder)


def quantize(values: np.ndarray, network_indices: Sequence[int]) -> np.ndarray:
    """vector_reduction.py:161-194: linear map min..max -> 0..n-1, np.rint (half to even), int."""
    scaled = remap_values_into_range(values, (min(values), max(values)), (0, len(network_indices) - 1))
    return np.rint(scaled).astype(int)


def rms_indices(audio: np.ndarray, vector_length: int, network_indices, savgol_window_length=7, savgol_polyorder=3) -> np.ndarray:
    """quantize_results_layers(reduce_vector_rms_rolling_average(...)) as both blend functions call it."""
    smoothed = smoothed_rolling_average(
        compute_raw_rms(audio, vector_length),
        savgol_window_length=savgol_window_length,
        savgol_polyorder=savgol_polyorder,
    )
    return quantize(smoothed, network_indices)


# -- a8: rotate_vectors_over_time -------------------------------------------------------------
def rotate_vectors_over_time(data: np.ndarray, vector_length: int, roll_values: np.ndarray) -> np.ndarray:
    """vector_sources_common.py:408-428: np.roll(frame_t, -cumsum(roll)[t])."""
    vectors = data.reshape(-1, vector_length)
    shift = np.cumsum(roll_values)
    columns = (np.arange(vector_length)[None, :] + shift[: len(vectors), None]) % vector_length
    return np.take_along_axis(vectors[: len(shift)], columns, axis=1).reshape(-1)


def create_spectrogram(audio: np.ndarray, vector_length: int, fft_amplitude_range, fft_roll_enabled: bool):
    """visualization_inputs.py:53-91. Returns (spectrogram, roll_values or None)."""
    spectrogram = compute_spectrogram_smooth_scale(audio, vector_length, fft_amplitude_range)
    roll_values = None
    if fft_roll_enabled:
        roll_values = rms_indices(audio, vector_length, list(np.arange(0, 3)))
        spectrogram = smooth_each_vector(
            rotate_vectors_over_time(spectrogram, vector_length, roll_values), vector_length
        )
    return spectrogram, roll_values


# -- a9: primatives.gaussian_data --------------------------------------------------------------
def gaussian_data(vector_length: int, num_vectors: int, sigma_across: float = 20, sigma_within: float = 0, seed: int = DEFAULT_RANDOM_SEED) -> np.ndarray:
    """primatives.py:49-74: RandomState(seed).randn -> float32 -> wrap-mode gaussian filter -> / global RMS."""
    latents = np.random.RandomState(seed).randn(num_vectors, 1, vector_length).astype(np.float32)
    latents = scipy.ndimage.gaussian_filter(input=latents, sigma=(sigma_across, 0, sigma_within), mode="wrap")
    latents /= np.sqrt(np.mean(np.square(latents)))
    return latents.reshape(vector_length * num_vectors)


# -- a10: alpha_blend_vectors_max_rms_power_audio ---------------------------------------------
def alpha_blend_vectors_max_rms_power_audio(alpha, fft_roll_enabled, fft_amplitude_range, audio, vector_length, network_indices) -> OracleInput:
    """visualization_inputs.py:94-166: combined = noise*(1-alpha) + spectrogram*alpha."""
    spectrogram, roll_values = create_spectrogram(audio, vector_length, fft_amplitude_range, fft_roll_enabled)
    num_vectors = int(spectrogram.shape[0] / vector_length)
    noise = minmax_scale_1d(gaussian_data(vector_length, num_vectors, sigma_across=50, sigma_within=0), (-4, 4))
    combined = noise * (1.0 - alpha) + spectrogram * alpha
    indices = rms_indices(audio, vector_length, network_indices)
    return OracleInput(spectrogram, combined, indices, roll_values)


# -- a11: alpha_blend_projection_file ---------------------------------------------------------
def duplicate_to_vector_count(data: np.ndarray, vector_length: int, target_vector_count: int) -> np.ndarray:
    """vector_sources_common.py:298-345 with divisor.divide_no_remainder (divisor.py:10-24)."""
    vectors = data.reshape(-1, vector_length)
    frac, repeats = np.modf(target_vector_count / len(vectors))
    if frac != 0:
        raise ValueError(
            f"Cannot duplicate the input vectors (count {len(vectors)}) to the desired count {target_vector_count}."
        )
    return np.repeat(vectors, int(repeats), axis=0).reshape(-1)


def alpha_blend_projection_file(final_latents: np.ndarray, alpha, fft_roll_enabled, fft_amplitude_range, blend_depth, audio, vector_length, network_indices) -> OracleInput:
    """
    visualization_inputs.py:169-270. `final_latents` is MatricesLabel.data, (18, F*512): row 0 is
    taken (224-226), each projected frame repeated T/F times, tiled to (18, T*512) (220-231);
    rows [0, blend_depth) = w*(1-alpha) + spectrogram*alpha (233-239); rows [blend_depth, 18) pure
    w (241-243); indices use savgol (3,2) (245-253).
    """
    spectrogram, roll_values = create_spectrogram(audio, vector_length, fft_amplitude_range, fft_roll_enabled)
    num_vectors = int(spectrogram.shape[0] / vector_length)
    depth = final_latents.shape[0]
    duplicated = duplicate_to_vector_count(final_latents[0], vector_length, num_vectors)
    projected = np.tile(duplicated, (depth, 1))
    blended = np.tile(projected[0] * (1.0 - alpha) + spectrogram * alpha, (blend_depth, 1))
    combined = np.concatenate((blended, projected[blend_depth:18]))
    indices = rms_indices(audio, vector_length, network_indices, savgol_window_length=3, savgol_polyorder=2)
    return OracleInput(spectrogram, combined, indices, roll_values)


# -- a12: frame splitting / ordering -----------------------------------------------------------
def sub_vectors(data: np.ndarray, vector_length: int) -> np.ndarray:
    """vector_sources_common.py:86-101: (T*L,) -> (T,L); (D,T*L) -> (T,D,L)."""
    if data.ndim >= 2:
        num_vectors = int(data.shape[-1] / vector_length)
        return np.array(np.split(data, num_vectors, axis=-1))
    return np.reshape(data, (-1, vector_length))


def render_order(network_indices: np.ndarray, num_networks: int, force_optimize: bool = True) -> List[int]:
    """
    network_visualization.py:653-674: with more than one network, frames are rendered in a STABLE
    sort by network index (sorted(), 659-663) and yielded back in frame order (671-674).
    Returns the order in which frame indices reach the network.
    """
    frames = list(range(len(network_indices)))
    if force_optimize and num_networks > 1:
        return sorted(frames, key=lambda i: int(network_indices[i]))
    return frames
