"""
ORACLE (test infrastructure, not product): restatements of the third-party arithmetic the
reference's audio path calls but which is absent from /root/reference and from this image
(SURVEY.md Appendix C). Each is pinned by a reference test:

  resampy==0.2.2  resample(x, sr_orig, sr_new, filter='kaiser_best')
        called at /root/reference/gance/vector_sources/music.py:223-227
        pinned by test/test_dynamic_model_switching.py:39 (first-frame RMS 0.00298562) and
        test/test_vector_source_music.py:24 (len_out == int(len * multiplier))
  librosa==0.8.1  feature.rms(y, frame_length, center=False)
        called at /root/reference/gance/vector_sources/vector_reduction.py:33-35
        pinned by the same known answer
  scikit-image    util.view_as_windows(x, (m,), step)
        called at /root/reference/gance/apply_spectrogram.py:69

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""

import numpy as np

try:  # numba is in the image; the pure-numpy fallback below is only for tiny inputs
    import numba
except Exception:  # pragma: no cover
    numba = None

KAISER_BEST_NUM_ZEROS = 64
KAISER_BEST_PRECISION = 9
KAISER_BEST_BETA = 14.769656459379492
KAISER_BEST_ROLLOFF = 0.9475937167399596


def kaiser_best_half_window():
    """
    resampy.filters.sinc_window(num_zeros=64, precision=9, window=kaiser(beta), rolloff):
    right half of a Kaiser-windowed sinc sampled 2**precision times per zero crossing.
    Returns (interp_win float64[32769], precision=512).
    """
    num_bits = 2**KAISER_BEST_PRECISION
    n = num_bits * KAISER_BEST_NUM_ZEROS
    sinc_win = KAISER_BEST_ROLLOFF * np.sinc(
        KAISER_BEST_ROLLOFF * np.linspace(0, KAISER_BEST_NUM_ZEROS, num=n + 1, endpoint=True)
    )
    taper = np.kaiser(2 * n + 1, KAISER_BEST_BETA)[n:]
    return taper * sinc_win, num_bits


def _resample_f_py(x, y, sample_ratio, interp_win, interp_delta, num_table):
    """resampy.interpn.resample_f (0.2.2), 1-D. Accumulates into y's dtype like the original."""
    scale = min(1.0, sample_ratio)
    time_increment = 1.0 / sample_ratio
    index_step = int(scale * num_table)
    time_register = 0.0
    nwin = interp_win.shape[0]
    n_orig = x.shape[0]
    n_out = y.shape[0]
    for t in range(n_out):
        n = int(time_register)
        frac = scale * (time_register - n)
        index_frac = frac * num_table
        offset = int(index_frac)
        eta = index_frac - offset
        i_max = min(n + 1, (nwin - offset) // index_step)
        for i in range(i_max):
            weight = interp_win[offset + i * index_step] + eta * interp_delta[offset + i * index_step]
            y[t] += weight * x[n - i]
        frac = scale - frac
        index_frac = frac * num_table
        offset = int(index_frac)
        eta = index_frac - offset
        k_max = min(n_orig - n - 1, (nwin - offset) // index_step)
        for k in range(k_max):
            weight = interp_win[offset + k * index_step] + eta * interp_delta[offset + k * index_step]
            y[t] += weight * x[n + k + 1]
        time_register += time_increment


_resample_f = numba.jit(nopython=True, nogil=True, cache=False)(_resample_f_py) if numba else _resample_f_py


def resample(x, sr_orig, sr_new, axis=-1, filter="kaiser_best", **kwargs):  # pylint: disable=redefined-builtin
    """resampy.core.resample (0.2.2) for 1-D input: output length int(n * sr_new / sr_orig)."""
    if filter != "kaiser_best" or kwargs:
        raise NotImplementedError("only the default kaiser_best filter is restated")
    if sr_orig <= 0 or sr_new <= 0:
        raise ValueError("Invalid sample rate")
    x = np.asarray(x)
    if x.ndim != 1:
        raise NotImplementedError("1-D input only")
    sample_ratio = float(sr_new) / sr_orig
    n_out = int(x.shape[0] * sample_ratio)
    if n_out < 1:
        raise ValueError("Input signal is too small to resample")
    y = np.zeros(n_out, dtype=x.dtype)
    interp_win, precision = kaiser_best_half_window()
    if sample_ratio < 1:
        interp_win = interp_win * sample_ratio
    interp_delta = np.zeros_like(interp_win)
    interp_delta[:-1] = np.diff(interp_win)
    _resample_f(np.ascontiguousarray(x), y, sample_ratio, interp_win, interp_delta, precision)
    return y


def librosa_rms(y, frame_length=2048, hop_length=512, center=False):
    """
    librosa.feature.rms (0.8.1) with center=False: frames y[i*hop : i*hop+frame_length],
    sqrt(mean(|frame|^2)) -> shape (1, n_frames). hop_length defaults to 512 whatever the
    frame_length.
    """
    if center:
        raise NotImplementedError("GANce always passes center=False")
    y = np.asarray(y)
    n_frames = 1 + (y.shape[-1] - frame_length) // hop_length
    if n_frames < 1:
        raise ValueError("Input is too short for frame_length")
    # librosa.util.frame: a strided (frame_length, n_frames) view, reduced over axis 0
    y = np.ascontiguousarray(y)
    frames = np.lib.stride_tricks.as_strided(
        y, shape=(frame_length, n_frames), strides=(y.itemsize, hop_length * y.itemsize), writeable=False
    )
    power = np.mean(np.abs(frames) ** 2, axis=0, keepdims=True)
    return np.sqrt(power)


def view_as_windows(arr_in, window_shape, step=1):
    """skimage.util.view_as_windows for 1-D input."""
    (m,) = window_shape if isinstance(window_shape, tuple) else (window_shape,)
    return np.lib.stride_tricks.sliding_window_view(np.asarray(arr_in), m)[::step]
