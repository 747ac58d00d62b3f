"""
Device side of the music vector source: every array-sized step of the reference's
audio -> spectrogram vectors -> blend path runs as one of the kernels in csrc/audio_kernels.cu,
in float64, on tensors that stay in HBM. The modules that
keep the reference's names and signatures (`gance_b200.apply_spectrogram`,
`gance_b200.vector_sources.music`, `...visualization_inputs`) are thin wrappers over this one.

Small host-built constants (cached): the Fourier-resampling matrix of `scipy.signal.resample`
(vector_sources_common.py:211-230), the Savitzky-Golay projection matrices of
`scipy.signal.savgol_filter(mode="interp")` (vector_sources_common.py:136-188) and resampy's
kaiser_best interpolation window (music.py:223-227).
"""

from functools import lru_cache
from typing import Optional, Tuple

import numpy as np
import torch
from scipy.signal import resample as _fourier_resample
from scipy.signal import savgol_coeffs

from gance_b200 import _lib

KAISER_BEST_NUM_ZEROS = 64
KAISER_BEST_PRECISION = 9
KAISER_BEST_BETA = 14.769656459379492
KAISER_BEST_ROLLOFF = 0.9475937167399596


def _device(device: Optional[torch.device] = None) -> torch.device:
    if not torch.cuda.is_available():
        raise _lib.GanceError("No CUDA device: the audio kernels have no CPU fallback")
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    _lib.init(device.index if device.index is not None else torch.cuda.current_device())
    return device


def _stream(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


@lru_cache(maxsize=8)
def resample_matrix_t(bins: int, vector_length: int) -> np.ndarray:
    """[bins, vector_length] float64: row k is scipy.signal.resample of the unit impulse at bin k."""
    return np.ascontiguousarray(_fourier_resample(np.eye(bins), vector_length, axis=1))


@lru_cache(maxsize=8)
def savgol_projection(window: int, polyorder: int) -> np.ndarray:
    """
    [window, window] float64. Row r holds the weights that give the degree-`polyorder` least-squares
    fit over a window, evaluated at window position r: the middle row is the interior FIR of
    savgol_filter, the other rows are its mode="interp" edge fits.
    """
    rows = [savgol_coeffs(window, polyorder, pos=r, use="dot") for r in range(window)]
    return np.ascontiguousarray(np.stack(rows))


@lru_cache(maxsize=2)
def kaiser_best_half_window() -> Tuple[np.ndarray, int]:
    """resampy.filters.sinc_window(64 zeros, precision 9, kaiser(beta), rolloff): right half, float64."""
    num_bits = 2**KAISER_BEST_PRECISION
    n = num_bits * KAISER_BEST_NUM_ZEROS
    sinc_win = KAISER_BEST_ROLLOFF * np.sinc(
        KAISER_BEST_ROLLOFF * np.linspace(0, KAISER_BEST_NUM_ZEROS, num=n + 1, endpoint=True)
    )
    taper = np.kaiser(2 * n + 1, KAISER_BEST_BETA)[n:]
    return taper * sinc_win, num_bits


_CONSTANTS = {}


def _constant(key, builder, device: torch.device) -> torch.Tensor:
    full_key = (key, device.index)
    tensor = _CONSTANTS.get(full_key)
    if tensor is None:
        tensor = torch.as_tensor(builder()).to(device)
        _CONSTANTS[full_key] = tensor
    return tensor


# ---- a1 / a2 ---------------------------------------------------------------------------------
def pcm_to_float(pcm: np.ndarray, device: Optional[torch.device] = None) -> torch.Tensor:
    """int16 / int32 / int8 PCM -> float32 in [-1, 1] on the device (music.py:190-205)."""
    device = _device(device)
    bits = {np.dtype(np.int16): 16, np.dtype(np.int32): 32, np.dtype(np.int8): 8}.get(pcm.dtype)
    if bits is None:
        raise ValueError(f"Cannot safely convert wav data to np.float32, unknown input format: {pcm.dtype}")
    flat = torch.as_tensor(np.ascontiguousarray(pcm)).to(device)
    out = torch.empty(flat.shape, dtype=torch.float32, device=device)
    _lib.check(
        _lib.load_library().gance_pcm_to_float(flat.data_ptr(), bits, flat.numel(), out.data_ptr(), _stream(device)),
        "gance_pcm_to_float",
    )
    return out


def sinc_resample(x: torch.Tensor, sr_orig: float, sr_new: float) -> torch.Tensor:
    """resampy.resample(x, sr_orig, sr_new) (kaiser_best) for a 1-D float32 device tensor."""
    device = _device(x.device)
    if sr_orig <= 0 or sr_new <= 0:
        raise ValueError("Invalid sample rate")
    x = x.to(torch.float32).contiguous()
    ratio = float(sr_new) / sr_orig
    n_out = int(x.shape[0] * ratio)
    if n_out < 1:
        raise ValueError("Input signal is too small to resample")
    win, num_table = kaiser_best_half_window()
    if ratio < 1:
        win = win * ratio
    delta = np.zeros_like(win)
    delta[:-1] = np.diff(win)
    # time_register += 1/ratio, accumulated in float64 exactly as the reference loop does
    times = np.zeros(n_out, dtype=np.float64)
    if n_out > 1:
        np.cumsum(np.full(n_out - 1, 1.0 / ratio, dtype=np.float64), out=times[1:])
    win_d = torch.as_tensor(win).to(device)
    delta_d = torch.as_tensor(delta).to(device)
    times_d = torch.as_tensor(times).to(device)
    y = torch.empty(n_out, dtype=torch.float32, device=device)
    _lib.check(
        _lib.load_library().gance_sinc_resample(
            x.data_ptr(), x.shape[0], times_d.data_ptr(), n_out, win_d.data_ptr(), delta_d.data_ptr(),
            win.shape[0], min(1.0, ratio), num_table, y.data_ptr(), _stream(device)),
        "gance_sinc_resample",
    )
    torch.cuda.current_stream(device).synchronize()  # the temporaries above die with this frame
    return y


# ---- a3 - a6, a8 -----------------------------------------------------------------------------
def spectrogram_vectors(
    audio: torch.Tensor,
    vector_length: int,
    amplitude_range: Optional[Tuple[float, float]],
    roll_shift: Optional[np.ndarray] = None,
) -> torch.Tensor:
    """
    compute_spectrogram_smooth_scale (apply_spectrogram.py:85-118), optionally followed by
    rotate_vectors_over_time + smooth_each_vector(51, 2) (visualization_inputs.py:82-89) when
    `roll_shift` (the running sum of the per-frame roll amounts) is given.
    audio: float32 [>= T*vector_length - 2] on the device -> float64 [T, vector_length].
    """
    device = _device(audio.device)
    lib, stream = _lib.load_library(), _stream(device)
    audio = audio.to(torch.float32).contiguous()
    window = vector_length - 2
    bins = window // 2
    num_frames = (audio.shape[0] - window) // vector_length + 1  # view_as_windows(step=vector_length)
    if num_frames < 1:
        raise ValueError("audio is shorter than one spectrogram window")
    mag = torch.empty((num_frames, bins), dtype=torch.float64, device=device)
    slots = torch.empty(3, dtype=torch.float64, device=device)  # [global max | min, max]
    _lib.check(lib.gance_stft_mag(audio.data_ptr(), num_frames, vector_length, mag.data_ptr(), slots.data_ptr(), stream),
               "gance_stft_mag")
    matrix = _constant(("resample", bins, vector_length), lambda: resample_matrix_t(bins, vector_length), device)
    vectors = torch.empty((num_frames, vector_length), dtype=torch.float64, device=device)
    _lib.check(
        lib.gance_db_resample(mag.data_ptr(), slots.data_ptr(), num_frames, bins, vector_length, matrix.data_ptr(),
                              vectors.data_ptr(), slots[1:].data_ptr(), stream),
        "gance_db_resample",
    )
    if amplitude_range is None:
        # no min-max scale: identity map (scale 1, offset 0) through the same kernel
        identity = torch.tensor([0.0, 1.0], dtype=torch.float64)
        encoded = _encode_ordered(identity).to(device)
        minmax_ptr, lo, hi = encoded.data_ptr(), 0.0, 1.0
    else:
        minmax_ptr, lo, hi = slots[1:].data_ptr(), float(amplitude_range[0]), float(amplitude_range[1])
    time_proj = _constant(("savgol", 7, 3), lambda: savgol_projection(7, 3), device)
    smoothed_time = torch.empty_like(vectors)
    _lib.check(
        lib.gance_minmax_savgol_time(vectors.data_ptr(), minmax_ptr, lo, hi, num_frames, vector_length,
                                     time_proj.data_ptr(), 7, smoothed_time.data_ptr(), stream),
        "gance_minmax_savgol_time",
    )
    vec_proj = _constant(("savgol", 5, 3), lambda: savgol_projection(5, 3), device)
    out = vectors  # reuse
    _lib.check(
        lib.gance_roll_savgol_vector(smoothed_time.data_ptr(), None, num_frames, vector_length, vec_proj.data_ptr(), 5,
                                     out.data_ptr(), stream),
        "gance_roll_savgol_vector",
    )
    if roll_shift is not None:
        shift = torch.as_tensor(np.asarray(roll_shift)[:num_frames] % vector_length, dtype=torch.int32).to(device)
        roll_proj = _constant(("savgol", 51, 2), lambda: savgol_projection(51, 2), device)
        rolled = smoothed_time  # reuse
        _lib.check(
            lib.gance_roll_savgol_vector(out.data_ptr(), shift.data_ptr(), num_frames, vector_length,
                                         roll_proj.data_ptr(), 51, rolled.data_ptr(), stream),
            "gance_roll_savgol_vector(roll)",
        )
        torch.cuda.current_stream(device).synchronize()
        return rolled
    return out


def _encode_ordered(values: torch.Tensor) -> torch.Tensor:
    """The order-preserving integer encoding the kernels use for their min/max slots, as float64 bits."""
    bits = values.view(torch.int64).clone()
    negative = bits < 0
    bits[negative] = ~bits[negative]
    bits[~negative] = bits[~negative] | torch.iinfo(torch.int64).min
    return bits.view(torch.float64)


# ---- a9 ------------------------------------------------------------------------------------------
def gaussian_kernel1d(sigma: float, truncate: float = 4.0):
    """scipy.ndimage._filters._gaussian_kernel1d(sigma, 0, radius) - (radius, weights from the centre outwards)."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (float(sigma) * float(sigma)) * x ** 2)
    phi = phi / phi.sum()
    return radius, np.ascontiguousarray(phi[radius:][::1], dtype=np.float64)


def gaussian_filter_time_wrap(latents: np.ndarray, sigma: float, device: Optional[torch.device] = None) -> np.ndarray:
    """
    scipy.ndimage.gaussian_filter(latents, sigma=(sigma, 0, 0), mode="wrap") for a float32 (T, 1, L) / (T, L) array
    (primatives.py:69), on the device, bit-exact with scipy; returns a host float32 array of the input's shape.
    """
    if device is None:
        if not torch.cuda.is_available():
            raise RuntimeError("gaussian_filter_time_wrap needs a CUDA device: the hot path has no CPU fallback")
        device = torch.device("cuda", torch.cuda.current_device())
    device = _device(device)
    array = np.ascontiguousarray(latents, dtype=np.float32)
    num_vectors, length = array.shape[0], int(np.prod(array.shape[1:]))
    radius, weights = gaussian_kernel1d(sigma)
    src = torch.as_tensor(array.reshape(num_vectors, length)).to(device)
    out = torch.empty_like(src)
    weights_d = torch.as_tensor(weights).to(device)
    _lib.check(
        _lib.load_library().gance_gaussian_filter_time_wrap(src.data_ptr(), num_vectors, length, weights_d.data_ptr(), radius,
                                                            out.data_ptr(), _stream(device)),
        "gance_gaussian_filter_time_wrap",
    )
    return out.cpu().numpy().reshape(array.shape)


# ---- a10 / a11 -------------------------------------------------------------------------------
def alpha_blend_noise(spec: torch.Tensor, noise: np.ndarray, alpha: float, want_f64: bool = True, want_f32: bool = True):
    """combined = noise*(1-alpha) + spectrogram*alpha (visualization_inputs.py:144). Returns (f64 | None, f32 | None)."""
    device = _device(spec.device)
    num_frames, length = spec.shape
    noise_d = torch.as_tensor(np.ascontiguousarray(noise, dtype=np.float32)).to(device)
    out64 = torch.empty((num_frames * length,), dtype=torch.float64, device=device) if want_f64 else None
    out32 = torch.empty((num_frames, length), dtype=torch.float32, device=device) if want_f32 else None
    _lib.check(
        _lib.load_library().gance_alpha_blend_noise(
            spec.data_ptr(), noise_d.data_ptr(), float(alpha), num_frames, length,
            None if out64 is None else out64.data_ptr(), None if out32 is None else out32.data_ptr(), _stream(device)),
        "gance_alpha_blend_noise",
    )
    torch.cuda.current_stream(device).synchronize()
    return out64, out32


def alpha_blend_wplus(spec: torch.Tensor, projected: np.ndarray, duplication: int, alpha: float, blend_depth: int,
                      depth: int, want_f64: bool = True, want_f32: bool = True):
    """
    alpha_blend_projection_file's array work (visualization_inputs.py:220-243): `projected` is row 0
    of the projection file as [F, L] float32; frame t uses projected[t // duplication].
    Returns ((depth, T*L) f64 | None, [T, depth, L] f32 | None).
    """
    device = _device(spec.device)
    num_frames, length = spec.shape
    projected_d = torch.as_tensor(np.ascontiguousarray(projected, dtype=np.float32)).to(device)
    out64 = torch.empty((depth, num_frames * length), dtype=torch.float64, device=device) if want_f64 else None
    out32 = torch.empty((num_frames, depth, length), dtype=torch.float32, device=device) if want_f32 else None
    _lib.check(
        _lib.load_library().gance_alpha_blend_wplus(
            spec.data_ptr(), projected_d.data_ptr(), int(duplication), float(alpha), int(blend_depth), int(depth),
            num_frames, length, None if out64 is None else out64.data_ptr(),
            None if out32 is None else out32.data_ptr(), _stream(device)),
        "gance_alpha_blend_wplus",
    )
    torch.cuda.current_stream(device).synchronize()
    return out64, out32
