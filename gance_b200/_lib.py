"""
ctypes binding of libgance_b200.so (include/gance_b200.h). This is the ONLY way the Python host
reaches the kernels; there is no fallback: if the library is missing or no CUDA device is
visible, loading/initialising raises.
"""

import ctypes
from ctypes import c_char_p, c_double, c_float, c_int, c_longlong, c_ulonglong, c_void_p
from pathlib import Path
from typing import Optional

LIBRARY_PATH = Path(__file__).resolve().parent / "csrc" / "libgance_b200.so"

GANCE_F32 = 0
GANCE_BF16 = 1
GANCE_F16 = 2

_SIGNATURES = {
    "gance_abi_version": (c_int, []),
    "gance_init": (c_int, [c_int]),
    "gance_last_error_string": (c_char_p, []),
    "gance_launch_count": (c_ulonglong, []),
    "gance_set_nonfinite_flag": (c_int, [c_void_p]),
    "gance_set_conv_pairing": (c_int, [c_int]),
    "gance_set_conv_two_rows": (c_int, [c_int]),
    "gance_mapping": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_float, c_int, c_void_p, c_void_p]),
    "gance_style_demod": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p],
    ),
    "gance_styles_all": (
        c_int,
        [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int,
         c_void_p, c_void_p],
    ),
    "gance_const_input": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "gance_scale_channels": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "gance_modconv3x3": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong,
         c_float, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_split_tf32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "gance_modconv3x3_tf32x3": (
        c_int,
        [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
         c_longlong, c_float, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_dual": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong,
         c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_upfir2x_pad": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "gance_modconv3x3_up2_fir": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_longlong,
         c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_up2_folded": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_up2_phase": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_up2_phase_static": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_up2_fused": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_modconv3x3_torgb": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "gance_torgb_skip": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p],
    ),
    "gance_to_uint8_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "gance_normal_noise": (c_int, [c_void_p, c_longlong, c_ulonglong, c_ulonglong, c_void_p]),
    "gance_normal_noise_next": (c_int, [c_void_p, c_longlong, c_ulonglong, c_void_p, c_void_p]),
    "gance_gaussian_filter_time_wrap": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "gance_resize_cubic_u8": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
         c_void_p],
    ),
    "gance_pcm_to_float": (c_int, [c_void_p, c_int, c_longlong, c_void_p, c_void_p]),
    "gance_sinc_resample": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_double, c_int, c_void_p, c_void_p],
    ),
    "gance_stft_mag": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "gance_db_resample": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "gance_minmax_savgol_time": (
        c_int,
        [c_void_p, c_void_p, c_double, c_double, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "gance_roll_savgol_vector": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "gance_alpha_blend_noise": (c_int, [c_void_p, c_void_p, c_double, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "gance_alpha_blend_wplus": (
        c_int,
        [c_void_p, c_void_p, c_int, c_double, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p],
    ),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_LIB: Optional[ctypes.CDLL] = None
_INITIALISED_DEVICES = set()


GANCE_E_NONFINITE = 7


class GanceError(RuntimeError):
    """A libgance_b200 entry returned non-zero."""


class NonFiniteFramesError(GanceError):
    """GANCE_E_NONFINITE: rendered frames held inf/NaN pixels - a 16-bit activation overflowed (include/gance_b200.h)."""


def load_library() -> ctypes.CDLL:
    """dlopen the in-tree library and declare every prototype. No compute happens here."""
    global _LIB  # pylint: disable=global-statement
    if _LIB is not None:
        return _LIB
    if not LIBRARY_PATH.exists():
        raise GanceError(
            f"{LIBRARY_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the CUDA path has no CPU fallback)"
        )
    lib = ctypes.CDLL(str(LIBRARY_PATH))
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _LIB = lib
    return lib


def last_error() -> str:
    return load_library().gance_last_error_string().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        kind = "argument rejected" if rc > 0 else f"CUDA error {-rc}"
        raise GanceError(f"{what}: {kind} (code {rc}): {last_error()}")


def init(device_index: int) -> ctypes.CDLL:
    """gance_init once per device; raises without a B200."""
    lib = load_library()
    if device_index not in _INITIALISED_DEVICES:
        check(lib.gance_init(device_index), "gance_init")
        _INITIALISED_DEVICES.add(device_index)
    return lib


def launch_count() -> int:
    return int(load_library().gance_launch_count())
