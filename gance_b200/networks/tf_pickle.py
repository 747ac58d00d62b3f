"""
Reading a REAL StyleGAN2 network pickle without TensorFlow or dnnlib.

The reference does `Gs = pickle.load(open(path, "rb"))[2]` (gance/network_interface/network_functions.py:103-110): the
file holds the tuple (G, D, Gs) of `dnnlib.tflib.network.Network` objects, and unpickling one imports dnnlib, builds a
TF graph from the source text stored in the pickle and needs a TF1 session. None of that exists here. What the pickle
stream itself contains is small and regular - `Network.__getstate__` (NVlabs/stylegan2 dnnlib/tflib/network.py, the
fork the reference pins at gance/stylegan2 @ 0f84feb) writes

    {"version": 4, "name": str, "static_kwargs": dict, "components": {name: Network}, "build_module_src": str,
     "build_func_name": str, "variables": [(name, ndarray), ...]}

with variable names relative to each network's own scope. Gs (`G_main`) owns `dlatent_avg` and `lod`; its components
`mapping` (`G_mapping`: `Dense{i}/weight|bias`) and `synthesis` (`G_synthesis_stylegan2`: `4x4/Const/const`,
`{res}x{res}/Conv0_up|Conv1|ToRGB/weight|mod_weight|mod_bias|noise_strength|bias`, `noise{i}`) own the rest.

`RestrictedUnpickler` walks that stream with an allow-list: numpy array reconstruction, plain containers, and stub
classes standing in for `dnnlib.tflib.network.Network` / `dnnlib.util.EasyDict` that only record their state. Anything
else in the stream (a pickle is a program: `os.system` is one GLOBAL opcode away) raises `pickle.UnpicklingError`; the
stored module source is never executed. `variables_from_tf_network` then renames the captured arrays to the full TF
names of SURVEY.md Appendix A.2, which is what `GeneratorOnDevice` and the oracle are keyed by.
"""

import io
import pickle
from collections import OrderedDict
from typing import Any, Dict

import numpy as np


class CapturedNetwork:
    """Stands in for dnnlib.tflib.network.Network while unpickling: keeps `__getstate__`'s dict, builds nothing."""

    def __init__(self) -> None:
        self.state: Dict[str, Any] = {}

    def __setstate__(self, state: Dict[str, Any]) -> None:
        if not isinstance(state, dict):
            raise pickle.UnpicklingError(f"Network state is {type(state).__name__}, expected dict")
        self.state = state

    @property
    def name(self) -> str:
        return str(self.state.get("name", ""))

    @property
    def static_kwargs(self) -> Dict[str, Any]:
        return dict(self.state.get("static_kwargs", {}))

    @property
    def components(self) -> Dict[str, "CapturedNetwork"]:
        return dict(self.state.get("components", {}))

    @property
    def variables(self) -> Dict[str, np.ndarray]:
        return OrderedDict((str(name), np.asarray(value)) for name, value in self.state.get("variables", []))


class CapturedEasyDict(dict):
    """Stands in for dnnlib.util.EasyDict (a dict with attribute access; pickled as a dict subclass)."""

    def __getattr__(self, name: str) -> Any:
        try:
            return self[name]
        except KeyError as error:
            raise AttributeError(name) from error

    def __setattr__(self, name: str, value: Any) -> None:
        self[name] = value


_NUMPY_MODULES = ("numpy", "numpy.core.multiarray", "numpy._core.multiarray", "numpy.core.numeric", "numpy._core.numeric")
_NUMPY_NAMES = ("_reconstruct", "ndarray", "dtype", "scalar", "_frombuffer")
_BUILTINS = ("dict", "list", "tuple", "set", "frozenset", "bytes", "bytearray", "str", "int", "float", "bool", "complex",
             "slice", "range")
_STUBS = {
    ("dnnlib.tflib.network", "Network"): CapturedNetwork,
    ("dnnlib.util", "EasyDict"): CapturedEasyDict,
    ("dnnlib", "EasyDict"): CapturedEasyDict,
}


class RestrictedUnpickler(pickle.Unpickler):
    """pickle.Unpickler whose GLOBAL / STACK_GLOBAL opcodes resolve only to the allow-list above."""

    def find_class(self, module: str, name: str) -> Any:
        # the reference vendors dnnlib under gance.stylegan2 (network_functions.py:28-30); accept both spellings
        short = module[len("gance.stylegan2."):] if module.startswith("gance.stylegan2.") else module
        if (short, name) in _STUBS:
            return _STUBS[(short, name)]
        if module in _NUMPY_MODULES and name in _NUMPY_NAMES:
            return super().find_class(module, name)
        if module == "collections" and name == "OrderedDict":
            return OrderedDict
        if module in ("builtins", "__builtin__") and name in _BUILTINS:
            return super().find_class("builtins", name)
        if module == "_codecs" and name == "encode":  # how protocol-2 pickles spell bytes objects (numpy array data)
            return super().find_class(module, name)
        raise pickle.UnpicklingError(f"network pickle references {module}.{name}, which is not on the allow-list")


def restricted_load(data: bytes) -> Any:
    return RestrictedUnpickler(io.BytesIO(data), encoding="latin1").load()


_SUPPORTED = {
    "architecture": ("skip",),          # config-e/f generators; 'resnet' / 'orig' synthesis networks are not built here
    "nonlinearity": ("lrelu",),
    "fused_modconv": (True, False),     # same function, different TF graph
    "resample_kernel": ([1, 3, 3, 1], (1, 3, 3, 1)),
    "dtype": ("float32",),
}


def _check_kwargs(network: CapturedNetwork) -> None:
    for holder in [network] + list(network.components.values()):
        kwargs = holder.static_kwargs
        for key, allowed in _SUPPORTED.items():
            if key in kwargs and kwargs[key] not in allowed:
                raise ValueError(f"{holder.name or 'network'}: {key}={kwargs[key]!r} is not supported (expected one of {allowed})")
        if int(kwargs.get("label_size", 0) or 0) != 0:
            raise ValueError(f"{holder.name or 'network'}: conditional networks (label_size > 0) are not supported")


def variables_from_tf_network(gs: CapturedNetwork) -> Dict[str, np.ndarray]:
    """
    Captured Gs -> {full TF name: float32 ndarray} (Appendix A.2 names: `dlatent_avg`, `G_mapping/Dense0/weight`,
    `G_synthesis/4x4/Conv/weight`, `G_synthesis/noise0`, ...). `lod` (progressive-growing level, unused by the
    'skip' architecture) is dropped.
    """
    components = gs.components
    if "synthesis" not in components or "mapping" not in components:
        raise ValueError(f"network '{gs.name}' has components {sorted(components)}; expected 'mapping' and 'synthesis' "
                         "(is tuple index 2 the generator Gs?)")
    _check_kwargs(gs)
    out: Dict[str, np.ndarray] = {}
    for name, value in gs.variables.items():
        if name != "lod":
            out[name] = np.asarray(value, dtype=np.float32)
    for prefix, component in (("G_mapping", components["mapping"]), ("G_synthesis", components["synthesis"])):
        for name, value in component.variables.items():
            if name != "lod":
                out[f"{prefix}/{name}"] = np.asarray(value, dtype=np.float32)
    if "dlatent_avg" not in out:
        raise ValueError("network has no dlatent_avg variable")
    return out


def fmap_config(gs: CapturedNetwork) -> str:
    """'config-f' / 'config-e' from the synthesis network's fmap_base (16384 / 8192; absent = the function default 16 << 10)."""
    kwargs = dict(gs.static_kwargs)
    kwargs.update(gs.components["synthesis"].static_kwargs if "synthesis" in gs.components else {})
    fmap_base = int(kwargs.get("fmap_base", 16 << 10))
    for config, base in (("config-f", 16 << 10), ("config-e", 8 << 10)):
        if fmap_base == base:
            return config
    raise ValueError(f"fmap_base={fmap_base} matches no known StyleGAN2 config")
