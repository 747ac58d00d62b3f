"""
ORACLE (test infrastructure, not product): import the reference's OWN audio/blend modules from
/root/reference (a Python reference) behind a dependency shim, so that the
numpy restatement in oracle/audio_ref.py can be pinned against it and golden vectors can be
generated (tests/golden/make_golden.py). /root/reference does not exist on the GPU box: nothing
that runs there may call `load()`; `available()` says whether it can be used.

Shimmed third-party modules (absent from this image):
  resampy.resample, librosa.feature.rms, skimage.util.view_as_windows -> oracle/third_party_ref.py
  matplotlib*, mpl_toolkits*, ffmpeg, vidgear, more_itertools, h5py, sentinels, lz, imagehash,
  dataclasses_json, click_option_group -> permissive dummies (never executed on the audio path)
  cv2.cv2 -> cv2
The reference's `remap_values_into_range` maps a scalar interp1d over every sample through a
multiprocessing.Pool (vector_sources_common.py:44-61); the shim leaves it untouched.
"""

import importlib
import importlib.abc
import importlib.machinery
import sys
import types
from pathlib import Path

from oracle import third_party_ref

REFERENCE_ROOT = Path("/root/reference")

_DUMMY_ROOTS = (
    "matplotlib",
    "mpl_toolkits",
    "ffmpeg",
    "vidgear",
    "more_itertools",
    "h5py",
    "sentinels",
    "lz",
    "imagehash",
    "dataclasses_json",
    "click_option_group",
    "face_recognition",
    "pympler",
    "objgraph",
)


class _Anything:
    """Attribute/callable sink standing in for plotting / video classes that are never run."""

    def __init__(self, *args, **kwargs):
        pass

    def __call__(self, *args, **kwargs):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __mro_entries__(self, bases):
        return (object,)

    def __iter__(self):
        return iter(())

    def __or__(self, other):
        return self

    def __ror__(self, other):
        return self

    def __getitem__(self, item):
        return _Anything()


class _DummyModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()


class _DummyFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in _DUMMY_ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        module = _DummyModule(spec.name)
        module.__path__ = []
        return module

    def exec_module(self, module):
        pass


_LOADED = {}


def available() -> bool:
    return (REFERENCE_ROOT / "gance" / "apply_spectrogram.py").exists()


def load():
    """
    Returns a dict of the reference's own modules: music, vector_sources_common, vector_reduction,
    apply_spectrogram, visualization_inputs, primatives, vector_types, visualization_common.
    """
    if _LOADED:
        return _LOADED
    if not available():
        raise RuntimeError("/root/reference is not present; the reference shim cannot be used here")

    if not any(isinstance(f, _DummyFinder) for f in sys.meta_path):
        sys.meta_path.append(_DummyFinder())

    def real(name, **attrs):
        module = types.ModuleType(name)
        module.__path__ = []
        for key, value in attrs.items():
            setattr(module, key, value)
        sys.modules[name] = module
        return module

    real("resampy", resample=third_party_ref.resample)
    feature = real("librosa.feature", rms=third_party_ref.librosa_rms)
    real("librosa", feature=feature)
    util = real("skimage.util", view_as_windows=third_party_ref.view_as_windows)
    real("skimage", util=util)

    import cv2  # the reference does `from cv2 import cv2` (opencv < 4.6 layout)

    if not hasattr(cv2, "cv2"):
        cv2.cv2 = cv2
        sys.modules["cv2.cv2"] = cv2

    if str(REFERENCE_ROOT) not in sys.path:
        sys.path.insert(0, str(REFERENCE_ROOT))

    names = {
        "music": "gance.vector_sources.music",
        "vector_sources_common": "gance.vector_sources.vector_sources_common",
        "vector_reduction": "gance.vector_sources.vector_reduction",
        "primatives": "gance.vector_sources.primatives",
        "vector_types": "gance.vector_sources.vector_types",
        "apply_spectrogram": "gance.apply_spectrogram",
        "visualization_common": "gance.data_into_network_visualization.visualization_common",
        "visualization_inputs": "gance.data_into_network_visualization.visualization_inputs",
    }
    for key, module_name in names.items():
        _LOADED[key] = importlib.import_module(module_name)
    return _LOADED
