"""
Real-pickle import without TensorFlow (gance_b200/networks/tf_pickle.py). The reference loads networks with
`pickle.load(f)[2]` on a (G, D, Gs) tuple of dnnlib Networks (gance/network_interface/network_functions.py:103-110);
no such pickle ships with it, so the test writes one the way `dnnlib.tflib.network.Network.__getstate__` does - from a
throw-away module named `dnnlib.tflib.network` that exists only while the file is written - and reads it back with
the module gone.
"""

import os
import pickle
import sys
import types

import numpy as np
import pytest

from gance_b200.networks import container, tf_pickle
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables


def _write_tf_style_pickle(path, variables, resolution, fmap_base=16 << 10, module_prefix="dnnlib", architecture="skip",
                           protocol=2):
    """(G, D, Gs) with Gs = Network(G_main){dlatent_avg, lod; components mapping + synthesis}, TF-relative variable names."""
    package_names = [module_prefix, f"{module_prefix}.tflib", f"{module_prefix}.tflib.network", f"{module_prefix}.util"]
    if module_prefix != "dnnlib":
        parts = module_prefix.split(".")
        package_names = [".".join(parts[: i + 1]) for i in range(len(parts))] + package_names[1:]
    created = {}
    for name in package_names:
        if name not in sys.modules:
            created[name] = sys.modules[name] = types.ModuleType(name)

    class Network:  # pylint: disable=too-few-public-methods
        def __init__(self, state):
            self._state = state

        def __getstate__(self):
            return self._state

    class EasyDict(dict):
        pass

    Network.__module__, Network.__qualname__ = f"{module_prefix}.tflib.network", "Network"
    EasyDict.__module__, EasyDict.__qualname__ = f"{module_prefix}.util", "EasyDict"
    sys.modules[f"{module_prefix}.tflib.network"].Network = Network
    sys.modules[f"{module_prefix}.util"].EasyDict = EasyDict
    try:
        def component(name, func, prefix, extra):
            own = [(k[len(prefix) + 1:], v) for k, v in variables.items() if k.startswith(prefix + "/")]
            return Network({"version": 4, "name": name, "static_kwargs": dict(EasyDict(extra)), "components": {},
                            "build_module_src": "raise RuntimeError('module source must never be executed')",
                            "build_func_name": func, "variables": own})

        synthesis = component("G_synthesis", "G_synthesis_stylegan2", "G_synthesis",
                              {"resolution": resolution, "fmap_base": fmap_base, "architecture": architecture,
                               "resample_kernel": [1, 3, 3, 1], "fused_modconv": True, "nonlinearity": "lrelu"})
        mapping = component("G_mapping", "G_mapping", "G_mapping", {"mapping_layers": 8, "label_size": 0})
        gs = Network({"version": 4, "name": "Gs", "static_kwargs": {"resolution": resolution, "num_channels": 3},
                      "components": dict(EasyDict(synthesis=synthesis, mapping=mapping)),
                      "build_module_src": "import os; os.system('false')", "build_func_name": "G_main",
                      "variables": [("lod", np.zeros((), np.float32)), ("dlatent_avg", variables["dlatent_avg"])]})
        discriminator = Network({"version": 4, "name": "D", "static_kwargs": {}, "components": {}, "build_module_src": "",
                                 "build_func_name": "D_stylegan2", "variables": [("4x4/Dense0/weight", np.ones((4, 4), np.float32))]})
        with open(path, "wb") as handle:
            pickle.dump((gs, discriminator, gs), handle, protocol=protocol)
    finally:
        for name in created:
            del sys.modules[name]
        for name in package_names:
            module = sys.modules.get(name)
            if module is not None:
                for attr in ("Network", "EasyDict"):
                    if hasattr(module, attr):
                        delattr(module, attr)


@pytest.mark.parametrize("module_prefix,protocol", [("dnnlib", 2), ("gance.stylegan2.dnnlib", 4)])
def test_real_pickle_round_trip_without_dnnlib(tmp_path, module_prefix, protocol):
    spec = GeneratorSpec(64)
    variables = random_init_variables(spec, seed=5, perturbed=True)
    path = tmp_path / "network-snapshot.pkl"
    _write_tf_style_pickle(path, variables, 64, module_prefix=module_prefix, protocol=protocol)
    assert "dnnlib" not in sys.modules and "dnnlib.tflib.network" not in sys.modules
    with pytest.raises(Exception):  # the stock unpickler needs dnnlib, as in the reference
        with open(path, "rb") as handle:
            pickle.load(handle)
    loaded = container.read_network(path)
    assert loaded.spec.resolution == 64 and loaded.spec.config == "config-f" and loaded.input_shape[1] == 512
    assert set(loaded.variables) == set(variables)
    for name, value in variables.items():
        assert loaded.variables[name].dtype == np.float32
        assert np.array_equal(loaded.variables[name], value), name


def test_config_e_is_recognised_by_fmap_base(tmp_path):
    spec = GeneratorSpec(32, "config-e")
    variables = random_init_variables(spec, seed=1)
    path = tmp_path / "e.pkl"
    _write_tf_style_pickle(path, variables, 32, fmap_base=8 << 10)
    loaded = container.read_network(path)
    assert loaded.spec.config == "config-e" and loaded.spec.resolution == 32


def test_unsupported_architecture_is_reported(tmp_path):
    variables = random_init_variables(GeneratorSpec(16), seed=1)
    path = tmp_path / "resnet.pkl"
    _write_tf_style_pickle(path, variables, 16, architecture="resnet")
    with pytest.raises(ValueError, match="architecture"):
        container.read_network(path)


def test_missing_variables_are_reported(tmp_path):
    variables = random_init_variables(GeneratorSpec(16), seed=1)
    del variables["G_synthesis/16x16/ToRGB/bias"]
    path = tmp_path / "broken.pkl"
    _write_tf_style_pickle(path, variables, 16)
    with pytest.raises(ValueError, match="missing variables"):
        container.read_network(path)


class _Evil:
    def __reduce__(self):
        return (os.system, ("echo pwned > /tmp/gance_b200_pwned",))


def test_pickles_that_reference_other_globals_are_refused(tmp_path):
    path = tmp_path / "evil.pkl"
    with open(path, "wb") as handle:
        pickle.dump((None, None, _Evil()), handle)
    with pytest.raises(ValueError, match="allow-list"):
        container.read_network(path)
    assert not os.path.exists("/tmp/gance_b200_pwned")
    with pytest.raises(pickle.UnpicklingError):
        tf_pickle.restricted_load(pickle.dumps(print))


def test_own_formats_still_load_through_the_restricted_unpickler(tmp_path):
    synthetic = container.write_synthetic_network(tmp_path / "s.pkl", 32, seed=2, perturbed=True)
    explicit = container.write_variables_network(tmp_path / "v.pkl", random_init_variables(GeneratorSpec(32), 2, True))
    a, b = container.read_network(synthetic), container.read_network(explicit)
    assert a.spec.resolution == b.spec.resolution == 32
    for name in a.variables:
        assert np.array_equal(a.variables[name], b.variables[name])
