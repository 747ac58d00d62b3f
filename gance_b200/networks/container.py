"""
Network files. The reference does `Gs = pickle.load(open(path, "rb"))[2]`
(network_functions.py:93-111): a `.pkl` holding the tuple (G, D, Gs) of dnnlib Networks. Those
pickles need TF1 + dnnlib to load and none ship with the reference, so this module
defines the two `.pkl` payloads this build can read, both kept at tuple index 2 so the file shape
matches the reference's:

  * "gance_b200/synthetic-v1": {resolution, config, seed, perturbed} - a random-init network of a
    named StyleGAN2 config, expanded deterministically at load time (a few hundred bytes on disk);
  * "gance_b200/variables-v1": {variables: {tf_name: ndarray}} - explicit weights keyed by the TF
    variable names of SURVEY.md Appendix A.2 (what a converted real pickle would contain);

and, third, a REAL StyleGAN2 pickle: tuple slot 2 is then a `dnnlib.tflib.network.Network`, whose pickled state
(static kwargs, components, (name, ndarray) variable list) is captured WITHOUT TensorFlow or dnnlib by the restricted
unpickler of `tf_pickle.py` and renamed to the same TF variable names. Every file is read through that allow-listed
unpickler: a network pickle from an untrusted source cannot run code here.
"""

import pickle
from pathlib import Path
from typing import Dict, NamedTuple

import numpy as np

from gance_b200.networks import tf_pickle
from gance_b200.networks.generator_spec import (
    GeneratorSpec,
    random_init_variables,
    variable_shapes,
)

SYNTHETIC_FORMAT = "gance_b200/synthetic-v1"
VARIABLES_FORMAT = "gance_b200/variables-v1"
NETWORK_SUFFIX = ".pkl"  # network_functions.py:38


class LoadedNetwork(NamedTuple):
    """What `load_network_network` hands to `wrap_loaded_network`."""

    spec: GeneratorSpec
    variables: Dict[str, np.ndarray]
    source: str

    @property
    def input_shape(self):
        """Mirrors `Gs.input_shape` ([None, 512]); network_functions.py:191 reads index 1."""
        return [None, 512]


def write_synthetic_network(
    path: Path, resolution: int, seed: int, perturbed: bool = False, config: str = "config-f"
) -> Path:
    """Write a random-init network description; the (G, D, Gs) tuple slots 0 and 1 stay None."""
    record = {
        "format": SYNTHETIC_FORMAT,
        "resolution": int(resolution),
        "config": config,
        "seed": int(seed),
        "perturbed": bool(perturbed),
    }
    path = Path(path)
    with open(path, "wb") as handle:
        pickle.dump((None, None, record), handle)
    return path


def write_variables_network(path: Path, variables: Dict[str, np.ndarray], config: str = "config-f") -> Path:
    record = {"format": VARIABLES_FORMAT, "config": config, "variables": dict(variables)}
    path = Path(path)
    with open(path, "wb") as handle:
        pickle.dump((None, None, record), handle)
    return path


def _spec_from_variables(variables: Dict[str, np.ndarray], config: str) -> GeneratorSpec:
    resolution = 4
    while f"G_synthesis/{resolution * 2}x{resolution * 2}/Conv1/weight" in variables:
        resolution *= 2
    spec = GeneratorSpec(resolution, config)
    expected = variable_shapes(spec)
    missing = sorted(set(expected) - set(variables))
    if missing:
        raise ValueError(f"network is missing variables, first few: {missing[:4]}")
    for name, shape in expected.items():
        if tuple(np.shape(variables[name])) != tuple(shape):
            raise ValueError(f"variable {name} has shape {np.shape(variables[name])}, expected {shape}")
    return spec


def read_network(path: Path) -> LoadedNetwork:
    """`pickle.load(f)[2]` and expansion into TF-named float32 variables."""
    path = Path(path)
    try:
        payload = tf_pickle.restricted_load(path.read_bytes())
    except pickle.UnpicklingError as error:
        raise ValueError(f"{path}: refused to unpickle: {error}") from error
    try:
        record = payload[2]
    except (TypeError, IndexError, KeyError) as error:
        raise ValueError(f"{path} is not a (G, D, Gs) network pickle") from error
    if isinstance(record, tf_pickle.CapturedNetwork):  # a real TF1 StyleGAN2 pickle
        variables = tf_pickle.variables_from_tf_network(record)
        return LoadedNetwork(_spec_from_variables(variables, tf_pickle.fmap_config(record)), variables, str(path))
    if not isinstance(record, dict) or "format" not in record:
        raise ValueError(f"{path}: unsupported network payload of type {type(record).__name__}")
    if record["format"] == SYNTHETIC_FORMAT:
        spec = GeneratorSpec(record["resolution"], record.get("config", "config-f"))
        variables = random_init_variables(spec, record["seed"], record.get("perturbed", False))
        return LoadedNetwork(spec, variables, str(path))
    if record["format"] == VARIABLES_FORMAT:
        variables = {k: np.asarray(v, dtype=np.float32) for k, v in record["variables"].items()}
        return LoadedNetwork(_spec_from_variables(variables, record.get("config", "config-f")), variables, str(path))
    raise ValueError(f"{path}: unknown network format {record['format']!r}")
