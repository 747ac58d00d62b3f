"""
StyleGAN2 generator description: channel schedule, TF variable names/shapes, and synthetic
(random-init) networks.

The reference loads `Gs = pickle.load(f)[2]` (network_functions.py:108-110) out of NVlabs/stylegan2
pickles that are not available offline, so every network this build runs is a
*random-init* network of a named config. Variable names and shapes follow the TF graph of
`G_main` / `G_mapping` / `G_synthesis_stylegan2` so that a real pickle's
`(name, ndarray)` list can be dropped into the same container later.

Two initialisers:
  * `tf_faithful` – what the TF initialisers give: weights N(0, 1/lrmul), biases 0, mod_bias 0,
    noise_strength 0, dlatent_avg 0, const N(0,1), noise buffers N(0,1).
  * `perturbed` – additionally biases N(0,0.1), noise_strength U(0,0.1), dlatent_avg N(0,0.5),
    mod_bias N(0,0.1), so the bias / noise / truncation paths are exercised by parity tests.
"""

from typing import Dict, List, NamedTuple, Tuple

import numpy as np

LATENT_SIZE = 512
DLATENT_SIZE = 512
MAPPING_LAYERS = 8
MAPPING_LRMUL = 0.01
NUM_CHANNELS = 3

# fmap_base per config (Appendix A.3): config-f doubles the feature maps of configs a-e.
FMAP_BASE = {"config-f": 16 << 10, "config-e": 8 << 10}


class GeneratorSpec(NamedTuple):
    """Static shape of one generator."""

    resolution: int
    config: str = "config-f"
    fmap_max: int = 512

    @property
    def resolution_log2(self) -> int:
        log2 = int(np.log2(self.resolution))
        if 2**log2 != self.resolution or self.resolution < 8:
            raise ValueError(f"resolution must be a power of two >= 8, got {self.resolution}")
        return log2

    @property
    def num_ws(self) -> int:
        """Number of rows of w+ the synthesis network consumes: 2*log2(R) - 2 (18 @ 1024)."""
        return 2 * self.resolution_log2 - 2

    def nf(self, stage: int) -> int:
        """Feature maps at a stage: clip(fmap_base / 2**stage, 1, fmap_max). (Appendix A.3)"""
        return int(np.clip(int(FMAP_BASE[self.config] / (2.0**stage)), 1, self.fmap_max))

    def block_channels(self, res_log2: int) -> int:
        """Output channels of the block working at resolution 2**res_log2."""
        return self.nf(res_log2 - 1)


class ConvLayerSpec(NamedTuple):
    """One modulated 3x3 convolution of the synthesis network, in execution order."""

    name: str  # TF scope, e.g. "G_synthesis/64x64/Conv0_up"
    w_index: int  # row of w+ that styles this layer
    noise_index: int  # index i of G_synthesis/noise{i}
    in_channels: int
    out_channels: int
    in_resolution: int
    out_resolution: int
    up: bool


class ToRGBSpec(NamedTuple):
    name: str
    w_index: int
    in_channels: int
    resolution: int


def synthesis_layers(spec: GeneratorSpec) -> Tuple[List[ConvLayerSpec], List[ToRGBSpec]]:
    """
    Layer list of G_synthesis_stylegan2 with architecture='skip' (Appendix A.4):
    4x4/Conv (w0), 4x4/ToRGB (w1), then per resolution r: Conv0_up (w[2r-5]), Conv1 (w[2r-4]),
    ToRGB (w[2r-3]).
    """
    convs = [
        ConvLayerSpec("G_synthesis/4x4/Conv", 0, 0, spec.nf(1), spec.nf(1), 4, 4, False)
    ]
    rgbs = [ToRGBSpec("G_synthesis/4x4/ToRGB", 1, spec.nf(1), 4)]
    for r in range(3, spec.resolution_log2 + 1):
        res = 2**r
        cin, cout = spec.block_channels(r - 1), spec.block_channels(r)
        scope = f"G_synthesis/{res}x{res}"
        convs.append(
            ConvLayerSpec(f"{scope}/Conv0_up", 2 * r - 5, 2 * r - 5, cin, cout, res // 2, res, True)
        )
        convs.append(ConvLayerSpec(f"{scope}/Conv1", 2 * r - 4, 2 * r - 4, cout, cout, res, res, False))
        rgbs.append(ToRGBSpec(f"{scope}/ToRGB", 2 * r - 3, cout, res))
    return convs, rgbs


def variable_shapes(spec: GeneratorSpec) -> Dict[str, Tuple[int, ...]]:
    """TF variable name -> shape for the whole generator (Appendix A.2). Conv weights are HWIO."""
    shapes: Dict[str, Tuple[int, ...]] = {}
    for i in range(MAPPING_LAYERS):
        fan_in = LATENT_SIZE if i == 0 else DLATENT_SIZE
        shapes[f"G_mapping/Dense{i}/weight"] = (fan_in, DLATENT_SIZE)
        shapes[f"G_mapping/Dense{i}/bias"] = (DLATENT_SIZE,)
    shapes["dlatent_avg"] = (DLATENT_SIZE,)
    shapes["G_synthesis/4x4/Const/const"] = (1, spec.nf(1), 4, 4)
    convs, rgbs = synthesis_layers(spec)
    for conv in convs:
        shapes[f"{conv.name}/weight"] = (3, 3, conv.in_channels, conv.out_channels)
        shapes[f"{conv.name}/mod_weight"] = (DLATENT_SIZE, conv.in_channels)
        shapes[f"{conv.name}/mod_bias"] = (conv.in_channels,)
        shapes[f"{conv.name}/noise_strength"] = ()
        shapes[f"{conv.name}/bias"] = (conv.out_channels,)
        side = conv.out_resolution
        shapes[f"G_synthesis/noise{conv.noise_index}"] = (1, 1, side, side)
    for rgb in rgbs:
        shapes[f"{rgb.name}/weight"] = (1, 1, rgb.in_channels, NUM_CHANNELS)
        shapes[f"{rgb.name}/mod_weight"] = (DLATENT_SIZE, rgb.in_channels)
        shapes[f"{rgb.name}/mod_bias"] = (rgb.in_channels,)
        shapes[f"{rgb.name}/bias"] = (NUM_CHANNELS,)
    return shapes


def random_init_variables(
    spec: GeneratorSpec, seed: int, perturbed: bool = False
) -> Dict[str, np.ndarray]:
    """
    Deterministic random-init variables (float32) for `spec`. Drawn in sorted-name order from
    `np.random.RandomState(seed)` so the oracle and the CUDA path get bit-identical arrays.
    """
    rng = np.random.RandomState(seed)
    out: Dict[str, np.ndarray] = {}
    for name, shape in sorted(variable_shapes(spec).items()):
        leaf = name.rsplit("/", 1)[-1]
        if leaf == "weight":
            std = 1.0 / MAPPING_LRMUL if name.startswith("G_mapping/") else 1.0
            value = rng.randn(*shape) * std
        elif leaf in ("mod_weight", "const") or (leaf.startswith("noise") and leaf != "noise_strength"):
            value = rng.randn(*shape)
        elif leaf == "noise_strength":
            value = rng.uniform(0.0, 0.1, size=shape) if perturbed else np.zeros(shape)
        elif leaf == "mod_bias":
            value = rng.randn(*shape) * 0.1 if perturbed else np.zeros(shape)
        elif leaf == "bias":
            if not perturbed:
                value = np.zeros(shape)
            elif name.startswith("G_mapping/"):
                # stored value is multiplied by lrmul=0.01 at run time
                value = rng.randn(*shape) * 0.1 / MAPPING_LRMUL
            else:
                value = rng.randn(*shape) * 0.1
        elif leaf == "dlatent_avg":
            value = rng.randn(*shape) * 0.5 if perturbed else np.zeros(shape)
        else:
            raise ValueError(f"no initialiser for variable {name}")
        out[name] = np.asarray(value, dtype=np.float32)
    return out
