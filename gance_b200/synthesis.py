"""
G_mapping + G_synthesis on one B200, driving libgance_b200.so through ctypes. PyTorch only owns
the device memory and the stream; every arithmetic step is one of our kernels.

Replaces the TF1 graph the reference runs in `wrap_loaded_network`
(/root/reference/gance/network_interface/network_functions.py:114-192):
  create_image_vector -> Gs.run(z, None, truncation_psi=1.2)                  -> `run_vectors`
  create_image_matrix -> Gs.components.synthesis.run(w+, randomize_noise=False) -> `run_matrices`
Both accept a whole batch of frames (the reference renders one frame per call).

Data flow per batch (NHWC activations, `mode` = "bf16" | "fp16" | "fp32"; DESIGN.md has the full table):
  styles/demod for every layer      gance_style_demod   (fp32)
  const * s0                        gance_const_input
  4x4 conv                          gance_modconv3x3            -> x (unscaled)
  per resolution:
      Conv0_up                      gance_modconv3x3_up2_fir    -> x * s_conv1
      Conv1                         gance_modconv3x3            -> x (unscaled)
      ToRGB + skip (+ uint8)        gance_torgb_skip
"""

import copy
import os
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np
import torch

from gance_b200 import _lib
from gance_b200.networks.generator_spec import (
    MAPPING_LAYERS,
    MAPPING_LRMUL,
    GeneratorSpec,
    synthesis_layers,
)

MODES = {
    "fp32": (_lib.GANCE_F32, torch.float32),
    # fp32-grade validation mode on the tensor core: tcgen05 kind::tf32 with hi/lo operand splitting ("3xTF32"); fp32
    # activations and fp32 bandwidth kernels exactly as "fp32", only the convolutions change (csrc/modconv_tf32.cu)
    "tf32": (_lib.GANCE_F32, torch.float32),
    "bf16": (_lib.GANCE_BF16, torch.bfloat16),
    "fp16": (_lib.GANCE_F16, torch.float16),
}
NOISE_MODES = ("const", "random", "provided", "none")


def fold_up_weight(hwio: torch.Tensor) -> torch.Tensor:
    """
    [3,3,Cin,Cout] (runtime-scaled HWIO) -> [4*Cout,3,3,Cin]: the four phase weights of the folded
    up-convolution (include/gance_b200.h, gance_modconv3x3_up2_folded), row n = (py*2+px)*Cout + co.
    Folded in float64, rounded to the operand type once.
    """
    w = hwio.double()
    taps = (
        torch.tensor([[0.75, 0.25, 0.0], [0.25, 0.75, 0.75], [0.0, 0.0, 0.25]], dtype=torch.float64, device=w.device),
        torch.tensor([[0.25, 0.0, 0.0], [0.75, 0.75, 0.25], [0.0, 0.25, 0.75]], dtype=torch.float64, device=w.device),
    )
    phases = [torch.einsum("yk,xl,klio->oyxi", taps[py], taps[px], w) for py in range(2) for px in range(2)]
    return torch.cat(phases, dim=0).contiguous().float()


PHASE_CHUNK = 64  # output channels per accumulator tile of gance_modconv3x3_up2_phase


def phase_up_weight(hwio: torch.Tensor, chunk: int = PHASE_CHUNK) -> torch.Tensor:
    """
    [3,3,Cin,Cout] (runtime-scaled HWIO) -> [9*Cout, Cin]: the B operand of the 4-phase transposed-convolution GEMM
    (include/gance_b200.h, gance_modconv3x3_up2_phase). The input shifted by (dy,dx) feeds phase (a,b) through tap
    w[iy(dy,a)][ix(dx,b)], iy(0,0)=2, iy(0,1)=1, iy(1,0)=0; rows are ordered shift -> `chunk`-channel chunk -> phase
    block (accumulator column blocks are [P01,P00,P10,P11]) -> channel.
    """
    tap = {(0, 0): 2, (0, 1): 1, (1, 0): 0}
    block_phase = [(0, 1), (0, 0), (1, 0), (1, 1)]
    shifts = [((0, 0), [0, 1, 2, 3]), ((1, 0), [0, 1]), ((0, 1), [1, 2]), ((1, 1), [1])]
    cout = hwio.shape[3]
    rows = []
    for (dy, dx), blocks in shifts:
        for nt in range(cout // chunk):
            for block in blocks:
                a, b = block_phase[block]
                rows.append(hwio[tap[(dy, a)], tap[(dx, b)]][:, nt * chunk: (nt + 1) * chunk].t())
    return torch.cat(rows, dim=0).contiguous()


_NONFINITE_FLAGS: Dict[int, torch.Tensor] = {}


def nonfinite_flag(device: torch.device) -> torch.Tensor:
    """
    The device's range-guard flag (include/gance_b200.h, gance_set_nonfinite_flag): an int32 the frame-writing kernels
    set when a pixel of the final fp32 image is inf/NaN, i.e. when an activation overflowed the 16-bit storage type
    anywhere upstream. One per device, registered with the library on first use.
    """
    index = device.index if device.index is not None else torch.cuda.current_device()
    flag = _NONFINITE_FLAGS.get(index)
    if flag is None:
        flag = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", index))
        with torch.cuda.device(index):
            _lib.check(_lib.init(index).gance_set_nonfinite_flag(flag.data_ptr()), "gance_set_nonfinite_flag")
        _NONFINITE_FLAGS[index] = flag
    return flag


def nonfinite_error(mode: str) -> _lib.NonFiniteFramesError:
    return _lib.NonFiniteFramesError(
        f"GANCE_E_NONFINITE (code {_lib.GANCE_E_NONFINITE}): rendered frames hold inf/NaN pixels - an activation left the "
        f"range of the {mode!r} storage type (IEEE half tops out at 65504; a style or weight of this network is too "
        "large for it). Load the network with mode='bf16' (fp32 range, 8-bit significand) or mode='fp32'.")


def _ptr(tensor: Optional[torch.Tensor]) -> Optional[int]:
    return None if tensor is None else tensor.data_ptr()


class GeneratorOnDevice:
    """One generator's weights resident in HBM in kernel layout, plus per-batch work buffers."""

    # smallest input side run as the fused up-convolution (`_up_form`). 4 = every up-convolution with Cout % 64 == 0: at 4^2 / 8^2
    # one 16 x 8-cell tile covers a whole sample (the TMA box is larger than the image, the rest is zero fill) and most of
    # the tile's MMAs are wasted, but those layers are bound by the weight bytes each CTA pulls through L2 - 9 Cin Cout
    # instead of the folded form's 36: 4->8 68 -> 26 us, 8->16 72 -> 45 us at batch 32 (tools/profile_layers.py --fused-min-side)
    fused_min_in_resolution = 4

    def __init__(self, spec: GeneratorSpec, variables: Dict[str, np.ndarray], device: torch.device, mode: str = "fp16",
                 noise_seed: Optional[int] = None, up_form: Optional[str] = None) -> None:
        """
        up_form: None picks the measured-best kernel per up-convolution (`_up_form`); "folded" | "phase" | "fused" forces
        one where the shape is eligible - an A/B switch for tools/profile_layers.py, not something a run needs.
        noise_seed: Philox key of the "random" noise mode (the vector path's fresh noise per call). None draws one from
        os.urandom, so two processes - or two ranks started from the same script - never share a noise stream; pass an
        integer for a reproducible run (mix the rank in yourself when every rank passes the same number).
        """
        if mode not in MODES:
            raise ValueError(f"mode must be one of {sorted(MODES)}, got {mode!r}")
        if device.type != "cuda":
            raise _lib.GanceError("GeneratorOnDevice needs a CUDA device: there is no CPU fallback")
        if up_form not in (None, "folded", "phase", "fused"):
            raise ValueError(f"up_form must be None, 'folded', 'phase' or 'fused', got {up_form!r}")
        self.forced_up_form = up_form or ""
        self.spec = spec
        self.device = device
        self.mode = mode
        self.dtype_code, self.torch_dtype = MODES[mode]
        self.lib = _lib.init(device.index if device.index is not None else torch.cuda.current_device())
        self.nonfinite = nonfinite_flag(device)
        self.num_ws = spec.num_ws
        self.resolution = spec.resolution
        self.convs, self.rgbs = synthesis_layers(spec)
        self._buffers: Dict[tuple, torch.Tensor] = {}
        self.profile_events: Optional[list] = None  # set to [] to collect (name, cuda event) marks
        if noise_seed is None:
            noise_seed = int.from_bytes(os.urandom(8), "little")
        self.noise_seed = int(noise_seed) & 0x7FFF_FFFF_FFFF_FFFF
        # Philox stream position, on the device: advances with every batch (also across CUDA-graph replays)
        self.noise_counter = torch.zeros(1, dtype=torch.int64, device=device)
        self._graphs: Dict[tuple, tuple] = {}
        self._renderers: Dict[tuple, "BatchRenderer"] = {}
        self._contexts: Dict[int, "GeneratorOnDevice"] = {}
        self.launches_per_replay = 0
        self._noise_stream: Optional[torch.cuda.Stream] = None  # side stream of the forked noise launch (run_vectors)
        self.graph_captures = 0  # how many CUDA graphs this generator has captured (tests: switching must not re-capture)
        self._upload(variables)

    # ---- weights ---------------------------------------------------------------------------------
    def _upload(self, v: Dict[str, np.ndarray]) -> None:
        dev = self.device

        def f32(array) -> torch.Tensor:
            return torch.as_tensor(np.ascontiguousarray(array, dtype=np.float32)).to(dev)

        # G_mapping: runtime coefficient lrmul/sqrt(fan_in) folded into W, lrmul into b (get_weight)
        weights = [v[f"G_mapping/Dense{i}/weight"] for i in range(MAPPING_LAYERS)]
        biases = [v[f"G_mapping/Dense{i}/bias"] for i in range(MAPPING_LAYERS)]
        self.map_w = f32(np.stack([w * (MAPPING_LRMUL / np.sqrt(w.shape[0])) for w in weights]))
        self.map_b = f32(np.stack([b * MAPPING_LRMUL for b in biases]))
        self.dlatent_avg = f32(v["dlatent_avg"])
        self.const = f32(np.transpose(v["G_synthesis/4x4/Const/const"][0], (1, 2, 0)))  # [4,4,C]

        self.conv_w: List[torch.Tensor] = []
        self.conv_wsq: List[torch.Tensor] = []
        self.conv_mod_w: List[torch.Tensor] = []
        self.conv_mod_b: List[torch.Tensor] = []
        self.conv_bias: List[torch.Tensor] = []
        self.conv_noise: List[torch.Tensor] = []
        self.conv_strength: List[float] = []
        for conv in self.convs:
            hwio = torch.as_tensor(v[f"{conv.name}/weight"], dtype=torch.float32).to(dev)
            hwio = hwio * float(1.0 / np.sqrt(9 * conv.in_channels))  # runtime_coef
            self.conv_wsq.append((hwio * hwio).sum(dim=(0, 1)).contiguous())  # [Cin,Cout]
            if self.mode == "fp32":
                self.conv_w.append(hwio.contiguous())  # [3,3,Cin,Cout]
            elif self.mode == "tf32":
                # K-major [Cout,3,3,Cin], split like the activations: hi = low 13 significand bits cleared, lo = w - hi
                kmajor = hwio.permute(3, 0, 1, 2).contiguous()
                hi = (kmajor.view(torch.int32) & -8192).view(torch.float32)
                self.conv_w.append(torch.stack([hi, kmajor - hi]).contiguous())  # [2,Cout,3,3,Cin]
            elif conv.up and self._up_form(conv) != "folded":
                chunk = PHASE_CHUNK if self._up_form(conv) == "phase" else min(PHASE_CHUNK, conv.out_channels)
                self.conv_w.append(phase_up_weight(hwio, chunk).to(self.torch_dtype))  # [9*Cout,Cin]
            elif conv.up:
                self.conv_w.append(fold_up_weight(hwio).to(self.torch_dtype))  # [4*Cout,3,3,Cin]
            else:
                self.conv_w.append(hwio.permute(3, 0, 1, 2).contiguous().to(self.torch_dtype))  # [Cout,3,3,Cin]
            self.conv_mod_w.append(f32(v[f"{conv.name}/mod_weight"] * (1.0 / np.sqrt(512.0))))
            self.conv_mod_b.append(f32(v[f"{conv.name}/mod_bias"]))
            self.conv_bias.append(f32(v[f"{conv.name}/bias"]))
            self.conv_noise.append(f32(v[f"G_synthesis/noise{conv.noise_index}"][0, 0]))
            self.conv_strength.append(float(v[f"{conv.name}/noise_strength"]))
        self.rgb_w: List[torch.Tensor] = []
        self.rgb_mod_w: List[torch.Tensor] = []
        self.rgb_mod_b: List[torch.Tensor] = []
        self.rgb_bias: List[torch.Tensor] = []
        for rgb in self.rgbs:
            w = v[f"{rgb.name}/weight"][0, 0] * (1.0 / np.sqrt(rgb.in_channels))  # [C,3]
            self.rgb_w.append(f32(w))
            self.rgb_mod_w.append(f32(v[f"{rgb.name}/mod_weight"] * (1.0 / np.sqrt(512.0))))
            self.rgb_mod_b.append(f32(v[f"{rgb.name}/mod_bias"]))
            self.rgb_bias.append(f32(v[f"{rgb.name}/bias"]))

        self._build_style_tables()

    def _up_form(self, conv) -> str:
        """
        Which kernel an up-convolution runs as (include/gance_b200.h):
          "fused"  - 4-phase GEMM with the FIR applied on chip (upconv_fused.cu): one launch, x read once, output written
                     once, 1.52x the algorithmic MACs. Default for every input side with Cout % 64 == 0 (measured on
                     B200, batch 32: 256->512 710 us vs 1280 phase / 1080 folded; 128->256 525 vs 794; 64->128 460 vs
                     595; 32->64 282 vs 282 - profiles/README.md).
          "phase"  - 4-phase GEMM + separate FIR pass through a 16-bit scratch tensor (upconv_phase.cu).
          "folded" - one N = 4*Cout SAME convolution at the input resolution, 4x the algorithmic MACs: 512->1024
                     (Cout = 32: the fused epilogue is issue-bound there, 1756 vs 1118 us).
        The constructor's `up_form` forces a form where the shape is eligible (A/B measurements).
        """
        forced = self.forced_up_form
        phase_ok = conv.in_resolution >= 32 and conv.in_channels % 64 == 0 and conv.out_channels % 64 == 0
        fused_ok = (conv.in_resolution >= self.fused_min_in_resolution and conv.in_channels % 64 == 0
                    and (conv.out_channels % 64 == 0 or conv.out_channels == 32))
        if forced == "folded":
            return "folded"
        if forced == "phase":
            return "phase" if phase_ok else "folded"
        if forced == "fused":
            return "fused" if fused_ok else "folded"
        if fused_ok and conv.out_channels % 64 == 0:  # (16->32: 100 us fused vs 132 us folded at batch 32)
            return "fused"
        return "folded"

    def _build_style_tables(self) -> None:
        """Concatenated affine weights + chunk descriptors for gance_styles_all (include/gance_b200.h)."""
        dev = self.device
        layers = [(c.w_index, c.in_channels, self.conv_mod_w[i], self.conv_mod_b[i]) for i, c in enumerate(self.convs)]
        layers += [(r.w_index, r.in_channels, self.rgb_mod_w[i], self.rgb_mod_b[i]) for i, r in enumerate(self.rgbs)]
        self.style_col0: List[int] = []
        style_chunks, col = [], 0
        for w_index, channels, _, _ in layers:
            self.style_col0.append(col)
            style_chunks += [[w_index, col, channels, first] for first in range(0, channels, 32)]
            col += channels
        self.style_cols = col
        self.mod_w_cat = torch.cat([w for _, _, w, _ in layers], dim=1).contiguous()  # [512, total]
        self.mod_b_cat = torch.cat([b for _, _, _, b in layers]).contiguous()
        self.style_chunks = torch.tensor(style_chunks, dtype=torch.int32).to(dev)
        demod_chunks, self.demod_col0, dcol, offset = [], [], 0, 0
        for i, conv in enumerate(self.convs):
            self.demod_col0.append(dcol)
            demod_chunks += [[offset, conv.in_channels, conv.out_channels, self.style_col0[i], first, dcol]
                             for first in range(0, conv.out_channels, 32)]
            dcol += conv.out_channels
            offset += conv.in_channels * conv.out_channels
        self.demod_cols = dcol
        self.wsq_cat = torch.cat([w.reshape(-1) for w in self.conv_wsq]).contiguous()
        self.demod_chunks = torch.tensor(demod_chunks, dtype=torch.int32).to(dev)

    def weight_bytes(self) -> int:
        tensors = [self.map_w, self.map_b, self.dlatent_avg, self.const]
        for group in (self.conv_w, self.conv_wsq, self.conv_mod_w, self.conv_mod_b, self.conv_bias,
                      self.conv_noise, self.rgb_w, self.rgb_mod_w, self.rgb_mod_b, self.rgb_bias):
            tensors.extend(group)
        return sum(t.numel() * t.element_size() for t in tensors)

    # ---- buffers ---------------------------------------------------------------------------------
    def _buffer(self, name: str, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        """
        Work buffer `name` viewed as `shape`. One flat allocation per (name, dtype) sized for the LARGEST request so
        far: batches of different sizes (the ragged per-network groups of a model-switching run) share it instead of
        each batch size owning a full set of activations (10.7 GB per set at 1024^2 x 32). Growing a buffer moves
        it, so the captured graphs (which hold addresses) are dropped and re-captured on next use; `reserve(batch)`
        sizes everything up front.
        """
        numel = 1
        for extent in shape:
            numel *= int(extent)
        key = (name, dtype)
        flat = self._buffers.get(key)
        if flat is None or flat.numel() < numel:
            if flat is not None:
                self._graphs.clear()
            flat = torch.empty((max(numel, 1),), dtype=dtype, device=self.device)
            self._buffers[key] = flat
        return flat[:numel].view(tuple(shape))

    def release_buffers(self) -> None:
        self._graphs.clear()  # captured graphs hold the work buffers' addresses
        self._renderers.clear()
        self._buffers.clear()
        for context in self._contexts.values():
            context.release_buffers()

    def graph_capture_count(self) -> int:
        """CUDA graphs captured by this generator and its contexts so far (tests: switching must not re-capture)."""
        return self.graph_captures + sum(context.graph_capture_count() for context in self._contexts.values())

    def context(self, index: int) -> "GeneratorOnDevice":
        """
        A view of this generator with its OWN work buffers, captured graphs and noise stream, sharing the resident
        weights: batches rendered through different contexts do not touch each other's memory, so they can run
        concurrently on different CUDA streams (the tail of one batch's kernels and its small latency-bound launches
        overlap the next batch's). Context 0 is the generator itself.
        """
        if index == 0:
            return self
        found = self._contexts.get(index)
        if found is None:
            found = copy.copy(self)
            found._buffers, found._graphs, found._renderers, found._contexts = {}, {}, {}, {}
            found._noise_stream = None  # its own side stream for the forked noise launch
            found.profile_events = None
            found.graph_captures, found.launches_per_replay = 0, 0
            found.noise_counter = torch.zeros(1, dtype=torch.int64, device=self.device)
            found.noise_seed = (self.noise_seed + index * 0x9E37_79B9_7F4A_7C15) & 0x7FFF_FFFF_FFFF_FFFF
            self._contexts[index] = found
        return found

    def renderer(self, max_batch: int = 32, truncation_psi: float = 1.2, output_side_length: Optional[int] = None,
                 noise_mode: Optional[str] = None) -> "BatchRenderer":
        """The BatchRenderer of this generator for a (batch, psi, output side, noise) combination, created once."""
        side = None if output_side_length in (None, self.resolution) else int(output_side_length)
        key = (int(max_batch), float(truncation_psi), side, noise_mode)
        found = self._renderers.get(key)
        if found is None:
            found = self._renderers[key] = BatchRenderer(self, max_batch=max_batch, truncation_psi=truncation_psi,
                                                         noise_mode=noise_mode, output_side_length=side)
        return found

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def check_finite(self) -> None:
        """
        Synchronising read of the device's range-guard flag: raises NonFiniteFramesError (GANCE_E_NONFINITE) when any
        frame rendered on this device since the last check held inf/NaN pixels, and clears the flag. `BatchRenderer`
        reads the flag back with each batch's frames instead, without an extra synchronisation.
        """
        if int(self.nonfinite.item()) != 0:
            self.nonfinite.zero_()
            raise nonfinite_error(self.mode)

    def _mark(self, name: str) -> None:
        """Per-layer timing for tools/profile_layers.py: an event after the step called `name`."""
        if self.profile_events is not None:
            event = torch.cuda.Event(enable_timing=True)
            event.record(torch.cuda.current_stream(self.device))
            self.profile_events.append((name, event))

    # ---- entry points ----------------------------------------------------------------------------
    def mapping(self, z: torch.Tensor, truncation_psi: float = 1.2) -> torch.Tensor:
        """G_mapping + truncation lerp: z [B,512] -> w+ [B,num_ws,512] (fp32, device)."""
        z = z.to(self.device, torch.float32).reshape(-1, 512).contiguous()
        batch = z.shape[0]
        wplus = torch.empty((batch, self.num_ws, 512), dtype=torch.float32, device=self.device)
        _lib.check(
            self.lib.gance_mapping(z.data_ptr(), batch, self.map_w.data_ptr(), self.map_b.data_ptr(),
                                   self.dlatent_avg.data_ptr(), float(truncation_psi), self.num_ws,
                                   wplus.data_ptr(), self._stream()),
            "gance_mapping",
        )
        return wplus

    def synthesis(
        self,
        wplus: torch.Tensor,
        noise_mode: str = "const",
        noise: Optional[Sequence[torch.Tensor]] = None,
        return_float_image: bool = False,
        out: Optional[torch.Tensor] = None,
        _prelaunched_noise=None,
    ):
        """
        G_synthesis + convert_images_to_uint8: w+ [B,num_ws,512] -> uint8 frames [B,R,R,3] on the
        device (`out` if given). noise_mode: "const" = the network's noise buffers
        (randomize_noise=False, the matrix path), "random" = fresh N(0,1) per call (the vector
        path's default), "provided" = `noise[layer]` of shape [B,H,W], "none" = no noise.
        With return_float_image the fp32 skip sum [B,R,R,4] (RGB + zero pad) is returned as well.
        """
        if noise_mode not in NOISE_MODES:
            raise ValueError(f"noise_mode must be one of {NOISE_MODES}")
        wplus = wplus.to(self.device, torch.float32).reshape(-1, self.num_ws, 512).contiguous()
        batch = wplus.shape[0]
        lib, stream, code, dt = self.lib, self._stream(), self.dtype_code, self.torch_dtype
        res = self.resolution
        frames = out if out is not None else torch.empty((batch, res, res, 3), dtype=torch.uint8, device=self.device)
        if tuple(frames.shape) != (batch, res, res, 3) or frames.dtype != torch.uint8 or not frames.is_contiguous():
            raise ValueError("out must be a contiguous uint8 [B,R,R,3] tensor")

        # -- styles and demodulation coefficients of every layer
        styles_all = self._buffer("styles_all", (batch * self.style_cols,), torch.float32)
        demod_all = self._buffer("demod_all", (batch * self.demod_cols,), torch.float32)
        _lib.check(
            lib.gance_styles_all(wplus.data_ptr(), batch, self.num_ws, self.mod_w_cat.data_ptr(), self.style_cols,
                                 self.mod_b_cat.data_ptr(), self.style_chunks.data_ptr(), self.style_chunks.shape[0],
                                 styles_all.data_ptr(), self.wsq_cat.data_ptr(), self.demod_chunks.data_ptr(),
                                 self.demod_chunks.shape[0], demod_all.data_ptr(), stream),
            "gance_styles_all",
        )

        def block(flat: torch.Tensor, col0: int, channels: int) -> torch.Tensor:
            return flat[batch * col0: batch * (col0 + channels)].view(batch, channels)

        styles = [block(styles_all, self.style_col0[i], c.in_channels) for i, c in enumerate(self.convs)]
        demods = [block(demod_all, self.demod_col0[i], c.out_channels) for i, c in enumerate(self.convs)]
        rgb_styles = [block(styles_all, self.style_col0[len(self.convs) + i], r.in_channels)
                      for i, r in enumerate(self.rgbs)]
        self._mark("styles")

        # fresh noise of every layer in one launch: already running on the side stream (run_vectors), or launched here
        random_noise: Dict[int, torch.Tensor] = {}
        if noise_mode == "random":
            random_noise, noise_stream = _prelaunched_noise if _prelaunched_noise is not None else \
                self._launch_random_noise(batch, fork=False)
            if noise_stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(noise_stream)  # join before the first layer reads it

        def layer_noise(i: int, side: int):
            if noise_mode == "none" or (noise_mode != "provided" and self.conv_strength[i] == 0.0):
                return None, 0
            if noise_mode == "const":
                return self.conv_noise[i], 0
            if noise_mode == "random":
                return random_noise[i], side * side
            tensor = noise[i].to(self.device, torch.float32).reshape(batch, side, side).contiguous()
            return tensor, side * side

        # -- 4x4 block
        c0 = self.convs[0]
        x_in = self._buffer("const_in", (batch, 4, 4, c0.in_channels), dt)
        _lib.check(lib.gance_const_input(self.const.data_ptr(), styles[0].data_ptr(), batch, c0.in_channels, code,
                                         x_in.data_ptr(), stream), "gance_const_input")
        x = self._buffer("act4", (batch, 4, 4, c0.out_channels), dt)
        nz, nz_stride = layer_noise(0, 4)
        last = len(self.rgbs) - 1

        def plain_conv(i: int, conv, side: int, x_src: torch.Tensor, x_dst: torch.Tensor, is_last: bool, nz, nz_stride):
            """
            A Conv layer whose ToRGB runs as its own kernel (Cout > 256, and the fp32 mode). In the 16-bit modes the
            epilogue also writes the next block's modulated input x * styles[i + 1] (returned; None otherwise), which
            used to be a gance_scale_channels pass over x.
            """
            if self.mode == "tf32":
                self._tf32_conv(i, conv, batch, side, side, 1, x_src, demods[i], nz, nz_stride, None, x_dst, stream)
                return None
            dual = self.mode != "fp32" and not is_last and conv.in_channels % 64 == 0 and conv.in_channels >= 128
            if not dual:
                _lib.check(
                    lib.gance_modconv3x3(x_src.data_ptr(), code, batch, side, side, conv.in_channels, conv.out_channels, 1,
                                         self.conv_w[i].data_ptr(), demods[i].data_ptr(), _ptr(nz), nz_stride,
                                         self.conv_strength[i], self.conv_bias[i].data_ptr(), None, x_dst.data_ptr(),
                                         stream),
                    f"gance_modconv3x3({side})",
                )
                return None
            scaled = self._buffer(f"scaled{side}", (batch, side, side, conv.out_channels), dt)
            _lib.check(
                lib.gance_modconv3x3_dual(x_src.data_ptr(), code, batch, side, side, conv.in_channels, conv.out_channels, 1,
                                          self.conv_w[i].data_ptr(), demods[i].data_ptr(), _ptr(nz), nz_stride,
                                          self.conv_strength[i], self.conv_bias[i].data_ptr(), styles[i + 1].data_ptr(),
                                          scaled.data_ptr(), x_dst.data_ptr(), stream),
                f"gance_modconv3x3_dual({side})",
            )
            return scaled

        x_scaled_next = plain_conv(0, c0, 4, x_in, x, last == 0, nz, nz_stride)
        self._mark("4/conv")
        want_float = return_float_image or last != 0
        y = self._buffer("skip4", (batch, 4, 4, 4), torch.float32) if want_float else None
        _lib.check(
            lib.gance_torgb_skip(x.data_ptr(), code, batch, 4, 4, c0.out_channels, rgb_styles[0].data_ptr(),
                                 self.rgb_w[0].data_ptr(), self.rgb_bias[0].data_ptr(), None, _ptr(y),
                                 frames.data_ptr() if last == 0 else None, stream),
            "gance_torgb_skip(4x4)",
        )
        self._mark("4/torgb")

        # -- one block per resolution
        x_prescaled = x_scaled_next is not None  # True when `x` already carries the next up-convolution's styles
        if x_prescaled:
            x = x_scaled_next
        for block in range(1, len(self.rgbs)):
            up_i, conv_i = 2 * block - 1, 2 * block
            up, conv = self.convs[up_i], self.convs[conv_i]
            side_in, side = up.in_resolution, up.out_resolution
            x_up = self._buffer(f"up{side}", (batch, side, side, up.out_channels), dt)
            nz, nz_stride = layer_noise(up_i, side)
            if self.mode == "tf32":  # FIR-first: upsampled + style-modulated input, then a VALID 3xTF32 convolution
                scratch = self._buffer("upfir", (batch, 2 * side_in + 2, 2 * side_in + 2, up.in_channels), dt)
                _lib.check(lib.gance_upfir2x_pad(x.data_ptr(), code, batch, side_in, side_in, up.in_channels,
                                                 styles[up_i].data_ptr(), scratch.data_ptr(), stream), "gance_upfir2x_pad")
                self._tf32_conv(up_i, up, batch, side, side, 0, scratch, demods[up_i], nz, nz_stride, styles[conv_i], x_up,
                                stream)
            elif self.mode == "fp32":  # validation mode: FIR-first form on CUDA cores
                scratch = self._buffer("upfir", (batch, 2 * side_in + 2, 2 * side_in + 2, up.in_channels), dt)
                _lib.check(
                    lib.gance_modconv3x3_up2_fir(x.data_ptr(), code, batch, side_in, side_in, up.in_channels,
                                                 up.out_channels, styles[up_i].data_ptr(),
                                                 self.conv_w[up_i].data_ptr(), demods[up_i].data_ptr(), _ptr(nz),
                                                 nz_stride, self.conv_strength[up_i],
                                                 self.conv_bias[up_i].data_ptr(), styles[conv_i].data_ptr(),
                                                 scratch.data_ptr(), x_up.data_ptr(), stream),
                    f"gance_modconv3x3_up2_fir({side})",
                )
            else:  # folded up-convolution: one implicit GEMM at the input resolution, N = 4*Cout
                x_scaled = x
                if not x_prescaled:
                    x_scaled = self._buffer(f"scaled{side_in}", (batch, side_in, side_in, up.in_channels), dt)
                    _lib.check(
                        lib.gance_scale_channels(x.data_ptr(), code, batch, side_in, side_in, up.in_channels,
                                                 styles[up_i].data_ptr(), x_scaled.data_ptr(), stream),
                        f"gance_scale_channels({side_in})",
                    )
                form = self._up_form(up)
                if form == "fused":
                    _lib.check(
                        lib.gance_modconv3x3_up2_fused(x_scaled.data_ptr(), code, batch, side_in, side_in, up.in_channels,
                                                       up.out_channels, self.conv_w[up_i].data_ptr(),
                                                       demods[up_i].data_ptr(), _ptr(nz), nz_stride,
                                                       self.conv_strength[up_i], self.conv_bias[up_i].data_ptr(),
                                                       styles[conv_i].data_ptr(), x_up.data_ptr(), stream),
                        f"gance_modconv3x3_up2_fused({side})",
                    )
                elif form == "phase":
                    scratch = self._buffer("phase_t", (batch * (side + 2) * (side + 2) * up.out_channels,), dt)
                    _lib.check(
                        lib.gance_modconv3x3_up2_phase(x_scaled.data_ptr(), code, batch, side_in, side_in, up.in_channels,
                                                       up.out_channels, self.conv_w[up_i].data_ptr(),
                                                       demods[up_i].data_ptr(), _ptr(nz), nz_stride,
                                                       self.conv_strength[up_i], self.conv_bias[up_i].data_ptr(),
                                                       styles[conv_i].data_ptr(), scratch.data_ptr(), x_up.data_ptr(),
                                                       stream),
                        f"gance_modconv3x3_up2_phase({side})",
                    )
                else:
                    _lib.check(
                        lib.gance_modconv3x3_up2_folded(x_scaled.data_ptr(), code, batch, side_in, side_in,
                                                        up.in_channels, up.out_channels, self.conv_w[up_i].data_ptr(),
                                                        demods[up_i].data_ptr(), _ptr(nz), nz_stride,
                                                        self.conv_strength[up_i], self.conv_bias[up_i].data_ptr(),
                                                        styles[conv_i].data_ptr(), x_up.data_ptr(), stream),
                        f"gance_modconv3x3_up2_folded({side})",
                    )
            self._mark(f"{side}/conv0_up")
            nz, nz_stride = layer_noise(conv_i, side)
            is_last = block == last
            want_float = return_float_image or not is_last
            y_out = self._buffer(f"skip{side}", (batch, side, side, 4), torch.float32) if want_float else None
            if self.mode not in ("fp32", "tf32") and conv.out_channels <= 256:
                # Conv1 + ToRGB + skip (+ uint8) in one kernel; its activation output is pre-multiplied by the
                # NEXT up-convolution's styles, and not written at all on the last block (nothing reads it).
                x = None if is_last else self._buffer(f"act{side}", (batch, side, side, conv.out_channels), dt)
                next_style = None if is_last else styles[conv_i + 1]
                _lib.check(
                    lib.gance_modconv3x3_torgb(x_up.data_ptr(), code, batch, side, side, conv.in_channels,
                                               conv.out_channels, self.conv_w[conv_i].data_ptr(),
                                               demods[conv_i].data_ptr(), _ptr(nz), nz_stride,
                                               self.conv_strength[conv_i], self.conv_bias[conv_i].data_ptr(),
                                               _ptr(next_style), _ptr(x), rgb_styles[block].data_ptr(),
                                               self.rgb_w[block].data_ptr(), self.rgb_bias[block].data_ptr(), _ptr(y),
                                               _ptr(y_out), frames.data_ptr() if is_last else None, stream),
                    f"gance_modconv3x3_torgb({side})",
                )
                x_prescaled = True
                self._mark(f"{side}/conv1")
            else:
                x = self._buffer(f"act{side}", (batch, side, side, conv.out_channels), dt)
                x_scaled_next = plain_conv(conv_i, conv, side, x_up, x, is_last, nz, nz_stride)
                x_prescaled = x_scaled_next is not None
                self._mark(f"{side}/conv1")
                _lib.check(
                    lib.gance_torgb_skip(x.data_ptr(), code, batch, side, side, conv.out_channels,
                                         rgb_styles[block].data_ptr(), self.rgb_w[block].data_ptr(),
                                         self.rgb_bias[block].data_ptr(), _ptr(y), _ptr(y_out),
                                         frames.data_ptr() if is_last else None, stream),
                    f"gance_torgb_skip({side})",
                )
                self._mark(f"{side}/torgb")
                if x_prescaled:
                    x = x_scaled_next
            y = y_out
        if return_float_image:
            return frames, y
        return frames

    def _tf32_conv(self, i: int, conv, batch: int, h_out: int, w_out: int, pad: int, x_src: torch.Tensor, demod, nz,
                   nz_stride: int, next_styles, x_dst: torch.Tensor, stream: int) -> None:
        """One convolution of the "tf32" mode: split the (already modulated) fp32 input into hi + lo, then 3xTF32."""
        h_in, w_in = h_out + 2 - 2 * pad, w_out + 2 - 2 * pad
        split = self._buffer("tf32_split", (2, batch, h_in, w_in, conv.in_channels), torch.float32)
        _lib.check(self.lib.gance_split_tf32(x_src.data_ptr(), None, batch, h_in, w_in, conv.in_channels,
                                             split[0].data_ptr(), split[1].data_ptr(), stream), "gance_split_tf32")
        weights = self.conv_w[i]
        _lib.check(
            self.lib.gance_modconv3x3_tf32x3(split[0].data_ptr(), split[1].data_ptr(), batch, h_out, w_out, conv.in_channels,
                                             conv.out_channels, pad, weights[0].data_ptr(), weights[1].data_ptr(),
                                             demod.data_ptr(), _ptr(nz), nz_stride, self.conv_strength[i],
                                             self.conv_bias[i].data_ptr(), _ptr(next_styles), x_dst.data_ptr(), stream),
            f"gance_modconv3x3_tf32x3({h_out})",
        )

    fork_noise = True  # A/B switch of the forked noise launch (bench.py --no-noise-fork)

    def run_vectors(self, z: torch.Tensor, truncation_psi: float = 1.2, noise_mode: str = "random", **kwargs):
        """
        Batched `Gs.run(z, None, truncation_psi=1.2)` (network_functions.py:151-158). The fresh noise of every layer does
        not depend on the latents: its one launch is forked onto a side stream before G_mapping, so it overlaps the eight
        latency-bound dense layers and the style / demodulation launches (a fork-join that a CUDA-graph capture keeps).
        """
        batch = int(z.reshape(-1, 512).shape[0])
        prelaunched = self._launch_random_noise(batch, fork=self.fork_noise) if noise_mode == "random" else None
        return self.synthesis(self.mapping(z, truncation_psi), noise_mode=noise_mode, _prelaunched_noise=prelaunched, **kwargs)

    def _launch_random_noise(self, batch: int, fork: bool):
        """
        N(0,1) noise of every layer with a non-zero strength for one batch, one launch (tf.random_normal per layer in the
        reference's graph). Returns ({layer: [batch,side,side] view}, side stream or None). With `fork` the launch goes to
        this generator's side stream, ordered after everything already queued on the current stream (the previous batch may
        still be reading the noise buffer); the consumer joins with `current.wait_stream(side)`.
        """
        sizes = [(i, c.out_resolution) for i, c in enumerate(self.convs) if self.conv_strength[i] != 0.0]
        total = sum(batch * side * side for _, side in sizes)
        views: Dict[int, torch.Tensor] = {}
        if not total:
            return views, None
        flat = self._buffer("noise_random", ((total + 3) // 4 * 4,), torch.float32)
        side_stream = None
        stream = self._stream()
        if fork:
            if self._noise_stream is None:
                self._noise_stream = torch.cuda.Stream(device=self.device)
            side_stream = self._noise_stream
            side_stream.wait_stream(torch.cuda.current_stream(self.device))
            stream = side_stream.cuda_stream
        _lib.check(self.lib.gance_normal_noise_next(flat.data_ptr(), flat.numel(), self.noise_seed,
                                                    self.noise_counter.data_ptr(), stream), "gance_normal_noise_next")
        start = 0
        for i, side in sizes:
            views[i] = flat[start: start + batch * side * side].view(batch, side, side)
            start += batch * side * side
        return views, side_stream

    def _replay(self, kind: str, latents: torch.Tensor, out: torch.Tensor, truncation_psi: float, noise_mode: str) -> torch.Tensor:
        """
        One batch replayed from a CUDA graph: the ~45 launches are captured once per (kind, batch size, output buffer,
        psi, noise mode) and replayed with one driver call - no per-launch CPU cost and back-to-back kernels on the
        device. `latents` is copied into the graph's static input; frames land in `out`.
        """
        key = (kind, latents.shape[0], out.data_ptr(), float(truncation_psi), noise_mode)
        entry = self._graphs.get(key)
        if entry is None:
            static_in = latents.clone()

            def run() -> None:
                if kind == "vectors":
                    self.run_vectors(static_in, truncation_psi=truncation_psi, noise_mode=noise_mode, out=out)
                else:
                    self.synthesis(static_in, noise_mode=noise_mode, out=out)

            run()  # warm-up: allocates the work buffers
            torch.cuda.current_stream(self.device).synchronize()
            graph = torch.cuda.CUDAGraph()
            launches_before = _lib.launch_count()
            with torch.cuda.graph(graph):
                run()
            self.launches_per_replay = _lib.launch_count() - launches_before  # kernels of ours inside one replay
            self.graph_captures += 1
            entry = (graph, static_in, self.launches_per_replay)
            self._graphs[key] = entry
        graph, static_in, self.launches_per_replay = entry
        static_in.copy_(latents, non_blocking=True)
        graph.replay()
        return out

    def run_vectors_graphed(self, z: torch.Tensor, out: torch.Tensor, truncation_psi: float = 1.2,
                            noise_mode: str = "random") -> torch.Tensor:
        """`run_vectors` as one CUDA-graph replay (see `_replay`)."""
        z = z.to(self.device, torch.float32).reshape(-1, 512)
        return self._replay("vectors", z, out, truncation_psi, noise_mode)

    def run_matrices_graphed(self, wplus: torch.Tensor, out: torch.Tensor, noise_mode: str = "const") -> torch.Tensor:
        """`run_matrices` as one CUDA-graph replay: w+ [B,num_ws,512] -> frames in `out`."""
        wplus = wplus.to(self.device, torch.float32).reshape(-1, self.num_ws, 512)
        return self._replay("matrices", wplus, out, 1.0, noise_mode)

    def run_matrices(self, wplus: torch.Tensor, **kwargs):
        """Batched `Gs.components.synthesis.run(w+, randomize_noise=False)` (network_functions.py:167-169)."""
        return self.synthesis(wplus, noise_mode="const", **kwargs)


class Arrival:
    """
    What `BatchRenderer.submit_into` returns: `synchronize()` blocks until the batch's frames are in the destination and
    raises NonFiniteFramesError (GANCE_E_NONFINITE) if the range-guard flag that travelled with them is set.
    """

    __slots__ = ("event", "flag", "generator")

    def __init__(self, event: torch.cuda.Event, flag: torch.Tensor, generator: "GeneratorOnDevice") -> None:
        self.event, self.flag, self.generator = event, flag, generator

    def synchronize(self) -> None:
        self.event.synchronize()
        if int(self.flag[0]) != 0:
            # batches still in flight on other streams would set the sticky device flag again after it is cleared
            torch.cuda.synchronize(self.generator.device)
            self.flag.zero_()
            self.generator.nonfinite.zero_()
            torch.cuda.synchronize(self.generator.device)
            raise nonfinite_error(self.generator.mode)

    def query(self) -> bool:
        return self.event.query()


class BatchRenderer:
    """
    Host latents in, host frames out, with the three legs overlapped: while batch i is computed,
    batch i-1's frames travel device -> host on a copy stream into pinned memory and batch i+1's
    latents are staged. This is what replaces the reference's per-frame
    `input_queue.put` / `output_queue.get` round trip through a TF child process
    (network_functions.py:289-301), which pickles 3 MB per frame.

    `render(batches)` yields one uint8 [n,R,R,3] array per input batch, in order. Each yielded
    array is a view of a pinned ring buffer and stays valid until the next batch is drawn from the
    iterator (copy it to keep it longer).
    """

    def __init__(self, generator: GeneratorOnDevice, max_batch: int = 32, depth: int = 2,
                 truncation_psi: float = 1.2, noise_mode: Optional[str] = None,
                 output_side_length: Optional[int] = None) -> None:
        """
        output_side_length: when given and different from the network's resolution, every batch is scaled on the
        device with the bicubic kernel (what `scale_square_source_duplicate` does per frame with cv2 in the reference,
        gance/projection_file_blend.py:184-199) before it leaves for the host, and `render` yields [n,S,S,3] arrays.
        """
        self.generator = generator
        self.max_batch = max_batch
        self.depth = max(2, depth)
        self.truncation_psi = truncation_psi
        self.noise_mode = noise_mode
        dev, res = generator.device, generator.resolution
        self.copy_stream = torch.cuda.Stream(device=dev)
        # one generator context and one compute stream per slot: consecutive batches overlap on the device
        self.contexts = [generator.context(slot) for slot in range(self.depth)]
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(self.depth)]
        self.frames_dev = [torch.empty((max_batch, res, res, 3), dtype=torch.uint8, device=dev) for _ in range(self.depth)]
        self.output_side = res if output_side_length is None else int(output_side_length)
        self.resizer = None
        self.scaled_dev: List[torch.Tensor] = []
        if self.output_side != res:
            from gance_b200.image_sources.video_common import FrameResizer

            self.resizer = FrameResizer(dev)
            self.scaled_dev = [torch.empty((max_batch, self.output_side, self.output_side, 3), dtype=torch.uint8, device=dev)
                               for _ in range(self.depth)]
        self.frames_host: List[torch.Tensor] = []  # pinned result ring of `render`, allocated on first use
        self.latents_host = [torch.empty((max_batch * generator.num_ws * 512,), dtype=torch.float32).pin_memory()
                             for _ in range(self.depth)]
        self.done = [torch.cuda.Event() for _ in range(self.depth)]
        self.staged = [torch.cuda.Event() for _ in range(self.depth)]  # H2D of the slot's staged latents has run
        self._next_slot = 0
        self.flags_host = [torch.zeros(1, dtype=torch.int32).pin_memory() for _ in range(self.depth)]
        self.bytes_h2d = 0
        self.bytes_d2h = 0

    def _compute(self, slot: int, latents, vector_noise_mode: str = "random"):
        """
        H2D of one batch's latents, the generator pass and the optional resize into this slot's device buffers, on this
        slot's compute stream and generator context (so that consecutive batches overlap on the device). Returns
        (n, device frames [n,S,S,3], event recorded when they are complete).

        (512,) / (N,512) are vectors, (N,rows,512) are matrices - the kind of a batch is its number of dimensions, never
        its length (a batch of exactly num_ws vectors is num_ws frames). A single (num_ws,512) matrix must arrive as
        (1,num_ws,512). `latents` may be a numpy array (staged through this slot's pinned buffer) or a CUDA tensor.
        """
        gen = self.contexts[slot]
        on_gpu = isinstance(latents, torch.Tensor) and latents.is_cuda
        shape = tuple(latents.shape)
        if len(shape) == 1:
            shape = (1,) + shape
        if len(shape) not in (2, 3) or shape[-1] != 512:
            raise ValueError(f"a batch is (N,512) vectors or (N,rows,512) matrices, got shape {tuple(latents.shape)}")
        is_vectors = len(shape) == 2
        n = shape[0]
        if n > self.max_batch:
            raise ValueError(f"batch of {n} exceeds max_batch {self.max_batch}")
        if not is_vectors and shape[1] < gen.num_ws:
            raise ValueError(f"latents have {shape[1]} rows, this network needs {gen.num_ws}")
        stream = self.streams[slot]
        stream.wait_stream(torch.cuda.current_stream(gen.device))  # device latents were produced on the caller's stream
        with torch.cuda.stream(stream):
            stream.wait_event(self.done[slot])  # the frames of the batch that last used this slot have left the device
            if on_gpu:
                latents.record_stream(stream)
                on_device = latents.reshape(shape).to(torch.float32)
                if not is_vectors:
                    on_device = on_device[:, : gen.num_ws]
                on_device = on_device.contiguous()
            else:
                array = np.asarray(latents, dtype=np.float32).reshape(shape)
                if not is_vectors:
                    array = array[:, : gen.num_ws]
                self.staged[slot].synchronize()  # host side: the H2D that last read this slot's staging buffer has run
                staged = self.latents_host[slot][: array.size].view(array.shape)
                staged.copy_(torch.from_numpy(np.ascontiguousarray(array)))
                on_device = staged.to(gen.device, non_blocking=True)
                self.staged[slot].record(stream)
                self.bytes_h2d += array.size * 4
            out = self.frames_dev[slot][:n]
            # full batches and multiples of 8 replay a captured graph; the odd-sized groups a model-switching window
            # leaves over run eagerly (same C-ABI calls), so a run never holds more than a handful of graphs per network
            graphed = n == self.max_batch or n % 8 == 0
            if is_vectors:
                noise_mode = self.noise_mode or vector_noise_mode
                if graphed:
                    gen.run_vectors_graphed(on_device, out, truncation_psi=self.truncation_psi, noise_mode=noise_mode)
                else:
                    gen.run_vectors(on_device, truncation_psi=self.truncation_psi, noise_mode=noise_mode, out=out)
            elif graphed:
                gen.run_matrices_graphed(on_device, out, noise_mode=self.noise_mode or "const")
            else:
                gen.synthesis(on_device, noise_mode=self.noise_mode or "const", out=out)
            if self.resizer is not None:
                out = self.resizer.resize(out, self.output_side, self.output_side, out=self.scaled_dev[slot][:n])
            computed = torch.cuda.Event()
            computed.record(stream)
        return n, out, computed

    def _submit(self, slot: int, latents, vector_noise_mode: str = "random") -> int:
        """One batch through `_compute`, then its frames to this slot's pinned host buffer on the copy stream."""
        gen = self.generator
        if not self.frames_host:
            side = self.output_side
            self.frames_host = [torch.empty((self.max_batch, side, side, 3), dtype=torch.uint8).pin_memory()
                                for _ in range(self.depth)]
        n, out, computed = self._compute(slot, latents, vector_noise_mode)
        self.copy_stream.wait_event(computed)
        with torch.cuda.stream(self.copy_stream):
            self.frames_host[slot][:n].copy_(out, non_blocking=True)
            self.flags_host[slot].copy_(gen.nonfinite, non_blocking=True)  # the range-guard flag travels with the frames
            self.done[slot].record(self.copy_stream)
        self.bytes_d2h += n * out[0].numel()
        return n

    def submit_into(self, latents, dest: torch.Tensor, dest_indices, vector_noise_mode: str = "random") -> Arrival:
        """
        Render one batch and deliver frame i of it to `dest[dest_indices[i]]`, without an intermediate host buffer: `dest`
        is a uint8 [T,S,S,3] tensor - pinned / CUDA-registered host memory (a window buffer, the encoder rank's shared
        frame sink) or device memory (the send buffer of a gather) - and every run of consecutive indices becomes one
        asynchronous copy on the copy stream. This is how frames rendered out of frame order (grouped by network, as the
        reference sorts them, network_visualization.py:659-674) land at their frame index: an indexed buffer instead of
        gzip-HDF5 spill files. Returns an `Arrival` (synchronize() = the frames have arrived and are finite); slots are
        used round-robin, so at most `depth` batches of this renderer are in flight.
        """
        gen = self.generator
        indices = np.asarray(dest_indices, dtype=np.int64).reshape(-1)
        slot = self._next_slot
        self._next_slot = (slot + 1) % self.depth
        n, out, computed = self._compute(slot, latents, vector_noise_mode)
        if len(indices) != n:
            raise ValueError(f"{n} latents but {len(indices)} destination indices")
        if tuple(dest.shape[1:]) != tuple(out.shape[1:]) or dest.dtype != torch.uint8:
            raise ValueError(f"dest must be uint8 [T,{out.shape[1]},{out.shape[2]},3], got {tuple(dest.shape)} {dest.dtype}")
        if n and (int(indices.min()) < 0 or int(indices.max()) >= dest.shape[0]):
            raise IndexError("destination index outside dest")
        self.copy_stream.wait_event(computed)
        arrived = torch.cuda.Event()
        with torch.cuda.stream(self.copy_stream):
            begin = 0
            for position in range(1, n + 1):  # one copy per run of consecutive destination indices
                if position == n or indices[position] != indices[position - 1] + 1:
                    first = int(indices[begin])
                    dest[first: first + position - begin].copy_(out[begin:position], non_blocking=True)
                    begin = position
            self.flags_host[slot].copy_(gen.nonfinite, non_blocking=True)
            self.done[slot].record(self.copy_stream)
            arrived.record(self.copy_stream)
        if not dest.is_cuda:
            self.bytes_d2h += n * out[0].numel()
        return Arrival(arrived, self.flags_host[slot], gen)

    def render(self, batches: Iterable[np.ndarray], check_alive=None, vector_noise_mode: str = "random") -> Iterator[np.ndarray]:
        pending: List[tuple] = []
        index = 0
        for latents in batches:
            if check_alive is not None:
                check_alive()  # the interface this iterator was drawn from may have been stopped meanwhile
            slot = index % self.depth
            n = self._submit(slot, latents, vector_noise_mode)
            pending.append((slot, n))
            index += 1
            if len(pending) == self.depth:  # keep depth-1 batches in flight
                yield self._collect(*pending.pop(0))
        for ready in pending:
            yield self._collect(*ready)

    def _collect(self, slot: int, n: int) -> np.ndarray:
        self.done[slot].synchronize()
        if int(self.flags_host[slot][0]) != 0:
            # later batches of this renderer are still running on the other slots' streams and would set the (device-wide,
            # sticky) flag again after it has been cleared: let everything finish first, so that the report leaves a clean flag
            torch.cuda.synchronize(self.generator.device)
            self.generator.nonfinite.zero_()
            torch.cuda.synchronize(self.generator.device)
            raise nonfinite_error(self.generator.mode)
        return self.frames_host[slot][:n].numpy()
