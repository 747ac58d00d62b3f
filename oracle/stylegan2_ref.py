"""
ORACLE (test infrastructure, not product): CPU restatement of the StyleGAN2 generator forward that
GANce calls through `Gs.run(...)` / `Gs.components.synthesis.run(...)`
(/root/reference/gance/network_interface/network_functions.py:151-158, 167-169).

PARITY UNPINNED. The algorithm lives in a third-party dependency that is absent from
/root/reference: git submodule `gance/stylegan2` -> github.com/esologic/stylegan2_gance @
0f84feb6fb89df8294f157de286d0a4cad15bb0d (.gitmodules:1-3, .SUBMODULES.json:9-15 "missing"), a fork
of NVlabs/stylegan2 (TF1). No network pickle and no reference test pins any pixel value
(only shape (1024,1024,3) and sum>0, test/test_network_functions.py:116-117). This
file restates the published algorithm of, function by function:

  training/networks_stylegan2.py : G_main, G_mapping, G_synthesis_stylegan2 (architecture='skip'),
                                   modulated_conv2d_layer (fused_modconv=True), get_weight,
                                   dense_layer, apply_bias_act
  dnnlib/tflib/ops/upfirdn_2d.py : _upfirdn_2d_ref, _setup_kernel, upsample_2d, upsample_conv_2d
  dnnlib/tflib/ops/fused_bias_act.py : _fused_bias_act_ref
  dnnlib/tflib/tfutil.py         : convert_images_to_uint8, lerp

(spec: SURVEY.md Appendix A). It is written the *reference's* way - per-sample modulated weights,
grouped convolution, transposed convolution followed by upfirdn - and NOT the way the CUDA path
computes (shared weights, pre-scaled activations, FIR-first up-convolution), so that agreement
between the two checks those identities instead of assuming them.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module.
"""

from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

LRELU_ALPHA = 0.2
LRELU_GAIN = float(np.sqrt(2.0))
RESAMPLE_KERNEL = (1, 3, 3, 1)


def _t(array: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    return torch.as_tensor(np.asarray(array), dtype=dtype)


def get_weight(var: torch.Tensor, lrmul: float = 1.0) -> torch.Tensor:
    """networks_stylegan2.get_weight with use_wscale=True: w * (gain/sqrt(fan_in)) * lrmul."""
    fan_in = int(np.prod(var.shape[:-1]))
    return var * (lrmul / np.sqrt(fan_in))


def dense_layer(x: torch.Tensor, weight_var: torch.Tensor, lrmul: float = 1.0) -> torch.Tensor:
    """networks_stylegan2.dense_layer: x @ get_weight([in, out])."""
    return x.reshape(x.shape[0], -1) @ get_weight(weight_var, lrmul)


def apply_bias_act(
    x: torch.Tensor, bias_var: Optional[torch.Tensor], act: str = "linear", lrmul: float = 1.0
) -> torch.Tensor:
    """
    networks_stylegan2.apply_bias_act -> fused_bias_act._fused_bias_act_ref:
    x + b*lrmul on axis 1; 'lrelu' = leaky_relu(0.2) * sqrt(2); 'linear' = gain 1.
    """
    if bias_var is not None:
        b = bias_var * lrmul
        x = x + b.reshape([1, -1] + [1] * (x.dim() - 2))
    if act == "lrelu":
        x = F.leaky_relu(x, LRELU_ALPHA) * LRELU_GAIN
    elif act != "linear":
        raise ValueError(act)
    return x


def _setup_kernel(k) -> torch.Tensor:
    """upfirdn_2d._setup_kernel: outer product of a separable kernel, normalised to sum 1."""
    k = np.asarray(k, dtype=np.float64)
    if k.ndim == 1:
        k = np.outer(k, k)
    k /= np.sum(k)
    return torch.as_tensor(k)


def upfirdn_2d(
    x: torch.Tensor, k: torch.Tensor, upx=1, upy=1, downx=1, downy=1, padx0=0, padx1=0, pady0=0, pady1=0
) -> torch.Tensor:
    """
    upfirdn_2d._upfirdn_2d_ref on an NCHW tensor (the reference reshapes NCHW to [N*C,H,W,1]):
    zero-insert (up-1 zeros AFTER every sample), pad (negative = crop), true convolution with k,
    keep every down-th sample.
    """
    n, c, in_h, in_w = x.shape
    x = x.reshape(n * c, 1, in_h, 1, in_w, 1)
    x = F.pad(x, [0, upx - 1, 0, 0, 0, upy - 1])
    x = x.reshape(n * c, 1, in_h * upy, in_w * upx)
    x = F.pad(x, [max(padx0, 0), max(padx1, 0), max(pady0, 0), max(pady1, 0)])
    x = x[:, :, max(-pady0, 0) : x.shape[2] - max(-pady1, 0), max(-padx0, 0) : x.shape[3] - max(-padx1, 0)]
    w = torch.flip(k.to(x.dtype), [0, 1])[None, None]
    x = F.conv2d(x, w)  # VALID cross-correlation with the flipped kernel = convolution
    x = x[:, :, ::downy, ::downx]
    return x.reshape(n, c, x.shape[2], x.shape[3])


def upsample_2d(x: torch.Tensor, k=RESAMPLE_KERNEL, factor: int = 2, gain: float = 1.0) -> torch.Tensor:
    """upfirdn_2d.upsample_2d: k * gain*factor^2, pad0 = (p+1)//2 + factor-1, pad1 = p//2."""
    kk = _setup_kernel(k) * (gain * factor**2)
    p = kk.shape[0] - factor
    pad0, pad1 = (p + 1) // 2 + factor - 1, p // 2
    return upfirdn_2d(x, kk, upx=factor, upy=factor, padx0=pad0, padx1=pad1, pady0=pad0, pady1=pad1)


def upsample_conv_2d(x: torch.Tensor, w: torch.Tensor, k=RESAMPLE_KERNEL, factor: int = 2) -> torch.Tensor:
    """
    upfirdn_2d.upsample_conv_2d for ONE sample group: `w` is HWIO [kh,kw,Cin,Cout].
    tf.nn.conv2d_transpose(x, flip_hw(w) as [h,w,out,in], stride=factor, VALID) then
    upfirdn with k*factor^2, p = (k-factor)-(kw-1), pad0 = (p+1)//2+factor-1, pad1 = p//2+1.
    torch conv_transpose2d scatters out[s*i+k] += x[i]*weight[in,out,k], same as the TF op.
    """
    conv_h, conv_w = w.shape[0], w.shape[1]
    assert conv_h == conv_w
    w_t = torch.flip(w, [0, 1]).permute(2, 3, 0, 1)  # [Cin, Cout, kh, kw]
    x = F.conv_transpose2d(x, w_t, stride=factor)  # VALID: (H-1)*factor + kh
    kk = _setup_kernel(k) * (factor**2)
    p = (kk.shape[0] - factor) - (conv_w - 1)
    pad0, pad1 = (p + 1) // 2 + factor - 1, p // 2 + 1
    return upfirdn_2d(x, kk, padx0=pad0, padx1=pad1, pady0=pad0, pady1=pad1)


def modulated_conv2d_layer(
    x: torch.Tensor,
    y: torch.Tensor,
    weight_var: torch.Tensor,
    mod_weight_var: torch.Tensor,
    mod_bias_var: torch.Tensor,
    up: bool = False,
    demodulate: bool = True,
) -> torch.Tensor:
    """
    networks_stylegan2.modulated_conv2d_layer, fused_modconv=True: per-sample weights
    ww = w[None] * s[:,None,None,:,None]; d = rsqrt(sum ww^2 over (kh,kw,in) + 1e-8); grouped conv.
    x: [B,Cin,H,W]; y: [B,512] (the row of w+ for this layer); weight_var HWIO.
    """
    kernel = weight_var.shape[0]
    w = get_weight(weight_var)
    ww = w[None]  # [B,kh,kw,I,O]
    s = dense_layer(y, mod_weight_var)
    s = apply_bias_act(s, mod_bias_var) + 1.0  # [B,I]
    ww = ww * s[:, None, None, :, None]
    if demodulate:
        d = torch.rsqrt(torch.sum(ww * ww, dim=(1, 2, 3)) + 1e-8)  # [B,O]
        ww = ww * d[:, None, None, None, :]
    outs = []
    for b in range(x.shape[0]):  # groups = batch
        xb, wb = x[b : b + 1], ww[b]
        if up:
            outs.append(upsample_conv_2d(xb, wb))
        else:
            outs.append(F.conv2d(xb, wb.permute(3, 2, 0, 1), padding=kernel // 2))  # SAME
    return torch.cat(outs, 0)


class ReferenceGenerator:
    """
    Holds one generator's variables (TF names) and evaluates
    G_mapping / G_main / G_synthesis on CPU in `dtype` (float64 = parity truth, float32 = timed
    CPU baseline).
    """

    def __init__(self, variables: Dict[str, np.ndarray], dtype: torch.dtype = torch.float64) -> None:
        self.dtype = dtype
        self.vars = {name: _t(value, dtype) for name, value in variables.items()}
        const = self.vars["G_synthesis/4x4/Const/const"]
        self.resolution = 4
        while f"G_synthesis/{self.resolution * 2}x{self.resolution * 2}/Conv1/weight" in self.vars:
            self.resolution *= 2
        self.resolution_log2 = int(np.log2(self.resolution))
        self.num_ws = 2 * self.resolution_log2 - 2
        self.const = const

    # -- G_mapping ---------------------------------------------------------------------------
    def mapping(self, z: np.ndarray) -> torch.Tensor:
        """G_mapping: normalize_latents, 8x dense(lrmul .01)+lrelu, broadcast to num_ws."""
        x = _t(z, self.dtype).reshape(-1, 512)
        x = x * torch.rsqrt(torch.mean(x * x, dim=1, keepdim=True) + 1e-8)
        for i in range(8):
            x = dense_layer(x, self.vars[f"G_mapping/Dense{i}/weight"], lrmul=0.01)
            x = apply_bias_act(x, self.vars[f"G_mapping/Dense{i}/bias"], act="lrelu", lrmul=0.01)
        return x[:, None, :].repeat(1, self.num_ws, 1)

    def truncate(self, dlatents: torch.Tensor, psi: Optional[float]) -> torch.Tensor:
        """G_main 'Truncation': lerp(dlatent_avg, w, psi) on all layers; skipped iff psi in (None, 1)."""
        if psi is None or psi == 1:
            return dlatents
        avg = self.vars["dlatent_avg"]
        return avg + (dlatents - avg) * psi

    # -- G_synthesis -------------------------------------------------------------------------
    def _layer(self, x, dlatents, scope: str, idx: int, up: bool, noise) -> torch.Tensor:
        v = self.vars
        x = modulated_conv2d_layer(
            x,
            dlatents[:, idx],
            v[f"{scope}/weight"],
            v[f"{scope}/mod_weight"],
            v[f"{scope}/mod_bias"],
            up=up,
        )
        nz = v[f"G_synthesis/noise{idx}"] if noise is None else _t(noise[idx], self.dtype)
        x = x + nz * v[f"{scope}/noise_strength"]
        return apply_bias_act(x, v[f"{scope}/bias"], act="lrelu")

    def _torgb(self, x, y, dlatents, scope: str, res_log2: int):
        v = self.vars
        t = modulated_conv2d_layer(
            x,
            dlatents[:, res_log2 * 2 - 3],
            v[f"{scope}/weight"],
            v[f"{scope}/mod_weight"],
            v[f"{scope}/mod_bias"],
            demodulate=False,
        )
        t = apply_bias_act(t, v[f"{scope}/bias"])
        return t if y is None else y + t

    def synthesis(self, dlatents, noise=None) -> torch.Tensor:
        """
        G_synthesis_stylegan2, architecture 'skip'. dlatents [B,num_ws,512]. `noise`: None ->
        the constant noise buffers (randomize_noise=False); else a list indexed by layer of
        [B,1,H,W] arrays (what tf.random_normal would have drawn). Returns [B,3,R,R].
        """
        dl = _t(dlatents, self.dtype).reshape(-1, self.num_ws, 512)
        x = self.const.repeat(dl.shape[0], 1, 1, 1)
        x = self._layer(x, dl, "G_synthesis/4x4/Conv", 0, False, noise)
        y = self._torgb(x, None, dl, "G_synthesis/4x4/ToRGB", 2)
        for r in range(3, self.resolution_log2 + 1):
            scope = f"G_synthesis/{2 ** r}x{2 ** r}"
            x = self._layer(x, dl, f"{scope}/Conv0_up", r * 2 - 5, True, noise)
            x = self._layer(x, dl, f"{scope}/Conv1", r * 2 - 4, False, noise)
            y = upsample_2d(y)
            y = self._torgb(x, y, dl, f"{scope}/ToRGB", r)
        return y

    # -- the two calls GANce makes -------------------------------------------------------------
    def run_vector(self, z: np.ndarray, truncation_psi: float = 1.2, noise=None) -> torch.Tensor:
        """Gs.run(z, None, truncation_psi=1.2) (network_functions.py:151-158): mapping, lerp, synthesis."""
        return self.synthesis(self.truncate(self.mapping(z), truncation_psi), noise=noise)

    def run_matrix(self, dlatents: np.ndarray) -> torch.Tensor:
        """Gs.components.synthesis.run(w+, randomize_noise=False) (network_functions.py:167-169)."""
        return self.synthesis(dlatents, noise=None)


def convert_images_to_uint8(images: torch.Tensor) -> np.ndarray:
    """
    tfutil.convert_images_to_uint8(drange=[-1,1], nchw_to_nhwc=True): NHWC; x*127.5 + 128
    (= 0.5 - (-1)*127.5); tf.saturate_cast(uint8) = clamp to [0,255] then truncate.
    The multiply-add is evaluated in float32 as TF does, whatever precision `images` carries.
    """
    x = images.permute(0, 2, 3, 1).to(torch.float32)
    x = x * np.float32(127.5) + np.float32(128.0)
    return torch.clamp(x, 0.0, 255.0).to(torch.uint8).numpy()
