"""
`-m gpu` parity tests, one kernel at a time, called THROUGH THE C ABI (ctypes) and compared with
the oracle (oracle/stylegan2_ref.py, float64) on the same seeded inputs.

Tolerances: fp32 validation mode - 2e-4 relative to the tensor's max (fp32 accumulation order);
bf16 mode - the inputs are rounded to bf16 first and the oracle evaluated on the ROUNDED values, so
what remains is fp32 accumulation order plus the final bf16 rounding of the output (2^-8 relative).
"""

import contextlib

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from gance_b200 import _lib
from oracle import stylegan2_ref as ref

pytestmark = pytest.mark.gpu

F64 = torch.float64


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    torch.cuda.set_device(0)
    return _lib.init(0)


def stream():
    return torch.cuda.current_stream().cuda_stream


_KEEP_ALIVE = []


@pytest.fixture(autouse=True)
def _release_device_tensors():
    yield
    torch.cuda.synchronize()
    _KEEP_ALIVE.clear()


def dev(array, dtype=torch.float32):
    """
    Host array -> device tensor that stays alive until the end of the test: the C ABI takes raw
    pointers, so a temporary freed right after `.data_ptr()` would be recycled by the caching
    allocator for the next upload while the kernel still reads it.
    """
    tensor = torch.as_tensor(np.asarray(array)).to("cuda", dtype).contiguous()
    _KEEP_ALIVE.append(tensor)
    return tensor


def rel_err(got: torch.Tensor, want: torch.Tensor) -> float:
    got, want = got.double().cpu(), want.double().cpu()
    return float((got - want).abs().max() / want.abs().max().clamp_min(1e-30))


def test_library_reports_version_and_counts_launches(lib):
    assert lib.gance_abi_version() == 1
    before = _lib.launch_count()
    z = dev(np.random.RandomState(0).randn(2, 512))
    w = dev(np.zeros((8, 512, 512)))
    b = dev(np.zeros((8, 512)))
    avg = dev(np.zeros(512))
    out = torch.empty(2, 4, 512, device="cuda")
    _lib.check(lib.gance_mapping(z.data_ptr(), 2, w.data_ptr(), b.data_ptr(), avg.data_ptr(), 1.0, 4, out.data_ptr(), stream()), "map")
    torch.cuda.synchronize()
    # one launch per dense layer plus the normalise / broadcast+lerp passes
    assert _lib.launch_count() > before


def test_argument_validation_returns_codes(lib):
    z = dev(np.zeros((1, 512)))
    assert lib.gance_mapping(None, 1, z.data_ptr(), z.data_ptr(), z.data_ptr(), 1.0, 4, z.data_ptr(), stream()) == 1
    assert "NULL" in _lib.last_error()
    x = torch.zeros(1, 8, 8, 48, device="cuda", dtype=torch.bfloat16)
    rc = lib.gance_modconv3x3(x.data_ptr(), 1, 1, 8, 8, 48, 48, 1, x.data_ptr(), z.data_ptr(), None, 0, 0.0,
                              z.data_ptr(), None, x.data_ptr(), stream())
    assert rc == 2  # GANCE_E_SHAPE: Cin % 32 != 0 on the tcgen05 path
    with pytest.raises(_lib.GanceError):
        _lib.check(rc, "modconv")


def test_mapping_matches_oracle(lib):
    from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables

    variables = random_init_variables(GeneratorSpec(8), seed=3, perturbed=True)
    gen = ref.ReferenceGenerator(variables, F64)
    z = np.random.RandomState(5).randn(5, 512)
    want = gen.truncate(gen.mapping(z), 1.2)
    w = dev(np.stack([variables[f"G_mapping/Dense{i}/weight"] * (0.01 / np.sqrt(512)) for i in range(8)]))
    b = dev(np.stack([variables[f"G_mapping/Dense{i}/bias"] * 0.01 for i in range(8)]))
    avg = dev(variables["dlatent_avg"])
    out = torch.empty(5, gen.num_ws, 512, device="cuda")
    _lib.check(lib.gance_mapping(dev(z).data_ptr(), 5, w.data_ptr(), b.data_ptr(), avg.data_ptr(), 1.2, gen.num_ws,
                                 out.data_ptr(), stream()), "gance_mapping")
    assert rel_err(out, want) < 2e-5


def _style_case(cin, cout, batch, seed):
    g = torch.Generator().manual_seed(seed)
    wplus = torch.randn(batch, 6, 512, dtype=F64, generator=g)
    weight = torch.randn(3, 3, cin, cout, dtype=F64, generator=g)
    mod_w = torch.randn(512, cin, dtype=F64, generator=g)
    mod_b = torch.randn(cin, dtype=F64, generator=g) * 0.1
    return wplus, weight, mod_w, mod_b


def run_style_demod(lib, wplus, w_index, weight, mod_w, mod_b):
    batch, cin, cout = wplus.shape[0], weight.shape[2], weight.shape[3]
    w_eff = ref.get_weight(weight)
    wsq = dev((w_eff * w_eff).sum((0, 1)))
    s = torch.empty(batch, cin, device="cuda")
    d = torch.empty(batch, cout, device="cuda")
    _lib.check(
        lib.gance_style_demod(dev(wplus).data_ptr(), batch, wplus.shape[1], w_index,
                              dev(mod_w / np.sqrt(512.0)).data_ptr(), dev(mod_b).data_ptr(), wsq.data_ptr(), cin, cout,
                              s.data_ptr(), d.data_ptr(), stream()),
        "gance_style_demod",
    )
    return s, d


@pytest.mark.parametrize("cin,cout", [(512, 512), (64, 32), (32, 32)])
def test_style_demod_matches_oracle(lib, cin, cout):
    wplus, weight, mod_w, mod_b = _style_case(cin, cout, 3, 0)
    s, d = run_style_demod(lib, wplus, 2, weight, mod_w, mod_b)
    s_ref = ref.dense_layer(wplus[:, 2], mod_w) + mod_b + 1
    w_eff = ref.get_weight(weight)
    d_ref = torch.rsqrt((s_ref * s_ref) @ (w_eff * w_eff).sum((0, 1)) + 1e-8)
    assert rel_err(s, s_ref) < 1e-5
    assert rel_err(d, d_ref) < 1e-5


@pytest.mark.parametrize("mode", ["fp32", "bf16", "fp16"])
def test_upfir2x_pad_matches_oracle(lib, mode):
    code, dt = {"fp32": (0, torch.float32), "bf16": (1, torch.bfloat16), "fp16": (2, torch.float16)}[mode]
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 16, 6, 10, dtype=F64, generator=g)  # NCHW for the oracle
    s = torch.randn(2, 16, dtype=F64, generator=g)
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)  # what the kernel actually saw
    k = ref._setup_kernel(ref.RESAMPLE_KERNEL) * 4
    want = ref.upfirdn_2d(x_used * s[:, :, None, None], k, upx=2, upy=2, padx0=3, padx1=2, pady0=3, pady1=2)
    u = torch.empty(2, 14, 22, 16, device="cuda", dtype=dt)
    _lib.check(lib.gance_upfir2x_pad(x_dev.data_ptr(), code, 2, 6, 10, 16, dev(s).data_ptr(), u.data_ptr(), stream()),
               "gance_upfir2x_pad")
    tol = {"fp32": 1e-5, "bf16": 2.0 ** -8, "fp16": 2.0 ** -11}[mode]
    assert rel_err(u.permute(0, 3, 1, 2), want) < tol


@pytest.mark.parametrize("mode", ["fp32", "bf16", "fp16"])
@pytest.mark.parametrize("channels,side,last", [(512, 4, False), (64, 16, False), (32, 32, True)])
def test_torgb_skip_matches_oracle(lib, mode, channels, side, last):
    code, dt = {"fp32": (0, torch.float32), "bf16": (1, torch.bfloat16), "fp16": (2, torch.float16)}[mode]
    g = torch.Generator().manual_seed(2)
    batch = 2
    x = torch.randn(batch, channels, side, side, dtype=F64, generator=g)
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    y_prev = torch.randn(batch, 3, side // 2, side // 2, dtype=F64, generator=g)
    y_prev_f32 = y_prev.float().double()
    wdl = torch.randn(batch, 512, dtype=F64, generator=g)
    weight = torch.randn(1, 1, channels, 3, dtype=F64, generator=g)
    mod_w = torch.randn(512, channels, dtype=F64, generator=g)
    mod_b = torch.randn(channels, dtype=F64, generator=g) * 0.1
    bias = torch.randn(3, dtype=F64, generator=g) * 0.1
    t = ref.modulated_conv2d_layer(x_used, wdl, weight, mod_w, mod_b, demodulate=False)
    want = ref.upsample_2d(y_prev_f32) + ref.apply_bias_act(t, bias)
    s_ref = (ref.dense_layer(wdl, mod_w) + mod_b + 1).float()
    y_prev_dev = torch.zeros(batch, side // 2, side // 2, 4, device="cuda")
    y_prev_dev[..., :3] = dev(y_prev.permute(0, 2, 3, 1))
    y_out = torch.empty(batch, side, side, 4, device="cuda")
    frames = torch.empty(batch, side, side, 3, device="cuda", dtype=torch.uint8)
    _lib.check(
        lib.gance_torgb_skip(x_dev.data_ptr(), code, batch, side, side, channels, dev(s_ref).data_ptr(),
                             dev(weight[0, 0] / np.sqrt(channels)).data_ptr(), dev(bias).data_ptr(),
                             y_prev_dev.data_ptr(), y_out.data_ptr(), frames.data_ptr() if last else None, stream()),
        "gance_torgb_skip",
    )
    assert rel_err(y_out[..., :3].permute(0, 3, 1, 2), want) < 2e-5
    assert float(y_out[..., 3].abs().max()) == 0.0
    if last:
        want_u8 = ref.convert_images_to_uint8(y_out[..., :3].permute(0, 3, 1, 2).cpu())
        assert np.array_equal(frames.cpu().numpy(), want_u8)  # bit-exact on identical fp32 input


CONV_CASES = [
    # cin, cout, side, batch, pad, name
    (512, 512, 4, 9, 1, "4x4 tile (4,4,8) + batch tail, N=128"),
    (512, 512, 8, 3, 1, "8x8 tile (8,8,2) + batch tail"),
    (512, 256, 16, 2, 1, "16x16 N=128"),
    (256, 256, 64, 5, 1, "N=256 path (m_tiles >= 148)"),
    (128, 64, 32, 2, 1, "N=64"),
    (64, 32, 32, 2, 1, "N=32 K=64"),
    (32, 32, 64, 2, 1, "Cin=32: SWIZZLE_64B"),
    (64, 64, 32, 3, 1, "halo kernel N=64 (512^2 Conv1 shape)"),
    (64, 128, 16, 2, 1, "halo kernel N=128"),
    (64, 64, 48, 1, 1, "halo kernel, 3x6 tiles"),
    (128, 128, 32, 2, 1, "128 -> 128 channels, halo-shaped image (CTA-pair halo kernel when pairing is on)"),
    (128, 128, 48, 1, 1, "128 -> 128 channels, 3x6 halo tiles"),
    (64, 64, 32, 2, 0, "VALID (up-conv second half)"),
    (512, 512, 8, 2, 0, "VALID 8x8"),
]


MODE_CODES = {"fp32": (0, torch.float32), "bf16": (1, torch.bfloat16), "fp16": (2, torch.float16)}


def _run_modconv(lib, mode, cin, cout, side, batch, pad, seed=0, with_next=True):
    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(seed)
    side_in = side + 2 - 2 * pad
    x = torch.randn(batch, cin, side_in, side_in, dtype=F64, generator=g)
    weight = torch.randn(3, 3, cin, cout, dtype=F64, generator=g) / np.sqrt(9 * cin)
    demod = torch.rand(batch, cout, dtype=F64, generator=g) + 0.5
    noise = torch.randn(side, side, dtype=F64, generator=g)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    nxt = torch.randn(batch, cout, dtype=F64, generator=g) if with_next else None
    strength = 0.07
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    if mode == "fp32":
        w_dev = dev(weight)  # HWIO
        w_used = w_dev.double().cpu()
    else:
        w_dev = dev(weight.permute(3, 0, 1, 2), dt)  # [Cout,3,3,Cin]
        w_used = w_dev.double().cpu().permute(1, 2, 3, 0)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    acc = F.conv2d(x_used, w_used.permute(3, 2, 0, 1), padding=pad)
    demod32, bias32, noise32 = demod.float().double(), bias.float().double(), noise.float().double()
    want = acc * demod32[:, :, None, None] + noise32[None, None] * np.float32(strength) + bias32[None, :, None, None]
    want = F.leaky_relu(want, 0.2) * np.sqrt(2.0)
    if with_next:
        want = want * nxt.float().double()[:, :, None, None]
    out = torch.full((batch, side, side, cout), float("nan"), device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3(x_dev.data_ptr(), code, batch, side, side, cin, cout, pad, w_dev.data_ptr(),
                             dev(demod).data_ptr(), dev(noise).data_ptr(), 0, strength, dev(bias).data_ptr(),
                             dev(nxt).data_ptr() if with_next else None, out.data_ptr(), stream()),
        "gance_modconv3x3",
    )
    torch.cuda.synchronize()
    return out.permute(0, 3, 1, 2), want


@pytest.mark.parametrize("cin,cout,side,batch,pad,name", CONV_CASES)
def test_modconv3x3_fp32_matches_oracle(lib, cin, cout, side, batch, pad, name):
    got, want = _run_modconv(lib, "fp32", cin, cout, side, batch, pad)
    assert not torch.isnan(got).any(), "kernel left output unwritten"
    assert rel_err(got, want) < 2e-5, name


@pytest.mark.parametrize("pairing", [0, 2])  # single-CTA kernels / CTA pairs (cta_group::2) wherever a shape is eligible
@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch,pad,name", CONV_CASES)
def test_modconv3x3_tcgen05_matches_oracle(lib, mode, cin, cout, side, batch, pad, name, pairing):
    with conv_pairing(lib, pairing):
        got, want = _run_modconv(lib, mode, cin, cout, side, batch, pad)
    assert not torch.isnan(got.float()).any(), "kernel left output unwritten"
    ulp = 2.0 ** -8 if mode == "bf16" else 2.0 ** -11  # half an output ulp, relative
    err = rel_err(got, want)
    assert err < 2 * ulp, f"{name}: rel err {err}"
    # element-wise: |got - want| <= output rounding + small absolute slack for accumulation order
    diff = (got.double().cpu() - want).abs()
    bound = want.abs() * ulp + 2e-5 * float(want.abs().max())
    assert bool((diff <= bound).all()), f"{name}: {int((diff > bound).sum())} elements out of bound"


@contextlib.contextmanager
def conv_pairing(lib, mode):
    """gance_set_conv_pairing(mode) for the duration of a block (0 single CTA, 1 default, 2 CTA pairs whenever eligible)."""
    before = lib.gance_set_conv_pairing(mode)
    try:
        yield
    finally:
        lib.gance_set_conv_pairing(before)


PAIR_CASES = [
    # cin, cout, side, batch, pad, name
    (512, 512, 4, 9, 1, "9 one-tile samples: odd m-tile count, the last pair holds a ghost tile; N=256 x 2 n-tiles"),
    (256, 512, 8, 1, 1, "ONE m-tile: the whole second CTA of the pair is a ghost"),
    (512, 256, 16, 2, 1, "N=256, 4 m-tiles"),
    (256, 256, 64, 5, 1, "N=256, 160 m-tiles: more pair items than co-resident clusters"),
    (128, 128, 24, 2, 1, "N=128, tiles that overhang the image (24 = 16 + 8)"),
    (128, 128, 64, 3, 1, "N=128, 96 m-tiles"),
    (512, 512, 8, 2, 0, "VALID 8x8"),
    (128, 128, 32, 3, 1, "CTA-pair halo kernel, two 64-channel K chunks per tile"),
    (128, 128, 64, 5, 1, "CTA-pair halo kernel, 320 m-tiles: more pair items than co-resident clusters"),
    (64, 64, 64, 5, 1, "CTA-pair halo kernel N=64"),
]


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch,pad,name", PAIR_CASES)
def test_modconv3x3_cta_pair_matches_oracle_and_the_single_cta_kernel(lib, mode, cin, cout, side, batch, pad, name):
    """The cta_group::2 flavour (two CTAs, one M=256 MMA) against float64 and against the cta_group::1 kernel."""
    with conv_pairing(lib, 2):
        launches = lib.gance_launch_count()
        got, want = _run_modconv(lib, mode, cin, cout, side, batch, pad)
        assert lib.gance_launch_count() == launches + 1
    with conv_pairing(lib, 0):
        single, _ = _run_modconv(lib, mode, cin, cout, side, batch, pad)
    assert not torch.isnan(got.float()).any(), "kernel left output unwritten"
    ulp = 2.0 ** -8 if mode == "bf16" else 2.0 ** -11
    diff = (got.double().cpu() - want).abs()
    bound = want.abs() * ulp + 2e-5 * float(want.abs().max())
    assert bool((diff <= bound).all()), f"{name}: {int((diff > bound).sum())} elements out of bound"
    # same K order (channel chunk, tap, 16-channel step), same fp32 accumulation: every flavour agrees to the last bit, so
    # a frame does not depend on which kernel the size of its batch selects
    assert torch.equal(got, single), f"{name}: {int((got != single).sum())} elements differ from the single-CTA kernel"


def _split_tf32(lib, tensor, styles=None):
    """gance_split_tf32 on an NHWC fp32 device tensor -> (hi, lo)."""
    batch, h, w, c = tensor.shape
    hi, lo = torch.empty_like(tensor), torch.empty_like(tensor)
    _lib.check(lib.gance_split_tf32(tensor.data_ptr(), None if styles is None else styles.data_ptr(), batch, h, w, c,
                                    hi.data_ptr(), lo.data_ptr(), stream()), "gance_split_tf32")
    return hi, lo


def test_split_tf32_is_exact_and_hi_is_a_tf32_value(lib):
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(2, 5, 7, 64, generator=g) * torch.logspace(-6, 6, 64)).cuda()
    styles = torch.randn(2, 64, generator=g).cuda()
    hi, lo = _split_tf32(lib, x)
    assert torch.equal(hi + lo, x)  # the split loses nothing
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0  # low 13 significand bits of hi are clear
    assert float((lo.abs() / x.abs().clamp_min(1e-30)).max()) < 2.0 ** -10
    hi_s, lo_s = _split_tf32(lib, x, styles)
    assert torch.equal(hi_s + lo_s, x * styles[:, None, None, :])


@pytest.mark.parametrize("cin,cout,side,batch,pad,name", CONV_CASES)
def test_modconv3x3_tf32x3_matches_oracle_at_fp32_grade(lib, cin, cout, side, batch, pad, name):
    """
    The tensor-core validation mode (tcgen05 kind::tf32, hi/lo split, 3 MMAs per K step) against float64 arithmetic, at
    the tolerance the CUDA-core fp32 kernel is held to - and against that kernel itself.
    """
    g = torch.Generator().manual_seed(0)
    side_in = side + 2 - 2 * pad
    x = torch.randn(batch, cin, side_in, side_in, dtype=F64, generator=g)
    weight = torch.randn(3, 3, cin, cout, dtype=F64, generator=g) / np.sqrt(9 * cin)
    demod = torch.rand(batch, cout, dtype=F64, generator=g) + 0.5
    noise = torch.randn(batch, side, side, dtype=F64, generator=g)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    strength = 0.07
    x_dev = dev(x.permute(0, 2, 3, 1))
    w_kmajor = dev(weight.permute(3, 0, 1, 2))  # [Cout,3,3,Cin]
    w_hi = (w_kmajor.view(torch.int32) & -8192).view(torch.float32)
    w_lo = w_kmajor - w_hi
    x_hi, x_lo = _split_tf32(lib, x_dev)
    out = torch.full((batch, side, side, cout), float("nan"), device="cuda")
    _lib.check(
        lib.gance_modconv3x3_tf32x3(x_hi.data_ptr(), x_lo.data_ptr(), batch, side, side, cin, cout, pad, w_hi.data_ptr(),
                                    w_lo.data_ptr(), dev(demod).data_ptr(), dev(noise).data_ptr(), side * side, strength,
                                    dev(bias).data_ptr(), dev(nxt).data_ptr(), out.data_ptr(), stream()),
        "gance_modconv3x3_tf32x3",
    )
    torch.cuda.synchronize()
    assert not torch.isnan(out).any(), "kernel left output unwritten"
    x_used, w_used = x_dev.double().cpu().permute(0, 3, 1, 2), w_kmajor.double().cpu().permute(1, 2, 3, 0)
    f32 = lambda t: t.float().double()  # noqa: E731
    want = F.conv2d(x_used, w_used.permute(3, 2, 0, 1), padding=pad) * f32(demod)[:, :, None, None]
    want = want + f32(noise)[:, None] * np.float32(strength) + f32(bias)[None, :, None, None]
    want = F.leaky_relu(want, 0.2) * np.sqrt(2.0) * f32(nxt)[:, :, None, None]
    err = rel_err(out.permute(0, 3, 1, 2), want)
    print(f"tf32x3 {name}: rel err {err:.3e}")
    assert err < 2e-5, name
    # the same layer on the CUDA-core fp32 kernel: the two validation modes agree far inside the parity tolerance
    plain = torch.empty_like(out)
    _lib.check(
        lib.gance_modconv3x3(x_dev.data_ptr(), 0, batch, side, side, cin, cout, pad, dev(weight).data_ptr(),
                             dev(demod).data_ptr(), dev(noise).data_ptr(), side * side, strength, dev(bias).data_ptr(),
                             dev(nxt).data_ptr(), plain.data_ptr(), stream()),
        "gance_modconv3x3",
    )
    assert rel_err(out.permute(0, 3, 1, 2), plain.permute(0, 3, 1, 2).double().cpu()) < 2e-5


def test_modconv3x3_tcgen05_without_next_style_and_noise_per_sample(lib):
    got, want = _run_modconv(lib, "bf16", 64, 64, 16, 3, 1, seed=4, with_next=False)
    assert rel_err(got, want) < 2.0 ** -7


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(512, 512, 4, 3), (512, 512, 16, 2), (256, 512, 8, 1), (128, 128, 24, 2), (128, 128, 32, 2)])
@pytest.mark.parametrize("pairing", [0, 2])
def test_modconv3x3_dual_writes_the_layer_output_and_the_next_modulated_input(lib, mode, cin, cout, side, batch, pairing):
    with conv_pairing(lib, pairing):
        _dual_case(lib, mode, cin, cout, side, batch)


def _dual_case(lib, mode, cin, cout, side, batch):
    """gance_modconv3x3_dual: `out` = the activated layer output (what ToRGB reads), `out_scaled` = out * next styles
    (what the next block's Conv0_up reads), both from one epilogue; each against the fp64 computation."""
    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(23)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    weight = torch.randn(3, 3, cin, cout, dtype=F64, generator=g) / np.sqrt(9 * cin)
    demod = torch.rand(batch, cout, dtype=F64, generator=g) + 0.5
    noise = torch.randn(batch, side, side, dtype=F64, generator=g)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    strength = 0.07
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    w_dev = dev(weight.permute(3, 0, 1, 2), dt)
    acc = F.conv2d(x_dev.double().cpu().permute(0, 3, 1, 2), w_dev.double().cpu().permute(0, 3, 1, 2), padding=1)
    want = acc * demod.float().double()[:, :, None, None] + noise.float().double()[:, None] * np.float32(strength)
    want = F.leaky_relu(want + bias.float().double()[None, :, None, None], 0.2) * np.sqrt(2.0)
    want_scaled = want * nxt.float().double()[:, :, None, None]
    out = torch.full((batch, side, side, cout), float("nan"), device="cuda", dtype=dt)
    out_scaled = torch.full_like(out, float("nan"))
    _lib.check(
        lib.gance_modconv3x3_dual(x_dev.data_ptr(), code, batch, side, side, cin, cout, 1, w_dev.data_ptr(),
                                  dev(demod).data_ptr(), dev(noise).data_ptr(), side * side, strength,
                                  dev(bias).data_ptr(), dev(nxt).data_ptr(), out_scaled.data_ptr(), out.data_ptr(),
                                  stream()),
        "gance_modconv3x3_dual",
    )
    torch.cuda.synchronize()
    tol = {"bf16": 2.0 ** -7, "fp16": 2.0 ** -10}[mode]  # one rounding of each output
    assert rel_err(out.permute(0, 3, 1, 2), want) < tol
    assert rel_err(out_scaled.permute(0, 3, 1, 2), want_scaled) < tol
    # arguments the entry refuses: the narrow layers run the halo kernel, which has no second output
    assert lib.gance_modconv3x3_dual(x_dev.data_ptr(), code, batch, side, side, 64, 64, 1, w_dev.data_ptr(),
                                     dev(demod).data_ptr(), None, 0, 0.0, dev(bias).data_ptr(), dev(nxt).data_ptr(),
                                     out_scaled.data_ptr(), out.data_ptr(), stream()) > 0
    assert lib.gance_modconv3x3_dual(x_dev.data_ptr(), code, batch, side, side, cin, cout, 1, w_dev.data_ptr(),
                                     dev(demod).data_ptr(), None, 0, 0.0, dev(bias).data_ptr(), dev(nxt).data_ptr(),
                                     out.data_ptr(), out.data_ptr(), stream()) > 0


@pytest.mark.parametrize("mode", ["fp32", "bf16", "fp16"])
def test_up_convolution_matches_oracle(lib, mode):
    """gance_modconv3x3_up2_fir against the reference's transposed-conv + upfirdn formulation."""
    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(7)
    batch, cin, cout, side = 2, 64, 32, 16
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    wplus, weight, mod_w, mod_b = _style_case(cin, cout, batch, 8)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    s, d = run_style_demod(lib, wplus, 1, weight, mod_w, mod_b)
    want = ref.modulated_conv2d_layer(x_used, wplus[:, 1], weight, mod_w, mod_b, up=True)
    want = ref.apply_bias_act(want, bias, act="lrelu")
    w_eff = ref.get_weight(weight)
    w_dev = dev(w_eff) if mode == "fp32" else dev(w_eff.permute(3, 0, 1, 2), dt)
    scratch = torch.empty(batch, 2 * side + 2, 2 * side + 2, cin, device="cuda", dtype=dt)
    out = torch.empty(batch, 2 * side, 2 * side, cout, device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3_up2_fir(x_dev.data_ptr(), code, batch, side, side, cin, cout, s.data_ptr(),
                                     w_dev.data_ptr(), d.data_ptr(), None, 0, 0.0, dev(bias).data_ptr(), None,
                                     scratch.data_ptr(), out.data_ptr(), stream()),
        "gance_modconv3x3_up2_fir",
    )
    # 16-bit modes: weights, upsampled activations and output are all rounded
    tol = {"fp32": 5e-5, "bf16": 2.0 ** -5, "fp16": 2.0 ** -8}[mode]
    assert rel_err(out.permute(0, 3, 1, 2), want) < tol


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(64, 32, 16, 2), (512, 512, 4, 3), (128, 64, 32, 1), (256, 128, 8, 5)])
@pytest.mark.parametrize("pairing", [0, 2])
def test_folded_up_convolution_matches_oracle(lib, mode, cin, cout, side, batch, pairing):
    with conv_pairing(lib, pairing):
        _folded_up_case(lib, mode, cin, cout, side, batch)


def _folded_up_case(lib, mode, cin, cout, side, batch):
    """gance_scale_channels + gance_modconv3x3_up2_folded against the transposed-conv + upfirdn oracle."""
    from gance_b200.synthesis import fold_up_weight

    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(11)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    wplus, weight, mod_w, mod_b = _style_case(cin, cout, batch, 12)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    noise = torch.randn(batch, 2 * side, 2 * side, dtype=F64, generator=g)
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    strength = 0.05
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    s, d = run_style_demod(lib, wplus, 1, weight, mod_w, mod_b)
    want = ref.modulated_conv2d_layer(x_used, wplus[:, 1], weight, mod_w, mod_b, up=True)
    want = want + noise.float().double()[:, None] * np.float32(strength)
    want = ref.apply_bias_act(want, bias, act="lrelu") * nxt.float().double()[:, :, None, None]
    w_fold = dev(fold_up_weight(ref.get_weight(weight).float()), dt)
    assert w_fold.shape == (4 * cout, 3, 3, cin)
    x_scaled = torch.empty_like(x_dev)
    _lib.check(lib.gance_scale_channels(x_dev.data_ptr(), code, batch, side, side, cin, s.data_ptr(),
                                        x_scaled.data_ptr(), stream()), "gance_scale_channels")
    out = torch.full((batch, 2 * side, 2 * side, cout), float("nan"), device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3_up2_folded(x_scaled.data_ptr(), code, batch, side, side, cin, cout, w_fold.data_ptr(),
                                        d.data_ptr(), dev(noise).data_ptr(), 4 * side * side, strength,
                                        dev(bias).data_ptr(), dev(nxt).data_ptr(), out.data_ptr(), stream()),
        "gance_modconv3x3_up2_folded",
    )
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "kernel left output unwritten"
    tol = {"bf16": 2.0 ** -5, "fp16": 2.0 ** -8}[mode]  # weights, scaled activations and output are each rounded once
    assert rel_err(out.permute(0, 3, 1, 2), want) < tol


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch,last", [(64, 64, 32, 2, False), (32, 32, 32, 3, True), (128, 128, 16, 2, False),
                                                     (256, 256, 16, 1, False), (64, 32, 16, 2, True),
                                                     (64, 64, 32, 2, True), (64, 64, 16, 3, True)])  # last block of a 512^2 network
@pytest.mark.parametrize("pairing", [0, 2])
def test_fused_conv_torgb_matches_oracle(lib, mode, cin, cout, side, batch, last, pairing):
    """gance_modconv3x3_torgb: Conv1 epilogue + ToRGB + skip upsample (+ uint8) against float64 arithmetic."""
    with conv_pairing(lib, pairing):
        _fused_conv_torgb_case(lib, mode, cin, cout, side, batch, last)


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(128, 128, 16, 2), (256, 256, 16, 1), (128, 128, 32, 3), (256, 256, 8, 3)])
def test_fused_conv_torgb_cta_pair_matches_oracle(lib, mode, cin, cout, side, batch):
    """The same through the cta_group::2 kernel (N = 128 / 256 with the fused ToRGB), incl. an odd m-tile count (8x8 x 3)."""
    with conv_pairing(lib, 2):
        _fused_conv_torgb_case(lib, mode, cin, cout, side, batch, False)


def _fused_conv_torgb_case(lib, mode, cin, cout, side, batch, last):
    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    weight = torch.randn(3, 3, cin, cout, dtype=F64, generator=g) / np.sqrt(9 * cin)
    demod = torch.rand(batch, cout, dtype=F64, generator=g) + 0.5
    noise = torch.randn(side, side, dtype=F64, generator=g)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    rgb_s = torch.randn(batch, cout, dtype=F64, generator=g)
    rgb_w = torch.randn(cout, 3, dtype=F64, generator=g) / np.sqrt(cout)
    rgb_b = torch.randn(3, dtype=F64, generator=g) * 0.1
    y_prev = torch.randn(batch, 3, side // 2, side // 2, dtype=F64, generator=g)
    strength = 0.07
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    w_dev = dev(weight.permute(3, 0, 1, 2), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    w_used = w_dev.double().cpu().permute(1, 2, 3, 0)
    f32 = lambda t: t.float().double()  # noqa: E731
    act = F.conv2d(x_used, w_used.permute(3, 2, 0, 1), padding=1) * f32(demod)[:, :, None, None]
    act = act + f32(noise)[None, None] * np.float32(strength) + f32(bias)[None, :, None, None]
    act = F.leaky_relu(act, 0.2) * np.sqrt(2.0)
    rgb = torch.einsum("bchw,bc,cj->bjhw", act, f32(rgb_s), f32(rgb_w)) + f32(rgb_b)[None, :, None, None]
    want_y = ref.upsample_2d(f32(y_prev)) + rgb
    want_x = act * f32(nxt)[:, :, None, None]
    y_prev_dev = torch.zeros(batch, side // 2, side // 2, 4, device="cuda")
    y_prev_dev[..., :3] = dev(y_prev.permute(0, 2, 3, 1))
    y_out = torch.full((batch, side, side, 4), float("nan"), device="cuda")
    frames = torch.zeros(batch, side, side, 3, device="cuda", dtype=torch.uint8)
    out = None if last else torch.full((batch, side, side, cout), float("nan"), device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3_torgb(x_dev.data_ptr(), code, batch, side, side, cin, cout, w_dev.data_ptr(),
                                   dev(demod).data_ptr(), dev(noise).data_ptr(), 0, strength, dev(bias).data_ptr(),
                                   None if last else dev(nxt).data_ptr(), None if last else out.data_ptr(),
                                   dev(rgb_s).data_ptr(), dev(rgb_w).data_ptr(), dev(rgb_b).data_ptr(),
                                   y_prev_dev.data_ptr(), y_out.data_ptr(), frames.data_ptr() if last else None, stream()),
        "gance_modconv3x3_torgb",
    )
    torch.cuda.synchronize()
    assert rel_err(y_out[..., :3].permute(0, 3, 1, 2), want_y) < 2e-5  # fp32 epilogue on the un-rounded activations
    assert float(y_out[..., 3].abs().max()) == 0.0
    if last:
        want_u8 = ref.convert_images_to_uint8(y_out[..., :3].permute(0, 3, 1, 2).cpu())
        assert np.array_equal(frames.cpu().numpy(), want_u8)
    else:
        ulp = 2.0 ** -8 if mode == "bf16" else 2.0 ** -11
        assert rel_err(out.permute(0, 3, 1, 2), want_x) < 2 * ulp
    return y_out, frames, out


@contextlib.contextmanager
def conv_two_rows(lib, enabled):
    """gance_set_conv_two_rows for the duration of a block (the 32 -> 32-channel halo kernel: N = 64 or N = 32 MMAs)."""
    before = lib.gance_set_conv_two_rows(enabled)
    try:
        yield
    finally:
        lib.gance_set_conv_two_rows(before)


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("side,batch,last", [(32, 3, True), (64, 2, True), (64, 1, False), (96, 2, True), (32, 5, False)])
def test_two_row_halo_kernel_matches_oracle_and_the_one_row_kernel(lib, mode, side, batch, last):
    """modconv3x3_halo2_kernel (32 -> 32 channels, two image rows per accumulator row: N = 64 / N = 32 MMAs over a
    (32+2) x (8+2) halo box, accumulators drained as two sub-tiles) with the fused ToRGB: against the oracle, and bit for
    bit against the one-row kernel (same accumulation order per output element). Side 96 = 3 x 12 tiles."""
    with conv_two_rows(lib, 1):
        y2, f2, o2 = _fused_conv_torgb_case(lib, mode, 32, 32, side, batch, last)
    with conv_two_rows(lib, 0):
        y1, f1, o1 = _fused_conv_torgb_case(lib, mode, 32, 32, side, batch, last)
    assert torch.equal(y2, y1) and torch.equal(f2, f1)
    if not last:
        assert torch.equal(o2, o1)


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("side,batch", [(64, 2), (32, 1), (160, 1)])
def test_two_row_halo_kernel_plain_convolution(lib, mode, side, batch):
    """The same kernel without ToRGB (gance_modconv3x3, 32 -> 32 channels): oracle + bit-identity with the one-row kernel."""
    ulp = 2.0 ** -8 if mode == "bf16" else 2.0 ** -11
    with conv_two_rows(lib, 1):
        got2, want = _run_modconv(lib, mode, 32, 32, side, batch, 1)
    with conv_two_rows(lib, 0):
        got1, _ = _run_modconv(lib, mode, 32, 32, side, batch, 1)
    assert not torch.isnan(got2.float()).any(), "kernel left output unwritten"
    assert rel_err(got2, want) < 2 * ulp
    assert torch.equal(got2, got1)


def _philox4x32_10(counter: np.ndarray, key: np.ndarray) -> np.ndarray:
    """numpy restatement of Philox4x32-10 (Salmon et al. 2011): counter [n,4] uint32, key [2] uint32."""
    ctr = counter.astype(np.uint64)
    k0, k1 = np.uint64(key[0]), np.uint64(key[1])
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(0xD2511F53) * ctr[:, 0]
        p1 = np.uint64(0xCD9E8D57) * ctr[:, 2]
        ctr = np.stack([(p1 >> np.uint64(32)) ^ ctr[:, 1] ^ k0, p1 & mask, (p0 >> np.uint64(32)) ^ ctr[:, 3] ^ k1, p0 & mask], axis=1)
        k0 = (k0 + np.uint64(0x9E3779B9)) & mask
        k1 = (k1 + np.uint64(0xBB67AE85)) & mask
    return ctr.astype(np.uint32)


def test_normal_noise_matches_philox_box_muller_restatement(lib):
    """gance_normal_noise against a numpy Philox4x32-10 + Box-Muller restatement, plus N(0,1) moments."""
    count, seed, offset = 1 << 20, 0x1234_5678_9ABC_DEF0, 77
    out = torch.empty(count, device="cuda", dtype=torch.float32)
    _lib.check(lib.gance_normal_noise(out.data_ptr(), count, seed, offset, stream()), "gance_normal_noise")
    got = out.cpu().numpy().astype(np.float64)
    groups = np.arange(count // 4, dtype=np.uint64) + np.uint64(offset)
    counter = np.stack([groups & np.uint64(0xFFFFFFFF), groups >> np.uint64(32), np.zeros_like(groups), np.zeros_like(groups)], axis=1)
    bits = _philox4x32_10(counter, np.array([seed & 0xFFFFFFFF, seed >> 32], dtype=np.uint64)).astype(np.float64)
    u1a, u2a, u1b, u2b = (bits[:, 0] + 1) * 2.0 ** -32, bits[:, 1] * 2.0 ** -32, (bits[:, 2] + 1) * 2.0 ** -32, bits[:, 3] * 2.0 ** -32
    ra, rb = np.sqrt(-2 * np.log(u1a)), np.sqrt(-2 * np.log(u1b))
    want = np.stack([ra * np.cos(2 * np.pi * u2a), ra * np.sin(2 * np.pi * u2a), rb * np.cos(2 * np.pi * u2b), rb * np.sin(2 * np.pi * u2b)], axis=1).reshape(-1)
    # float32 uniforms + fast-math log/sincos on the device: absolute agreement to ~1e-3 of a unit-variance value
    assert np.abs(got - want).max() < 5e-3 and np.abs(got - want).mean() < 2e-5
    assert abs(got.mean()) < 4e-3 and abs(got.var() - 1) < 6e-3
    kurt = ((got - got.mean()) ** 4).mean() / got.var() ** 2
    assert abs(kurt - 3) < 0.03
    again = torch.empty(count, device="cuda", dtype=torch.float32)
    _lib.check(lib.gance_normal_noise(again.data_ptr(), count, seed, offset, stream()), "gance_normal_noise")
    assert torch.equal(out, again)
    _lib.check(lib.gance_normal_noise(again.data_ptr(), count, seed, offset + count // 4, stream()), "gance_normal_noise")
    assert not torch.equal(out, again) and abs(float((out * again).mean())) < 4e-3


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(64, 64, 16, 2), (128, 64, 32, 1), (256, 128, 24, 3), (512, 256, 16, 2)])
def test_phase_up_convolution_matches_oracle(lib, mode, cin, cout, side, batch):
    """gance_scale_channels + gance_modconv3x3_up2_phase (4-phase transposed-conv GEMM, then FIR pass) against the
    transposed-conv + upfirdn oracle; side 24 exercises ragged cell tiles (25 cells = 2 x 4 tiles of 16 x 8)."""
    from gance_b200.synthesis import phase_up_weight

    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(13)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    wplus, weight, mod_w, mod_b = _style_case(cin, cout, batch, 14)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    noise = torch.randn(batch, 2 * side, 2 * side, dtype=F64, generator=g)
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    strength = 0.05
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    s, d = run_style_demod(lib, wplus, 1, weight, mod_w, mod_b)
    want = ref.modulated_conv2d_layer(x_used, wplus[:, 1], weight, mod_w, mod_b, up=True)
    want = want + noise.float().double()[:, None] * np.float32(strength)
    want = ref.apply_bias_act(want, bias, act="lrelu") * nxt.float().double()[:, :, None, None]
    w_phase = dev(phase_up_weight(ref.get_weight(weight).float()), dt)
    assert w_phase.shape == (9 * cout, cin)
    x_scaled = torch.empty_like(x_dev)
    _lib.check(lib.gance_scale_channels(x_dev.data_ptr(), code, batch, side, side, cin, s.data_ptr(),
                                        x_scaled.data_ptr(), stream()), "gance_scale_channels")
    scratch = torch.full((batch, 2 * side + 2, 2 * side + 2, cout), float("nan"), device="cuda", dtype=dt)
    out = torch.full((batch, 2 * side, 2 * side, cout), float("nan"), device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3_up2_phase(x_scaled.data_ptr(), code, batch, side, side, cin, cout, w_phase.data_ptr(),
                                       d.data_ptr(), dev(noise).data_ptr(), 4 * side * side, strength,
                                       dev(bias).data_ptr(), dev(nxt).data_ptr(), scratch.data_ptr(), out.data_ptr(),
                                       stream()),
        "gance_modconv3x3_up2_phase",
    )
    torch.cuda.synchronize()
    assert not torch.isnan(scratch.float()).any(), "phase GEMM left part of t unwritten"
    assert float(scratch[:, -1].float().abs().max()) == 0.0 and float(scratch[:, :, -1].float().abs().max()) == 0.0
    assert not torch.isnan(out.float()).any(), "FIR pass left output unwritten"
    tol = {"bf16": 2.0 ** -5, "fp16": 2.0 ** -8}[mode]  # weights, scaled activations, t and the output are each rounded once
    assert rel_err(out.permute(0, 3, 1, 2), want) < tol


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(64, 64, 16, 2), (128, 64, 32, 1), (256, 128, 24, 3), (512, 256, 16, 2),
                                                 (64, 32, 40, 2), (128, 64, 64, 1), (512, 512, 4, 3), (512, 512, 8, 2),
                                                 (128, 64, 4, 1)])
def test_fused_up_convolution_matches_oracle(lib, mode, cin, cout, side, batch):
    """Single-CTA kernel against the oracle, then the CTA-pair flavour of the same 16 x 8-cell kernel (cta_group::2, half of
    the weight rows per CTA; side 24 x batch 3 and side 40 give odd tile counts = a ghost tile) bit for bit against it."""
    with conv_pairing(lib, 0):
        single = _fused_up_case(lib, mode, cin, cout, side, batch)
    with conv_pairing(lib, 2):
        paired = _fused_up_case(lib, mode, cin, cout, side, batch)
    assert torch.equal(single, paired)


@pytest.mark.parametrize("mode", ["bf16", "fp16"])
@pytest.mark.parametrize("cin,cout,side,batch", [(256, 128, 32, 2), (512, 256, 32, 1), (512, 512, 40, 1), (256, 64, 64, 3),
                                                 (512, 128, 48, 2)])
def test_fused_up_convolution_cta_pair_matches_oracle(lib, mode, cin, cout, side, batch):
    """The Cin >= 256 form (upconv_fused_pair.cu: CTA pairs, one activation halo box per K chunk, 8 x 16-cell tiles with
    12 x 28 output pixels). Side 40 gives 7 x 3 = 21 tiles - an odd count, the last pair holds a ghost tile - and ragged
    tiles on both axes; side 48 ragged rows."""
    with conv_pairing(lib, 2):
        paired = _fused_up_case(lib, mode, cin, cout, side, batch)
    with conv_pairing(lib, 0):
        single = _fused_up_case(lib, mode, cin, cout, side, batch)
    assert torch.equal(paired, single)  # same accumulation order, same filter arithmetic: the two forms agree bit for bit


def _fused_up_case(lib, mode, cin, cout, side, batch):
    """gance_scale_channels + gance_modconv3x3_up2_fused (4-phase transposed-conv GEMM with the FIR applied from a
    shared-memory patch of t) against the transposed-conv + upfirdn oracle. Sides 24 and 40 give ragged tiles (output
    48 / 80 pixels against 28 x 12-pixel tiles); Cout = 32 is the 1024^2 layer's single 32-channel chunk; sides 4 and 8 are
    the 4^2 -> 8^2 / 8^2 -> 16^2 layers (one tile per sample or two, TMA boxes larger than the image)."""
    from gance_b200.synthesis import phase_up_weight

    code, dt = MODE_CODES[mode]
    g = torch.Generator().manual_seed(17)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    wplus, weight, mod_w, mod_b = _style_case(cin, cout, batch, 14)
    bias = torch.randn(cout, dtype=F64, generator=g) * 0.1
    noise = torch.randn(batch, 2 * side, 2 * side, dtype=F64, generator=g)
    nxt = torch.randn(batch, cout, dtype=F64, generator=g)
    strength = 0.05
    x_dev = dev(x.permute(0, 2, 3, 1), dt)
    x_used = x_dev.double().cpu().permute(0, 3, 1, 2)
    s, d = run_style_demod(lib, wplus, 1, weight, mod_w, mod_b)
    want = ref.modulated_conv2d_layer(x_used, wplus[:, 1], weight, mod_w, mod_b, up=True)
    want = want + noise.float().double()[:, None] * np.float32(strength)
    want = ref.apply_bias_act(want, bias, act="lrelu") * nxt.float().double()[:, :, None, None]
    w_phase = dev(phase_up_weight(ref.get_weight(weight).float(), min(64, cout)), dt)
    assert w_phase.shape == (9 * cout, cin)
    x_scaled = torch.empty_like(x_dev)
    _lib.check(lib.gance_scale_channels(x_dev.data_ptr(), code, batch, side, side, cin, s.data_ptr(),
                                        x_scaled.data_ptr(), stream()), "gance_scale_channels")
    out = torch.full((batch, 2 * side, 2 * side, cout), float("nan"), device="cuda", dtype=dt)
    _lib.check(
        lib.gance_modconv3x3_up2_fused(x_scaled.data_ptr(), code, batch, side, side, cin, cout, w_phase.data_ptr(),
                                       d.data_ptr(), dev(noise).data_ptr(), 4 * side * side, strength,
                                       dev(bias).data_ptr(), dev(nxt).data_ptr(), out.data_ptr(), stream()),
        "gance_modconv3x3_up2_fused",
    )
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "fused up-convolution left output unwritten"
    tol = {"bf16": 2.0 ** -5, "fp16": 2.0 ** -8}[mode]  # weights, scaled activations, t and the output are each rounded once
    got = out.double().cpu().permute(0, 3, 1, 2)
    assert float((got - want).abs().max() / want.abs().max()) < tol
    return out
