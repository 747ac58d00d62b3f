"""
Self-consistency tests for oracle/stylegan2_ref.py. The reference pins no pixel values for the
generator (only shape and sum > 0, test/test_network_functions.py:116-117), so the
restatement is protected by the algebraic identities of Appendix A.4 instead - the same identities
the CUDA path relies on (shared weights + pre-scaled activations; FIR-first up-convolution).
"""

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from gance_b200.networks.generator_spec import (
    GeneratorSpec,
    random_init_variables,
    synthesis_layers,
    variable_shapes,
)
from oracle import stylegan2_ref as ref

F64 = torch.float64


def test_channel_schedule_config_f():
    spec = GeneratorSpec(1024)
    assert spec.num_ws == 18
    assert [spec.block_channels(r) for r in range(2, 11)] == [512, 512, 512, 512, 512, 256, 128, 64, 32]
    assert GeneratorSpec(512).num_ws == 16 and GeneratorSpec(256).num_ws == 14
    e = GeneratorSpec(1024, "config-e")
    assert [e.block_channels(r) for r in range(2, 11)] == [512, 512, 512, 512, 256, 128, 64, 32, 16]


def test_layer_indices_and_macs():
    """w indices of Appendix A.4 and the 74.260 GMAC / frame of Appendix A.5."""
    convs, rgbs = synthesis_layers(GeneratorSpec(1024))
    assert [c.w_index for c in convs] == list(range(0, 17))
    assert [t.w_index for t in rgbs] == [1, 3, 5, 7, 9, 11, 13, 15, 17]
    macs = 0
    for c in convs:
        pixels = c.in_resolution**2 if c.up else c.out_resolution**2
        macs += 9 * c.in_channels * c.out_channels * pixels
    macs += sum(t.in_channels * 3 * t.resolution**2 for t in rgbs)
    assert abs(macs / 1e9 - 74.260) < 0.01
    shapes = variable_shapes(GeneratorSpec(1024))
    n_params = sum(int(np.prod(s)) for n, s in shapes.items() if "/noise" not in n or n.endswith("noise_strength"))
    assert abs(n_params - 30.37e6) < 0.02e6  # the published config-f generator size


def test_fir_kernel_dc_gain_and_shapes():
    k = ref._setup_kernel(ref.RESAMPLE_KERNEL)
    assert float(k.sum()) == pytest.approx(1.0)
    y = torch.ones(1, 3, 8, 8, dtype=F64)
    up = ref.upsample_2d(y)
    assert up.shape == (1, 3, 16, 16)
    np.testing.assert_allclose(up[:, :, 2:-2, 2:-2].numpy(), 1.0, atol=1e-15)  # DC gain 1


def test_upsample_2d_explicit_taps():
    """upsample_2d: y[2a] = .25 x[a-1] + .75 x[a]; y[2a+1] = .75 x[a] + .25 x[a+1] per axis."""
    x = torch.randn(1, 1, 1, 6, dtype=F64)
    row = x[0, 0, 0]
    padded = torch.cat([torch.zeros(1, dtype=F64), row, torch.zeros(1, dtype=F64)])
    want = torch.empty(12, dtype=F64)
    for a in range(6):
        want[2 * a] = 0.25 * padded[a] + 0.75 * padded[a + 1]
        want[2 * a + 1] = 0.75 * padded[a + 1] + 0.25 * padded[a + 2]
    k1 = torch.tensor([[1.0, 3.0, 3.0, 1.0]], dtype=F64) / 4.0
    got = ref.upfirdn_2d(x, k1, upx=2, padx0=2, padx1=1)
    np.testing.assert_allclose(got[0, 0, 0].numpy(), want.numpy(), atol=1e-15)


def test_transposed_conv_equals_zero_insert_correlate():
    """Appendix A.4: conv2d_transpose(x, flip(w), stride 2, VALID) == zero-insert, pad 2, correlate."""
    x = torch.randn(1, 4, 5, 5, dtype=F64)
    w = torch.randn(3, 3, 4, 6, dtype=F64)
    t = F.conv_transpose2d(x, torch.flip(w, [0, 1]).permute(2, 3, 0, 1), stride=2)
    assert t.shape == (1, 6, 11, 11)
    z = torch.zeros(1, 4, 9, 9, dtype=F64)
    z[:, :, ::2, ::2] = x
    want = F.conv2d(F.pad(z, [2, 2, 2, 2]), w.permute(3, 2, 0, 1))
    np.testing.assert_allclose(t.numpy(), want.numpy(), atol=1e-13)


def _mod_inputs(cin=16, cout=24, batch=2, side=8):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(batch, cin, side, side, dtype=F64, generator=g)
    y = torch.randn(batch, 512, dtype=F64, generator=g)
    w = torch.randn(3, 3, cin, cout, dtype=F64, generator=g)
    mw = torch.randn(512, cin, dtype=F64, generator=g)
    mb = torch.randn(cin, dtype=F64, generator=g) * 0.1
    return x, y, w, mw, mb


def test_modconv_equals_prescaled_activation_form():
    """conv(x*s, W) * d with d = rsqrt(s^2 @ sum_kk W^2 + 1e-8) == per-sample modulated weights."""
    x, y, w, mw, mb = _mod_inputs()
    want = ref.modulated_conv2d_layer(x, y, w, mw, mb)
    s = ref.dense_layer(y, mw) + mb + 1
    w_eff = ref.get_weight(w)
    d = torch.rsqrt((s * s) @ (w_eff * w_eff).sum((0, 1)) + 1e-8)
    got = F.conv2d(x * s[:, :, None, None], w_eff.permute(3, 2, 0, 1), padding=1) * d[:, :, None, None]
    np.testing.assert_allclose(got.numpy(), want.numpy(), atol=1e-12)


def test_upconv_equals_fir_first_then_valid_conv():
    """
    upsample_conv_2d (transposed conv then FIR) == VALID 3x3 correlation of the (2H+2)^2 tensor
    u[2a] = .75 x[a-1] + .25 x[a], u[2a+1] = .25 x[a-1] + .75 x[a] (per axis, x[-1] = x[H] = 0).
    This is the form the CUDA up-convolution uses.
    """
    x, y, w, mw, mb = _mod_inputs()
    want = ref.modulated_conv2d_layer(x, y, w, mw, mb, up=True)
    assert want.shape[-1] == 16
    s = ref.dense_layer(y, mw) + mb + 1
    w_eff = ref.get_weight(w)
    d = torch.rsqrt((s * s) @ (w_eff * w_eff).sum((0, 1)) + 1e-8)
    xs = F.pad(x * s[:, :, None, None], [1, 1, 1, 1])  # x[-1] and x[H] = 0

    def up_axis(t, dim):
        lo, hi = t.narrow(dim, 0, t.shape[dim] - 1), t.narrow(dim, 1, t.shape[dim] - 1)
        even, odd = 0.75 * lo + 0.25 * hi, 0.25 * lo + 0.75 * hi
        return torch.stack([even, odd], dim=dim + 1).flatten(dim, dim + 1)

    u = up_axis(up_axis(xs, 2), 3)
    assert u.shape[-1] == 18
    got = F.conv2d(u, w_eff.permute(3, 2, 0, 1)) * d[:, :, None, None]
    np.testing.assert_allclose(got.numpy(), want.numpy(), atol=1e-12)


@pytest.mark.parametrize("perturbed", [False, True])
def test_full_forward_shapes_and_paths(perturbed):
    spec = GeneratorSpec(32)
    variables = random_init_variables(spec, seed=0, perturbed=perturbed)
    gen = ref.ReferenceGenerator(variables, F64)
    assert gen.resolution == 32 and gen.num_ws == 8
    z = np.random.RandomState(1).randn(2, 512)
    w = gen.mapping(z)
    assert w.shape == (2, 8, 512)
    assert torch.equal(w[:, 0], w[:, 7])
    img = gen.run_vector(z)
    assert img.shape == (2, 3, 32, 32)
    u8 = ref.convert_images_to_uint8(img)
    assert u8.shape == (2, 32, 32, 3) and u8.dtype == np.uint8 and u8.sum() > 0
    # matrix path = synthesis only, constant noise, no truncation (network_functions.py:167-169)
    img_m = gen.run_matrix(gen.truncate(w, 1.2).numpy())
    np.testing.assert_allclose(img_m.numpy(), img.numpy(), atol=1e-12)
    if perturbed:  # truncation and noise paths are live only with the perturbed init
        assert not np.allclose(gen.run_vector(z, truncation_psi=1.0).numpy(), img.numpy())
        noise = [np.zeros((2, 1, 1, 1))] * 8
        assert not np.allclose(gen.synthesis(w.numpy(), noise=noise).numpy(), gen.synthesis(w.numpy()).numpy())
    # fp32 evaluation of the same network stays within 1e-3 of fp64 (the north star's fp32 bar)
    img32 = ref.ReferenceGenerator(variables, torch.float32).run_vector(z)
    assert float((img32.double() - img).abs().max()) < 1e-3


def test_uint8_conversion_truncates_and_saturates():
    x = torch.tensor([-2.0, -1.0, -0.999, 0.0, 0.0039, 0.999, 1.0, 3.0], dtype=F64).reshape(1, 1, 1, 8).repeat(1, 3, 1, 1)
    got = ref.convert_images_to_uint8(x)[0, 0, :, 0]
    assert got.tolist() == [0, 0, 0, 128, 128, 255, 255, 255]


def test_phase_weight_matrix_reproduces_the_transposed_convolution():
    """
    gance_b200.synthesis.phase_up_weight: the [9*Cout, Cin] B operand of the 4-phase GEMM, emulated here in float64
    (cells x shifts x column blocks exactly as the kernel walks them) + the [1,3,3,1] FIR, equals upsample_conv_2d.
    """
    from gance_b200.synthesis import PHASE_CHUNK, phase_up_weight

    g = torch.Generator().manual_seed(5)
    batch, cin, cout, side = 2, 8, 2 * PHASE_CHUNK, 5
    x = torch.randn(batch, cin, side, side, dtype=torch.float64, generator=g)
    w = torch.randn(3, 3, cin, cout, dtype=torch.float64, generator=g)
    want = torch.stack([ref.upsample_conv_2d(x[i: i + 1], w)[0] for i in range(batch)])
    wp = phase_up_weight(w)
    assert wp.shape == (9 * cout, cin)
    xp = F.pad(x, (1, 1, 1, 1))
    t = torch.zeros(batch, cout, 2 * side + 2, 2 * side + 2, dtype=torch.float64)
    block_phase = [(0, 1), (0, 0), (1, 0), (1, 1)]
    bases, counts, firsts = [0, 4, 6, 8], [4, 2, 2, 1], [0, 0, 1, 1]
    for sh in range(4):
        dy, dx = sh & 1, sh >> 1
        xs = xp[:, :, 1 - dy: 2 - dy + side, 1 - dx: 2 - dx + side]  # x[m-dy, n-dx] for cells m,n in [0, side]
        for chunk in range(cout // PHASE_CHUNK):
            row0 = bases[sh] * cout + chunk * counts[sh] * PHASE_CHUNK
            for j in range(counts[sh]):
                a, b = block_phase[firsts[sh] + j]
                rows = wp[row0 + j * PHASE_CHUNK: row0 + (j + 1) * PHASE_CHUNK]  # [64, Cin]
                t[:, chunk * PHASE_CHUNK: (chunk + 1) * PHASE_CHUNK, a::2, b::2] += torch.einsum("bimn,oi->bomn", xs, rows)
    assert float(t[:, :, -1].abs().max()) == 0.0 and float(t[:, :, :, -1].abs().max()) == 0.0
    f = torch.tensor([0.25, 0.75, 0.75, 0.25], dtype=torch.float64)
    tp = F.pad(t, (1, 0, 1, 0))
    out = sum(f[a] * f[b] * tp[:, :, a: a + 2 * side, b: b + 2 * side] for a in range(4) for b in range(4))
    assert float((out - want).abs().max()) < 1e-12
