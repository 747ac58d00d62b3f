"""
`-m gpu` test of gance_modconv3x3_up2_phase_static (include/gance_b200.h): the up-convolution whose intermediate t lives
in a file-scope `static __device__` array of the library instead of a caller-allocated scratch. No counterpart in the
reference; the check is the existing entry point gance_modconv3x3_up2_phase (itself checked against the oracle in
tests/test_gpu_kernels.py::test_phase_up_convolution_matches_oracle) on the same inputs, through the C ABI.

Collected last on purpose (the file name sorts after every other test file).
Self-contained: it passes alone in a fresh process and after the other tests.
"""

import pytest
import torch

from gance_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 4 << 20  # GANCE_PHASE_STATIC_SCRATCH_BYTES
MODES = {"bf16": (_lib.GANCE_BF16, torch.bfloat16, 2.0 ** -5), "fp16": (_lib.GANCE_F16, torch.float16, 2.0 ** -8)}


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.fail("this test needs a CUDA device; there is no CPU fallback to test")
    torch.cuda.set_device(0)
    return _lib.init(0)


def _case(seed, dt, batch, side, cin, cout):
    g = torch.Generator(device="cuda").manual_seed(seed)
    rand = lambda *shape: torch.randn(*shape, device="cuda", dtype=torch.float32, generator=g)
    return {
        "x": rand(batch, side, side, cin).to(dt).contiguous(),                       # already multiplied by the styles
        "w": (rand(9 * cout, cin) * (9 * cin) ** -0.5).to(dt).contiguous(),            # phase-ordered rows
        "demod": (1.0 + 0.1 * rand(batch, cout)).contiguous(),
        "noise": rand(batch, 2 * side, 2 * side).contiguous(),
        "bias": (0.1 * rand(cout)).contiguous(),
        "next": rand(batch, cout).contiguous(),
    }


def _blank_output(dt, batch, side, cout):
    return torch.full((batch, 2 * side, 2 * side, cout), float("nan"), device="cuda", dtype=dt)


def _run(lib, static, code, dt, batch, side, cin, cout, t, out, strength=0.05):
    stream = torch.cuda.current_stream().cuda_stream
    head = (t["x"].data_ptr(), code, batch, side, side, cin, cout, t["w"].data_ptr(), t["demod"].data_ptr(),
            t["noise"].data_ptr(), 4 * side * side, strength, t["bias"].data_ptr(), t["next"].data_ptr())
    if static:
        _lib.check(lib.gance_modconv3x3_up2_phase_static(*head, out.data_ptr(), stream), "gance_modconv3x3_up2_phase_static")
        scratch = None
    else:
        scratch = torch.full((batch, 2 * side + 2, 2 * side + 2, cout), float("nan"), device="cuda", dtype=dt)
        _lib.check(lib.gance_modconv3x3_up2_phase(*head, scratch.data_ptr(), out.data_ptr(), stream), "gance_modconv3x3_up2_phase")
    torch.cuda.synchronize()
    del scratch
    return out


@pytest.mark.parametrize("mode", ["fp16", "bf16"])
def test_static_scratch_up_convolution_equals_the_caller_scratch_entry_on_two_calls_in_a_row(lib, mode):
    code, dt, tol = MODES[mode]
    # call 1 fills 3.4 MB of the array; call 2 (other shape, ragged cell tiles, fewer channels) finds it full of call 1's t
    shapes = [(2, 40, 128, 128), (3, 24, 64, 64)]
    assert all(2 * b * (2 * s + 2) ** 2 * co <= STATIC_SCRATCH_BYTES for b, s, _, co in shapes)
    cases = [(_case(31 + i, dt, b, s, ci, co), (b, s, ci, co)) for i, (b, s, ci, co) in enumerate(shapes)]
    want = [_run(lib, False, code, dt, *shape, tensors, _blank_output(dt, shape[0], shape[1], shape[3])) for tensors, shape in cases]
    blanks = [_blank_output(dt, b, s, co) for b, s, _, co in shapes]   # torch allocates here, outside the window below
    got = []
    for index, (tensors, shape) in enumerate(cases):
        if index == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        got.append(_run(lib, True, code, dt, *shape, tensors, blanks[index]))
    free_after = torch.cuda.mem_get_info()[0]
    for (tensors, shape), w, g in zip(cases, want, got):
        assert not torch.isnan(g.float()).any(), f"{shape}: the static-scratch entry left output unwritten"
        err = float((g.double() - w.double()).abs().max() / w.double().abs().max())
        assert err < tol, f"{shape}: {err}"
        assert torch.equal(g, w), f"{shape}: same kernels, same inputs - only the address of t differs"
    # the second call ran with every kernel already loaded and its output allocated beforehand: the device's free memory
    # must not have moved (a scratch of its own would be 960 KB)
    assert free_before == free_after, (free_before, free_after)


def test_static_scratch_entry_rejects_sizes_beyond_its_array(lib):
    dt = torch.float16
    batch, side, cin, cout = 2, 48, 64, 128   # t = 2 * 2 * 98 * 98 * 128 B = 4.9 MB > 4 MiB
    assert 2 * batch * (2 * side + 2) ** 2 * cout > STATIC_SCRATCH_BYTES
    tensors = _case(5, dt, batch, side, cin, cout)
    out = torch.zeros((batch, 2 * side, 2 * side, cout), device="cuda", dtype=dt)
    rc = lib.gance_modconv3x3_up2_phase_static(
        tensors["x"].data_ptr(), _lib.GANCE_F16, batch, side, side, cin, cout, tensors["w"].data_ptr(),
        tensors["demod"].data_ptr(), tensors["noise"].data_ptr(), 4 * side * side, 0.05, tensors["bias"].data_ptr(),
        tensors["next"].data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert rc == 2  # GANCE_E_SHAPE
    message = _lib.last_error()
    assert "static scratch holds 4194304" in message and "gance_modconv3x3_up2_phase" in message
    assert float(out.float().abs().max()) == 0.0, "a rejected call must not launch"
