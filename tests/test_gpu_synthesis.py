"""
`-m gpu` whole-generator parity: gance_b200.synthesis.GeneratorOnDevice (CUDA, through the C ABI)
against oracle/stylegan2_ref.py (float64 CPU) on the same random-init / perturbed-init networks.

Tolerances are BASELINE.json's: fp32 mode images within 1e-3 max-abs in [-1,1]; 16-bit mode uint8
frames within +-1 LSB on >= 99.9 % of pixels. The 16-bit mode that meets that bar is "fp16"
(11 significand bits, same tcgen05 kind::f16 rate as bf16) and it is the default everywhere.
"bf16" is a NON-CONFORMING fallback for networks whose activations overflow half precision (the range
guard, GANCE_E_NONFINITE, says when): 8 significand bits per stored activation over 2*log2(R)-2
layers put it at 3-7 LSB worst case and 83-97 % of pixels within 1 LSB. Its conformance test is an
`xfail(strict=True)` that states the gap; `test_bf16_fallback_error_does_not_grow` holds it to the
error it was measured at (BF16_* below) - a regression guard, not a parity claim ("Precision").
(Pixel parity is "unpinned": the oracle is a restatement, the reference's StyleGAN2 code is absent
- see oracle/stylegan2_ref.py.)
"""

import numpy as np
import pytest
import torch

from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
from gance_b200.synthesis import GeneratorOnDevice
from oracle import stylegan2_ref as ref

pytestmark = pytest.mark.gpu

BF16_MAX_ABS = 0.12    # image max-abs error in [-1,1] units (measured 2e-2 .. 6e-2 on 32^2 .. 256^2)
BF16_WORST_LSB = 12    # worst uint8 deviation (measured 3 .. 7)
BF16_WITHIN_2_LSB = 0.90


def assert_16bit_tolerance(mode, got_u8, want_u8, max_abs=None):
    within, worst, _ = lsb_stats(got_u8, want_u8)
    if mode == "bf16":
        diff = np.abs(got_u8.astype(np.int32) - want_u8.astype(np.int32))
        assert worst <= BF16_WORST_LSB and float((diff <= 2).mean()) >= BF16_WITHIN_2_LSB, (within, worst)
        if max_abs is not None:
            assert max_abs < BF16_MAX_ABS
    else:
        assert within >= 0.999, (within, worst)


def _device():
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    torch.cuda.set_device(0)
    return torch.device("cuda", 0)


def _case(resolution, seed, perturbed, batch):
    spec = GeneratorSpec(resolution)
    variables = random_init_variables(spec, seed=seed, perturbed=perturbed)
    oracle = ref.ReferenceGenerator(variables, torch.float64)
    z = np.random.RandomState(100 + seed).randn(batch, 512)
    return spec, variables, oracle, z


def fp32_grade_bound(mode: str, want: torch.Tensor) -> float:
    """
    north_star: "fp32-mode images within 1e-3 max-abs in [-1,1]". Both validation modes are held to exactly that,
    ABSOLUTE, although the random-init networks of these tests produce images spanning +-5..7 rather than the +-1 of a
    trained network: the CUDA-core fp32 mode measures 2-3e-5, the tensor-core 3xTF32 mode 1.5-5e-4.
    """
    del mode, want
    return 1e-3


def lsb_stats(got_u8: np.ndarray, want_u8: np.ndarray):
    diff = np.abs(got_u8.astype(np.int32) - want_u8.astype(np.int32))
    return float((diff <= 1).mean()), int(diff.max()), float((diff == 0).mean())


@pytest.mark.parametrize("mode", ["fp32", "tf32"])
@pytest.mark.parametrize("resolution,perturbed", [(32, False), (64, True)])
def test_fp32_mode_matches_oracle(resolution, perturbed, mode):
    """Both validation modes: CUDA-core fp32 FMA ("fp32") and 3xTF32 on the tensor core ("tf32")."""
    spec, variables, oracle, z = _case(resolution, 0, perturbed, 3)
    gen = GeneratorOnDevice(spec, variables, _device(), mode=mode)
    wplus = gen.mapping(torch.as_tensor(z), truncation_psi=1.2)
    want_w = oracle.truncate(oracle.mapping(z), 1.2)
    assert float((wplus.double().cpu() - want_w).abs().max() / want_w.abs().max()) < 2e-5
    frames, image = gen.synthesis(wplus, noise_mode="const", return_float_image=True)
    want = oracle.synthesis(want_w.numpy())  # [B,3,R,R] float64
    got = image[..., :3].permute(0, 3, 1, 2).double().cpu()
    max_abs = float((got - want).abs().max())
    print(f"{mode} mode {resolution}^2 perturbed={perturbed}: max-abs {max_abs:.3e} (|image| max {float(want.abs().max()):.2f})")
    assert max_abs < fp32_grade_bound(mode, want)
    within, worst, exact = lsb_stats(frames.cpu().numpy(), ref.convert_images_to_uint8(want))
    print(f"   uint8: exact {exact:.5f}, within 1 LSB {within:.5f}, worst {worst}")
    assert within >= 0.999


BF16_GAP = ("bf16 operands (8-bit significand) do not reach the north_star tolerance of +-1 LSB on >= 99.9 % of pixels: "
            "measured 83-97 % (3-7 LSB worst case). fp16 - same tcgen05 kind::f16 instruction, same rate - is the conforming "
            "16-bit mode and the default; bf16 is the fallback for networks that overflow half's range.")


def _sixteen_bit_case(mode, resolution, perturbed):
    spec, variables, oracle, z = _case(resolution, 1, perturbed, 2)
    gen = GeneratorOnDevice(spec, variables, _device(), mode=mode)
    want_w = oracle.truncate(oracle.mapping(z), 1.2)
    frames, image = gen.synthesis(torch.as_tensor(want_w.numpy()), noise_mode="const", return_float_image=True)
    want = oracle.synthesis(want_w.numpy())
    got = image[..., :3].permute(0, 3, 1, 2).double().cpu()
    within, worst, exact = lsb_stats(frames.cpu().numpy(), ref.convert_images_to_uint8(want))
    max_abs = float((got - want).abs().max())
    print(f"{mode} mode {resolution}^2 perturbed={perturbed}: float max-abs {max_abs:.3e}; "
          f"uint8 exact {exact:.5f}, within 1 LSB {within:.5f}, worst {worst}")
    return frames.cpu().numpy(), ref.convert_images_to_uint8(want), max_abs


@pytest.mark.parametrize("mode", ["fp16", pytest.param("bf16", marks=pytest.mark.xfail(strict=True, reason=BF16_GAP))])
@pytest.mark.parametrize("resolution,perturbed", [(32, False), (64, True), (128, True)])
def test_16bit_modes_match_oracle(mode, resolution, perturbed):
    """The north_star bar for the 16-bit mode: uint8 frames within +-1 LSB on >= 99.9 % of pixels. fp16 meets it."""
    got, want, _ = _sixteen_bit_case(mode, resolution, perturbed)
    within, worst, _ = lsb_stats(got, want)
    assert within >= 0.999, (within, worst)


@pytest.mark.parametrize("resolution,perturbed", [(32, False), (64, True), (128, True)])
def test_bf16_fallback_error_does_not_grow(resolution, perturbed):
    """Regression guard of the non-conforming bf16 fallback: it stays within the error it was measured at."""
    got, want, max_abs = _sixteen_bit_case("bf16", resolution, perturbed)
    assert_16bit_tolerance("bf16", got, want, max_abs)


def test_vector_and_matrix_paths_agree_with_oracle_semantics():
    """Vector path = mapping + psi 1.2 + synthesis; matrix path = synthesis only, constant noise."""
    spec, variables, oracle, z = _case(32, 2, True, 2)
    gen = GeneratorOnDevice(spec, variables, _device(), mode="fp32")
    frames_vec = gen.run_vectors(torch.as_tensor(z), noise_mode="const")
    want = ref.convert_images_to_uint8(oracle.run_vector(z))
    within, worst, _ = lsb_stats(frames_vec.cpu().numpy(), want)
    assert within >= 0.999, worst
    wplus = oracle.truncate(oracle.mapping(z), 1.2).numpy()
    frames_mat = gen.run_matrices(torch.as_tensor(wplus))
    assert lsb_stats(frames_mat.cpu().numpy(), ref.convert_images_to_uint8(oracle.run_matrix(wplus)))[0] >= 0.999
    # provided noise = what tf.random_normal would have drawn, handed to both sides
    rng = np.random.RandomState(9)
    noise = [rng.randn(2, 1, c.out_resolution, c.out_resolution) for c in gen.convs]
    got = gen.synthesis(torch.as_tensor(wplus), noise_mode="provided",
                        noise=[torch.as_tensor(n[:, 0]) for n in noise])
    want = ref.convert_images_to_uint8(oracle.synthesis(wplus, noise=noise))
    assert lsb_stats(got.cpu().numpy(), want)[0] >= 0.999
    # random noise only changes the image when noise_strength != 0 (perturbed init)
    a = gen.synthesis(torch.as_tensor(wplus), noise_mode="random")
    assert not np.array_equal(a.cpu().numpy(), frames_mat.cpu().numpy())


def test_batch_independence_and_ordering():
    """Frame i of a batch equals the same latent rendered alone: ordering is bit-exact."""
    spec, variables, _, z = _case(32, 3, True, 5)
    gen = GeneratorOnDevice(spec, variables, _device(), mode="bf16")
    wplus = gen.mapping(torch.as_tensor(z))
    together = gen.synthesis(wplus).cpu().numpy()
    for i in (0, 3, 4):
        alone = gen.synthesis(wplus[i : i + 1]).cpu().numpy()
        assert np.array_equal(alone[0], together[i])
    reversed_frames = gen.synthesis(torch.flip(wplus, [0])).cpu().numpy()
    assert np.array_equal(reversed_frames[::-1], together)


def test_config1_256_frames_match_oracle():
    """BASELINE config 1 shape: 256^2 config-f generator; 2 of its frames against the fp64 oracle."""
    spec, variables, oracle, z = _case(256, 0, False, 2)
    for mode in ("fp32", "bf16", "fp16"):
        gen = GeneratorOnDevice(spec, variables, _device(), mode=mode)
        frames = gen.run_vectors(torch.as_tensor(z), noise_mode="const")
        assert frames.shape == (2, 256, 256, 3) and int(frames.sum()) > 0
        want = ref.convert_images_to_uint8(oracle.run_vector(z))
        within, worst, exact = lsb_stats(frames.cpu().numpy(), want)
        print(f"256^2 {mode}: exact {exact:.5f} within1 {within:.5f} worst {worst}")
        assert_16bit_tolerance("fp16" if mode == "fp32" else mode, frames.cpu().numpy(), want)
        gen.release_buffers()


def test_config2_1024_frame_matches_oracle():
    """
    BASELINE config 2 at its full size: one frame of the 1024^2 config-f generator (perturbed init: noise, bias and
    truncation paths active) in the fp16 mode against the fp64 oracle, through the same call the bench times.
    """
    spec, variables, oracle, z = _case(1024, 4, True, 1)
    gen = GeneratorOnDevice(spec, variables, _device(), mode="fp16")
    frames = gen.run_vectors(torch.as_tensor(z), noise_mode="const")
    assert frames.shape == (1, 1024, 1024, 3) and int(frames.sum()) > 0  # the reference's own check (test_network_functions.py:116-117)
    want = ref.convert_images_to_uint8(oracle.run_vector(z))
    within, worst, exact = lsb_stats(frames.cpu().numpy(), want)
    print(f"1024^2 fp16: exact {exact:.5f} within1 {within:.5f} worst {worst}")
    assert within >= 0.999, (within, worst)


def test_full_size_batch_properties():
    """
    Size-independent properties at the bench's size (batch 32, 1024^2), where the oracle is too slow: every frame of a
    batch equals the same latent rendered in a batch of 1 (bit-exact: no cross-sample leakage through the tiling, the
    coefficient registers or the accumulator buffers), a permuted batch gives permuted frames, and two runs agree.
    """
    spec = GeneratorSpec(1024)
    gen = GeneratorOnDevice(spec, random_init_variables(spec, seed=5, perturbed=True), _device(), mode="fp16")
    z = torch.as_tensor(np.random.RandomState(7).randn(32, 512))
    wplus = gen.mapping(z)
    together = gen.synthesis(wplus, noise_mode="const").clone()
    again = gen.synthesis(wplus, noise_mode="const")
    assert torch.equal(together, again)
    for i in (0, 13, 31):
        alone = gen.synthesis(wplus[i: i + 1], noise_mode="const")
        assert torch.equal(alone[0], together[i]), f"frame {i} differs when rendered alone"
    perm = torch.randperm(32, generator=torch.Generator().manual_seed(3))
    permuted = gen.synthesis(wplus[perm.to(wplus.device)], noise_mode="const")
    assert torch.equal(permuted, together[perm.to(together.device)])
    gen.release_buffers()


def test_graph_replay_equals_eager_and_draws_fresh_noise():
    """run_vectors_graphed (CUDA-graph replay of the same C-ABI calls) against the eager path."""
    spec, variables, _, z = _case(64, 6, True, 4)
    gen = GeneratorOnDevice(spec, variables, _device(), mode="fp16")
    z_dev = torch.as_tensor(z, dtype=torch.float32, device=gen.device)
    eager = gen.run_vectors(z_dev, noise_mode="const").clone()
    out = torch.empty_like(eager)
    for _ in range(3):  # capture, then two replays
        gen.run_vectors_graphed(z_dev, out, noise_mode="const")
        assert torch.equal(out, eager)
    other = torch.as_tensor(z[::-1].copy(), dtype=torch.float32, device=gen.device)
    gen.run_vectors_graphed(other, out, noise_mode="const")  # new input through the static buffer
    assert torch.equal(out, torch.flip(eager, [0]))
    assert gen.launches_per_replay > 10
    # random noise: the Philox position lives on the device, so replays never repeat it
    first = gen.run_vectors_graphed(z_dev, out, noise_mode="random").clone()
    second = gen.run_vectors_graphed(z_dev, out, noise_mode="random").clone()
    third = gen.run_vectors_graphed(z_dev, out, noise_mode="random").clone()
    assert not torch.equal(first, second) and not torch.equal(second, third)


@pytest.mark.parametrize("mode", ["fp16", "fp32", "tf32"])
def test_512_generator_matches_oracle(mode):
    """
    A 512^2 config-f generator end to end (BASELINE config 5 names 512^2 networks). Its last block runs the 64-channel
    halo kernel with ToRGB + uint8 fused and NO activation output, a flavour no other size exercises.
    """
    spec, variables, oracle, z = _case(512, 7, True, 2)
    gen = GeneratorOnDevice(spec, variables, _device(), mode=mode)
    frames, image = gen.run_vectors(torch.as_tensor(z), noise_mode="const", return_float_image=True)
    assert frames.shape == (2, 512, 512, 3)
    want = oracle.run_vector(z)
    got = image[..., :3].permute(0, 3, 1, 2).double().cpu()
    max_abs = float((got - want).abs().max())
    within, worst, exact = lsb_stats(frames.cpu().numpy(), ref.convert_images_to_uint8(want))
    print(f"512^2 {mode}: float max-abs {max_abs:.3e}; uint8 exact {exact:.5f} within1 {within:.5f} worst {worst}")
    if mode in ("fp32", "tf32"):
        assert max_abs < fp32_grade_bound(mode, want)
    assert within >= 0.999, (within, worst)
    # the production call (no float image requested: the last block writes uint8 only) gives the same frames
    assert torch.equal(gen.run_vectors(torch.as_tensor(z), noise_mode="const"), frames)
    gen.release_buffers()


@pytest.mark.parametrize("mode", ["fp32", "tf32"])
@pytest.mark.parametrize("resolution", [256, 1024])
def test_fp32_mode_max_abs_at_the_sizes_the_configs_name(resolution, mode):
    """north_star: fp32-mode images within 1e-3 max-abs in [-1,1] - at 256^2 (config 1) and 1024^2 (config 2), one frame."""
    spec, variables, oracle, z = _case(resolution, 8, True, 1)
    gen = GeneratorOnDevice(spec, variables, _device(), mode=mode)
    frames, image = gen.run_vectors(torch.as_tensor(z), noise_mode="const", return_float_image=True)
    want = oracle.run_vector(z)
    got = image[..., :3].permute(0, 3, 1, 2).double().cpu()
    max_abs = float((got - want).abs().max())
    within, worst, exact = lsb_stats(frames.cpu().numpy(), ref.convert_images_to_uint8(want))
    print(f"{mode} mode {resolution}^2: max-abs {max_abs:.3e} (|image| max {float(want.abs().max()):.2f}); "
          f"uint8 exact {exact:.5f} within1 {within:.5f} worst {worst}")
    assert max_abs < fp32_grade_bound(mode, want)
    assert within >= 0.999
    gen.release_buffers()


def test_config2_batch32_provided_noise_frames_match_oracle():
    """
    BASELINE config 2 the way the bench runs it - batch 32, 1024^2, per-sample noise on every layer - with the noise
    handed to both sides (`noise_mode="provided"` = what tf.random_normal would have drawn): frames 0, 13 and 31 of the
    batch against the fp64 oracle, and the graph-replayed const-noise batch against the eager one.
    """
    spec, variables, oracle, z = _case(1024, 9, True, 32)
    gen = GeneratorOnDevice(spec, variables, _device(), mode="fp16")
    generator = torch.Generator().manual_seed(17)
    noise = [torch.randn(32, c.out_resolution, c.out_resolution, generator=generator) for c in gen.convs]
    frames = gen.run_vectors(torch.as_tensor(z), noise_mode="provided", noise=noise).cpu().numpy()
    for index in (0, 13, 31):
        sample_noise = [n[index: index + 1, None].numpy() for n in noise]
        want = ref.convert_images_to_uint8(oracle.run_vector(z[index: index + 1], noise=sample_noise))
        within, worst, exact = lsb_stats(frames[index: index + 1], want)
        print(f"1024^2 fp16 batch-32 frame {index}: exact {exact:.5f} within1 {within:.5f} worst {worst}")
        assert within >= 0.999, (index, within, worst)
    z_dev = torch.as_tensor(z, dtype=torch.float32, device=gen.device)
    eager = gen.run_vectors(z_dev, noise_mode="const").clone()
    out = torch.empty_like(eager)
    gen.run_vectors_graphed(z_dev, out, noise_mode="const")
    gen.run_vectors_graphed(z_dev, out, noise_mode="const")
    assert torch.equal(out, eager)
    wplus = gen.mapping(z_dev)
    gen.run_matrices_graphed(wplus, out)  # the matrix path's replay: synthesis only, constant noise
    gen.run_matrices_graphed(wplus, out)
    assert torch.equal(out, eager)
    gen.release_buffers()


def test_fp16_range_guard_reports_overflow_instead_of_black_pixels():
    """
    Activations are stored in IEEE half already multiplied by the next layer's style. A network whose style is large
    (here: one mod_bias of 3e5 - harmless in exact arithmetic, the demodulation divides it out again) overflows 65504:
    the frames would come out as plausible black pixels (`fmaxf(NaN, 0) = 0`). The sticky device flag set by the
    frame-writing epilogues surfaces it as GANCE_E_NONFINITE; bf16 (fp32 range) and fp32 render the same network fine.
    """
    from gance_b200 import _lib
    from gance_b200.synthesis import BatchRenderer

    spec, variables, oracle, z = _case(64, 10, True, 2)
    large = dict(variables)
    large["G_synthesis/64x64/Conv0_up/mod_bias"] = variables["G_synthesis/64x64/Conv0_up/mod_bias"] + np.float32(3e5)
    want = ref.convert_images_to_uint8(ref.ReferenceGenerator(large, torch.float64).run_vector(z))
    gen = GeneratorOnDevice(spec, large, _device(), mode="fp16")
    frames = gen.run_vectors(torch.as_tensor(z), noise_mode="const")
    with pytest.raises(_lib.NonFiniteFramesError, match="GANCE_E_NONFINITE"):
        gen.check_finite()
    gen.check_finite()  # the flag was cleared by the report
    assert lsb_stats(frames.cpu().numpy(), want)[0] < 0.9  # what the caller would otherwise have been handed
    renderer = BatchRenderer(gen, max_batch=2, noise_mode="const")
    with pytest.raises(_lib.NonFiniteFramesError):
        list(renderer.render(iter([z, z])))
    gen.check_finite()
    # the same network in the wide-range modes
    for mode in ("bf16", "fp32"):
        wide = GeneratorOnDevice(spec, large, _device(), mode=mode)
        got = wide.run_vectors(torch.as_tensor(z), noise_mode="const").cpu().numpy()
        wide.check_finite()
        assert_16bit_tolerance("bf16" if mode == "bf16" else "fp16", got, want)
    # an ordinary network never trips the guard, through the per-frame boundary call either
    normal = GeneratorOnDevice(spec, variables, _device(), mode="fp16")
    normal.run_vectors(torch.as_tensor(z), noise_mode="random")
    normal.check_finite()
