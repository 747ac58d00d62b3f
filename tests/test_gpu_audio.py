"""
`-m gpu` parity tests of the audio -> latents -> blend half of the path:
the CUDA kernels, called through the C ABI by the reference-named host functions, against
  * the golden vectors generated from the reference's OWN Python (tests/golden/make_golden.py), and
  * the pinned numpy restatement (oracle/audio_ref.py) on inputs that have no golden.

Tolerances (BASELINE.json): integer/byte/index work bit-exact - PCM conversion, resampling (same
operation order as the numba loop), roll amounts, network indices; float64 latents within 1e-5
relative (measured: ~1e-12, the DFT / resampling sums differ from numpy only in summation order).
"""

import numpy as np
import pytest
import torch

from gance_b200 import apply_spectrogram, audio_device, synthetic
from gance_b200.data_into_network_visualization import visualization_inputs
from gance_b200.vector_sources import music
from gance_b200.vector_sources.vector_types import MatricesLabel
from oracle import audio_ref

pytestmark = pytest.mark.gpu

LATENT_RTOL = 1e-5


@pytest.fixture(scope="module", autouse=True)
def _cuda():
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    torch.cuda.set_device(0)


def assert_latents_close(got, want, what=""):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    scale = np.abs(want).max()
    err = np.abs(got - want).max() / scale
    assert err < LATENT_RTOL, f"{what}: max error {err:.3e} of the value range"


def test_pcm_to_float_is_bit_exact(chirp_golden):
    got = audio_device.pcm_to_float(chirp_golden["wav_int16"]).cpu().numpy()
    assert got.dtype == np.float32
    assert np.array_equal(got, chirp_golden["wav_float32"])
    extremes = np.array([-32768, -1, 0, 1, 32767], dtype=np.int16)
    assert np.array_equal(audio_device.pcm_to_float(extremes).cpu().numpy(), audio_ref.wav_to_float32(extremes))
    wide = np.array([-2147483648, -7, 0, 123456789, 2147483647], dtype=np.int32)
    assert np.array_equal(audio_device.pcm_to_float(wide).cpu().numpy(), audio_ref.wav_to_float32(wide))
    with pytest.raises(ValueError):
        audio_device.pcm_to_float(np.zeros(4, dtype=np.uint8))


def test_resample_is_bit_exact_with_reference_golden(chirp_golden, claps_golden):
    wav = music.WavFileProperties(44100, chirp_golden["wav_float32"], "chirp")
    scaled = music.read_wavs_scale_for_video([wav], 512, frames_per_second=60.0)
    assert scaled.wav_data.dtype == np.float32 and scaled.wav_data.shape == (60 * 512,)
    assert np.array_equal(scaled.wav_data, chirp_golden["scaled"])
    assert scaled.sample_rate == 44100 and scaled.name == "chirp_mono_scaled_padded"
    # target-vector-count mode and the un-padded variant (projection_file_blend.py:150-157)
    by_count = music.read_wavs_scale_for_video([wav], 512, target_num_vectors=60)
    want = audio_ref.scale_wav_for_video(chirp_golden["wav_float32"], 44100, 512, target_num_vectors=60)
    assert np.array_equal(by_count.wav_data, want)
    raw = music.read_wavs_scale_for_video([wav], 512, frames_per_second=60.0, pad_to_length=False)
    assert len(raw.wav_data) == int(44100 * (30720 / 44100))


@pytest.mark.parametrize("multiplier", [2.0, 1.5, 0.3, 0.1, 10])
def test_resample_length_law_and_values(claps_golden, multiplier):
    """test/test_vector_source_music.py:13-24 (length == int(len*multiplier)) + values against the oracle."""
    rng = np.random.RandomState(3)
    data = (rng.randn(7000) * 0.1).astype(np.float32)
    wav = music.WavFileProperties(44100, data, "noise")
    scaled = music._scale_wav_to_sample_rate(wav, int(44100 * multiplier))  # pylint: disable=protected-access
    assert len(scaled.wav_data) == int(len(data) * (int(44100 * multiplier) / 44100))
    from oracle import third_party_ref

    want = third_party_ref.resample(data, 44100, int(44100 * multiplier))
    assert np.array_equal(scaled.wav_data, want)


def test_read_wavs_argument_errors(chirp_golden):
    wav = music.WavFileProperties(44100, chirp_golden["wav_float32"], "chirp")
    with pytest.raises(ValueError):
        music.read_wavs_scale_for_video([wav], 512, frames_per_second=60.0, target_num_vectors=60)
    with pytest.raises(ValueError):
        music.read_wavs_scale_for_video([wav], 512)
    other = music.WavFileProperties(48000, chirp_golden["wav_float32"], "other")
    with pytest.raises(ValueError):
        music.read_wavs_scale_for_video([wav, other], 512, frames_per_second=60.0)


def test_read_wav_file_roundtrip(tmp_path, chirp_golden):
    from scipy.io import wavfile

    path = tmp_path / "chirp.wav"
    wavfile.write(str(path), 44100, chirp_golden["wav_int16"])
    loaded = music.read_wav_file(path)
    assert loaded.sample_rate == 44100 and loaded.name == "chirp"
    assert np.array_equal(loaded.wav_data, chirp_golden["wav_float32"])
    stereo = np.stack([chirp_golden["wav_int16"], chirp_golden["wav_int16"]], axis=1)
    wavfile.write(str(path), 44100, stereo)
    both = music.read_wavs_scale_for_video([path], 512, frames_per_second=60.0)
    assert np.array_equal(both.wav_data, chirp_golden["scaled"])  # mono mean of two equal channels


def test_spectrogram_smooth_scale_matches_oracle(chirp_golden):
    scaled = chirp_golden["scaled"]
    for amplitude_range in ((-5, 5), (-1, 1), None):
        got = apply_spectrogram.compute_spectrogram_smooth_scale(scaled, 512, amplitude_range)
        want = audio_ref.compute_spectrogram_smooth_scale(scaled, 512, amplitude_range)
        assert got.dtype == np.float64
        assert_latents_close(got, want, f"range {amplitude_range}")


@pytest.mark.parametrize("roll", [False, True])
def test_blends_match_reference_golden(chirp_golden, roll):
    tag = "roll" if roll else "noroll"
    scaled = chirp_golden["scaled"]
    noise_blend = visualization_inputs.alpha_blend_vectors_max_rms_power_audio(
        alpha=0.25, fft_roll_enabled=roll, fft_amplitude_range=(-5, 5), time_series_audio_vectors=scaled,
        vector_length=512, network_indices=[0, 1, 2])
    assert_latents_close(noise_blend.a_vectors.data, chirp_golden[f"{tag}_spectrogram"], "spectrogram")
    assert_latents_close(noise_blend.combined.data, chirp_golden[f"{tag}_noise_blend_combined"], "noise blend")
    assert np.array_equal(noise_blend.b_vectors.data, chirp_golden["noise"])
    assert np.array_equal(noise_blend.network_indices.result.data, chirp_golden[f"{tag}_noise_blend_indices"])

    projection = MatricesLabel(synthetic.projection_latents(30), 512, "synthetic")
    blend = visualization_inputs.alpha_blend_projection_file(
        final_latents_matrices_label=projection, alpha=0.25, fft_roll_enabled=roll, fft_amplitude_range=(-5, 5),
        blend_depth=12, time_series_audio_vectors=scaled, vector_length=512, network_indices=[0, 1, 2])
    combined = blend.combined.data
    assert combined.shape == (18, 60 * 512) and combined.dtype == np.float64
    for row in range(12):
        assert_latents_close(combined[row], chirp_golden[f"{tag}_projection_blend_row0"], f"blended row {row}")
    for row in range(12, 18):  # rows at and below blend_depth are the projected w, untouched: bit-exact
        assert np.array_equal(combined[row], chirp_golden[f"{tag}_projection_blend_row12"])
    assert np.array_equal(blend.network_indices.result.data, chirp_golden[f"{tag}_projection_blend_indices"])
    assert blend.b_vectors.data.shape == (18, 60 * 512)


def test_device_twins_return_generator_layout(chirp_golden):
    scaled = chirp_golden["scaled"]
    host = visualization_inputs.alpha_blend_vectors_max_rms_power_audio(0.25, True, (-5, 5), scaled, 512, [0, 1])
    dev = visualization_inputs.alpha_blend_vectors_max_rms_power_audio_device(0.25, True, (-5, 5), scaled, 512, [0, 1])
    assert dev.latents.shape == (60, 512) and dev.latents.dtype == torch.float32 and dev.latents.is_cuda
    assert np.array_equal(dev.latents.cpu().numpy().reshape(-1), host.combined.data.astype(np.float32))
    assert np.array_equal(dev.network_indices, host.network_indices.result.data)

    projection = MatricesLabel(synthetic.projection_latents(30), 512, "synthetic")
    host = visualization_inputs.alpha_blend_projection_file(projection, 0.25, True, (-5, 5), 12, scaled, 512, [0])
    dev = visualization_inputs.alpha_blend_projection_file_device(projection, 0.25, True, (-5, 5), 12, scaled, 512, [0])
    assert dev.latents.shape == (60, 18, 512)
    want = np.transpose(host.combined.data.reshape(18, 60, 512), (1, 0, 2)).astype(np.float32)
    assert np.array_equal(dev.latents.cpu().numpy(), want)


def test_projection_blend_edge_cases(chirp_golden):
    scaled = chirp_golden["scaled"]
    # blend depth 0 and 18; a projection file whose frame count does not divide the video's
    for depth in (0, 18):
        projection = MatricesLabel(synthetic.projection_latents(60), 512, "synthetic")
        got = visualization_inputs.alpha_blend_projection_file(projection, 0.5, False, (-1, 1), depth, scaled, 512, [0])
        want = audio_ref.alpha_blend_projection_file(projection.data, 0.5, False, (-1, 1), depth, scaled, 512, [0])
        assert got.combined.data.shape == want.combined.shape
        assert_latents_close(got.combined.data, want.combined, f"blend depth {depth}")
    with pytest.raises(ValueError):
        bad = MatricesLabel(synthetic.projection_latents(7), 512, "synthetic")
        visualization_inputs.alpha_blend_projection_file(bad, 0.5, False, (-1, 1), 12, scaled, 512, [0])


def test_short_and_ragged_audio():
    lib_error = audio_device._lib.GanceError  # pylint: disable=protected-access
    with pytest.raises(ValueError):  # shorter than one 510-sample window
        apply_spectrogram.compute_spectrogram_smooth_scale(np.zeros(100, dtype=np.float32), 512, (-1, 1))
    with pytest.raises(lib_error):  # 3 frames < the 7-frame Savitzky-Golay window: scipy raises as well
        apply_spectrogram.compute_spectrogram_smooth_scale(np.random.RandomState(0).randn(3 * 512).astype(np.float32), 512, (-1, 1))
    # un-padded audio: the last partial frame still yields a window when >= 510 samples remain
    data = (np.random.RandomState(1).randn(9 * 512 + 510) * 0.1).astype(np.float32)
    got = apply_spectrogram.compute_spectrogram_smooth_scale(data, 512, (-2, 2))
    want = audio_ref.compute_spectrogram_smooth_scale(data, 512, (-2, 2))
    assert got.shape == (10 * 512,)
    assert_latents_close(got, want, "ragged")
    # other vector lengths (the reference's golden test uses 1000)
    data = (np.random.RandomState(2).randn(12 * 1000) * 0.1).astype(np.float32)
    assert_latents_close(apply_spectrogram.compute_spectrogram_smooth_scale(data, 1000, (-1, 1)),
                         audio_ref.compute_spectrogram_smooth_scale(data, 1000, (-1, 1)), "vector_length 1000")


def test_album_length_track_matches_oracle_and_is_frame_sharded_consistently():
    """BASELINE config 4: 3-minute track at 60 fps = 10 800 frames, full size, against the oracle."""
    wav = synthetic.chirp_wav(180.0)
    as_float = audio_device.pcm_to_float(wav)
    assert np.array_equal(as_float.cpu().numpy(), audio_ref.wav_to_float32(wav))
    scaled = music.read_wavs_scale_for_video([music.WavFileProperties(44100, as_float.cpu().numpy(), "chirp")], 512,
                                             frames_per_second=60.0).wav_data
    assert scaled.shape == (10800 * 512,)
    want_scaled = audio_ref.scale_wav_for_video(as_float.cpu().numpy(), 44100, 512, frames_per_second=60.0)
    assert np.array_equal(scaled, want_scaled)
    dev = visualization_inputs.alpha_blend_vectors_max_rms_power_audio_device(0.25, True, (-5, 5), scaled, 512, [0, 1, 2])
    want = audio_ref.alpha_blend_vectors_max_rms_power_audio(0.25, True, (-5, 5), scaled, 512, [0, 1, 2])
    assert dev.latents.shape == (10800, 512)
    assert_latents_close(dev.latents.cpu().numpy().reshape(-1), want.combined, "album noise blend")
    assert np.array_equal(dev.network_indices, want.network_indices)


@pytest.mark.parametrize("num_vectors,sigma", [(60, 50), (150, 50), (450, 50), (1000, 20), (7, 3)])
def test_gaussian_filter_along_time_is_bit_exact_with_scipy(num_vectors, sigma):
    """gance_gaussian_filter_time_wrap against scipy.ndimage.gaussian_filter(sigma=(s,0,0), mode='wrap') (primatives.py:69),
    including tracks shorter than the kernel radius (the periodic extension wraps more than once)."""
    import scipy.ndimage

    latents = np.random.RandomState(num_vectors).randn(num_vectors, 1, 512).astype(np.float32)
    want = scipy.ndimage.gaussian_filter(input=latents, sigma=(sigma, 0, 0), mode="wrap")
    got = audio_device.gaussian_filter_time_wrap(latents, float(sigma))
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.array_equal(got, want)


def test_gaussian_noise_source_equals_the_oracle():
    """gaussian_data (host MT19937 stream -> device filter -> host RMS) bit for bit against oracle/audio_ref.py."""
    from gance_b200.vector_sources.primatives import Sigmas, gaussian_data

    for frames in (60, 333):
        got = gaussian_data(vector_length=512, num_vectors=frames, sigmas=Sigmas(across_vectors=50, within_vectors=0))
        want = audio_ref.gaussian_data(512, frames, sigma_across=50)
        assert got.dtype == np.float32 and np.array_equal(np.asarray(got), np.asarray(want))
