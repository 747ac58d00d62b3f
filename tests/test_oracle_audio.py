"""
Pins oracle/audio_ref.py + oracle/third_party_ref.py: against the golden vectors generated from
the reference's own Python (tests/golden/make_golden.py), against the reference's known answers
(test/test_dynamic_model_switching.py:39, test/test_vector_source_music.py:24,
test/test_vector_sources_common.py:66-97), and - where /root/reference exists - against the
reference itself on fresh inputs.
"""

import hashlib
from pathlib import Path

import numpy as np
import pytest

from gance_b200 import synthetic
from oracle import audio_ref, reference_shim, third_party_ref

CLAPS = Path("/root/reference/test/assets/claps.wav")
needs_reference = pytest.mark.skipif(not reference_shim.available(), reason="/root/reference absent")


def test_synthetic_inputs_match_golden(chirp_golden):
    assert np.array_equal(synthetic.chirp_wav(1.0), chirp_golden["wav_int16"])


def test_wav_to_float_and_resample_bit_exact(chirp_golden):
    as_float = audio_ref.wav_to_float32(chirp_golden["wav_int16"])
    assert as_float.dtype == np.float32
    assert np.array_equal(as_float, chirp_golden["wav_float32"])
    scaled = audio_ref.scale_wav_for_video(as_float, 44100, 512, frames_per_second=60.0)
    assert scaled.shape == (60 * 512,)
    assert np.array_equal(scaled, chirp_golden["scaled"])


def test_spectrogram_db(chirp_golden):
    db = audio_ref.compute_spectrogram(chirp_golden["scaled"], 512)
    assert db.shape == (255, 60)
    np.testing.assert_allclose(db, chirp_golden["spectrogram_db"], rtol=0, atol=1e-9)


@pytest.mark.parametrize("roll", [False, True])
def test_blends_match_reference_golden(chirp_golden, roll):
    tag = "roll" if roll else "noroll"
    scaled = chirp_golden["scaled"]
    noise_blend = audio_ref.alpha_blend_vectors_max_rms_power_audio(0.25, roll, (-5, 5), scaled, 512, [0, 1, 2])
    np.testing.assert_allclose(noise_blend.spectrogram, chirp_golden[f"{tag}_spectrogram"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(noise_blend.combined, chirp_golden[f"{tag}_noise_blend_combined"], rtol=0, atol=1e-12)
    assert np.array_equal(noise_blend.network_indices, chirp_golden[f"{tag}_noise_blend_indices"])
    projection = synthetic.projection_latents(30)
    blend = audio_ref.alpha_blend_projection_file(projection, 0.25, roll, (-5, 5), 12, scaled, 512, [0, 1, 2])
    assert blend.combined.shape == (18, 60 * 512)
    np.testing.assert_allclose(blend.combined[0], chirp_golden[f"{tag}_projection_blend_row0"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(blend.combined[11], chirp_golden[f"{tag}_projection_blend_row0"], rtol=0, atol=1e-12)
    for row in range(12, 18):  # rows at and below blend_depth are the projected w untouched
        assert np.array_equal(blend.combined[row], chirp_golden[f"{tag}_projection_blend_row12"])
    assert np.array_equal(blend.network_indices, chirp_golden[f"{tag}_projection_blend_indices"])
    if not roll:  # without the roll the whole array is reproduced bit for bit
        digest = hashlib.sha256(np.ascontiguousarray(blend.combined).tobytes()).hexdigest()
        assert digest == str(chirp_golden[f"{tag}_projection_blend_sha256"])


def test_noise_source(chirp_golden):
    noise = audio_ref.minmax_scale_1d(audio_ref.gaussian_data(512, 60, sigma_across=50), (-4, 4))
    assert noise.dtype == np.float32
    np.testing.assert_allclose(noise, chirp_golden["noise"], rtol=0, atol=1e-6)
    assert noise.min() == pytest.approx(-4) and noise.max() == pytest.approx(4)


def test_known_answer_fixture_values(claps_golden):
    # test/test_dynamic_model_switching.py:39
    assert np.isclose(0.00298562, claps_golden["first_frame_rms"][0])
    # test/test_vector_source_music.py:24
    n = int(claps_golden["num_samples"])
    for multiplier, length in zip(claps_golden["multipliers"], claps_golden["scaled_lengths"]):
        assert int(n * multiplier) == int(length)


@needs_reference
@pytest.mark.reference
def test_known_answer_rms_from_claps(claps_golden):
    """The reference's own known answer, re-derived with the restated resampy / librosa."""
    from scipy.io import wavfile

    sample_rate, raw = wavfile.read(str(CLAPS))
    as_float = audio_ref.wav_to_float32(raw)
    assert hashlib.sha256(as_float.tobytes()).hexdigest() == str(claps_golden["claps_float32_sha256"])
    audio = audio_ref.scale_wav_for_video(as_float, sample_rate, 1000, frames_per_second=60.0)
    assert len(audio) == int(claps_golden["scaled_1000_60fps_len"])
    assert np.array_equal(audio[:2000], claps_golden["scaled_1000_60fps_head"])
    rms = third_party_ref.librosa_rms(audio[:1000], frame_length=1000, center=False)[0]
    assert rms.shape[0] == 1
    assert np.isclose(0.00298562, rms[0])


@needs_reference
@pytest.mark.reference
@pytest.mark.parametrize("multiplier", [2.0, 1.5, 0.3, 0.1, 10])
def test_resample_length_law(multiplier):
    from scipy.io import wavfile

    sample_rate, raw = wavfile.read(str(CLAPS))
    out = third_party_ref.resample(raw.astype(np.float32), sample_rate, int(sample_rate * multiplier))
    assert int(len(raw) * multiplier) == len(out)


def test_sub_vectors_shapes():
    """test/test_vector_sources_common.py:66-83."""
    assert audio_ref.sub_vectors(np.zeros(5120), 512).shape == (10, 512)
    assert audio_ref.sub_vectors(np.zeros((18, 5120)), 512).shape == (10, 18, 512)
    assert audio_ref.sub_vectors(np.zeros((18, 512)), 512).shape == (1, 18, 512)


def test_fourier_resample_sanity():
    """test/test_vector_sources_common.py:16-63: up-then-down resampling returns the signal (atol 0.5)."""
    x = np.sin(np.linspace(0, 10, 255))
    up = audio_ref.reshape_spectrogram_to_vectors(x[:, None], 512, None)
    assert up.shape == (512,)
    from scipy.signal import resample

    np.testing.assert_allclose(resample(up, 255), x, atol=0.5)


def test_render_order_is_stable_sort():
    idx = np.array([2, 0, 1, 0, 2, 1])
    assert audio_ref.render_order(idx, 3) == [1, 3, 2, 5, 0, 4]
    assert audio_ref.render_order(idx, 1) == [0, 1, 2, 3, 4, 5]


@needs_reference
@pytest.mark.reference
def test_against_reference_on_fresh_input():
    """A second, differently seeded input straight through the reference (2 s, target-count mode)."""
    ref = reference_shim.load()
    rng = np.random.RandomState(7)
    wav = (rng.randn(22050) * 3000).astype(np.int16)
    as_float = audio_ref.wav_to_float32(wav)
    music = ref["music"]
    theirs = music.read_wavs_scale_for_video(
        [music.WavFileProperties(44100, as_float, "n")], 512, target_num_vectors=24
    ).wav_data
    mine = audio_ref.scale_wav_for_video(as_float, 44100, 512, target_num_vectors=24)
    assert np.array_equal(theirs, mine)
    projection = synthetic.projection_latents(12, seed=3)
    types = ref["vector_types"]
    theirs_blend = ref["visualization_inputs"].alpha_blend_projection_file(
        final_latents_matrices_label=types.MatricesLabel(projection, 512, "p"),
        alpha=0.4,
        fft_roll_enabled=True,
        fft_amplitude_range=(-1, 1),
        blend_depth=10,
        time_series_audio_vectors=theirs,
        vector_length=512,
        network_indices=[0, 1],
    )
    mine_blend = audio_ref.alpha_blend_projection_file(projection, 0.4, True, (-1, 1), 10, mine, 512, [0, 1])
    np.testing.assert_allclose(mine_blend.combined, theirs_blend.combined.data, rtol=0, atol=1e-12)
    assert np.array_equal(mine_blend.network_indices, theirs_blend.network_indices.result.data)


def test_eight_bit_pcm_behaves_as_the_reference_does(tmp_path):
    """
    `read_wav_file` (gance/vector_sources/music.py:190-205) maps int8 data through interp1d([0, 255], [-1, 1]): samples in
    [0, 127] land in [-1, -0.004], a NEGATIVE sample is outside the interpolation range and raises ValueError; uint8 (what
    scipy.io.wavfile actually returns for 8-bit files) is refused as an unknown format. Oracle and drop-in both do that -
    the host-side checks run before any GPU work, so this is a CPU test.
    """
    from unittest import mock

    from gance_b200.vector_sources import music

    ok = np.array([0, 1, 64, 127], dtype=np.int8)
    want = (2.0 / 255.0 * ok.astype(np.float64) - 1.0).astype(np.float32)
    np.testing.assert_allclose(audio_ref.wav_to_float32(ok), want, rtol=0, atol=1e-7)
    with pytest.raises(ValueError):
        audio_ref.wav_to_float32(np.array([3, -1], dtype=np.int8))
    with pytest.raises(ValueError):
        audio_ref.wav_to_float32(np.array([3, 200], dtype=np.uint8))
    for data, message in ((np.array([3, -1], dtype=np.int8), "interpolation range"), (np.array([3, 200], dtype=np.uint8), "unknown input format")):
        with mock.patch.object(music.wavfile, "read", return_value=(44100, data)):
            with pytest.raises(ValueError, match=message):
                music.read_wav_file(tmp_path / "eight_bit.wav")


@needs_reference
def test_eight_bit_pcm_reference_errors(tmp_path):
    """The reference's own `read_wav_file` on the same arrays (scipy's reader patched to hand them over)."""
    from unittest import mock

    reference_music = reference_shim.load()["music"]
    with mock.patch.object(reference_music.wavfile, "read", return_value=(44100, np.array([0, 64, 127], dtype=np.int8))):
        got = reference_music.read_wav_file(tmp_path / "x.wav").wav_data
    np.testing.assert_array_equal(got, audio_ref.wav_to_float32(np.array([0, 64, 127], dtype=np.int8)))
    for data in (np.array([3, -1], dtype=np.int8), np.array([3, 200], dtype=np.uint8)):
        with mock.patch.object(reference_music.wavfile, "read", return_value=(44100, data)):
            with pytest.raises(ValueError):
                reference_music.read_wav_file(tmp_path / "x.wav")


def test_prefetched_noise_values_are_the_serial_draw():
    """`NoisePrefetch` (the RNG draw moved to a background thread, started from an upper bound of the track length) hands
    `gaussian_data` bit for bit what `RandomState(1234).randn(T, 1, L)` returns - for T below, at and above the bound."""
    from gance_b200.vector_sources import primatives

    length = 64
    for bound, count in ((40, 33), (40, 40), (40, 57), (0, 5)):
        want = np.random.RandomState(primatives.DEFAULT_RANDOM_SEED).randn(count, 1, length).astype(np.float32)  # pylint: disable=no-member
        got = primatives.NoisePrefetch(length, bound).take(count)
        assert got.dtype == np.float32 and got.shape == want.shape and np.array_equal(got, want)
        got[...] = 0  # writable, and not aliased with a later take
    no_filter = primatives.Sigmas(across_vectors=0, within_vectors=0)
    plain = primatives.gaussian_data(length, 33, no_filter)
    ahead = primatives.gaussian_data(length, 33, no_filter, prefetched=primatives.NoisePrefetch(length, 48))
    assert np.array_equal(plain, ahead)
    other_length = primatives.gaussian_data(length, 33, no_filter, prefetched=primatives.NoisePrefetch(length * 2, 48))
    assert np.array_equal(plain, other_length)  # a prefetch of the wrong shape is ignored
