"""
Generates tests/golden/*.npz by running the REFERENCE'S OWN Python (imported from
/root/reference through oracle/reference_shim.py) on the synthetic inputs of
gance_b200/synthetic.py. Only runs in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The fixtures are what the `-m "not gpu"` and `-m gpu` tests compare the oracle and the CUDA path
against on machines that do not have the reference.
"""

import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from gance_b200 import synthetic  # noqa: E402
from oracle import reference_shim  # noqa: E402

GOLDEN_DIR = Path(__file__).resolve().parent
CLAPS = Path("/root/reference/test/assets/claps.wav")


def sha(array: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(array).tobytes()).hexdigest()


def main() -> None:
    ref = reference_shim.load()
    music, common, reduction = ref["music"], ref["vector_sources_common"], ref["vector_reduction"]
    inputs, types = ref["visualization_inputs"], ref["vector_types"]

    # ---- known answers the reference's own tests hold -------------------------
    claps = music.read_wav_file(CLAPS)
    audio_1000 = music.read_wavs_scale_for_video(
        wavs=[CLAPS], vector_length=1000, frames_per_second=60.0
    ).wav_data
    first = common.sub_vectors(data=audio_1000, vector_length=1000)[0]
    rms_first = reduction.reduce_vector_rms_rolling_max(first, 1000).result.data
    lengths = {}
    for multiplier in (2.0, 1.5, 0.3, 0.1, 10):
        scaled = music._scale_wav_to_sample_rate(  # pylint: disable=protected-access
            wav_file=claps, new_sample_rate=int(claps.sample_rate * multiplier)
        )
        lengths[str(multiplier)] = len(scaled.wav_data)
    np.savez_compressed(
        GOLDEN_DIR / "claps_known_answers.npz",
        sample_rate=claps.sample_rate,
        num_samples=len(claps.wav_data),
        claps_float32_sha256=sha(claps.wav_data),
        claps_float32_head=claps.wav_data[:4096],
        scaled_1000_60fps_len=len(audio_1000),
        scaled_1000_60fps_head=audio_1000[:2000],
        first_frame_rms=rms_first,
        multipliers=np.array([float(k) for k in lengths]),
        scaled_lengths=np.array(list(lengths.values())),
    )

    # ---- 1 s chirp, T = 60 (BASELINE config 1 audio) -----------------------------------------
    wav = synthetic.chirp_wav(1.0)
    as_float = np.array(
        list(common.remap_values_into_range(wav, (-32768, 32767), (-1, 1)))
    ).astype(np.float32)
    scaled = music.read_wavs_scale_for_video(
        [music.WavFileProperties(44100, as_float, "chirp")], 512, frames_per_second=60.0
    ).wav_data
    spectrogram_db = ref["apply_spectrogram"].compute_spectrogram(scaled, 512)
    out = dict(wav_int16=wav, wav_float32=as_float, scaled=scaled, spectrogram_db=spectrogram_db)
    projection = synthetic.projection_latents(30)
    for roll in (False, True):
        tag = "roll" if roll else "noroll"
        noise_blend = inputs.alpha_blend_vectors_max_rms_power_audio(
            alpha=0.25,
            fft_roll_enabled=roll,
            fft_amplitude_range=(-5, 5),
            time_series_audio_vectors=scaled,
            vector_length=512,
            network_indices=[0, 1, 2],
        )
        out[f"{tag}_spectrogram"] = noise_blend.a_vectors.data
        out[f"{tag}_noise_blend_combined"] = noise_blend.combined.data
        out[f"{tag}_noise_blend_indices"] = noise_blend.network_indices.result.data
        if not roll:
            out["noise"] = noise_blend.b_vectors.data
        blend = inputs.alpha_blend_projection_file(
            final_latents_matrices_label=types.MatricesLabel(projection, 512, "synthetic"),
            alpha=0.25,
            fft_roll_enabled=roll,
            fft_amplitude_range=(-5, 5),
            blend_depth=12,
            time_series_audio_vectors=scaled,
            vector_length=512,
            network_indices=[0, 1, 2],
        )
        combined = blend.combined.data
        assert combined.shape == (18, 60 * 512)
        # rows [0,12) are identical, rows [12,18) are identical: keep one of each + a digest of all
        assert all(np.array_equal(combined[0], combined[i]) for i in range(12))
        assert all(np.array_equal(combined[12], combined[i]) for i in range(12, 18))
        out[f"{tag}_projection_blend_row0"] = combined[0]
        out[f"{tag}_projection_blend_row12"] = combined[12]
        out[f"{tag}_projection_blend_sha256"] = sha(combined)
        out[f"{tag}_projection_blend_indices"] = blend.network_indices.result.data
    np.savez_compressed(GOLDEN_DIR / "chirp_1s_60fps.npz", **out)
    for name in ("claps_known_answers.npz", "chirp_1s_60fps.npz"):
        print(name, (GOLDEN_DIR / name).stat().st_size, "bytes")


if __name__ == "__main__":
    main()
