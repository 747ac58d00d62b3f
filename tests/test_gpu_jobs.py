"""
`-m gpu`: the callers of the path end to end (SURVEY.md §8 "next" rows either side of the kernels), checked against the
oracles: `projection_file_blend_frames` / `_api` (gance/projection_file_blend.py:56-199) from a projection FILE and a WAV
FILE on disk, `noise_blend_frames` (music_into_networks.py:338-378), and the frame-sharded whole-track jobs that bench.py
reports (tools/track_jobs.py) at a short track length.
"""

import json
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
from scipy.io import wavfile

from gance_b200 import noise_blend, projection_file_blend, synthetic
from gance_b200.network_interface import network_functions
from gance_b200.networks import container
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
from gance_b200.projection import projector_file_writer
from oracle import audio_ref, resize_ref
from oracle import stylegan2_ref as ref

pytestmark = pytest.mark.gpu

NETWORKS = [(64, 0), (128, 1)]
FPS, SECONDS, SIDE = 30.0, 1.0, 96


@pytest.fixture(scope="module")
def inputs(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    directory = tmp_path_factory.mktemp("jobs")
    paths = [container.write_synthetic_network(directory / f"net{i}.pkl", resolution, seed, perturbed=True)
             for i, (resolution, seed) in enumerate(NETWORKS)]
    wav = synthetic.chirp_wav(SECONDS).astype(np.float64)
    wav = np.round(wav * (0.15 + 0.85 * np.sin(np.linspace(0.0, np.pi, len(wav))) ** 2)).astype(np.int16)
    wav_path = directory / "track.wav"
    wavfile.write(str(wav_path), 44100, wav)
    frames = int(SECONDS * FPS)
    walk = synthetic.projection_latents(frames // 2)
    projection_path = projector_file_writer.write_projection_file(
        directory / "projection.hdf5", projector_file_writer.matrices_to_frames(walk),
        projector_file_writer.synthetic_attributes(frames // 2, projection_fps=FPS / 2))
    return directory, paths, wav_path, wav, projection_path, walk, frames


def _within(got, want):
    return float((np.abs(got.astype(np.int32) - want.astype(np.int32)) <= 1).mean())


def test_projection_file_blend_frames_match_the_oracles(inputs):
    _, paths, wav_path, wav, projection_path, walk, frames = inputs
    source = projection_file_blend.projection_file_blend_frames(
        wav=[str(wav_path)], network_paths=paths, frames_to_visualize=None, output_fps=FPS, output_side_length=SIDE,
        alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5), projection_file_path=str(projection_path),
        blend_depth=12, batch_size=8)
    got = projection_file_blend.frames_of(source)
    assert got.shape == (frames, SIDE, SIDE, 3) and got.dtype == np.uint8
    # the oracle's view of the same job: audio scaled to `frames` vectors, blended w+, per-frame network index
    audio = audio_ref.scale_wav_for_video(audio_ref.wav_to_float32(wav), 44100, 512, target_num_vectors=frames)
    want = audio_ref.alpha_blend_projection_file(walk, 0.25, True, (-5, 5), 12, audio, 512, [0, 1])
    indices = np.asarray(want.network_indices)
    assert set(indices.tolist()) == {0, 1}
    matrices = np.transpose(np.asarray(want.combined).reshape(18, frames, 512), (1, 0, 2))
    for frame in range(0, frames, 4):
        resolution, seed = NETWORKS[int(indices[frame])]
        spec = GeneratorSpec(resolution)
        oracle = ref.ReferenceGenerator(random_init_variables(spec, seed, perturbed=True), torch.float64)
        native = ref.convert_images_to_uint8(oracle.run_matrix(matrices[frame][None, : spec.num_ws]))[0]
        scaled = resize_ref.resize_cubic_u8(native, SIDE, SIDE)
        # +-1 LSB of the 16-bit mode before the resize can move a resized pixel by a little more than 1
        diff = np.abs(got[frame].astype(np.int32) - scaled.astype(np.int32))
        assert float((diff <= 2).mean()) >= 0.999 and float((diff <= 1).mean()) >= 0.99, (frame, int(diff.max()))


def test_projection_file_blend_api_writes_a_raw_frame_file_and_keeps_the_references_errors(inputs):
    directory, paths, wav_path, _, projection_path, walk, frames = inputs
    output = directory / "frames.u8"
    arguments = dict(wav=[str(wav_path)], output_path=str(output), network_paths=paths[:1], frames_to_visualize=10,
                     output_fps=FPS, output_side_length=64, debug_path=None, debug_window=None, debug_side_length=None,
                     alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5), projection_file_path=str(projection_path),
                     blend_depth=12, complexity_change_rolling_sum_window=None, complexity_change_threshold=None,
                     phash_distance=None, bbox_distance=None, track_length=None)
    assert projection_file_blend.projection_file_blend_api(**arguments) == 10
    assert output.stat().st_size == 10 * 64 * 64 * 3
    assert json.loads(Path(str(output) + ".json").read_text())["shape"] == [10, 64, 64, 3]
    with pytest.raises(ValueError, match="Overlay music mask"):
        projection_file_blend.projection_file_blend_api(**{**arguments, "complexity_change_rolling_sum_window": 3,
                                                           "complexity_change_threshold": 1})
    with pytest.raises(ValueError):  # 25 fps over a 15 fps projection does not divide (gance/divisor.py:10-24)
        projection_file_blend.projection_file_blend_api(**{**arguments, "output_fps": 25.0})
    incomplete = projector_file_writer.write_projection_file(
        directory / "incomplete.hdf5", projector_file_writer.matrices_to_frames(walk),
        projector_file_writer.synthetic_attributes(frames // 2, projection_fps=FPS / 2, complete=False))
    with pytest.raises(ValueError, match="Invalid Projection File"):
        projection_file_blend.projection_file_blend_api(**{**arguments, "projection_file_path": str(incomplete)})
    with pytest.raises(NotImplementedError):
        projection_file_blend.projection_file_blend_api(**{**arguments, "phash_distance": 1, "bbox_distance": 1.0, "track_length": 3})


def test_noise_blend_frames_route_every_frame_to_its_network(inputs):
    _, paths, wav_path, wav, _, _, _ = inputs
    previous = network_functions.VECTOR_NOISE_MODE
    network_functions.VECTOR_NOISE_MODE = "const"
    try:
        got = projection_file_blend.frames_of(noise_blend.noise_blend_frames(
            wav=[str(wav_path)], network_paths=paths, frames_to_visualize=None, output_fps=FPS, output_side_length=SIDE,
            alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5), batch_size=8))
    finally:
        network_functions.VECTOR_NOISE_MODE = previous
    audio = audio_ref.scale_wav_for_video(audio_ref.wav_to_float32(wav), 44100, 512, frames_per_second=FPS)
    want = audio_ref.alpha_blend_vectors_max_rms_power_audio(0.25, True, (-5, 5), audio, 512, [0, 1])
    latents = np.asarray(want.combined).reshape(-1, 512)
    indices = np.asarray(want.network_indices)
    assert got.shape == (len(latents), SIDE, SIDE, 3)
    for frame in (0, len(latents) // 2, len(latents) - 1):
        resolution, seed = NETWORKS[int(indices[frame])]
        oracle = ref.ReferenceGenerator(random_init_variables(GeneratorSpec(resolution), seed, perturbed=True), torch.float64)
        native = ref.convert_images_to_uint8(oracle.run_vector(latents[frame][None]))[0]
        diff = np.abs(got[frame].astype(np.int32) - resize_ref.resize_cubic_u8(native, SIDE, SIDE).astype(np.int32))
        assert float((diff <= 2).mean()) >= 0.999, (frame, int(diff.max()))


@pytest.mark.parametrize("kind,seconds", [("album", 2.0), ("projection", 2.0), ("switching", 8.0)])
def test_whole_track_jobs_deliver_every_frame_in_order(kind, seconds, tmp_path):
    """The jobs bench.py reports, at a short track length: frames in the sink are verified against re-rendered samples."""
    sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
    import track_jobs

    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    report = track_jobs.run_track_job(kind, seconds, device, directory=tmp_path,
                                      output_side=540 if kind == "switching" else None)
    assert report["frames"] == int(seconds * 60) and report["frame_order_ok"], report
    assert report["frames_per_s"] > 0 and len(report["frames_checked_for_order"]) >= 2
    if kind == "switching":
        assert len(report["networks_used"]) == 3, report["networks_used"]  # the swell visits every network
    network_functions.evict_resident_networks()
