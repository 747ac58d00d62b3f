"""
End to end through the reference-facing calls, the way gance/projection_file_blend.py:122-199 chains them (BASELINE
configs 3 and 5 at test size): WAV -> read_wavs_scale_for_video -> alpha_blend_projection_file -> vector_synthesis over a
MultiNetwork of THREE networks of different resolutions -> scale_square_source_duplicate. Checked frame by frame against
the oracles: oracle/audio_ref.py for the latents and the per-frame network index, oracle/stylegan2_ref.py for the pixels
of the network each frame was routed to, oracle/resize_ref.py for the scaled frame.
"""

import numpy as np
import pytest
import torch

from gance_b200 import synthetic
from gance_b200.data_into_network_visualization import network_visualization, visualization_inputs
from gance_b200.image_sources import video_common
from gance_b200.network_interface import network_functions
from gance_b200.networks import container
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
from gance_b200.vector_sources import music
from gance_b200.vector_sources.vector_types import MatricesLabel
from oracle import audio_ref, resize_ref
from oracle import stylegan2_ref as ref

pytestmark = pytest.mark.gpu

NETWORKS = [(64, 0), (128, 1), (64, 2)]  # (resolution, seed): two resolutions, three networks, as in config 5
FPS, SECONDS, SIDE = 30.0, 1.0, 150


def _swelling_wav() -> np.ndarray:
    """A chirp whose loudness swells, so the RMS-driven network index visits all three networks."""
    wav = synthetic.chirp_wav(SECONDS).astype(np.float64)
    envelope = 0.15 + 0.85 * np.sin(np.linspace(0.0, np.pi, len(wav))) ** 2
    return np.round(wav * envelope).astype(np.int16)


@pytest.fixture(scope="module")
def pipeline(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    directory = tmp_path_factory.mktemp("networks")
    paths = []
    for position, (resolution, seed) in enumerate(NETWORKS):
        paths.append(container.write_synthetic_network(directory / f"net{position}.pkl", resolution, seed, perturbed=True))
    wav = _swelling_wav()
    as_float = audio_ref.wav_to_float32(wav)
    audio = music.read_wavs_scale_for_video([music.WavFileProperties(44100, as_float, "swell")], 512, frames_per_second=FPS)
    num_frames = len(audio.wav_data) // 512
    # the projection arrives the way gance/projection_file_blend.py:122-126 gets it: from an HDF5 projection file on disk
    from gance_b200.projection import projection_file_reader, projector_file_writer

    walk = synthetic.projection_latents(num_frames // 2)  # projection_fps = fps / 2
    projection_path = projector_file_writer.write_projection_file(
        directory / "projection.hdf5", projector_file_writer.matrices_to_frames(walk),
        projector_file_writer.synthetic_attributes(num_frames // 2, projection_fps=FPS / 2))
    projection_file_reader.verify_projection_file_assumptions(projection_path)
    with projection_file_reader.load_projection_file(projection_path) as reader:
        projection = projection_file_reader.final_latents_matrices_label(reader)
        assert reader.projection_attributes.projection_fps * 2 == FPS
    assert isinstance(projection, MatricesLabel) and np.array_equal(projection.data, walk)
    data = visualization_inputs.alpha_blend_projection_file(
        final_latents_matrices_label=projection, alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5),
        blend_depth=12, time_series_audio_vectors=audio.wav_data, vector_length=512, network_indices=[0, 1, 2])
    return paths, audio, projection, data, num_frames


def test_latents_and_network_indices_match_the_audio_oracle(pipeline):
    _, audio, projection, data, num_frames = pipeline
    want = audio_ref.alpha_blend_projection_file(projection.data, 0.25, True, (-5, 5), 12, audio.wav_data, 512, [0, 1, 2])
    got, expected = np.asarray(data.combined.data), np.asarray(want.combined)
    assert got.shape == expected.shape == (18, num_frames * 512)
    assert np.abs(got - expected).max() <= 1e-5 * np.abs(expected).max()
    indices = np.asarray(data.network_indices.result.data)
    assert np.array_equal(indices, np.asarray(want.network_indices))
    assert set(indices.tolist()) == {0, 1, 2}, "the test track must exercise all three networks"


def test_model_switching_frames_match_the_oracle_in_frame_order(pipeline):
    paths, _, _, data, num_frames = pipeline
    indices = np.asarray(data.network_indices.result.data)
    with network_functions.MultiNetwork(paths) as networks:
        output = network_visualization.vector_synthesis(data=data, networks=networks, batch_size=8)
        frames = [np.array(frame) for frame in output.synthesized_images]
    assert len(frames) == num_frames and output.visualization_images is None
    matrices = np.transpose(np.asarray(data.combined.data).reshape(18, num_frames, 512), (1, 0, 2))
    oracles = {}
    for frame_index in range(0, num_frames, 3):  # every third frame through the float64 oracle (seconds each)
        network = int(indices[frame_index])
        resolution, seed = NETWORKS[network]
        if network not in oracles:
            spec = GeneratorSpec(resolution)
            oracles[network] = (spec, ref.ReferenceGenerator(random_init_variables(spec, seed, perturbed=True), torch.float64))
        spec, oracle = oracles[network]
        assert frames[frame_index].shape == (resolution, resolution, 3) and frames[frame_index].dtype == np.uint8
        want = ref.convert_images_to_uint8(oracle.run_matrix(matrices[frame_index][None, : spec.num_ws]))[0]
        diff = np.abs(frames[frame_index].astype(np.int32) - want.astype(np.int32))
        assert float((diff <= 1).mean()) >= 0.999, (frame_index, network, int(diff.max()))
    for frame_index in range(num_frames):  # every frame has its own network's resolution: routing is per frame
        assert frames[frame_index].shape[0] == NETWORKS[int(indices[frame_index])][0]


def test_scaled_output_matches_the_resize_oracle_and_duplicates_frames(pipeline):
    paths, _, _, data, num_frames = pipeline
    with network_functions.MultiNetwork(paths) as networks:
        native = [np.array(f) for f in network_visualization.vector_synthesis(data=data, networks=networks, batch_size=8,
                                                                                frames_to_visualize=12).synthesized_images]
        again = network_visualization.vector_synthesis(data=data, networks=networks, batch_size=8, frames_to_visualize=12)
        scaled = list(video_common.scale_square_source_duplicate(again.synthesized_images, SIDE, frame_multiplier=2))
    assert len(native) == 12 and len(scaled) == 24
    for frame_index, frame in enumerate(native):
        want = resize_ref.resize_cubic_u8(frame, SIDE, SIDE)
        assert np.array_equal(scaled[2 * frame_index], want) and np.array_equal(scaled[2 * frame_index + 1], want)


def test_grouping_by_network_does_not_change_the_frames(pipeline):
    """force_optimize_synthesis_order only reorders work (network_visualization.py:653-674): frames are identical."""
    paths, _, _, data, _ = pipeline
    with network_functions.MultiNetwork(paths) as networks:
        grouped = [np.array(f) for f in network_visualization.vector_synthesis(
            data=data, networks=networks, batch_size=6, frames_to_visualize=18, force_optimize_synthesis_order=True).synthesized_images]
        in_order = [np.array(f) for f in network_visualization.vector_synthesis(
            data=data, networks=networks, batch_size=6, frames_to_visualize=18, force_optimize_synthesis_order=False).synthesized_images]
    assert len(grouped) == len(in_order) == 18
    for a, b in zip(grouped, in_order):
        assert np.array_equal(a, b)
