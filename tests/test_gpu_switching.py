"""
`-m gpu`: the vector path through the reference-facing calls when a per-network group has EXACTLY num_ws frames (14 on a
256^2 network, 16 on 512^2, 18 on 1024^2), which round 1 rendered as one matrix frame and padded with None; and model
switching on resident generators (no re-allocation / re-capture after the first visit of each network).

Reference behaviour being matched: one latent -> one frame, yielded in frame-index order
(gance/data_into_network_visualization/network_visualization.py:598-681), vectors rendered by `Gs.run(z, None,
truncation_psi=1.2)` (gance/network_interface/network_functions.py:151-158). The reference draws fresh noise per call;
the parity checks here set the module's VECTOR_NOISE_MODE to "const" so both sides use the network's noise buffers.
"""

import numpy as np
import pytest
import torch

from gance_b200.data_into_network_visualization import network_visualization
from gance_b200.data_into_network_visualization.visualization_common import DataLabel, ResultLayers, VisualizationInput
from gance_b200.network_interface import network_functions
from gance_b200.networks import container
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
from gance_b200.vector_sources.vector_types import VectorsLabel
from oracle import stylegan2_ref as ref

pytestmark = pytest.mark.gpu

NETWORKS = [(256, 0), (512, 1), (1024, 2)]  # (resolution, seed): num_ws 14 / 16 / 18


@pytest.fixture(scope="module")
def network_paths(tmp_path_factory):
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    directory = tmp_path_factory.mktemp("switching")
    return [container.write_synthetic_network(directory / f"net{i}.pkl", resolution, seed, perturbed=True)
            for i, (resolution, seed) in enumerate(NETWORKS)]


@pytest.fixture()
def const_noise():
    previous = network_functions.VECTOR_NOISE_MODE
    network_functions.VECTOR_NOISE_MODE = "const"
    yield
    network_functions.VECTOR_NOISE_MODE = previous


def _oracle_frame(network: int, z: np.ndarray) -> np.ndarray:
    resolution, seed = NETWORKS[network]
    spec = GeneratorSpec(resolution)
    oracle = ref.ReferenceGenerator(random_init_variables(spec, seed, perturbed=True), torch.float64)
    return ref.convert_images_to_uint8(oracle.run_vector(z[None]))[0]


def _within_one_lsb(got: np.ndarray, want: np.ndarray) -> float:
    return float((np.abs(got.astype(np.int32) - want.astype(np.int32)) <= 1).mean())


def _schedule():
    """48 frames, batch 32: window 0 holds 14 frames of network 0 interleaved with 18 of network 2, window 1 holds 16 of network 1."""
    window0 = np.array([0, 2] * 14 + [2] * 4)
    np.random.RandomState(5).shuffle(window0)
    indices = np.concatenate([window0, np.full(16, 1)])
    assert (window0 == 0).sum() == 14 and (window0 == 2).sum() == 18 and len(indices) == 48
    return indices


def test_num_ws_sized_vector_groups_render_every_frame_in_order(network_paths, const_noise):
    indices = _schedule()
    z = np.random.RandomState(8).randn(len(indices), 512)
    data = VisualizationInput(
        a_vectors=VectorsLabel(z.reshape(-1), 512, "a"), b_vectors=VectorsLabel(z.reshape(-1), 512, "b"),
        combined=VectorsLabel(z.reshape(-1), 512, "combined"),
        network_indices=ResultLayers(result=DataLabel(indices, "indices")))
    with network_functions.MultiNetwork(network_paths) as networks:
        output = network_visualization.vector_synthesis(data=data, networks=networks, batch_size=32)
        frames = [None if frame is None else np.array(frame) for frame in output.synthesized_images]
        assert len(frames) == len(indices)
        for frame_index, frame in enumerate(frames):
            side = NETWORKS[int(indices[frame_index])][0]
            assert frame is not None and frame.shape == (side, side, 3) and frame.dtype == np.uint8, frame_index
        # order + content: a frame equals the per-frame call the reference would have made for it (same noise: const)
        for frame_index in (0, 5, 31, 32, 47):
            alone = networks.indexed_create_image_vector(int(indices[frame_index]), z[frame_index])
            assert np.array_equal(alone, frames[frame_index]), frame_index
    # oracle parity of one frame per network, first and last member of its num_ws-sized group
    for network in (0, 1, 2):
        members = np.flatnonzero(indices == network)
        assert len(members) == GeneratorSpec(NETWORKS[network][0]).num_ws
        for frame_index in (members[0], members[-1]) if network != 2 else (members[-1],):
            want = _oracle_frame(network, z[frame_index])
            assert _within_one_lsb(frames[frame_index], want) >= 0.999, (network, frame_index)
    # all frames of a group differ from one another (a mis-routed batch repeats or drops frames)
    group = [frames[i] for i in np.flatnonzero(indices == 1)]
    assert len({frame.tobytes() for frame in group}) == len(group)


@pytest.mark.parametrize("network", [0, 1, 2])
def test_num_ws_sized_batches_through_indexed_render_batches(network_paths, const_noise, network):
    num_ws = GeneratorSpec(NETWORKS[network][0]).num_ws
    z = np.random.RandomState(20 + network).randn(2 * num_ws + 3, 512)
    batches = [z[:num_ws], z[num_ws: 2 * num_ws], z[2 * num_ws:]]
    with network_functions.MultiNetwork(network_paths) as networks:
        rendered = [np.array(frames) for frames in networks.indexed_render_batches(network, iter(batches), max_batch=32)]
        assert [len(frames) for frames in rendered] == [num_ws, num_ws, 3]
        together = np.concatenate(rendered)
        direct = networks.indexed_create_images_batch(network, z)
        assert np.array_equal(together, direct)
        for position in (0, num_ws - 1, num_ws, len(z) - 1):
            assert np.array_equal(networks.indexed_create_image_vector(network, z[position]), together[position])
    assert _within_one_lsb(together[num_ws - 1], _oracle_frame(network, z[num_ws - 1])) >= 0.999


def test_switching_back_and_forth_reuses_buffers_renderers_and_graphs(network_paths, const_noise):
    """After each network has been visited once, alternating between them allocates nothing and captures nothing."""
    z = np.random.RandomState(31).randn(16, 512)
    with network_functions.MultiNetwork(network_paths[:2]) as networks:
        generators = [network_functions.resident_generator(path) for path in network_paths[:2]]

        def visit(index):
            return [np.array(f) for f in networks.indexed_render_batches(index, iter([z, z[:8]]), max_batch=16)]

        first = [visit(0), visit(1)]
        captures = [g.graph_capture_count() for g in generators]
        pointers = [{key: buf.data_ptr() for key, buf in g._buffers.items()} for g in generators]  # pylint: disable=protected-access
        renderers = [dict(g._renderers) for g in generators]  # pylint: disable=protected-access
        assert all(count >= 2 for count in captures)  # batch 16 and batch 8, one capture per renderer slot (the resident generators may hold graphs of earlier tests too)
        torch.cuda.synchronize()
        allocated = torch.cuda.memory_allocated()
        for _ in range(3):
            again = [visit(0), visit(1)]
            for a, b in zip(first, again):
                assert all(np.array_equal(x, y) for x, y in zip(a, b))
        assert [g.graph_capture_count() for g in generators] == captures
        assert [{key: buf.data_ptr() for key, buf in g._buffers.items()} for g in generators] == pointers  # pylint: disable=protected-access
        assert all(dict(g._renderers) == r for g, r in zip(generators, renderers))  # pylint: disable=protected-access
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == allocated
    assert all(not g._buffers for g in generators)  # pylint: disable=protected-access  (unload released the idle buffers)


def test_two_interfaces_of_one_file_do_not_disturb_each_other(network_paths, const_noise):
    z = np.random.RandomState(32).randn(4, 512)
    one = network_functions.create_network_interface_process(network_paths[0])
    two = network_functions.create_network_interface_process(network_paths[0])
    want = one.network_interface.create_images_batch(z)
    stream = two.network_interface.render_batches(iter([z, z]), 4)
    first = np.array(next(stream))
    one.stop_function()  # must not pull buffers or graphs from under `two`
    second = np.array(next(stream))
    assert np.array_equal(first, want) and np.array_equal(second, want)
    with pytest.raises(RuntimeError):
        one.network_interface.create_images_batch(z)
    two.stop_function()
    network_functions.release_idle_buffers()


def test_ragged_batch_sizes_share_one_set_of_work_buffers(network_paths):
    """Batches of 1..9 frames on a 512^2 network: the work buffers are sized by the largest batch, not one set per size."""
    generator = network_functions.resident_generator(network_paths[1])
    generator.release_buffers()
    z = torch.as_tensor(np.random.RandomState(33).randn(9, 512))
    reference = generator.run_vectors(z, noise_mode="const").clone()
    torch.cuda.synchronize()
    allocated = torch.cuda.memory_allocated()
    for count in (1, 5, 8, 3, 9):
        frames = generator.run_vectors(z[:count], noise_mode="const")
        assert torch.equal(frames, reference[:count])
        del frames
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() <= allocated + (1 << 20)
    generator.release_buffers()
