"""
CPU tests of the multi-GPU layout: contiguous frame ranges and the ONE collective
on the path, the gather of uint8 frames to the encoder rank - run with world_size 2 and 3 over
`gloo` (NCCL on the GPUs uses the same code).
"""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gance_b200.data_into_network_visualization.network_visualization import frame_schedule
from gance_b200.frame_sharding import all_frame_ranges, frame_range, gather_frames


@pytest.mark.parametrize("world,frames", [(1, 60), (2, 60), (4, 10800), (8, 10800), (8, 61), (3, 2), (8, 0)])
def test_frame_ranges_partition_the_track(world, frames):
    ranges = all_frame_ranges(world, frames)
    assert ranges[0][0] == 0 and ranges[-1][1] == frames
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a0 <= a1 and b0 <= b1
    sizes = [b - a for a, b in ranges]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_frame_range_rejects_bad_rank():
    with pytest.raises(ValueError):
        frame_range(2, 2, 10)
    with pytest.raises(ValueError):
        frame_range(-1, 2, 10)


def _free_port() -> int:
    with socket.socket() as sock:
        sock.bind(("127.0.0.1", 0))
        return sock.getsockname()[1]


def _frame(index: int) -> np.ndarray:
    """A 4x5x3 uint8 frame that encodes its global frame index in every pixel."""
    return np.full((4, 5, 3), index % 251, dtype=np.uint8) + np.arange(3, dtype=np.uint8)


def _worker(rank: int, world: int, port: int, total_frames: int, results) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ranges = all_frame_ranges(world, total_frames)
        start, stop = ranges[rank]
        local = torch.from_numpy(np.stack([_frame(i) for i in range(start, stop)])) if stop > start else \
            torch.zeros((0, 4, 5, 3), dtype=torch.uint8)
        counts = [b - a for a, b in ranges]
        gathered = gather_frames(local, counts, dst=0)
        if len(set(counts)) == 1:  # equal shares: chunks are received straight into a caller-provided buffer
            out = torch.zeros((total_frames, 4, 5, 3), dtype=torch.uint8) if rank == 0 else None
            again = gather_frames(local, counts, dst=0, out=out)
            if rank == 0:
                results.put(("out", again is out and bool(torch.equal(out, gathered)), ()))
                try:
                    gather_frames(local, counts, dst=0, out=torch.zeros((1, 4, 5, 3), dtype=torch.uint8))
                    results.put(("out-shape", False, ()))
                except ValueError:
                    results.put(("out-shape", True, ()))
            else:
                results.put(("out", again is None, ()))
                results.put(("out-shape", True, ()))
        if rank == 0:
            want = np.stack([_frame(i) for i in range(total_frames)])
            results.put(("ok", bool(np.array_equal(gathered.numpy(), want)), tuple(gathered.shape)))
        else:
            results.put(("none", gathered is None, ()))
        # a wrong `counts` is an argument error on every rank, before any communication
        try:
            gather_frames(local, [1] * (world + 1), dst=0)
            results.put(("raise", False, ()))
        except ValueError:
            results.put(("raise", True, ()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,total_frames", [(2, 8), (2, 7), (3, 2)])
def test_gather_frames_keeps_global_frame_order_gloo(world, total_frames):
    """Ragged (7 over 2) and empty (2 over 3) shares included; rank 0 gets every frame in order."""
    ctx = mp.get_context("spawn")
    results = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(rank, world, port, total_frames, results)) for rank in range(world)]
    for proc in procs:
        proc.start()
    per_rank = 4 if total_frames % world == 0 else 2
    got = [results.get(timeout=120) for _ in range(per_rank * world)]
    for proc in procs:
        proc.join(timeout=60)
        assert proc.exitcode == 0
    assert all(flag for _, flag, _ in got), got
    shapes = [shape for kind, _, shape in got if kind == "ok"]
    assert shapes == [(total_frames, 4, 5, 3)]


def test_gather_is_identity_without_a_process_group():
    frames = torch.zeros((3, 2, 2, 3), dtype=torch.uint8)
    assert gather_frames(frames, [3]) is frames


def test_sharded_schedule_equals_the_unsharded_one():
    """Every rank schedules only its range; concatenated in rank order that is the whole track, once."""
    rng = np.random.RandomState(0)
    indices = rng.randint(0, 3, size=101)
    seen = []
    for start, stop in all_frame_ranges(4, len(indices)):
        for network_index, frames in frame_schedule(indices[start:stop], stop - start, 32, True):
            assert np.all(indices[start:stop][frames] == network_index) and len(frames) <= 32
            seen.extend((frames + start).tolist())
    assert sorted(seen) == list(range(len(indices)))


# ---- the indexed frame buffer: batches grouped by network land at their frame index ----------------------------------
from gance_b200 import track_renderer  # noqa: E402
from gance_b200.frame_sharding import SharedFrameSink, open_shared_frame_sink  # noqa: E402


class _FakeNetworks:
    """Duck-typed MultiNetwork without a GPU: frame = (network index, first latent element) in every pixel."""

    def __init__(self, side: int = 2) -> None:
        self.side = side
        self.calls = []

    def indexed_render_into(self, index, latents, dest, dest_indices, max_batch=32, output_side_length=None):
        latents = np.asarray(latents)
        assert len(latents) == len(dest_indices) <= max_batch
        self.calls.append((int(index), len(latents)))
        for item, position in zip(latents, dest_indices):
            dest[int(position)] = torch.tensor([index, int(item.reshape(-1)[0]) % 251, 7], dtype=torch.uint8)
        return None


@pytest.mark.parametrize("group", [True, False])
def test_network_batches_cover_every_frame_once(group):
    rng = np.random.RandomState(1)
    indices = rng.randint(0, 3, size=200)
    frames = np.arange(37, 170)
    batches = track_renderer.network_batches(indices, frames, 32, group)
    seen = np.concatenate([members for _, members in batches])
    assert sorted(seen.tolist()) == frames.tolist()
    for network, members in batches:
        assert 1 <= len(members) <= 32 and np.all(indices[members] == network) and np.all(np.diff(members) > 0)
    if group:  # a network's frames fill whole batches: at most one ragged batch per network
        assert sum(1 for _, members in batches if len(members) < 32) <= 3
        assert [network for network, _ in batches] == sorted(network for network, _ in batches)
    assert track_renderer.network_batches(indices, np.arange(0), 32, group) == []
    with pytest.raises(ValueError):
        track_renderer.network_batches(indices, frames, 0, group)


def test_render_frames_into_puts_each_frame_at_its_index():
    rng = np.random.RandomState(2)
    total = 150
    indices = rng.randint(0, 3, size=total)
    latents = np.arange(total, dtype=np.float64)[:, None] * np.ones((1, 512))
    networks = _FakeNetworks()
    dest = torch.zeros((total, 2, 2, 3), dtype=torch.uint8)
    batches = track_renderer.render_frames_into(networks, latents, indices, dest, batch_size=32)
    assert batches == len(networks.calls) and sum(count for _, count in networks.calls) == total
    for frame in range(total):
        assert dest[frame, 0, 0].tolist() == [indices[frame], frame % 251, 7]
    # a rank's contiguous share, written at an offset (the rank's slice of the shared sink)
    start, stop = frame_range(1, 4, total)
    share = torch.zeros((stop - start, 2, 2, 3), dtype=torch.uint8)
    track_renderer.render_frames_into(networks, latents, indices, share, frames=np.arange(start, stop), dest_offset=-start)
    assert torch.equal(share, dest[start:stop])


def _sink_worker(rank: int, world: int, port: int, total_frames: int, name: str, directory: str, results) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sink = open_shared_frame_sink(name, total_frames, 2, register=False, directory=directory)
        start, stop = sink.frame_range
        indices = np.arange(total_frames) % 2
        latents = np.arange(total_frames, dtype=np.float64)[:, None] * np.ones((1, 512))
        # every rank renders ITS frames, grouped by network, straight to their global positions
        track_renderer.render_frames_into(_FakeNetworks(), latents, indices, sink.frames, frames=np.arange(start, stop), batch_size=3)
        dist.barrier()
        if rank == 0:  # the encoder rank sees every frame, in order, in its own mapping
            ok = all(sink.frames[f, 1, 1].tolist() == [f % 2, f % 251, 7] for f in range(total_frames))
            results.put(("sink", ok, tuple(sink.frames.shape)))
        dist.barrier()
        path = sink.path
        sink.close()
        dist.barrier()
        results.put(("closed", not os.path.exists(path), ()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,total_frames", [(2, 11), (3, 4)])
def test_shared_frame_sink_collects_every_ranks_frames_in_order_gloo(tmp_path, world, total_frames):
    ctx = mp.get_context("spawn")
    results = ctx.Queue()
    port = _free_port()
    name = f"gance_b200_test_sink_{os.getpid()}_{world}"
    procs = [ctx.Process(target=_sink_worker, args=(rank, world, port, total_frames, name, str(tmp_path), results))
             for rank in range(world)]
    for proc in procs:
        proc.start()
    got = [results.get(timeout=120) for _ in range(world + 1)]
    for proc in procs:
        proc.join(timeout=60)
        assert proc.exitcode == 0
    assert all(flag for _, flag, _ in got), got
    assert [shape for kind, _, shape in got if kind == "sink"] == [(total_frames, 2, 2, 3)]


def test_shared_frame_sink_single_process(tmp_path):
    with open_shared_frame_sink("solo", 5, 3, register=False, directory=str(tmp_path)) as sink:
        assert sink.frame_range == (0, 5) and tuple(sink.frames.shape) == (5, 3, 3, 3)
        sink.local_frames()[2] = 9
        other = SharedFrameSink("solo", 5, 3, (0, 0), create=False, register=False, directory=str(tmp_path))
        assert int(other.frames[2].sum()) == 9 * 27  # a second mapping of the same file sees the write
        other.close()
        with pytest.raises(ValueError):
            SharedFrameSink("solo", 6, 3, (0, 0), create=False, register=False, directory=str(tmp_path))
    assert not (tmp_path / "solo").exists()


def _projection_worker(rank: int, world: int, port: int, path: str, results) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gance_b200.projection import projection_file_reader

        with projection_file_reader.load_projection_file(path) as reader:
            label = projection_file_reader.final_latents_matrices_label_sharded(reader)
        results[rank] = (label.data.copy(), label.vector_length, label.label)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,frames", [(2, 7), (3, 2), (2, 1)])
def test_projection_file_decoding_shared_over_ranks_equals_the_whole_read_gloo(tmp_path, world, frames):
    """Every rank inflates its share of the projection file's frames; one all-gather gives each the array a whole read gives
    (ragged and empty shares included)."""
    from pathlib import Path

    from gance_b200.projection import projection_file_reader, projector_file_writer

    path = Path(tmp_path) / "projection.hdf5"
    latents = [np.random.RandomState(i).randn(18, 512).astype(np.float32) for i in range(frames)]
    projector_file_writer.write_projection_file(path, latents, projector_file_writer.synthetic_attributes(frames, projection_fps=30.0))
    whole = projection_file_reader.load_final_latents_matrices_label(path)
    with projection_file_reader.load_projection_file(path) as reader:
        assert reader.final_latents_count() == frames and reader.final_latents_shape() == (18, 512)
        assert np.array_equal(np.concatenate(list(reader.final_latents_slice(1, 3)), axis=-1) if frames > 1 else whole.data[:, :0],
                              whole.data[:, 512: 3 * 512])
        single = projection_file_reader.final_latents_matrices_label_sharded(reader)  # no process group: the plain read
    assert np.array_equal(single.data, whole.data)
    manager = mp.Manager()
    results = manager.dict()
    mp.spawn(_projection_worker, args=(world, _free_port(), str(path), results), nprocs=world, join=True)
    for rank in range(world):
        data, vector_length, label = results[rank]
        assert np.array_equal(data, whole.data) and data.dtype == whole.data.dtype
        assert vector_length == whole.vector_length and label == whole.label
