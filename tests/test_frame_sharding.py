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
