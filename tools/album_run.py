"""
BASELINE config 4: an album-length track (default 180 s at 60 fps = 10 800 frames) rendered at 1024^2, frame-sharded over
the ranks of one box with the NCCL gather of uint8 frames to the encoder rank (rank 0), which drains every gathered
batch to pinned host memory (the hand-over point to a video encoder). Wall clock covers everything after process
start-up (buffers, one warm-up batch and gather): synthetic WAV -> latents on the audio kernels (every rank,
deterministic), synthesis, gather, D2H.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/album_run.py

Frame order is verified, not assumed: noise is the network's constant buffers, so a frame is a pure function of its
latent; rank 0 re-renders a sample of global frame indices from its own copy of the latent sequence and compares them
bit for bit with what arrived through the gather.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gance_b200.frame_sharding import all_frame_ranges, gather_frames  # noqa: E402
from gance_b200.network_interface import network_functions  # noqa: E402
from gance_b200.networks import container  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=180.0)
    ap.add_argument("--resolution", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=32)
    args = ap.parse_args()
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=device)
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    import bench  # track_latents: the bench's WAV -> latents helper (our CUDA audio path)

    path = Path(f"/tmp/gance_b200_album_{args.resolution}_{rank}.pkl")
    container.write_synthetic_network(path, args.resolution, seed=0, perturbed=True)
    generator = network_functions.resident_generator(path, "fp16")
    res, batch = args.resolution, args.batch
    # ---- process start-up (untimed): buffers, lazy imports / CUDA modules of the audio path, NCCL channel set-up ----
    slots = [torch.zeros((batch, res, res, 3), dtype=torch.uint8, device=device) for _ in range(2)]
    gathered = [torch.empty((world * batch, res, res, 3), dtype=torch.uint8, device=device) if rank == 0 else None for _ in range(2)]
    pinned = [torch.empty((world * batch, res, res, 3), dtype=torch.uint8).pin_memory() if rank == 0 else None for _ in range(2)]
    warm = bench.track_latents(2.0, device)
    generator.run_vectors(warm[:batch], truncation_psi=1.2, noise_mode="const", out=slots[0])
    if world > 1:
        gather_frames(slots[0], [batch] * world, dst=0, out=gathered[0])
        dist.barrier()
    torch.cuda.synchronize()
    # ---- the job: WAV -> latents -> frames -> encoder rank's host memory ----
    t_start = time.perf_counter()
    latents_all = bench.track_latents(args.seconds, device)  # [T,512], identical on every rank
    torch.cuda.synchronize()
    t_latents = time.perf_counter()
    total = latents_all.shape[0]
    ranges = all_frame_ranges(world, total)
    start, stop = ranges[rank]
    steps = max((b - a + batch - 1) // batch for a, b in ranges)
    comm = torch.cuda.Stream(device=device)
    slot_free = [None, None]
    host_ready = [None, None]
    kept = {}  # global frame index -> host copy, for the order check
    probe = sorted({0, total // 3, total // 2, total - 1, min(total - 1, ranges[-1][0] + 5)})
    checksum = 0

    def drain(step: int) -> None:
        nonlocal checksum
        slot = step % 2
        host_ready[slot].synchronize()
        host = pinned[slot].numpy()
        for r, (a, b) in enumerate(ranges):
            first = a + step * batch
            count = max(0, min(batch, b - first))
            chunk = host[r * batch: r * batch + count]
            if count:
                checksum += int(chunk[:, 0, 0, 0].astype(np.int64).sum())
            for g in probe:
                if first <= g < first + count:
                    kept[g] = chunk[g - first].copy()

    main_stream = torch.cuda.current_stream(device)
    for step in range(steps):
        slot = step % 2
        if slot_free[slot] is not None:
            main_stream.wait_event(slot_free[slot])
        first = start + step * batch
        count = max(0, min(batch, stop - first))
        if count:
            generator.run_vectors(latents_all[first: first + count], truncation_psi=1.2, noise_mode="const", out=slots[slot][:count])
        if world > 1:
            computed = torch.cuda.Event()
            computed.record(main_stream)
            comm.wait_event(computed)
            with torch.cuda.stream(comm):
                if rank == 0 and step >= 2:
                    drain(step - 2)  # the pinned slot is about to be overwritten
                gather_frames(slots[slot], [batch] * world, dst=0, out=gathered[slot])
                if rank == 0:
                    pinned[slot].copy_(gathered[slot], non_blocking=True)
                    host_ready[slot] = torch.cuda.Event()
                    host_ready[slot].record(comm)
                slot_free[slot] = torch.cuda.Event()
                slot_free[slot].record(comm)
        else:
            with torch.cuda.stream(comm):
                comm.wait_stream(main_stream)
                if step >= 2:
                    drain(step - 2)
                pinned[slot].copy_(slots[slot], non_blocking=True)
                host_ready[slot] = torch.cuda.Event()
                host_ready[slot].record(comm)
                slot_free[slot] = host_ready[slot]
    if rank == 0:
        for step in range(max(0, steps - 2), steps):
            drain(step)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_end = time.perf_counter()
    if rank == 0:
        order_ok = True
        for g, frame in kept.items():
            again = generator.run_vectors(latents_all[g: g + 1], truncation_psi=1.2, noise_mode="const")
            order_ok = order_ok and bool(np.array_equal(again[0].cpu().numpy(), frame))
        wall = t_end - t_start
        print(json.dumps({
            "workload": f"album-length track: {args.seconds:.0f} s at 60 fps = {total} frames at {res}^2, config-f random-init, "
                        f"frame-sharded over {world} GPU(s), NCCL gather to rank 0, D2H to pinned host memory",
            "n_gpus": world, "frames": int(total), "wall_s": wall, "frames_per_s": total / wall,
            "realtime_multiple_at_60fps": total / wall / 60.0, "latents_s": t_latents - t_start,
            "frames_checked_for_order": sorted(kept), "frame_order_ok": order_ok, "checksum": checksum}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
