"""
BASELINE configs 3 / 4 / 5 as stand-alone jobs (the same code bench.py runs as its `album` / `projection` / `switching`
legs - tools/track_jobs.py has the description):

    python tools/album_run.py --job album                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        tools/album_run.py --job album --seconds 180

Prints one JSON report on rank 0: wall time from the input files on disk to every frame in the encoder rank's host
buffer, the stage breakdown (inputs / latents / synthesis), device -> host GB/s per rank and the frame-order check.
"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import track_jobs  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--job", default="album", choices=sorted(track_jobs.JOBS))
    ap.add_argument("--seconds", type=float, default=180.0)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--mode", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--output-side", type=int, default=None)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=device)
    report = track_jobs.run_track_job(args.job, args.seconds, device, mode=args.mode, batch=args.batch,
                                      output_side=args.output_side)
    if report is not None:
        print(json.dumps(report), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
