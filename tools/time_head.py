"""Device time of G_mapping and of one graphed vector batch (mapping, styles, demodulation, synthesis), for A/B
runs of two builds of the library (`--library`). CUDA events around CUDA-graph replays; prints the median of `--iters`."""
import argparse
import statistics
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gance_b200 import _lib  # noqa: E402
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables  # noqa: E402
from gance_b200.synthesis import GeneratorOnDevice  # noqa: E402


def timed_graph(fn, iters: int, repeat: int) -> float:
    """Median ms of one graph replay holding `repeat` calls of fn, divided by repeat."""
    fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(repeat):
            fn()
    times = []
    for _ in range(iters):
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        graph.replay()
        stop.record()
        torch.cuda.synchronize()
        times.append(start.elapsed_time(stop) / repeat)
    return statistics.median(times)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--library", default=None)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    if args.library:
        _lib.LIBRARY_PATH = Path(args.library).resolve()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    spec = GeneratorSpec(1024)
    gen = GeneratorOnDevice(spec, random_init_variables(spec, 0, perturbed=True), device, mode="fp16", noise_seed=1)
    z = torch.randn(args.batch, 512, device=device)
    side = torch.cuda.Stream(device)
    with torch.cuda.stream(side):
        mapping_ms = timed_graph(lambda: gen.mapping(z), args.iters, 10)
        out = torch.empty((args.batch, 1024, 1024, 3), dtype=torch.uint8, device=device)
        gen.run_vectors_graphed(z, out)
        torch.cuda.synchronize()
        times = []
        for _ in range(args.iters):
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
            gen.run_vectors_graphed(z, out)
            stop.record()
            torch.cuda.synchronize()
            times.append(start.elapsed_time(stop))
    print(f"{_lib.LIBRARY_PATH.name}: G_mapping {1000 * mapping_ms:.1f} us per batch of {args.batch}; "
          f"graphed vector batch median {statistics.median(times):.3f} ms, min {min(times):.3f} ms")


if __name__ == "__main__":
    main()
