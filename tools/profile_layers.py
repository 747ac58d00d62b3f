"""Per-layer device times of one synthesis batch (CUDA events), next to each layer's roofline."""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gance_b200 import _lib  # noqa: E402
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables  # noqa: E402
from gance_b200.synthesis import GeneratorOnDevice  # noqa: E402

HBM_GBS, TENSOR_TFLOPS = 6531.9, 1614.0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--resolution", type=int, default=1024)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--mode", default="bf16")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--perturbed", action="store_true", help="non-zero noise strengths / biases (what bench.py runs)")
    ap.add_argument("--noise", default="const", choices=["const", "random", "none"])
    ap.add_argument("--library", default=None, help="an alternative build of the library (measurement variants of csrc/build.py)")
    ap.add_argument("--pairing", type=int, default=None, choices=[0, 1, 2], help="gance_set_conv_pairing")
    ap.add_argument("--two-rows", type=int, default=None, choices=[0, 1], help="gance_set_conv_two_rows")
    ap.add_argument("--fused-min-side", type=int, default=None, help="GeneratorOnDevice.fused_min_in_resolution (A/B)")
    ap.add_argument("--only", default=None, help="print only layers whose name contains this")
    args = ap.parse_args()
    if args.library:
        _lib.LIBRARY_PATH = Path(args.library).resolve()
    if args.pairing is not None:
        _lib.load_library().gance_set_conv_pairing(args.pairing)
    if args.two_rows is not None:
        _lib.load_library().gance_set_conv_two_rows(args.two_rows)
    if args.fused_min_side is not None:
        GeneratorOnDevice.fused_min_in_resolution = args.fused_min_side
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    spec = GeneratorSpec(args.resolution)
    gen = GeneratorOnDevice(spec, random_init_variables(spec, 0, perturbed=args.perturbed), device, mode=args.mode)
    wplus = torch.randn(args.batch, spec.num_ws, 512, device=device)
    for _ in range(2):
        gen.synthesis(wplus, noise_mode=args.noise)
    torch.cuda.synchronize()
    totals = {}
    for _ in range(args.iters):
        gen.profile_events = []
        start = torch.cuda.Event(enable_timing=True)
        start.record()
        gen.synthesis(wplus, noise_mode=args.noise)
        torch.cuda.synchronize()
        prev = start
        for name, event in gen.profile_events:
            totals.setdefault(name, []).append(prev.elapsed_time(event))
            prev = event
    gen.profile_events = None
    bytes_per = 2 if args.mode != "fp32" else 4
    layer_info = {}
    for conv in gen.convs:
        res = conv.out_resolution
        key = f"{res}/conv0_up" if conv.up else (f"{res}/conv1" if res > 4 else "4/conv")
        macs = 9 * conv.in_channels * conv.out_channels * (conv.in_resolution ** 2 if conv.up else res ** 2)
        traffic = (conv.in_resolution ** 2 * conv.in_channels + res ** 2 * conv.out_channels) * bytes_per
        layer_info[key] = (2 * macs, traffic)
    for rgb in gen.rgbs:
        layer_info[f"{rgb.resolution}/torgb"] = (2 * 3 * rgb.in_channels * rgb.resolution ** 2,
                                                 rgb.resolution ** 2 * (rgb.in_channels * bytes_per + 16 + 4))
    total_ms = 0.0
    print(f"{'layer':>16} {'ms':>9} {'us/img':>9} {'TFLOP/s':>9} {'GB/s(alg)':>10} {'roof us/img':>12}")
    for name, samples in totals.items():
        ms = float(np.median(samples))
        total_ms += ms
        if args.only and args.only not in name:
            continue
        flops, traffic = layer_info.get(name, (0, 0))
        tf = flops * args.batch / (ms * 1e-3) / 1e12
        gbs = traffic * args.batch / (ms * 1e-3) / 1e9
        roof = max(flops / (TENSOR_TFLOPS * 1e12), traffic / (HBM_GBS * 1e9)) * 1e6
        print(f"{name:>16} {ms:9.3f} {ms * 1e3 / args.batch:9.1f} {tf:9.1f} {gbs:10.1f} {roof:12.2f}")
    print(f"total {total_ms:.3f} ms per batch of {args.batch} -> {args.batch / total_ms * 1e3:.1f} frames/s "
          f"({args.mode}, {args.resolution}^2)")


if __name__ == "__main__":
    main()
