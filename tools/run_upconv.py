"""One up-convolution launch on random data (for ncu captures and A/B timing of the fused / phase / folded forms)."""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gance_b200 import _lib  # noqa: E402
from gance_b200.synthesis import fold_up_weight, phase_up_weight  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--form", default="fused", choices=["fused", "phase", "folded"])
    ap.add_argument("--cin", type=int, default=128)
    ap.add_argument("--cout", type=int, default=64)
    ap.add_argument("--side", type=int, default=256)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _lib.init(0)
    dt, code = torch.float16, _lib.GANCE_F16
    b, s, cin, cout = args.batch, args.side, args.cin, args.cout
    x = torch.randn(b, s, s, cin, device="cuda").to(dt)
    hwio = torch.randn(3, 3, cin, cout, device="cuda") / (9 * cin) ** 0.5
    demod = torch.rand(b, cout, device="cuda") + 0.5
    noise = torch.randn(b, 2 * s, 2 * s, device="cuda")
    bias = torch.randn(cout, device="cuda") * 0.1
    nxt = torch.rand(b, cout, device="cuda") + 0.5
    out = torch.empty(b, 2 * s, 2 * s, cout, device="cuda", dtype=dt)
    stream = torch.cuda.current_stream().cuda_stream
    if args.form == "folded":
        w = fold_up_weight(hwio).to(dt)
        call = lambda: lib.gance_modconv3x3_up2_folded(x.data_ptr(), code, b, s, s, cin, cout, w.data_ptr(), demod.data_ptr(),
                                                       noise.data_ptr(), 4 * s * s, 0.05, bias.data_ptr(), nxt.data_ptr(),
                                                       out.data_ptr(), stream)
    elif args.form == "phase":
        w = phase_up_weight(hwio, 64).to(dt)
        scratch = torch.empty(b * (2 * s + 2) * (2 * s + 2) * cout, device="cuda", dtype=dt)
        call = lambda: lib.gance_modconv3x3_up2_phase(x.data_ptr(), code, b, s, s, cin, cout, w.data_ptr(), demod.data_ptr(),
                                                      noise.data_ptr(), 4 * s * s, 0.05, bias.data_ptr(), nxt.data_ptr(),
                                                      scratch.data_ptr(), out.data_ptr(), stream)
    else:
        w = phase_up_weight(hwio, min(64, cout)).to(dt)
        call = lambda: lib.gance_modconv3x3_up2_fused(x.data_ptr(), code, b, s, s, cin, cout, w.data_ptr(), demod.data_ptr(),
                                                      noise.data_ptr(), 4 * s * s, 0.05, bias.data_ptr(), nxt.data_ptr(),
                                                      out.data_ptr(), stream)
    _lib.check(call(), args.form)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(args.iters):
        _lib.check(call(), args.form)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / args.iters
    flops = 2 * 9 * cin * cout * s * s * b
    traffic = (s * s * cin + 4 * s * s * cout) * 2 * b
    print(f"{args.form} Cin={cin} Cout={cout} {s}->{2 * s} B={b}: {ms * 1e3:.1f} us, {flops / ms / 1e9:.1f} TFLOP/s (algorithmic), "
          f"{traffic / ms / 1e6:.1f} GB/s (algorithmic), checksum {float(out.float().abs().mean()):.4f}")


if __name__ == "__main__":
    main()
