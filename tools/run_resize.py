"""Times gance_resize_cubic_u8 on a batch of 1024^2 -> 2160^2 frames (CUDA events) and reports algorithmic HBM GB/s."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from gance_b200.image_sources.video_common import FrameResizer  # noqa: E402


def main() -> None:
    batch, src, dst = 32, 1024, 2160
    if len(sys.argv) > 1:
        batch, src, dst = (int(v) for v in sys.argv[1:4])
    resizer = FrameResizer(torch.device("cuda", 0))
    frames = torch.randint(0, 256, (batch, src, src, 3), dtype=torch.uint8, device="cuda")
    out = torch.empty((batch, dst, dst, 3), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        resizer.resize(frames, dst, dst, out=out)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 10
    start.record()
    for _ in range(iters):
        resizer.resize(frames, dst, dst, out=out)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / iters
    nbytes = batch * 3 * (src * src + dst * dst)
    print(f"resize {src}^2 -> {dst}^2 x{batch}: {ms * 1e3:.1f} us per batch, {ms * 1e3 / batch:.2f} us/frame, "
          f"{nbytes / ms / 1e6:.0f} GB/s algorithmic (read {src}^2*3 + write {dst}^2*3 bytes per frame)")


if __name__ == "__main__":
    main()
