"""
Where do the single-CTA halo kernels spend their time? Runs one eager batch of the 1024^2 generator through the
INSTRUMENTED library (GANCE_PIPE_STATS=1 build: cycle counters in the TMA producer, the MMA issuer and the epilogue)
and prints, per halo launch, the mean over CTAs of: MMA-warp loop cycles, cycles it waited for a drained accumulator,
cycles it waited for TMA data, cycles the producer waited for a free stage, cycles the epilogue groups waited for a
finished accumulator. Run under gpurun:  GANCE_PIPE_STATS=1 python gance_b200/csrc/build.py && python tools/pipe_stats.py
"""
import ctypes
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from gance_b200 import _lib  # noqa: E402

_lib.LIBRARY_PATH = ROOT / "gance_b200" / "csrc" / "libgance_b200_stats.so"

from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables  # noqa: E402
from gance_b200.synthesis import GeneratorOnDevice  # noqa: E402


def main() -> None:
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    lib = _lib.load_library()
    lib.gance_set_conv_pairing(0)
    spec = GeneratorSpec(1024)
    gen = GeneratorOnDevice(spec, random_init_variables(spec, seed=0, perturbed=True), device, mode="fp16")
    z = torch.as_tensor(np.random.RandomState(0).randn(batch, 512), device=device)
    gen.run_vectors(z, noise_mode="random")  # warm-up (allocations)
    torch.cuda.synchronize()
    slots = 16
    stats = torch.zeros((slots, 256, 8), dtype=torch.int64, device=device)
    lib.gance_debug_pipe_stats.argtypes = [ctypes.c_void_p]
    lib.gance_debug_pipe_stats(stats.data_ptr())
    gen.run_vectors(z, noise_mode="random")
    torch.cuda.synchronize()
    lib.gance_debug_pipe_stats(None)
    host = stats.cpu().numpy().astype(np.float64)
    names = ["mma_loop", "mma_wait_acc", "mma_wait_data", "-", "tma_wait_stage", "epi0_wait_acc", "epi1_wait_acc", "epi0_loop"]
    for slot in range(slots):
        rows = host[slot]
        used = rows[rows[:, 0] > 0]
        if len(used) == 0:
            continue
        mean = used.mean(axis=0)
        print(f"halo launch {slot}: {len(used)} CTAs; " + ", ".join(f"{n} {v / 1e3:.0f}k" for n, v in zip(names, mean) if n != "-"))
        print(f"    fractions of the MMA loop: wait_acc {mean[1] / mean[0]:.2f}, wait_data {mean[2] / mean[0]:.2f}, "
              f"issue {1 - (mean[1] + mean[2]) / mean[0]:.2f}; epilogue group 0 waits {mean[5] / mean[7]:.2f} of its loop")


if __name__ == "__main__":
    main()
