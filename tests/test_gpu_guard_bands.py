"""
`-m gpu` guard-band tests: every device buffer the renderer's kernels WRITE is placed between two sentinel bands, the
whole path runs, and the bands must come back untouched. This is the out-of-bounds-store check of the suite (the
kernels use tiles that overhang ragged edges: 28 x 12-pixel fused up-convolution tiles on 32..512-pixel planes, 8 x 16
halo-kernel tiles, 16 x 128 resize tiles), written as a test because compute-sanitizer is not available on the GPU
pool. Each guarded run is also compared bit for bit with an unguarded run of the same inputs, so a kernel that
depended on the placement of its buffers would show up as well.
"""

import math

import numpy as np
import pytest
import torch

from gance_b200.image_sources.video_common import FrameResizer
from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
from gance_b200.synthesis import GeneratorOnDevice

pytestmark = pytest.mark.gpu

GUARD_BYTES = 8192  # each side; a multiple of 512 so the guarded view keeps the allocator's alignment
SENTINEL = 0xA5


def _device():
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device; there is no CPU fallback to test")
    torch.cuda.set_device(0)
    return torch.device("cuda", 0)


class GuardedAllocations:
    """Hands out tensors that sit between two sentinel bands and checks the bands afterwards."""

    def __init__(self, device):
        self.device = device
        self.raw = []  # (name, raw uint8 tensor, payload bytes)

    def empty(self, name, shape, dtype):
        payload = math.prod(shape) * torch.empty((), dtype=dtype).element_size()
        raw = torch.full((GUARD_BYTES + payload + GUARD_BYTES,), SENTINEL, dtype=torch.uint8, device=self.device)
        self.raw.append((name, raw, payload))
        return raw[GUARD_BYTES:GUARD_BYTES + payload].view(dtype).view(tuple(shape))

    def check(self):
        torch.cuda.synchronize()
        assert self.raw, "nothing was allocated through the guarded allocator"
        for name, raw, payload in self.raw:
            before = raw[:GUARD_BYTES]
            after = raw[GUARD_BYTES + payload:]
            assert bool((before == SENTINEL).all()), f"{name}: bytes BEFORE the buffer were overwritten"
            assert bool((after == SENTINEL).all()), f"{name}: bytes AFTER the buffer were overwritten"


@pytest.mark.parametrize("mode,resolution,batch", [("fp16", 256, 3), ("bf16", 128, 5), ("fp32", 64, 2), ("fp16", 512, 1),
                                                   ("fp16", 1024, 2)])
@pytest.mark.parametrize("noise_mode", ["const", "random"])
def test_renderer_kernels_stay_inside_their_buffers(mode, resolution, batch, noise_mode):
    device = _device()
    spec = GeneratorSpec(resolution)
    variables = random_init_variables(spec, seed=5, perturbed=True)
    z = torch.as_tensor(np.random.RandomState(3).randn(batch, 512))

    plain = GeneratorOnDevice(spec, variables, device, mode=mode)
    wplus = plain.mapping(z, truncation_psi=1.2)
    want = plain.synthesis(wplus, noise_mode="const").cpu() if noise_mode == "const" else None

    guarded = GeneratorOnDevice(spec, variables, device, mode=mode)
    guards = GuardedAllocations(device)
    cache = {}

    def guarded_buffer(name, shape, dtype):
        key = (name, tuple(shape), dtype)
        if key not in cache:
            cache[key] = guards.empty(name, shape, dtype)
        return cache[key]

    guarded._buffer = guarded_buffer  # every intermediate: styles, activations, skip images, noise, scratch
    out = guards.empty("frames", (batch, resolution, resolution, 3), torch.uint8)
    got = guarded.synthesis(wplus, noise_mode=noise_mode, out=out)
    guards.check()
    assert got.data_ptr() == out.data_ptr()
    assert len(guards.raw) > 2 * int(math.log2(resolution))  # the activations of every block went through the guards
    if want is not None:
        assert torch.equal(got.cpu(), want), "frames depend on where the work buffers sit"
    else:
        assert int(got.max()) > int(got.min())  # fresh noise: just a rendered image


@pytest.mark.parametrize("src_hw,dst_hw,batch", [((100, 70), (37, 53), 3), ((64, 64), (135, 135), 2),
                                                 ((33, 129), (200, 17), 1), ((256, 256), (540, 540), 2)])
def test_resize_kernel_stays_inside_its_output(src_hw, dst_hw, batch):
    device = _device()
    rng = np.random.RandomState(11)
    frames = torch.as_tensor(rng.randint(0, 256, size=(batch, *src_hw, 3), dtype=np.uint8)).to(device)
    resizer = FrameResizer(device)
    want = resizer.resize(frames, *dst_hw).cpu()
    guards = GuardedAllocations(device)
    out = guards.empty("resized", (batch, *dst_hw, 3), torch.uint8)
    got = resizer.resize(frames, *dst_hw, out=out)
    guards.check()
    assert torch.equal(got.cpu(), want)
