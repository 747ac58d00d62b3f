"""
Multi-GPU layout of a synthesis run: frames are independent once the latents exist,
so rank r of R renders the contiguous frame range [r*T/R, (r+1)*T/R) with the weights of every
network replicated; every rank recomputes the (cheap, deterministic) latent sequence and slices its
range, so there is no collective on the data path except ONE gather of uint8 frames to the encoder
rank, in global frame order (rank-major, then local index). The reference has no counterpart: it is
single-process, single-GPU (docker-compose.yml:14).

Works with any torch.distributed backend: NCCL on the GPUs, gloo in the CPU tests.
"""
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def frame_range(rank: int, world_size: int, num_frames: int) -> Tuple[int, int]:
    """[start, stop) of rank's contiguous share; shares differ by at most one frame, earlier ranks first."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside [0, {world_size})")
    base, extra = divmod(num_frames, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def all_frame_ranges(world_size: int, num_frames: int) -> List[Tuple[int, int]]:
    return [frame_range(rank, world_size, num_frames) for rank in range(world_size)]


def gather_frames(
    local_frames: torch.Tensor, counts: List[int], dst: int = 0, group: Optional[dist.ProcessGroup] = None,
    out: Optional[torch.Tensor] = None,
) -> Optional[torch.Tensor]:
    """
    Gather uint8 [n_r,H,W,3] chunks (n_r = counts[r], possibly ragged or zero) to rank `dst`, which
    gets [sum(counts),H,W,3] in global frame order; other ranks get None. One collective:
    `dist.gather` (grouped ncclSend/ncclRecv under NCCL) over chunks padded to max(counts). When every rank sends the
    same number of frames the chunks are received straight into `out` ([sum(counts),H,W,3], allocated if not given):
    no staging copies on the encoder rank.
    """
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local_frames
    world_size, rank = dist.get_world_size(group), dist.get_rank(group)
    if len(counts) != world_size or counts[rank] != local_frames.shape[0]:
        raise ValueError("counts must list every rank's frame count, including this rank's")
    longest = max(counts)
    frame_shape = tuple(local_frames.shape[1:])
    send = local_frames
    if local_frames.shape[0] != longest:
        send = torch.zeros((longest,) + frame_shape, dtype=local_frames.dtype, device=local_frames.device)
        send[: local_frames.shape[0]] = local_frames
    send = send.contiguous()
    if rank == dst and min(counts) == longest:
        total = (longest * world_size,) + frame_shape
        if out is None:
            out = torch.empty(total, dtype=send.dtype, device=send.device)
        elif tuple(out.shape) != total or out.dtype != send.dtype or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous {send.dtype} tensor of shape {total}")
        dist.gather(send, list(out.split(longest, dim=0)), dst=dst, group=group)
        return out
    if rank == dst:
        received = [torch.empty_like(send) for _ in range(world_size)]
        dist.gather(send, received, dst=dst, group=group)
        return torch.cat([chunk[:count] for chunk, count in zip(received, counts)], dim=0)
    dist.gather(send, None, dst=dst, group=group)
    return None


class SharedFrameSink:
    """
    The encoder rank's host frame buffer, written by every rank over its own PCIe link.

    The reference's consumer is ONE process that needs every frame, in frame order, in host memory
    (gance/projection_file_blend.py:280-286, `write_source_to_disk_forward`). Gathering the ranks' frames to rank 0's
    HBM and copying them down from there puts all of them on one PCIe link (48 GB/s = 15 k 1024^2 frames/s - the
    ceiling the round-1 album job ran into at N = 4). Here the buffer is one file under /dev/shm: uint8
    [num_frames, side, side, 3] in global frame order, mapped by every rank; each rank registers ITS OWN frame range with
    CUDA (`cudaHostRegister`), so its device -> host copies are true DMA straight into the frames' final positions, and
    the eight links of an 8-GPU box carry an eighth of the frames each. Nothing is copied twice on the host; rank 0 (or
    any process that maps the file) reads frames [0, T) in order. The NCCL gather to rank 0's HBM remains available
    (`gather_frames`) - that is the path BASELINE.json names for `value`; this sink is the host hand-over.

    Life cycle: every rank constructs it with the same `name` (rank `create_rank` creates and sizes the file; call a
    barrier between creation and the other ranks' construction - `open_shared_frame_sink` does), writes through
    `.frames[a:b].copy_(device_frames, non_blocking=True)` or `BatchRenderer.submit_into`, then `close()`; the creating
    rank also unlinks the file.
    """

    def __init__(self, name: str, num_frames: int, side: int, frame_range: Tuple[int, int], create: bool,
                 register: Optional[bool] = None, directory: str = "/dev/shm") -> None:
        import mmap
        import os

        self.path = os.path.join(directory, name)
        self.num_frames, self.side = int(num_frames), int(side)
        self.frame_bytes = self.side * self.side * 3
        self.created = create
        total = max(1, self.num_frames * self.frame_bytes)
        flags = os.O_RDWR | (os.O_CREAT | os.O_TRUNC if create else 0)
        descriptor = os.open(self.path, flags, 0o600)
        try:
            if create:
                os.ftruncate(descriptor, total)
            elif os.fstat(descriptor).st_size != total:
                raise ValueError(f"{self.path} holds {os.fstat(descriptor).st_size} bytes, expected {total}")
            self._map = mmap.mmap(descriptor, total, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        finally:
            os.close(descriptor)
        import numpy as np

        flat = np.frombuffer(self._map, dtype=np.uint8, count=self.num_frames * self.frame_bytes)
        self.frames = torch.from_numpy(flat).view(self.num_frames, self.side, self.side, 3)
        self.frame_range = (int(frame_range[0]), int(frame_range[1]))
        self._registered: Optional[Tuple[int, int]] = None
        if register is None:
            register = torch.cuda.is_available()
        if register and self.frame_range[1] > self.frame_range[0]:
            page = mmap.PAGESIZE
            base = self.frames.data_ptr()
            first = base + self.frame_range[0] * self.frame_bytes
            last = base + self.frame_range[1] * self.frame_bytes
            first, last = first // page * page, min(-(-last // page) * page, base + -(-total // page) * page)
            status = int(torch.cuda.cudart().cudaHostRegister(first, last - first, 0))
            if status != 0:
                raise RuntimeError(f"cudaHostRegister of {last - first} bytes of {self.path} failed (cudaError {status})")
            self._registered = (first, last - first)

    def local_frames(self) -> torch.Tensor:
        """This rank's own (CUDA-registered) slice [start:stop] of the frame buffer."""
        return self.frames[self.frame_range[0]: self.frame_range[1]]

    def close(self) -> None:
        import os

        if self._registered is not None:
            torch.cuda.cudart().cudaHostUnregister(self._registered[0])
            self._registered = None
        self.frames = None
        if self._map is not None:
            try:
                self._map.close()
            except BufferError:  # a numpy view is still alive somewhere; the mapping goes with the process
                pass
            self._map = None
        if self.created:
            try:
                os.unlink(self.path)
            except FileNotFoundError:
                pass
            self.created = False

    def __enter__(self) -> "SharedFrameSink":
        return self

    def __exit__(self, *exc) -> None:
        self.close()


def open_shared_frame_sink(name: str, num_frames: int, side: int, group: Optional[dist.ProcessGroup] = None,
                           register: Optional[bool] = None, directory: str = "/dev/shm") -> SharedFrameSink:
    """
    Collective: rank 0 creates the sink, everyone else opens it after a barrier; each rank registers the contiguous
    frame range `frame_range(rank, world, num_frames)` gives it. Works without torch.distributed (one process).
    """
    if dist.is_initialized():
        world_size, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world_size, rank = 1, 0
    own = frame_range(rank, world_size, num_frames)
    sink = SharedFrameSink(name, num_frames, side, own, create=True, register=register, directory=directory) if rank == 0 else None
    if world_size > 1:
        dist.barrier(group)
    if sink is None:
        sink = SharedFrameSink(name, num_frames, side, own, create=False, register=register, directory=directory)
    return sink
