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
