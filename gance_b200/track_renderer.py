"""
Rendering a whole frame range into an INDEXED frame buffer: the scheduling half of `create_network_frames`
(/root/reference/gance/data_into_network_visualization/network_visualization.py:598-681), shared by `vector_synthesis`,
the album-length job (tools/album_run.py) and bench.py.

The reference renders one frame per call and, because switching networks costs it seconds, sorts the frames by network
index, renders them in that order, spills each to a gzip-HDF5 temp file and reads them back in frame order (631-674).
Here every network is resident, a batch renders many frames, and the "re-sort" is an indexed store: frames are grouped by
network (stable, the reference's sort), cut into batches, and each batch's frames are copied device -> host straight to
their FRAME INDEX in `dest` (`BatchRenderer.submit_into`) - a pinned window buffer, or the encoder rank's shared sink
(`frame_sharding.SharedFrameSink`). Frame <-> latent assignment and the order in which a consumer sees frames are the
reference's; nothing touches the disk, and the copies of batch i overlap the rendering of batch i+1.
"""

from collections import deque
from typing import Deque, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

Latents = Union[np.ndarray, torch.Tensor]


def network_batches(network_indices: np.ndarray, frames: np.ndarray, batch_size: int, group_by_network: bool
                    ) -> List[Tuple[int, np.ndarray]]:
    """
    (network index, frame indices) batches covering `frames` exactly once, none longer than `batch_size`. With
    `group_by_network` the frames are stably grouped by network index first (the reference's `sorted(..., key=network_index)`,
    network_visualization.py:659-663), so a network's frames fill whole batches; otherwise runs of equal index in frame
    order.
    """
    if batch_size < 1:
        raise ValueError("batch_size must be positive")
    frames = np.asarray(frames, dtype=np.int64)
    if len(frames) == 0:
        return []
    indices = np.asarray(network_indices)[frames].astype(np.int64)
    batches: List[Tuple[int, np.ndarray]] = []
    if group_by_network:
        for network_index in np.unique(indices):
            members = frames[indices == network_index]
            batches += [(int(network_index), members[i: i + batch_size]) for i in range(0, len(members), batch_size)]
        return batches
    run_start = 0
    for position in range(1, len(frames) + 1):
        if position == len(frames) or indices[position] != indices[run_start]:
            run = frames[run_start:position]
            batches += [(int(indices[run_start]), run[i: i + batch_size]) for i in range(0, len(run), batch_size)]
            run_start = position
    return batches


def _take(latents: Latents, frames: np.ndarray) -> Latents:
    if isinstance(latents, torch.Tensor):
        if len(frames) and int(frames[-1]) - int(frames[0]) + 1 == len(frames):
            return latents[int(frames[0]): int(frames[-1]) + 1]
        return latents[torch.as_tensor(frames, device=latents.device)]
    return latents[frames]


def submit_frames(networks, latents: Latents, network_indices: np.ndarray, dest: torch.Tensor, frames: Sequence[int],
                  dest_offset: int = 0, batch_size: int = 32, output_side_length: Optional[int] = None,
                  group_by_network: bool = True, in_flight: Optional[Deque] = None, max_in_flight: int = 8) -> List:
    """
    Queue the rendering of `frames` (indices into `latents` / `network_indices`): frame f is delivered to
    `dest[f + dest_offset]`. Returns the arrivals (one `synthesis.Arrival` per batch; None entries for interfaces without a CUDA
    renderer, whose frames are written synchronously). At most `max_in_flight` batches are left unsynchronised, so the
    host never runs further ahead of the device than that.
    """
    events = []
    waiting: Deque = in_flight if in_flight is not None else deque()
    for network_index, members in network_batches(network_indices, np.asarray(frames), batch_size, group_by_network):
        event = networks.indexed_render_into(network_index, _take(latents, members), dest, members + dest_offset,
                                             max_batch=batch_size, output_side_length=output_side_length)
        events.append(event)
        if event is not None:
            waiting.append(event)
            while len(waiting) > max_in_flight:
                waiting.popleft().synchronize()
    return events


def wait_for(events: Sequence) -> None:
    """Block until every batch has arrived; an arrival whose range-guard flag is set raises GANCE_E_NONFINITE."""
    for event in events:
        if event is not None:
            event.synchronize()


def render_frames_into(networks, latents: Latents, network_indices: np.ndarray, dest: torch.Tensor,
                       frames: Optional[Sequence[int]] = None, dest_offset: int = 0, batch_size: int = 32,
                       output_side_length: Optional[int] = None, group_by_network: bool = True) -> int:
    """
    Render `frames` (default: every latent) and return once they are all in `dest`: the whole-range form used by the
    album job and the bench, where `dest` is the rank's slice of the encoder's frame sink. Returns the number of batches.
    """
    if frames is None:
        frames = np.arange(min(len(latents), len(network_indices)))
    events = submit_frames(networks, latents, network_indices, dest, frames, dest_offset, batch_size, output_side_length,
                           group_by_network)
    wait_for(events)
    return len(events)


_WINDOW_RINGS = {}


def _window_ring(frames_per_window: int, side: int, slots: int) -> List[torch.Tensor]:
    """Pinned [frames_per_window, side, side, 3] window buffers, kept for the life of the process (pinning is slow)."""
    key = (frames_per_window, side, slots)
    ring = _WINDOW_RINGS.get(key)
    if ring is None:
        _WINDOW_RINGS.clear()  # one geometry at a time: a 2160^2 ring is gigabytes of pinned memory
        ring = [torch.empty((frames_per_window, side, side, 3), dtype=torch.uint8).pin_memory() for _ in range(slots)]
        _WINDOW_RINGS[key] = ring
    return ring


def release_window_buffers() -> None:
    _WINDOW_RINGS.clear()


def windowed_frames(networks, latents: Latents, network_indices: np.ndarray, side: int, batch_size: int = 32,
                    output_side_length: Optional[int] = None, group_by_network: bool = True, window_batches: int = 2,
                    reuse_frame_buffers: bool = False) -> Iterator[np.ndarray]:
    """
    Frames 0..T-1 in frame order, lazily, `window_batches * batch_size` frames at a time: window w is rendered (grouped
    by network) into one of three pinned window buffers while the consumer is still drawing window w-1, so rendering,
    device -> host copies and consumption overlap. Yields (side,side,3) uint8 arrays: copies by default (the reference's
    iterator yields independent arrays; `list(frames)` must stay safe), views of the window buffer with
    `reuse_frame_buffers` (valid until two windows later - for consumers that are done with a frame before they draw the
    next, as a video encoder is).
    """
    total = min(len(latents), len(network_indices))
    per_window = max(1, window_batches) * batch_size
    ring = _window_ring(per_window, side, 3)
    pending: Deque[Tuple[torch.Tensor, int, List]] = deque()
    in_flight: Deque = deque()

    def drain(entry) -> Iterator[np.ndarray]:
        buffer, count, events = entry
        wait_for(events)
        host = buffer.numpy()
        for position in range(count):
            yield host[position] if reuse_frame_buffers else host[position].copy()

    for window, start in enumerate(range(0, total, per_window)):
        stop = min(start + per_window, total)
        buffer = ring[window % len(ring)]
        events = submit_frames(networks, latents, network_indices, buffer, np.arange(start, stop), -start, batch_size,
                               output_side_length, group_by_network, in_flight=in_flight)
        pending.append((buffer, stop - start, events))
        if len(pending) == 2:
            yield from drain(pending.popleft())
    while pending:
        yield from drain(pending.popleft())
