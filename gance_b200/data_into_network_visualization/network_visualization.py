"""
Feeding latents into networks (mirrors the synthesis half of
/root/reference/gance/data_into_network_visualization/network_visualization.py: `vector_synthesis`
(462-690), `SynthesisOutput` (403-409), `compute_force_synthesis_order` (431-446)). The matplotlib
debug plots that module also draws are outside this build's scope:
`visualization_images` is always None.

What changes underneath. The reference renders ONE frame per call and, because switching networks
costs it seconds, first renders every frame sorted by network index, spills each to a gzip-HDF5
temp file, then reads them back in frame order (631-674). Here all networks stay resident in HBM,
so frames are walked in frame order in windows of `batch_size`; inside a window the frames are
grouped by network index (stable, like the reference's sort), each group is rendered as one batch,
and the results are scattered back to their frame slots. Frame order and frame <-> latent
assignment are exactly the reference's; nothing touches the disk.
"""
from typing import Iterator, List, NamedTuple, Optional, Tuple

import numpy as np

from gance_b200.data_into_network_visualization.visualization_common import VisualizationInput
from gance_b200.gance_types import ImageSourceType, RGBInt8ImageType
from gance_b200.logger_common import LOGGER
from gance_b200.network_interface.network_functions import MultiNetwork
from gance_b200.vector_sources.vector_sources_common import sub_vectors

DEFAULT_BATCH_SIZE = 32


class SynthesisOutput(NamedTuple):
    """The two image sources of a synthesis run (network_visualization.py:403-409)."""

    synthesized_images: Optional[ImageSourceType]
    visualization_images: Optional[ImageSourceType]


def compute_force_synthesis_order(force_optimize_synthesis_order: bool, num_network_indices: Optional[int]) -> bool:
    """network_visualization.py:431-446: ordering only matters with more than one network."""
    return force_optimize_synthesis_order if num_network_indices is not None and num_network_indices > 1 else False


def frame_schedule(
    network_indices: np.ndarray, num_frames: int, batch_size: int, group_by_network: bool
) -> List[Tuple[int, np.ndarray]]:
    """
    The order in which frames reach the networks: a list of (network index, frame indices) batches.
    Windows of `batch_size` consecutive frames; inside a window a stable grouping by network index
    when `group_by_network` (the reference's `sorted(..., key=network_index)`, 659-663), else runs of
    equal index in frame order. Every frame appears exactly once; batches never exceed batch_size.
    """
    if batch_size < 1:
        raise ValueError("batch_size must be positive")
    schedule: List[Tuple[int, np.ndarray]] = []
    indices = np.asarray(network_indices[:num_frames]).astype(np.int64)
    for start in range(0, num_frames, batch_size):
        window = np.arange(start, min(start + batch_size, num_frames))
        window_indices = indices[window]
        if group_by_network:
            for network_index in np.unique(window_indices):  # ascending, stable within a group
                schedule.append((int(network_index), window[window_indices == network_index]))
        else:
            run_start = 0
            for position in range(1, len(window) + 1):
                if position == len(window) or window_indices[position] != window_indices[run_start]:
                    schedule.append((int(window_indices[run_start]), window[run_start:position]))
                    run_start = position
    return schedule


def _uniform_output_side(networks: MultiNetwork, network_indices: np.ndarray, output_side_length: Optional[int]) -> Optional[int]:
    """The one side length every frame of the run will have, or None when it cannot be known / is not uniform."""
    if getattr(networks, "indexed_render_into", None) is None or getattr(networks, "indexed_output_resolution", None) is None:
        return None
    sides = {networks.indexed_output_resolution(int(index)) for index in np.unique(network_indices)}
    if None in sides:
        return None  # an interface built from the reference's four fields only (no CUDA renderer behind it)
    if output_side_length is not None:
        return int(output_side_length)
    return sides.pop() if len(sides) == 1 else None


def create_network_frames(
    latents: np.ndarray, network_indices: np.ndarray, networks: MultiNetwork, batch_size: int, group_by_network: bool,
    output_side_length: Optional[int] = None, reuse_frame_buffers: bool = False,
) -> Iterator[RGBInt8ImageType]:
    """
    Batched `create_network_frames` (network_visualization.py:598-681): `latents` is (T,512) or
    (T,num_ws,512); yields T frames in frame-index order.

    When every frame of the run has the same size - one network, networks of one resolution, or `output_side_length`
    given (then frames are scaled on the device, the reference's `scale_square_source_duplicate` step,
    gance/projection_file_blend.py:184-199) - the run is pipelined through pinned window buffers
    (gance_b200/track_renderer.py): rendering, device -> host copies and the consumer overlap. Networks of different
    resolutions without a common output size are rendered batch by batch, synchronously.
    """
    num_frames = len(latents)
    if num_frames == 0:
        return
    side = _uniform_output_side(networks, np.asarray(network_indices[:num_frames]), output_side_length)
    if side is not None:
        from gance_b200 import track_renderer

        for image in track_renderer.windowed_frames(networks, latents, np.asarray(network_indices), side, batch_size,
                                                    output_side_length, group_by_network,
                                                    reuse_frame_buffers=reuse_frame_buffers):
            yield RGBInt8ImageType(image)
        return
    schedule = frame_schedule(network_indices, num_frames, batch_size, group_by_network)
    position = 0
    while position < len(schedule):
        # all batches of one window
        window_start = (int(schedule[position][1][0]) // batch_size) * batch_size
        window_end = min(window_start + batch_size, num_frames)
        window_frames: List[Optional[np.ndarray]] = [None] * (window_end - window_start)
        while position < len(schedule) and int(schedule[position][1][0]) < window_end:
            network_index, frame_indices = schedule[position]
            LOGGER.debug(f"Rendering frames {frame_indices[0]}..{frame_indices[-1]} on network {network_index}")
            images = networks.indexed_create_images_batch(network_index, latents[frame_indices])
            if len(images) != len(frame_indices):  # one latent -> one frame, always (frame indexing is bit-exact)
                raise RuntimeError(f"network {network_index} returned {len(images)} frames for {len(frame_indices)} latents")
            for image, frame_index in zip(images, frame_indices):
                window_frames[int(frame_index) - window_start] = image
            position += 1
        for image in window_frames:
            yield RGBInt8ImageType(image)


def vector_synthesis(  # pylint: disable=too-many-arguments,unused-argument
    data: VisualizationInput,
    networks: Optional[MultiNetwork],
    default_vector_length: Optional[int] = 1024,
    visualization_height: Optional[int] = None,
    enable_3d: bool = False,
    enable_2d: bool = True,
    frames_to_visualize: Optional[int] = None,
    network_index_window_width: Optional[int] = None,
    force_optimize_synthesis_order: bool = True,
    unload_networks_when_complete: bool = False,
    batch_size: int = DEFAULT_BATCH_SIZE,
    output_side_length: Optional[int] = None,
    reuse_frame_buffers: bool = False,
) -> SynthesisOutput:
    """
    network_visualization.py:462-690 without the plots: every latent in `data.combined` goes through
    the network its `network_indices` entry names; `synthesized_images` is a lazy iterator of frames in
    frame-index order at the network's native resolution. The frame count is the number of latents
    (or network indices, whichever is fewer - `zip`, 245-249), truncated to `frames_to_visualize`.

    Additions (keyword-only in spirit, defaults keep the reference's behaviour): `batch_size` frames are rendered per
    launch sequence; `output_side_length` scales every frame on the device before it leaves for the host (the
    `scale_square_source_duplicate` step of gance/projection_file_blend.py:184-199, which then passes the frames through
    untouched); `reuse_frame_buffers` yields views of pinned window buffers instead of copies (valid until two windows
    later).
    """
    if networks is None:
        if not enable_3d and not enable_2d:
            raise ValueError("Nothing to render!")
        return SynthesisOutput(synthesized_images=None, visualization_images=None)
    vector_length = networks.expected_vector_length
    latents = sub_vectors(data=data.combined.data, vector_length=vector_length)
    indices = np.asarray(data.network_indices.result.data)
    total = min(len(latents), len(indices))
    if frames_to_visualize is not None:
        LOGGER.warning(f"Truncating output to the first {frames_to_visualize} frames!")
        total = min(total, frames_to_visualize)
    group = compute_force_synthesis_order(force_optimize_synthesis_order, len(networks.network_indices))

    def frames() -> Iterator[RGBInt8ImageType]:
        yield from create_network_frames(latents[:total], indices[:total], networks, batch_size, group,
                                         output_side_length=output_side_length, reuse_frame_buffers=reuse_frame_buffers)
        if unload_networks_when_complete:
            networks.unload()

    return SynthesisOutput(synthesized_images=frames(), visualization_images=None)
