"""
The hot-path half of `projection_file_blend_api` (/root/reference/gance/projection_file_blend.py:56-199): projection
file -> audio scaled to the projection's frame count -> alpha / blend-depth mix -> synthesis on the listed networks ->
frames at `output_side_length`. Same argument names and meaning, same checks and errors (incomplete projection file,
frame multiplier that does not divide, overlay mask without overlay) in the same order.

What is not here: the eye-tracking overlay (:223-278), the matplotlib debug video (:288-341) and the ffmpeg/x264 writer
(:280-286) - CPU post-effects and a host codec, outside this build's scope. The frames are handed over
instead: `projection_file_blend_frames` is the lazy frame iterator the reference passes to
`write_source_to_disk_forward`; `projection_file_blend_api` consumes it and, when `output_path` is given, writes the
frames as one headless raw file (uint8 [T,S,S,3] + a JSON sidecar) where the reference writes an .mp4;
`projection_file_blend_into` renders a rank's contiguous share straight into the encoder rank's indexed frame buffer
(the multi-GPU form).
"""

import json
from pathlib import Path
from typing import Iterator, List, NamedTuple, Optional, Sequence, Tuple, Union, cast

import numpy as np
import torch

from gance_b200 import divisor, track_renderer
from gance_b200.data_into_network_visualization import visualization_inputs
from gance_b200.data_into_network_visualization.network_visualization import vector_synthesis
from gance_b200.gance_types import ImageSourceType
from gance_b200.image_sources.video_common import scale_square_source_duplicate
from gance_b200.logger_common import LOGGER
from gance_b200.network_interface.network_functions import MultiNetwork
from gance_b200.projection import projection_file_reader
from gance_b200.vector_sources import music
from gance_b200.vector_sources.vector_sources_common import underlying_length
from gance_b200.vector_sources.vector_types import ConcatenatedVectors, MatricesLabel

WavInput = Union[str, Path, music.WavFileProperties]


class ProjectionBlendInputs(NamedTuple):
    """Everything `projection_file_blend_api` derives before synthesis starts (reference :124-157)."""

    final_latents: MatricesLabel
    frame_multiplier: int
    num_output_frames: int
    time_series_audio_vectors: ConcatenatedVectors


def _wavs(wav: Sequence[WavInput]) -> list:
    return [item if isinstance(item, music.WavFileProperties) else Path(item) for item in wav]


def read_projection_blend_inputs(wav: Sequence[WavInput], projection_file_path: Union[str, Path], output_fps: float,
                                 vector_length: int, share_decoding: bool = False) -> ProjectionBlendInputs:
    """
    Reference :124-157: read the final latents, refuse an incomplete file, derive the frame multiplier
    (`output_fps / projection_fps`, which must divide exactly) and scale the audio to `multiplier * frames_in_file` vectors.
    `share_decoding` (an addition, for frame-sharded runs; every rank must make the call): each rank inflates its share of
    the file's frames and one all-gather completes the array on every rank (`final_latents_matrices_label_sharded`).
    """
    with projection_file_reader.load_projection_file(Path(projection_file_path)) as reader:
        if share_decoding:
            final_latents = projection_file_reader.final_latents_matrices_label_sharded(reader)
        else:
            final_latents = projection_file_reader.final_latents_matrices_label(reader)
        final_latents_in_file = underlying_length(final_latents.data) / vector_length
        processed_frames_in_file = reader.projection_attributes.projection_frame_count
        projection_complete = reader.projection_attributes.complete
        LOGGER.info(
            f"Reading projection file. Complete: {projection_complete}, "
            f"Final Latent Count: {final_latents_in_file}, Processed Frames: {processed_frames_in_file}"
        )
        if not projection_complete or abs(final_latents_in_file - processed_frames_in_file) > 2:
            raise ValueError("Invalid Projection File, cannot continue.")
        frame_multiplier = divisor.divide_no_remainder(
            numerator=output_fps, denominator=reader.projection_attributes.projection_fps
        )
    num_output_frames = int(frame_multiplier * final_latents_in_file)
    time_series_audio_vectors = cast(
        ConcatenatedVectors,
        music.read_wavs_scale_for_video(wavs=_wavs(wav), vector_length=vector_length, target_num_vectors=num_output_frames).wav_data,
    )
    return ProjectionBlendInputs(final_latents, frame_multiplier, num_output_frames, time_series_audio_vectors)


def _check_overlay_arguments(complexity_change_rolling_sum_window, complexity_change_threshold, phash_distance,
                             bbox_distance, track_length, debug_path) -> None:
    overlay_enabled = all(param is not None for param in (phash_distance, bbox_distance, track_length))
    overlay_music_mask_enabled = all(
        param is not None for param in (complexity_change_rolling_sum_window, complexity_change_threshold)
    )
    if overlay_music_mask_enabled and not overlay_enabled:
        raise ValueError("Overlay music mask without overlay being enabled is not supported!")
    if overlay_enabled or debug_path is not None:
        raise NotImplementedError(
            "The eye-tracking overlay and the debug visualization are CPU post-effects outside this build's scope "
            "(DESIGN.md §8): pass None for phash_distance / bbox_distance / track_length / debug_path."
        )


def projection_file_blend_frames(  # pylint: disable=too-many-arguments
    wav: Sequence[WavInput],
    network_paths: List[Path],
    frames_to_visualize: Optional[int],
    output_fps: float,
    output_side_length: int,
    alpha: float,
    fft_roll_enabled: bool,
    fft_amplitude_range: Tuple[int, int],
    projection_file_path: Union[str, Path],
    blend_depth: int,
    batch_size: int = 32,
    reuse_frame_buffers: bool = False,
) -> ImageSourceType:
    """
    Reference :122-199 as a lazy frame source: what it hands to `write_source_to_disk_forward`. Frames are scaled to
    `output_side_length` on the device (the `scale_square_source_duplicate` of :192-195 then passes them through).
    """
    multi_networks = MultiNetwork(network_paths=network_paths, load=True)
    vector_length = multi_networks.expected_vector_length
    inputs = read_projection_blend_inputs(wav, projection_file_path, output_fps, vector_length)
    synthesis_output = vector_synthesis(
        networks=multi_networks,
        data=visualization_inputs.alpha_blend_projection_file(
            final_latents_matrices_label=inputs.final_latents,
            alpha=alpha,
            fft_roll_enabled=fft_roll_enabled,
            fft_amplitude_range=fft_amplitude_range,
            blend_depth=blend_depth,
            time_series_audio_vectors=inputs.time_series_audio_vectors,
            vector_length=vector_length,
            network_indices=multi_networks.network_indices,
        ),
        default_vector_length=vector_length,
        enable_3d=False,
        enable_2d=False,
        frames_to_visualize=frames_to_visualize,
        unload_networks_when_complete=True,
        batch_size=batch_size,
        output_side_length=output_side_length,
        reuse_frame_buffers=reuse_frame_buffers,
    )
    return scale_square_source_duplicate(source=synthesis_output.synthesized_images, output_side_length=output_side_length)


def projection_file_blend_api(  # pylint: disable=too-many-arguments,too-many-locals
    wav: List[str],
    output_path: Optional[str],
    network_paths: List[Path],
    frames_to_visualize: Optional[int],
    output_fps: float,
    output_side_length: int,
    debug_path: Optional[str],
    debug_window: Optional[int],
    debug_side_length: Optional[int],
    alpha: float,
    fft_roll_enabled: bool,
    fft_amplitude_range: Tuple[int, int],
    projection_file_path: str,
    blend_depth: int,
    complexity_change_rolling_sum_window: Optional[int],
    complexity_change_threshold: Optional[int],
    phash_distance: Optional[int],
    bbox_distance: Optional[float],
    track_length: Optional[int],
) -> int:
    """
    The reference's entry point with its argument list (:56-76). Renders every frame; with `output_path` the frames go to
    a raw frame file (uint8 [T,S,S,3] back to back, shape in `output_path + ".json"`) instead of the reference's x264 video. Returns the number of frames rendered.
    """
    del debug_window, debug_side_length
    _check_overlay_arguments(complexity_change_rolling_sum_window, complexity_change_threshold, phash_distance,
                             bbox_distance, track_length, debug_path)
    frames = projection_file_blend_frames(
        wav=wav, network_paths=network_paths, frames_to_visualize=frames_to_visualize, output_fps=output_fps,
        output_side_length=output_side_length, alpha=alpha, fft_roll_enabled=fft_roll_enabled,
        fft_amplitude_range=fft_amplitude_range, projection_file_path=projection_file_path, blend_depth=blend_depth,
        reuse_frame_buffers=True,  # every frame is consumed before the next is drawn
    )
    count = 0
    if output_path is None:
        for _ in frames:
            count += 1
        return count
    with open(output_path, "wb") as handle:  # headless artefact: raw RGB frames back to back + a sidecar describing them
        for frame in frames:
            handle.write(np.ascontiguousarray(frame).tobytes())
            count += 1
    with open(str(output_path) + ".json", "w") as sidecar:
        json.dump({"shape": [count, output_side_length, output_side_length, 3], "dtype": "uint8", "layout": "THWC RGB",
                   "fps": output_fps}, sidecar)
    return count


def projection_blend_latents(inputs: ProjectionBlendInputs, networks: MultiNetwork, alpha: float, fft_roll_enabled: bool,
                             fft_amplitude_range: Tuple[int, int], blend_depth: int):
    """w+ [T,18,512] float32 on the device + network index per frame, for the whole track (every rank computes the same)."""
    from gance_b200.noise_blend import TrackLatents

    result = visualization_inputs.alpha_blend_projection_file_device(
        final_latents_matrices_label=inputs.final_latents, alpha=alpha, fft_roll_enabled=fft_roll_enabled,
        fft_amplitude_range=fft_amplitude_range, blend_depth=blend_depth,
        time_series_audio_vectors=inputs.time_series_audio_vectors, vector_length=networks.expected_vector_length,
        network_indices=networks.network_indices)
    return TrackLatents(result.latents, result.network_indices)


def projection_file_blend_into(  # pylint: disable=too-many-arguments
    dest: torch.Tensor,
    frame_range: Tuple[int, int],
    networks: MultiNetwork,
    track,
    output_side_length: Optional[int] = None,
    frames_to_visualize: Optional[int] = None,
    batch_size: int = 32,
    dest_offset: int = 0,
) -> int:
    """
    The multi-GPU form: the w+ latents of the WHOLE track (`projection_blend_latents`: cheap, deterministic, identical
    on every rank - no collective) are on this rank's device, and the frames of `frame_range` are rendered, grouped by
    network, straight to `dest[frame + dest_offset]` - the rank's registered slice of the encoder's `SharedFrameSink`,
    or a device buffer. Returns the number of frames this call rendered.
    """
    from gance_b200.noise_blend import noise_blend_into

    return noise_blend_into(dest, frame_range, networks, track, output_side_length, frames_to_visualize, batch_size, dest_offset)


def frames_of(source: Iterator[np.ndarray]) -> np.ndarray:
    """Materialise a frame source (tests)."""
    return np.stack([np.array(frame) for frame in source])
