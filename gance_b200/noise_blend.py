"""
The hot-path body of the `noise-blend` command (/root/reference/music_into_networks.py:287-401): audio scaled so that one
video frame spans `vector_length` samples -> spectrogram latents blended with smoothed Gaussian noise -> the VECTOR path
of every listed network (G_mapping, psi = 1.2, fresh noise) -> frames at `output_side_length`. Argument names and
meaning are the command's; the click decoration, the run-config dump and the ffmpeg writer stay with the caller
(DESIGN.md §8). `noise_blend_frames` is the frame source the command hands to `write_source_to_disk_forward`;
`noise_blend_into` renders a rank's contiguous share into the encoder rank's indexed frame buffer.
"""

from pathlib import Path
from typing import List, NamedTuple, Optional, Sequence, Tuple, Union, cast

import numpy as np
import torch

from gance_b200 import track_renderer
from gance_b200.data_into_network_visualization import visualization_inputs
from gance_b200.data_into_network_visualization.network_visualization import vector_synthesis
from gance_b200.gance_types import ImageSourceType
from gance_b200.image_sources.video_common import scale_square_source_duplicate
from gance_b200.network_interface.network_functions import MultiNetwork
from gance_b200.vector_sources import music
from gance_b200.vector_sources.vector_types import ConcatenatedVectors

WavInput = Union[str, Path, music.WavFileProperties]


def scaled_audio(wav: Sequence[WavInput], vector_length: int, output_fps: float) -> ConcatenatedVectors:
    """music_into_networks.py:344-351: `read_wavs_scale_for_video(..., frames_per_second=output_fps).wav_data`."""
    wavs = [item if isinstance(item, music.WavFileProperties) else Path(item) for item in wav]
    return cast(ConcatenatedVectors, music.read_wavs_scale_for_video(wavs=wavs, vector_length=vector_length,
                                                                     frames_per_second=output_fps).wav_data)


def noise_blend_frames(  # pylint: disable=too-many-arguments
    wav: Sequence[WavInput],
    network_paths: List[Path],
    frames_to_visualize: Optional[int],
    output_fps: float,
    output_side_length: int,
    alpha: float,
    fft_roll_enabled: bool,
    fft_amplitude_range: Tuple[int, int],
    batch_size: int = 32,
    reuse_frame_buffers: bool = False,
) -> ImageSourceType:
    """music_into_networks.py:338-378 as a lazy frame source. Raises RuntimeError when no network can be loaded."""
    multi_networks = MultiNetwork(network_paths=network_paths)
    if multi_networks.__enter__() is None:  # the reference's `with` would hand None to the body and fail on attribute access
        raise RuntimeError("Couldn't load network into GPU")
    vector_length = multi_networks.expected_vector_length
    time_series_audio_vectors = scaled_audio(wav, vector_length, output_fps)
    synthesis_output = vector_synthesis(
        networks=multi_networks,
        data=visualization_inputs.alpha_blend_vectors_max_rms_power_audio(
            alpha=alpha,
            fft_roll_enabled=fft_roll_enabled,
            fft_amplitude_range=fft_amplitude_range,
            time_series_audio_vectors=time_series_audio_vectors,
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


class TrackLatents(NamedTuple):
    """The latent sequence of a whole track on the device plus the per-frame network index."""

    latents: torch.Tensor
    network_indices: np.ndarray


def prefetch_noise(wav: Sequence[WavInput], vector_length: int, output_fps: float):
    """
    Start drawing the noise source's random values (`primatives.NoisePrefetch`) from the WAV HEADERS alone - an upper bound
    of the frame count is all the draw needs - so that it overlaps reading and resampling the audio. Returns None when a
    header cannot be read (the values are then drawn where the reference draws them).
    """
    import wave

    from gance_b200.vector_sources.primatives import NoisePrefetch

    seconds = 0.0
    try:
        for item in wav:
            if isinstance(item, music.WavFileProperties):
                return None
            with wave.open(str(item), "rb") as handle:
                seconds += handle.getnframes() / float(handle.getframerate())
    except (wave.Error, OSError, EOFError):
        return None
    return NoisePrefetch(vector_length, int(np.ceil(seconds * output_fps)) + 8)


def noise_blend_latents(time_series_audio_vectors: ConcatenatedVectors, networks: MultiNetwork, alpha: float,
                        fft_roll_enabled: bool, fft_amplitude_range: Tuple[int, int], noise_prefetch=None) -> TrackLatents:
    """z [T,512] float32 on the device + network index per frame, for the whole track (every rank computes the same)."""
    result = visualization_inputs.alpha_blend_vectors_max_rms_power_audio_device(
        alpha=alpha, fft_roll_enabled=fft_roll_enabled, fft_amplitude_range=fft_amplitude_range,
        time_series_audio_vectors=time_series_audio_vectors, vector_length=networks.expected_vector_length,
        network_indices=networks.network_indices, noise_prefetch=noise_prefetch)
    return TrackLatents(result.latents, result.network_indices)


def noise_blend_into(  # pylint: disable=too-many-arguments
    dest: torch.Tensor,
    frame_range: Tuple[int, int],
    networks: MultiNetwork,
    track: TrackLatents,
    output_side_length: Optional[int] = None,
    frames_to_visualize: Optional[int] = None,
    batch_size: int = 32,
    dest_offset: int = 0,
) -> int:
    """Render frames [start, stop) of the track, grouped by network, to `dest[frame + dest_offset]`; returns their count."""
    total = min(int(track.latents.shape[0]), len(track.network_indices))
    if frames_to_visualize is not None:
        total = min(total, frames_to_visualize)
    start, stop = max(0, frame_range[0]), min(frame_range[1], total)
    if stop <= start:
        return 0
    track_renderer.render_frames_into(networks, track.latents, track.network_indices, dest, frames=np.arange(start, stop),
                                      dest_offset=dest_offset, batch_size=batch_size,
                                      output_side_length=output_side_length, group_by_network=True)
    return stop - start
