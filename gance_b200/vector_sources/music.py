"""
Reads music into the shape the vector sources want (mirrors
/root/reference/gance/vector_sources/music.py: `WavFileProperties`, `read_wav_file`,
`read_wavs_scale_for_video`, `_scale_wav_to_sample_rate` with the same arguments and errors).

The two slow host steps of the reference run as CUDA kernels here: the PCM -> [-1,1] float map,
which the reference evaluates one sample at a time through a multiprocessing Pool
(vector_sources_common.py:44-61), and the band-limited sinc resampling, a single-threaded numba loop
in resampy (music.py:223-227). Both are bit-exact with the reference (tests/test_gpu_audio.py).
"""
import pickle
from pathlib import Path
from typing import List, NamedTuple, Optional, Union

import numpy as np
import torch
from scipy.io import wavfile

from gance_b200 import audio_device
from gance_b200.logger_common import LOGGER
from gance_b200.vector_sources.vector_sources_common import pad_array

WavDataType = Union["np.ndarray[np.int16]", "np.ndarray[np.float32]"]


class WavFileProperties(NamedTuple):
    """Sample rate, samples and a name for logging (music.py:20-33)."""

    sample_rate: int
    wav_data: WavDataType
    name: str


def read_wav_file(wav_path: Path, convert_to_32bit_float: bool = True) -> WavFileProperties:
    """music.py:172-209: integer PCM is scaled to float32 in [-1, 1]."""
    sample_rate, wav_data = wavfile.read(str(wav_path))
    if convert_to_32bit_float and wav_data.dtype != np.float32:
        if wav_data.dtype not in (np.int32, np.int16, np.int8):
            raise ValueError(f"Cannot safely convert wav data to np.float32, unknown input format: {wav_data.dtype}")
        if wav_data.dtype == np.int8 and wav_data.size and int(wav_data.min()) < 0:
            # the reference maps int8 through interp1d([0, 255], [-1, 1]) (music.py:197-198), which refuses values
            # outside its x range: a negative sample is a ValueError there, so it is one here (scipy itself hands
            # 8-bit WAVs over as uint8, which the line above already refuses, exactly like the reference)
            raise ValueError("A value in x_new is below the interpolation range.")
        wav_data = audio_device.pcm_to_float(wav_data).cpu().numpy()
    return WavFileProperties(sample_rate=sample_rate, wav_data=wav_data, name=Path(wav_path).with_suffix("").name)


def _scale_wav_to_sample_rate(wav_file: WavFileProperties, new_sample_rate: Union[int, float]) -> WavFileProperties:
    """music.py:212-230: resampy.resample(kaiser_best); the NT keeps the input sample rate."""
    device = audio_device._device()  # pylint: disable=protected-access
    data = torch.as_tensor(np.ascontiguousarray(wav_file.wav_data, dtype=np.float32)).to(device)
    scaled = audio_device.sinc_resample(data, sr_orig=wav_file.sample_rate, sr_new=new_sample_rate)
    return WavFileProperties(
        wav_data=scaled.cpu().numpy(), sample_rate=wav_file.sample_rate, name=f"{wav_file.name}_scaled"
    )


def read_wavs_scale_for_video(  # pylint: disable=too-many-locals
    wavs: Union[List[Path], List[WavFileProperties]],
    vector_length: int,
    frames_per_second: Optional[float] = None,
    target_num_vectors: Optional[int] = None,
    cache_path: Optional[Path] = None,
    pad_to_length: bool = True,
) -> WavFileProperties:
    """
    music.py:60-169: mono mix, concatenate, resample so that one video frame spans `vector_length`
    samples (FPS mode) or so that the track yields `target_num_vectors` vectors, zero-pad to a
    multiple of `vector_length`. Exactly one of the two modes must be chosen.
    """
    if frames_per_second is not None and target_num_vectors is not None:
        raise ValueError("Can't use both FPS mode and target vector count mode.")
    if frames_per_second is None and target_num_vectors is None:
        raise ValueError("Need to use FPS mode or target vector count mode.")
    if cache_path is not None:
        if cache_path.exists():
            with open(str(cache_path), "rb") as read_file:
                LOGGER.info("Cached audio found. Loading.")
                return pickle.load(read_file)
        LOGGER.warning("No audio cache found, will be created.")

    input_wavs = [read_wav_file(wav) if isinstance(wav, Path) else wav for wav in wavs]
    sample_rates = {input_wav.sample_rate for input_wav in input_wavs}
    if len(sample_rates) != 1:
        raise ValueError("Multiple sample rates for input audio files is unsupported.")
    input_wav = WavFileProperties(
        wav_data=np.concatenate(
            [w.wav_data.mean(axis=1) if len(w.wav_data.shape) > 1 else w.wav_data for w in input_wavs]
        ),
        sample_rate=next(iter(sample_rates)),
        name="_".join([w.name for w in input_wavs]) + "_mono",
    )
    num_wav_samples = input_wav.wav_data.shape[0]
    if frames_per_second is not None:
        scaled_sample_rate = int(
            input_wav.sample_rate
            * (vector_length * (frames_per_second * (num_wav_samples / input_wav.sample_rate)))
            / num_wav_samples
        )
    else:
        original_num_vectors = num_wav_samples / vector_length
        scaled_sample_rate = float(input_wav.sample_rate) * (target_num_vectors / original_num_vectors)
    scaled_wav = _scale_wav_to_sample_rate(input_wav, scaled_sample_rate)
    wav_data = (
        pad_array(scaled_wav.wav_data, int(np.ceil(scaled_wav.wav_data.shape[0] / vector_length) * vector_length))
        if pad_to_length
        else scaled_wav.wav_data
    )
    output = WavFileProperties(wav_data=wav_data, sample_rate=input_wav.sample_rate, name=f"{scaled_wav.name}_padded")
    if cache_path is not None:
        with open(str(cache_path), "wb") as write_file:
            pickle.dump(output, write_file)
    return output
