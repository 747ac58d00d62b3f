"""
The Gaussian noise source of the noise-blend path (mirrors `gaussian_data` of
/root/reference/gance/vector_sources/primatives.py:49-74). The random values stay on the host - they
come from numpy's legacy MT19937 + polar Box-Muller stream (`RandomState(1234).randn`), which has to be
reproduced bit for bit - and so do the two global float32 reductions (RMS here, min-max in the caller).
The wrap-around Gaussian filter along time, 61 % of the whole audio -> latents time for an album-length
track when scipy ran it, is `gance_gaussian_filter_time_wrap` (bit-exact with scipy's NI_Correlate1D).
"""
from typing import NamedTuple, Optional

import numpy as np
import scipy.ndimage

from gance_b200.vector_sources.vector_types import ConcatenatedVectors

DEFAULT_RANDOM_SEED = 1234


class NoisePrefetch:
    """
    The random values of `gaussian_data`, drawn on a background thread while the caller reads and resamples the audio.

    numpy's legacy `RandomState.randn` stream is strictly sequential (MT19937 + polar Box-Muller with rejection: no
    jump-ahead) and costs ~90 ms for an album-length track - the longest serial host step between the input files and the
    first frame, repeated on every rank. It only needs the NUMBER of vectors, and a longer draw starts with the same values
    (C order), so it can begin from an upper bound of the track length as soon as the WAV header has been read;
    `take(num_vectors)` joins the thread and returns the first `num_vectors` rows - bit for bit what
    `RandomState(seed).randn(num_vectors, 1, vector_length)` returns. numpy releases the GIL inside the draw.
    """

    def __init__(self, vector_length: int, num_vectors_upper_bound: int, seed: int = DEFAULT_RANDOM_SEED) -> None:
        import threading

        self.vector_length, self.bound, self.seed = int(vector_length), max(0, int(num_vectors_upper_bound)), seed
        self._values: Optional[np.ndarray] = None
        self._thread = threading.Thread(target=self._draw, daemon=True)
        self._thread.start()

    def _draw(self) -> None:
        state = np.random.RandomState(self.seed)  # pylint: disable=no-member
        self._values = state.randn(self.bound, 1, self.vector_length).astype(np.float32)

    def take(self, num_vectors: int) -> np.ndarray:
        """float32 [num_vectors, 1, vector_length], a fresh writable array."""
        self._thread.join()
        if self._values is None or num_vectors > self.bound:  # the bound was too small: draw again, serially
            state = np.random.RandomState(self.seed)  # pylint: disable=no-member
            return state.randn(num_vectors, 1, self.vector_length).astype(np.float32)
        return np.ascontiguousarray(self._values[:num_vectors]).copy() if num_vectors < self.bound else self._values


class Sigmas(NamedTuple):
    """Smoothing across consecutive vectors / within a vector (primatives.py:36-46)."""

    across_vectors: int
    within_vectors: int


def gaussian_data(
    vector_length: int,
    num_vectors: int,
    sigmas: Sigmas = Sigmas(20, 0),
    random_state: Optional["np.random.RandomState"] = None,
    prefetched: Optional[NoisePrefetch] = None,
) -> ConcatenatedVectors:
    """
    float32 (T*L,): randn -> wrap-around gaussian filter along time -> divided by its RMS. `prefetched` (an addition)
    supplies the default-seed random values drawn ahead of time on a background thread; the result is the same.
    """
    if prefetched is not None and random_state is None and prefetched.vector_length == vector_length \
            and prefetched.seed == DEFAULT_RANDOM_SEED:
        latents = prefetched.take(num_vectors)
    else:
        if random_state is None:
            random_state = np.random.RandomState(DEFAULT_RANDOM_SEED)  # pylint: disable=no-member
        latents = random_state.randn(num_vectors, 1, vector_length).astype(np.float32)
    if sigmas.across_vectors > 1e-15:  # scipy skips axes whose sigma is zero
        from gance_b200 import audio_device  # needs the CUDA library: no CPU fallback on the hot path

        latents = audio_device.gaussian_filter_time_wrap(latents, float(sigmas.across_vectors))
    if sigmas.within_vectors > 1e-15:  # never set by the hot path (visualization_inputs.py:135-142 passes 0): demo shapes only
        latents = scipy.ndimage.gaussian_filter(input=latents, sigma=(0, 0, sigmas.within_vectors), mode="wrap")
    latents /= np.sqrt(np.mean(np.square(latents)))
    return ConcatenatedVectors(latents.reshape(vector_length * num_vectors))
