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


class Sigmas(NamedTuple):
    """Smoothing across consecutive vectors / within a vector (primatives.py:36-46)."""

    across_vectors: int
    within_vectors: int


def gaussian_data(
    vector_length: int,
    num_vectors: int,
    sigmas: Sigmas = Sigmas(20, 0),
    random_state: Optional["np.random.RandomState"] = None,
) -> ConcatenatedVectors:
    """float32 (T*L,): randn -> wrap-around gaussian filter along time -> divided by its RMS."""
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
