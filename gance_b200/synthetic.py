"""
Synthetic inputs for tests and benchmarks: no WAVs, projection files or network
pickles ship with the reference (`.gitignore:107-110`, `.MISSING_LARGE_BLOBS`), so every run here
is driven by these deterministic generators.
"""

from typing import Optional

import numpy as np


def chirp_wav(
    seconds: float, sample_rate: int = 44100, noise_floor_db: Optional[float] = -60.0
) -> np.ndarray:
    """
    Mono int16 linear chirp 100 Hz -> 8 kHz at amplitude 0.8. The optional white-noise floor
    (RandomState(0)) keeps spectrogram bins away from log10(0).
    """
    n = int(round(seconds * sample_rate))
    t = np.arange(n) / sample_rate
    f0, f1 = 100.0, 8000.0
    phase = 2 * np.pi * (f0 * t + 0.5 * (f1 - f0) / seconds * t * t)
    signal = 0.8 * np.sin(phase)
    if noise_floor_db is not None:
        signal = signal + (10.0 ** (noise_floor_db / 20.0)) * np.random.RandomState(0).randn(n)
    return np.clip(np.round(signal * 32767.0), -32768, 32767).astype(np.int16)


def projection_latents(
    num_frames: int, depth: int = 18, vector_length: int = 512, seed: int = 1
) -> np.ndarray:
    """
    Stand-in for `ProjectionFileReader.final_latents_matrices_label().data`
    (projection_file_reader.py:263-300): (depth, F*512) float32, a smooth random walk in w whose
    depth rows are identical (what `verify_projection_file_assumptions` checks, 236-260).
    """
    rng = np.random.RandomState(seed)
    base = rng.randn(vector_length)
    steps = rng.randn(num_frames, vector_length) * 0.05
    walk = (base[None, :] + np.cumsum(steps, axis=0)).astype(np.float32)
    return np.tile(walk.reshape(-1), (depth, 1))
