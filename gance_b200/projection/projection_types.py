"""Types of the projection files (mirrors /root/reference/gance/projection/projection_types.py)."""

from typing import NewType, Tuple

import numpy as np

from gance_b200.vector_sources.vector_types import SingleMatrix

CompleteLatentsType = NewType("CompleteLatentsType", np.ndarray)  # (1, num_ws, 512) float32
FlattenedNoisesType = NewType("FlattenedNoisesType", np.ndarray)
NoisesShapesType = NewType("NoisesShapesType", Tuple[Tuple[int, int, int, int], ...])


def complete_latents_to_matrix(complete_latents: CompleteLatentsType) -> SingleMatrix:
    """(1, num_ws, 512) as stored -> (num_ws, 512) (projection_types.py:22-28)."""
    return SingleMatrix(complete_latents[0])
