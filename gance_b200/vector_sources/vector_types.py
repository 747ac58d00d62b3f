"""
Shapes that cross the boundary (mirrors /root/reference/gance/vector_sources/vector_types.py):
  SingleVector (512,)            ConcatenatedVectors (T*512,)         DividedVectors (T,512)
  SingleMatrix (num_ws,512)      ConcatenatedMatrices (num_ws,T*512)  DividedMatrices (T,num_ws,512)
"""
from typing import NamedTuple, NewType, Union

import numpy as np

SingleVector = NewType("SingleVector", np.ndarray)
ConcatenatedVectors = NewType("ConcatenatedVectors", np.ndarray)
DividedVectors = NewType("DividedVectors", np.ndarray)
SingleMatrix = NewType("SingleMatrix", np.ndarray)
ConcatenatedMatrices = NewType("ConcatenatedMatrices", np.ndarray)
DividedMatrices = NewType("DividedMatrices", np.ndarray)


class VectorsLabel(NamedTuple):
    """A concatenated vector array and the name it is plotted under (vector_types.py:37-44)."""

    data: ConcatenatedVectors
    vector_length: int
    label: str


class MatricesLabel(NamedTuple):
    """A concatenated matrix array and its name (vector_types.py:47-55)."""

    data: ConcatenatedMatrices
    vector_length: int
    label: str


def is_vector(data: Union[SingleVector, SingleMatrix, np.ndarray]) -> bool:
    """vector_types.py:58-68: anything with fewer than two dimensions is a vector."""
    return len(data.shape) < 2
