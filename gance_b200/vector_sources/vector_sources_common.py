"""
Host-side helpers on concatenated vector arrays that the callers of the hot path use directly
(mirrors the function names of /root/reference/gance/vector_sources/vector_sources_common.py).
The smoothing / rotation / resampling functions of that module run on the GPU here: see
gance_b200/apply_spectrogram.py.
"""
from typing import Tuple, Union

import numpy as np

from gance_b200 import divisor
from gance_b200.vector_sources.vector_types import (
    ConcatenatedMatrices,
    ConcatenatedVectors,
    DividedMatrices,
    DividedVectors,
    SingleMatrix,
    SingleVector,
    is_vector,
)


def pad_array(array: np.ndarray, size: int) -> np.ndarray:
    """Zero-pad the end of a 1-D array to `size` (vector_sources_common.py:33-41)."""
    return np.pad(array, pad_width=(0, size - len(array)), mode="constant", constant_values=0)


def remap_values_into_range(
    data, input_range: Tuple[float, float], output_range: Tuple[float, float]
) -> np.ndarray:
    """
    Linear map of input_range onto output_range (vector_sources_common.py:44-61). The reference
    evaluates a scipy `interp1d` once per value through a multiprocessing Pool; this is the same
    arithmetic - slope = (y1-y0)/(x1-x0); slope*(x-x0)+y0 in float64 - on the whole array.
    """
    x0, x1 = float(input_range[0]), float(input_range[1])
    y0, y1 = float(output_range[0]), float(output_range[1])
    slope = (y1 - y0) / (x1 - x0)
    return slope * (np.asarray(data, dtype=np.float64) - x0) + y0


def sub_vectors(
    data: Union[ConcatenatedMatrices, ConcatenatedVectors], vector_length: int
) -> Union[DividedMatrices, DividedVectors]:
    """(T*L,) -> (T,L); (D,T*L) -> (T,D,L) (vector_sources_common.py:86-101)."""
    if len(data.shape) >= 2:
        num_vectors = int(data.shape[-1] / vector_length)
        return DividedMatrices(np.array(np.split(data, num_vectors, axis=-1)))
    return DividedVectors(np.reshape(data, (-1, vector_length)))


def underlying_length(data: Union[SingleVector, SingleMatrix, ConcatenatedVectors, ConcatenatedMatrices]) -> int:
    """Length of a vector, or of the vectors inside a matrix (vector_sources_common.py:125-133)."""
    return int(data.shape[0] if is_vector(data) else data.shape[1])


def duplication_factor(original_count: int, target_vector_count: int) -> int:
    """
    How often each projected frame repeats (duplicate_to_vector_count,
    vector_sources_common.py:298-345): the division must be exact or ValueError is raised.
    """
    try:
        return divisor.divide_no_remainder(numerator=target_vector_count, denominator=original_count)
    except ValueError as error:
        raise ValueError(
            f"Cannot duplicate the input vectors (count {original_count}) "
            f"to the desired count {target_vector_count}."
        ) from error


def duplicate_to_vector_count(data: ConcatenatedVectors, vector_length: int, target_vector_count: int) -> ConcatenatedVectors:
    """Repeat every sub-vector the same number of times (vector_sources_common.py:298-345)."""
    split = sub_vectors(data=data, vector_length=vector_length)
    factor = duplication_factor(len(split), target_vector_count)
    return ConcatenatedVectors(np.repeat(split, factor, axis=0).reshape(-1))


def promote_to_matrix_duplicate(data: ConcatenatedVectors, target_depth: int) -> ConcatenatedMatrices:
    """(N,) -> (target_depth, N) by tiling (vector_sources_common.py:348-365)."""
    if len(data.shape) != 1:
        raise ValueError("Undefined behavior!")
    return ConcatenatedMatrices(np.tile(data, (target_depth, 1)))


def demote_to_vector_select(data: Union[SingleMatrix, ConcatenatedMatrices], index_to_take: int = 0):
    """Row `index_to_take` of a matrix (vector_sources_common.py:380-391)."""
    return ConcatenatedVectors(data[index_to_take])
