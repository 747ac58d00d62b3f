"""
The NamedTuples that carry latents from the vector sources to the synthesis loop (mirrors
/root/reference/gance/data_into_network_visualization/visualization_common.py:30-116, without
the matplotlib helpers that module also holds).
"""
from typing import List, NamedTuple, Union

import numpy as np

from gance_b200.vector_sources.vector_types import MatricesLabel, SingleMatrix, SingleVector, VectorsLabel

DataLabelDataType = Union[np.ndarray, SingleVector, SingleMatrix]


class DataLabel(NamedTuple):
    """A piece of data and the label it is plotted under (visualization_common.py:30-37)."""

    data: DataLabelDataType
    label: str


class ResultLayers(NamedTuple):
    """A result plus the intermediate steps that led to it (visualization_common.py:40-47)."""

    result: DataLabel
    layers: List[DataLabel] = []


class VisualizationInput(NamedTuple):
    """
    visualization_common.py:65-86. `combined.data` is what reaches the networks: (T*512,) for the
    vector path or (18, T*512) for the matrix path; `network_indices.result.data` is int (T,).
    """

    a_vectors: Union[VectorsLabel, MatricesLabel]
    b_vectors: Union[VectorsLabel, MatricesLabel]
    combined: Union[VectorsLabel, MatricesLabel]
    network_indices: ResultLayers


class FrameInput(NamedTuple):
    """Everything needed to render one frame (visualization_common.py:89-116)."""

    frame_index: int
    a_sample: DataLabel
    b_sample: DataLabel
    combined_sample: DataLabel
    network_index: int
    surrounding_network_indices: np.ndarray
    network_index_layers: List[DataLabel]
