"""
Types shared by the image sources (mirror of gance/image_sources/image_sources_common.py:1-27).
"""

from typing import NamedTuple

import numpy as np


class ImageResolution(NamedTuple):
    """Creators are responsible for the order (reference: image_sources_common.py:10-17)."""

    width: int
    height: int


def image_resolution(image: np.ndarray) -> ImageResolution:
    """Resolution of an (H,W,3) image (reference: image_sources_common.py:20-27)."""
    return ImageResolution(height=image.shape[0], width=image.shape[1])
