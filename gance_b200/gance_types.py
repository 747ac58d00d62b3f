"""Image types at the boundary (mirrors /root/reference/gance/gance_types.py:31-36)."""
from typing import Iterator, NewType, Optional

import numpy as np  # noqa: F401

# dimensions are (Height, Width, Colors), dtype uint8, RGB order
RGBInt8ImageType = NewType("RGBInt8ImageType", "np.ndarray")

ImageSourceType = Iterator[RGBInt8ImageType]
OptionalImageSourceType = Iterator[Optional[RGBInt8ImageType]]
