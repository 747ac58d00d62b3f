"""Exact division (mirrors /root/reference/gance/divisor.py:10-24)."""
import math
from typing import Union


def divide_no_remainder(numerator: Union[int, float], denominator: Union[int, float]) -> int:
    """numerator / denominator as an int; ValueError when the division leaves a remainder."""
    frac, whole = math.modf(numerator / denominator)
    if frac != 0:
        raise ValueError(f"Cannot evenly divide {numerator} into {denominator}")
    return int(whole)
