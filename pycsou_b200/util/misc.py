"""Shape helpers used by the map algebra (mirror of ``pycsou/util/misc.py:15-88``)."""
from typing import Tuple


def is_range_broadcastable(shape1: Tuple[int, int], shape2: Tuple[int, int]) -> bool:
    """Two maps can be summed iff their domains agree and their ranges agree or one of them is scalar-valued."""
    same_domain = shape1[1] == shape2[1]
    ranges_ok = shape1[0] == shape2[0] or 1 in (shape1[0], shape2[0])
    return bool(same_domain and ranges_ok)


def range_broadcast_shape(shape1: Tuple[int, int], shape2: Tuple[int, int]) -> Tuple[int, int]:
    if not is_range_broadcastable(shape1, shape2):
        raise ValueError('Shapes are not (range) broadcastable.')
    return (max(shape1[0], shape2[0]), max(shape1[1], shape2[1]))
