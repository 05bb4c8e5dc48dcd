from .base import CompositeTransform, Transform, InverseNotAvailable  # noqa: F401
from .permutations import Permutation, ReversePermutation  # noqa: F401
from .autoregressive import (  # noqa: F401
    AutoregressiveTransform,
    MaskedPiecewiseRationalQuadraticAutoregressiveTransform,
)
from . import made, splines  # noqa: F401
