from .base import Distribution  # noqa: F401
from .normal import StandardNormal  # noqa: F401
