from . import torchutils  # noqa: F401
