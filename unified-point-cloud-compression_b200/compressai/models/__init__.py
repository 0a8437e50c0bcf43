from . import base  # noqa: F401
from .base import CompressionModel  # noqa: F401
