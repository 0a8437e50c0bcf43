from . import ops  # noqa: F401
from .bound_ops import LowerBound  # noqa: F401
from .ops import quantize_ste  # noqa: F401
from .parametrizers import NonNegativeParametrizer  # noqa: F401
