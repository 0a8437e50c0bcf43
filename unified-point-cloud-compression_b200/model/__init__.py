"""Host-side mirror of the reference's model/*.py (same class names, constructor configs, parameter
names and state-dict keys), written against the fused libpcg.so paths instead of one ME call per op."""
from .blocks import MinkowskiGDN  # noqa: F401
from .entropy_models import MeanScaleHyperprior  # noqa: F401
from .model import UnifiedModel  # noqa: F401
from .transforms import AnalysisTransform, SparseSynthesisTransform  # noqa: F401
