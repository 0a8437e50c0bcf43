"""Drop-in for the `compressai` names the reference imports (compressai==1.2.4, requirements.txt:9).

Same class names, parameters, buffers and state-dict keys; the likelihood / quantise / CDF-index
arithmetic of EntropyBottleneck and GaussianConditional runs in libpcg.so kernels, the range coder
is the host coder in csrc/host_rans.cpp (format-compatible with compressai's `ans` extension).
"""
from . import entropy_models, layers, models, ops  # noqa: F401

__version__ = "1.2.4+pcg.b200"
