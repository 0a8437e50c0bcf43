"""B200-native graft of the learned point-cloud codec hot path (ikt-luh/Unified-Point-Cloud-Compression).

Layout
  csrc/              hand-written sm_100a CUDA kernels + the C ABI (include/pcg.h) -> libpcg.so
  _lib.py            ctypes binding (no torch types cross the ABI)
  sparse_ops.py      coordinate sets / kernel maps / sparse conv / prune / lookup on top of the ABI
  entropy_ops.py     entropy-model arithmetic on top of the ABI
  MinkowskiEngine/   drop-in for the `MinkowskiEngine` names the reference imports
  compressai/        drop-in for the `compressai` names the reference imports
  model/             host-side mirror of the reference's model/*.py (same classes, same state-dict keys)
  codec.py           frame / block sharding driver (one process per GPU, no collective)

`install_shims()` registers the two drop-in packages under their upstream names so the reference's
unmodified `model/*.py` can `import MinkowskiEngine as ME` / `from compressai... import ...`.
"""
import sys

from . import _lib, sparse_ops  # noqa: F401


def install_shims():
    from . import MinkowskiEngine as _me
    from . import compressai as _cai
    sys.modules["MinkowskiEngine"] = _me
    sys.modules["MinkowskiEngine.utils"] = _me.utils
    sys.modules["MinkowskiEngine.utils.collation"] = _me.utils.collation
    sys.modules["compressai"] = _cai
    for sub in ("ops", "layers", "entropy_models", "models"):
        sys.modules[f"compressai.{sub}"] = getattr(_cai, sub)
    sys.modules["compressai.ops.ops"] = _cai.ops.ops
    sys.modules["compressai.models.base"] = _cai.models.base
    return _me, _cai
