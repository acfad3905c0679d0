"""concrete_b200 — B200-native (sm_100a) TFHE hot path behind the tfhe-cuda-backend C ABI.

The product is ``libconcrete_b200.so`` (``csrc/``, C ABI in ``include/concrete_b200.h``).  This
package is the host-side harness around it: it loads the library, fails loudly if it is missing,
and exposes the entry points with torch tensors standing in for raw device buffers.
There is no CPU fallback and nothing here imports ``oracle``.
"""
from .lib import load_library, library_path  # noqa: F401
from .engine import Engine, as_i64, as_u64  # noqa: F401
from .params import PARAMS, TfheParams  # noqa: F401

__all__ = ["load_library", "library_path", "Engine", "PARAMS", "TfheParams", "as_i64", "as_u64"]
