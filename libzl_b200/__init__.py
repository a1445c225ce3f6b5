"""libzl_b200 — B200-native (sm_100a CUDA) sampler voice-render path of zynthbox/libzl.

The compute lives in ``libzl_b200/libzl_b200.so`` (hand-written CUDA kernels behind the C ABI of
``include/zl_b200.h``).  This package is the thin Python host view of that ABI.  There is no CPU
fallback: if the shared library is missing the import of :func:`load` fails loudly.
"""
import os

from . import _capi
from ._capi import *  # noqa: F401,F403  (constants + PODs)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ZL_B200_LIB") or os.path.join(_HERE, "libzl_b200.so")   # env: kernel-variant experiments
_lib = None


def load():
    """Load the CUDA product library (never the oracle)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc -gencode arch=compute_100a,code=sm_100a). There is no CPU fallback.")
        _lib = _capi.Lib(LIB_PATH, "zl_")
    return _lib


from .engine import Engine, ZlError  # noqa: E402,F401
