"""gpu_fhe_b200 -- host side of the B200-native NTT / RNS poly-mul path.

The directory is named ``gpu-fhe_b200`` (not importable by a plain ``import``); the drop-in
modules at the repo root (``arithmetic.py``, ``primitive.py``, ``polynomial.py``) load it under
the module name ``gpu_fhe_b200``.  It holds only what the hot path needs:

  csrc/         hand-written sm_100a kernels + the C ABI (include/gpufhe_b200.h)
  _capi.py      ctypes binding of libgpufhe_b200.so -- fails loudly if the library is missing
  ring.py       roots of unity, RNS ring context, plan (twiddle table) cache
  dispatch.py   tensor / ndarray marshalling for the seven reference-level functions
  shard.py      limb/batch partition across GPUs + the final all-gather
"""
from . import _capi, ring, dispatch, shard  # noqa: F401

__all__ = ["_capi", "ring", "dispatch", "shard"]
