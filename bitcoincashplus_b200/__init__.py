"""bitcoincashplus_b200 — B200 (sm_100a) engine for the block-validation hot path of the
grospy/bitcoincashplus node: batched secp256k1 ECDSA verification, SHA-256d txid / Merkle root and
Equihash solution verification.

The product is the C-ABI shared library ``libbcp_b200.so`` (include/bcp_gpu.h) built by nvcc from
``csrc/``; the reference's host language is C++, and the C++ drop-in shims live in
``include/compat/``.  This Python package is only a thin ctypes face over the same C ABI, used by the
tests and bench.py.  There is no CPU fallback anywhere: every call raises ``BcpError`` when the
library or a CUDA device is missing.
"""
from ._lib import BcpError, lib, lib_path  # noqa: F401
from .api import (  # noqa: F401
    ECDSA_FLAG_NORMALIZE_S,
    Engine,
    block_index,
    der_lax_parse,
    device_count,
)

__all__ = ["BcpError", "Engine", "lib", "lib_path", "device_count", "block_index", "der_lax_parse",
           "ECDSA_FLAG_NORMALIZE_S"]
