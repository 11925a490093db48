"""One call of each secondary path at a representative size, for an ncu launch list of the non-ECDSA kernels:
Equihash (200,9) x 10,000 headers, SHA-256d over 1,000,000 80-byte messages, a Merkle root over 1,000,000 leaves."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bitcoincashplus_b200 as bcp  # noqa: E402

eng = bcp.Engine(0)
hdrs = np.load(os.path.join(ROOT, "tests", "golden", "equihash_200_9.npz"))["headers"]
allh = np.tile(hdrs, ((10000 + len(hdrs) - 1) // len(hdrs), 1))[:10000].copy()
rng = np.random.default_rng(1)
for rep in range(2):
    ok, _ = eng.equihash_verify_batch(200, 9, allh)
    print("equihash valid", int(ok.sum()))
    data = rng.integers(0, 256, size=80 * 1000000, dtype=np.uint8)
    off = np.arange(0, 80 * 1000000 + 1, 80, dtype=np.uint32)
    d = eng.sha256d_batch(data, off)
    root, mut = eng.merkle_root(d)
    print("sha256d/merkle", root.hex()[:16], mut)
