"""Probe for the window call: W copies of one synthetic 8 MB block through bcp_gpu_blocks_connect_p2pkh (for ncu launch
lists and quick timing).  usage: python tools/window_probe.py [W] [reps]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bitcoincashplus_b200 as bcp  # noqa: E402
import synth  # noqa: E402

W = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
eng = bcp.Engine(0)
L = bcp.lib()
blk, spk, amt = synth.synthetic_block(41650, True, with_prevouts=True)


def pinned(a):
    p = L.bcp_gpu_host_alloc(max(a.nbytes, 1))
    out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(a.nbytes, 1),))[: a.nbytes].view(a.dtype)
    out[...] = a
    return out


n = len(spk)
off0 = np.zeros(n * W + 1, dtype=np.uint32)
off0[1:] = np.cumsum(np.array([len(x) for x in spk] * W, dtype=np.uint64))
blks = [pinned(np.frombuffer(blk, dtype=np.uint8)) for _ in range(W)]
cat, off, am = pinned(np.frombuffer(b"".join(spk) * W, dtype=np.uint8)), pinned(off0), pinned(np.asarray(list(amt) * W, dtype=np.int64))
lanes = int(sys.argv[3]) if len(sys.argv) > 3 else 1
if lanes > 1:
    import threading
    engs = [eng] + [bcp.Engine(0, lane=k) for k in range(1, lanes)]
    for e in engs:
        e.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)

    def work(e):
        for _ in range(reps):
            e.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)

    th = [threading.Thread(target=work, args=(e,)) for e in engs]
    t0 = time.perf_counter()
    [t.start() for t in th]
    [t.join() for t in th]
    dt = time.perf_counter() - t0
    print("W=%d lanes=%d: %.2f ms per window, %.0f blocks/s" % (W, lanes, dt * 1e3 / (reps * lanes), W * reps * lanes / dt))
    sys.exit(0)
for r in range(reps):
    t0 = time.perf_counter()
    res = eng.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)
    print("window %d: %.2f ms  all_ok=%s" % (r, (time.perf_counter() - t0) * 1e3, all(res["all_ok"])), flush=True)
