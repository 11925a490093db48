"""CPU, world_size 2 over gloo: the multi-GPU plan of the hot path.  Signature batches shard by
index range across ranks with NO data-path collective; each rank produces verdict bytes for its range, the
block verdict is the AND over ranks (`fAllOk &= fOk`, src/checkqueue.h:77).  Here each rank stands in for a GPU
and verifies its shard with the oracle; the test checks the shard arithmetic (ragged sizes, empty shards), that
concatenated shard verdicts equal the unsharded verdicts, and that the only cross-rank step is the final
reduction bench.py also uses (max over ranks of the timing, AND of the verdict)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def shard_range(n, world, rank):
    per = (n + world - 1) // world
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import synth
    from oracle import pyoracle as po

    msg, pub, sig, kinds = synth.sigset(n, nkeys=16, corrupt_pct=10 if n > 20 else 0)
    lo, hi = shard_range(n, world, rank)
    ok = po.oracle().ecdsa_verify_batch(msg[lo:hi].copy(), pub[lo:hi].copy(), sig[lo:hi].copy(), 1) if hi > lo else np.zeros(0, np.uint8)
    # the block-level verdict: AND over ranks (a MIN all-reduce of one byte) — the only cross-rank step
    flag = torch.tensor([int(ok.all()) if len(ok) else 1], dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    # timing reduction used by bench.py: max over ranks
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, lo, hi, ok.tobytes(), int(flag.item()), float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1, 2, 7, 200])
def test_two_rank_sharding_matches_unsharded(n):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import synth
    from oracle import pyoracle as po

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + n) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    msg, pub, sig, _ = synth.sigset(n, nkeys=16, corrupt_pct=10 if n > 20 else 0)
    want = po.oracle().ecdsa_verify_batch(msg, pub, sig, 1)
    got = np.frombuffer(b"".join(r[3] for r in res), dtype=np.uint8)
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == n  # contiguous, disjoint, complete
    assert (got == want).all()
    assert all(r[4] == int(want.all()) for r in res)
    assert all(r[5] == 2.0 for r in res)


def test_shard_range_properties():
    for n in [0, 1, 2, 3, 7, 8, 9, 1000, 41650, 1000000]:
        for world in [1, 2, 4, 8]:
            cover = []
            for r in range(world):
                lo, hi = shard_range(n, world, r)
                assert 0 <= lo <= hi <= n
                cover.extend(range(lo, hi)) if n <= 1000 else None
                if r:
                    assert lo == shard_range(n, world, r - 1)[1]
            assert shard_range(n, world, world - 1)[1] == n
            if n <= 1000:
                assert cover == list(range(n))
