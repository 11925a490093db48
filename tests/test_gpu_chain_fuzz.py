"""-m gpu: randomized differential test of both block connectors (include/compat/bcp_chain.h: sharded UTXO table on the
host, worker team, two blocks in flight, scan-ahead, rollback; include/compat/bcp_chain_gpu.h: coin table in HBM, blocks
dealt out to check threads, several blocks in flight, rollback by kernel) against a small Python model of ConnectBlock's coin rules
(dict of coins; per transaction: every input's coin must exist — coins created earlier in the same block included —
and is spent; then the outputs are added) with the oracle deciding the script checks."""
import ctypes as C
import hashlib
import os
import random

import numpy as np
import pytest

import synth
from test_abi import _py_scan
from test_gpu_chain import MAGIC, chain_lib, new_chain

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


def model_connect(coins, blk, oracle):
    """returns (ok, reason-prefix) and updates `coins` only when ok"""
    tx_off, tx_out, ins, outs = _py_scan(blk, False)
    txids = [hashlib.sha256(hashlib.sha256(blk[tx_off[t]:tx_off[t + 1]]).digest()).digest() for t in range(len(tx_off) - 1)]
    n, otx, root, mutated = oracle.block_txids(blk, False)
    if root != blk[36:68]:
        return False, b"bad-txnmrklroot"
    work = dict(coins)
    spk, amt = [], []
    by_tx = {}
    for pos, t, i in ins:
        by_tx.setdefault(t, []).append(pos)
    outs_by_tx = {}
    for pos, t, v in outs:
        outs_by_tx.setdefault(t, []).append((pos, v))
    for t in range(len(txids)):
        for pos in by_tx.get(t, []):
            key = blk[pos:pos + 36]
            if key not in work:
                return False, b"bad-txns-inputs-missingorspent"
            s, a = work.pop(key)
            spk.append(s)
            amt.append(a)
        for pos, v in outs_by_tx.get(t, []):
            a = int.from_bytes(blk[pos:pos + 8], "little")
            sl = blk[pos + 8]
            assert sl < 253
            work[txids[t] + v.to_bytes(4, "little")] = (blk[pos + 9:pos + 9 + sl], a)
    if spk:
        err = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, NT)
        bad = np.nonzero(err)[0]
        if len(bad):
            return False, b"blk-bad-inputs (input %d, script error %d)" % (bad[0], err[bad[0]])
    coins.clear()
    coins.update(work)
    return True, b""


@pytest.mark.parametrize("impl", ["host", "device"])
@pytest.mark.parametrize("seed", range(12))
def test_random_chains_against_the_python_model(oracle, seed, impl):
    rng = random.Random(100 + seed)
    L = chain_lib(impl=impl)
    nblocks, ntx = 6, 26
    inblock = tuple(sorted(rng.sample(range(1, ntx), rng.randrange(0, 9))))
    kinds = [None, "bad_sig", "double_spend", "missing_input", "bad_merkle", "inblock_reversed", "inblock_double"]
    kind = rng.choice(kinds)
    if kind in ("inblock_reversed", "inblock_double"):
        inblock = ()  # those generators place their own in-block spends at fixed positions
    corrupt = (rng.randrange(1, nblocks), kind) if kind else None
    blocks, seeds, _ = synth.synthetic_chain(nblocks, ntx, corrupt=corrupt, inblock=inblock)
    # model
    coins = {txid + vout.to_bytes(4, "little"): (spk, amount) for txid, vout, spk, amount in seeds}
    want_n, want_reason = 0, b""
    for b in blocks:
        ok, why = model_connect(coins, b, oracle)
        if not ok:
            want_reason = why
            break
        want_n += 1
    assert (want_n == nblocks) == (kind is None), (kind, want_n, want_reason)
    # connector: whole file (two blocks in flight), several thread counts; then block by block
    data = b"".join(MAGIC + len(b).to_bytes(4, "little") + b for b in blocks)
    reason = C.create_string_buffer(256)
    for threads in (1, 2, 8):
        h = new_chain(L, seeds, walk_threads=threads)
        failed = C.c_long(-2)
        n = L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed))
        assert n == want_n and reason.value.startswith(want_reason), (kind, threads, n, want_n, reason.value, want_reason)
        assert failed.value == (-1 if want_n == nblocks else want_n)
        assert L.bcp_chain_utxo_size(h) == len(coins), (kind, threads)
        L.bcp_chain_free(h)
    h = new_chain(L, seeds, walk_threads=4)
    for i, b in enumerate(blocks):
        ok = L.bcp_chain_connect_block(h, b, len(b), reason, 256)
        assert ok == (1 if i < want_n else 0), (kind, i, reason.value)
        if not ok:
            assert reason.value.startswith(want_reason)
            break
    assert L.bcp_chain_utxo_size(h) == len(coins)
    L.bcp_chain_free(h)
