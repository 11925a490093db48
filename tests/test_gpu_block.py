"""-m gpu: the whole-block fast path (txids + Merkle + P2PKH script checks + batched ECDSA) vs the oracle."""
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


def test_block_with_every_mutation(engine, oracle):
    kinds = list(synth.INPUT_MUTATIONS)
    mut = {5 + 7 * i: k for i, k in enumerate(kinds)}
    blk, spk, amt = synth.synthetic_block(400, True, nkeys=32, with_prevouts=True, mutations=mut)
    res = engine.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS, want_txids=True)
    want = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, NT)
    assert (res["script_err"] == want).all(), np.nonzero(res["script_err"] != want)[0]
    for i, k in mut.items():
        exp = synth.INPUT_MUTATIONS[k]
        assert res["script_err"][i] == (255 if exp is None else exp), k
    assert not res["all_ok"]
    n, txids, root, mutated = oracle.block_txids(blk, False)
    assert res["root"] == root == blk[36:68] and res["mutated"] == mutated
    assert (res["txids"][:n] == txids).all()


def test_all_valid_block_and_single_bad_input(engine, oracle):
    blk, spk, amt = synth.synthetic_block(3000, True, nkeys=64, with_prevouts=True)
    res = engine.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS)
    assert res["all_ok"] and not res["script_err"].any()
    blk2, spk2, amt2 = synth.synthetic_block(3000, True, nkeys=64, with_prevouts=True, bad_inputs=(1777,))
    res2 = engine.block_connect_p2pkh(blk2, spk2, amt2, synth.POSTFORK_FLAGS)
    assert not res2["all_ok"]
    assert list(np.nonzero(res2["script_err"])[0]) == [1777] and res2["script_err"][1777] == 31
    # wrong amount in the UTXO side table changes the sighash: NULLFAIL exactly there
    amt3 = list(amt)
    amt3[5] += 1
    res3 = engine.block_connect_p2pkh(blk, spk, amt3, synth.POSTFORK_FLAGS)
    assert list(np.nonzero(res3["script_err"])[0]) == [5] and res3["script_err"][5] == 31
    want = oracle.block_script_errors(blk, False, spk, amt3, synth.POSTFORK_FLAGS, NT)
    assert (res3["script_err"] == want).all()


def test_flag_sets_outside_the_fast_path_are_refused(engine):
    import bitcoincashplus_b200 as bcp

    blk, spk, amt = synth.synthetic_block(3, True, nkeys=2, with_prevouts=True)
    for flags in (0, synth.POSTFORK_FLAGS & ~(1 << 14), synth.POSTFORK_FLAGS | (1 << 17)):
        with pytest.raises(bcp.BcpError):
            engine.block_connect_p2pkh(blk, spk, amt, flags)
    with pytest.raises(bcp.BcpError):
        engine.block_connect_p2pkh(blk, spk[:-1], amt[:-1], synth.POSTFORK_FLAGS)


def test_empty_block_only_coinbase(engine, oracle):
    blk = synth.synthetic_block(0, True)
    res = engine.block_connect_p2pkh(blk, [], [], synth.POSTFORK_FLAGS)
    n, txids, root, mutated = oracle.block_txids(blk, False)
    assert n == 1 and res["root"] == root == txids[0].tobytes() and res["all_ok"]


def test_multi_input_output_transactions_all_hashtypes(engine, oracle):
    """1..5 inputs, 1..4 outputs per transaction, all six FORKID hashtypes, compressed and uncompressed keys."""
    blk, spk, amt = synth.synthetic_block_multi(500, seed=21)
    res = engine.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS, want_txids=True)
    assert res["all_ok"] and not res["script_err"].any()
    n, txids, root, mutated = oracle.block_txids(blk, False)
    assert res["root"] == root == blk[36:68] and (res["txids"][:n] == txids).all()
    amt2 = list(amt)
    bad = [0, 17, len(amt) - 1]
    for i in bad:
        amt2[i] ^= 1
    res2 = engine.block_connect_p2pkh(blk, spk, amt2, synth.POSTFORK_FLAGS)
    want = oracle.block_script_errors(blk, False, spk, amt2, synth.POSTFORK_FLAGS, NT)
    assert (res2["script_err"] == want).all() and list(np.nonzero(want)[0]) == bad


def test_block_index_inputs_matches_oracle(engine, oracle):
    import ctypes as C

    import bitcoincashplus_b200 as bcp

    blk, spk, amt = synth.synthetic_block_multi(80, seed=5)
    L = bcp.lib()
    b = np.frombuffer(blk, dtype=np.uint8)
    tx_off = np.zeros(300, np.uint32)
    arrs = [np.zeros(len(spk), np.uint32) for _ in range(3)]
    ntx = C.c_long(0)
    nin = L.bcp_block_index_inputs(b.ctypes.data, len(blk), 0, tx_off.ctypes.data, 299, arrs[0].ctypes.data, arrs[1].ctypes.data,
                                   arrs[2].ctypes.data, len(spk), C.byref(ntx))
    assert nin == len(spk) and ntx.value == 81
    n, ooff = oracle.block_index(blk, False)
    assert (tx_off[: n + 1] == ooff).all()
    assert (arrs[1][:-1] <= arrs[1][1:]).all() and arrs[1].min() == 1  # coinbase (tx 0) skipped, block order
    # capacity error
    assert L.bcp_block_index_inputs(b.ctypes.data, len(blk), 0, tx_off.ctypes.data, 299, arrs[0].ctypes.data, arrs[1].ctypes.data,
                                    arrs[2].ctypes.data, 3, C.byref(ntx)) == -2


def _window(blocks):
    """marshals [(blk, spk list, amounts)] for the window call"""
    cat = np.frombuffer(b"".join(b"".join(s) for _, s, _ in blocks) or b"\x00", dtype=np.uint8)
    lens = [len(x) for _, s, _ in blocks for x in s]
    off = np.zeros(len(lens) + 1, dtype=np.uint32)
    off[1:] = np.cumsum(lens)
    am = np.array([a for _, _, aa in blocks for a in aa], dtype=np.int64)
    return [np.frombuffer(b, dtype=np.uint8) for b, _, _ in blocks], cat, off, am, [len(s) for _, s, _ in blocks]


def test_window_of_blocks_matches_per_block_calls(engine, oracle):
    """bcp_gpu_blocks_connect_p2pkh over a window of different blocks (mutated inputs, multi-input transactions, an
    empty block, a block with a broken Merkle root, a duplicated last transaction) == the single-block call per block
    == the oracle."""
    kinds = list(synth.INPUT_MUTATIONS)
    b0 = synth.synthetic_block(300, True, nkeys=32, with_prevouts=True, mutations={3 + 5 * i: k for i, k in enumerate(kinds)})
    b1 = synth.synthetic_block_multi(200, seed=5)
    b2 = (synth.synthetic_block(0, True), [], [])
    b3 = synth.synthetic_block(1000, True, nkeys=64, with_prevouts=True, bad_inputs=(999,))
    blk4, spk4, amt4 = synth.synthetic_block(64, True, nkeys=8, with_prevouts=True)
    blk4 = bytearray(blk4)
    blk4[40] ^= 1  # header Merkle root no longer matches: the call reports the real root, the caller compares
    b4 = (bytes(blk4), spk4, amt4)
    b5 = synth.synthetic_block(777, True, nkeys=16, with_prevouts=True)
    blocks = [b0, b1, b2, b3, b4, b5]
    arrs, cat, off, am, nin = _window(blocks)
    res = engine.blocks_connect_p2pkh(arrs, cat, off, am, nin, synth.POSTFORK_FLAGS, want_txids=True)
    pos = tpos = 0
    for b, (blk, spk, amt) in enumerate(blocks):
        one = engine.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS, want_txids=True)
        n, txids, root, mutated = oracle.block_txids(blk, False)
        assert res["ntx"][b] == n
        assert bytes(res["roots"][b]) == one["root"] == root, b
        assert res["mutated"][b] == one["mutated"] == mutated
        assert (res["txids"][tpos:tpos + n] == txids).all()
        err = res["script_err"][pos:pos + len(spk)]
        assert (err == one["script_err"]).all(), b
        if spk:
            want = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, NT)
            assert (err == want).all(), b
        assert res["all_ok"][b] == one["all_ok"] == (not err.any())
        pos += len(spk)
        tpos += n
    assert res["all_ok"] == [False, True, True, False, True, True]


def test_window_rejects_malformed_block_and_wrong_counts(engine):
    import bitcoincashplus_b200 as bcp

    good = synth.synthetic_block(50, True, nkeys=4, with_prevouts=True)
    arrs, cat, off, am, nin = _window([good, good])
    bad = arrs[1].copy()[:-7]  # truncated
    with pytest.raises(bcp.BcpError):
        engine.blocks_connect_p2pkh([arrs[0], bad], cat, off, am, nin, synth.POSTFORK_FLAGS)
    with pytest.raises(bcp.BcpError):  # input counts swapped between blocks of different size
        other = synth.synthetic_block(20, True, nkeys=4, with_prevouts=True)
        a2, c2, o2, m2, n2 = _window([good, other])
        engine.blocks_connect_p2pkh(a2, c2, o2, m2, n2[::-1], synth.POSTFORK_FLAGS)


def test_two_lanes_of_one_gpu_run_concurrently_and_agree(engine, oracle):
    """BCP_DEVICE_LANE: a second context on the same GPU; both lanes validate different blocks from two host threads at
    the same time and each returns what the oracle says."""
    import threading

    import bitcoincashplus_b200 as bcp

    lane1 = bcp.Engine(0, lane=1)
    blocks = [synth.synthetic_block(600 + 50 * i, True, nkeys=16, with_prevouts=True, bad_inputs=(17 * i + 3,)) for i in range(4)]
    want = [oracle.block_script_errors(b, False, s, a, synth.POSTFORK_FLAGS, NT) for b, s, a in blocks]
    got = {}

    def run(e, ids):
        for _ in range(5):
            for i in ids:
                b, s, a = blocks[i]
                got[(e.device, i)] = e.block_connect_p2pkh(b, s, a, synth.POSTFORK_FLAGS)

    th = [threading.Thread(target=run, args=(engine, [0, 1, 2, 3])), threading.Thread(target=run, args=(lane1, [3, 2, 1, 0]))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for (dev, i), r in got.items():
        assert (r["script_err"] == want[i]).all() and r["root"] == blocks[i][0][36:68], (dev, i)
    assert len(got) == 8
    with pytest.raises(bcp.BcpError):
        bcp.Engine(0, lane=9)


def test_huge_transaction_takes_the_global_memory_path(engine, oracle):
    """The per-transaction kernels parse from a shared-memory copy of their thread block's span of the block; a span that
    does not fit (here: a transaction with a 150,000-byte output script in the middle of ordinary ones) must fall back
    to global memory for that thread block only — same txids, Merkle root, sighashes and verdicts as the oracle."""
    import hashlib

    o = oracle
    nkeys = 8
    keys = [synth.seckey(9000 + j) for j in range(nkeys)]
    pubs = [o.pubkey_create(kb, True) for kb in keys]
    spks = [synth.p2pkh_script(pb) for pb in pubs]
    big_script = b"\x6a" + bytes(range(256)) * 586  # OP_RETURN + 150,016 bytes
    cb = (2).to_bytes(4, "little") + b"\x01" + bytes(32) + b"\xff\xff\xff\xff" + b"\x04" + (700123).to_bytes(4, "little")
    cb += b"\xff\xff\xff\xff" + b"\x01" + (1250000000).to_bytes(8, "little") + bytes([25]) + spks[0] + bytes(4)
    txs, spent_spk, spent_amt = [cb], [], []
    ntx = 300
    for j in range(ntx):
        kj = j % nkeys
        prev = hashlib.sha256(b"huge-tx-test" + j.to_bytes(4, "little")).digest()
        amount = 10**7 + j
        outs = (amount - 1000).to_bytes(8, "little") + bytes([25]) + spks[(kj + 1) % nkeys]
        nout = 1
        if j == 137:  # the big one: a second output carrying the long script
            outs += (0).to_bytes(8, "little") + synth._compact(len(big_script)) + big_script
            nout = 2

        def build(ss):
            return ((2).to_bytes(4, "little") + b"\x01" + prev + (0).to_bytes(4, "little") + synth._compact(len(ss)) + ss
                    + b"\xff\xff\xff\xff" + bytes([nout]) + outs + bytes(4))
        sh = o.sighash_forkid(spks[kj], build(b""), 0, 0x41, amount)
        s64 = o.sign(keys[kj], sh)
        der = synth._der(int.from_bytes(s64[:32], "big"), int.from_bytes(s64[32:], "big")) + b"\x41"
        txs.append(build(bytes([len(der)]) + der + bytes([33]) + pubs[kj]))
        spent_spk.append(spks[kj])
        spent_amt.append(amount)
    root, _ = o.merkle_root(b"".join(o.sha256d(t) for t in txs))
    blk = synth.synthetic_header(height=700123, merkle_root=root, n=48, k=5) + synth._compact(len(txs)) + b"".join(txs)
    want = o.block_script_errors(blk, False, spent_spk, spent_amt, synth.POSTFORK_FLAGS, NT)
    assert not want.any()
    res = engine.block_connect_p2pkh(blk, spent_spk, spent_amt, synth.POSTFORK_FLAGS, want_txids=True)
    n, txids, root2, mutated = o.block_txids(blk, False)
    assert res["root"] == root == root2 and (res["txids"][:n] == txids).all()
    assert (res["script_err"] == want).all() and res["all_ok"]
    # the same block inside a window, and with a wrong amount for the big transaction's input
    arrs, cat, off, am, nin = _window([(blk, spent_spk, spent_amt), synth.synthetic_block(200, True, nkeys=8, with_prevouts=True)])
    w = engine.blocks_connect_p2pkh(arrs, cat, off, am, nin, synth.POSTFORK_FLAGS)
    assert w["all_ok"] == [True, True] and bytes(w["roots"][0]) == root
    bad_amt = list(spent_amt)
    bad_amt[137] += 1
    res2 = engine.block_connect_p2pkh(blk, spent_spk, bad_amt, synth.POSTFORK_FLAGS)
    assert list(np.nonzero(res2["script_err"])[0]) == [137] and res2["script_err"][137] == 31
