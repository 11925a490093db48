"""-m gpu: the device-resident coin table (bcp_gpu_utxo_*, bitcoincashplus_b200/csrc/bcp_utxo.cu) through the C ABI
against a Python dict model of ConnectBlock's coin rules: per transaction, every input's coin must exist — coins created
by EARLIER transactions of the same block included — and is spent once; then the outputs become coins
(src/validation.cpp:2030-2100, UpdateCoins :1327-1347)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import synth
from test_abi import _py_scan

pytestmark = pytest.mark.gpu


class Table:
    def __init__(self, L, log2_slots, seeds, arena=1 << 20):
        self.L = L
        self.h = L.bcp_gpu_utxo_create(0, log2_slots, arena)
        assert self.h, L.bcp_gpu_last_error()
        if seeds:
            ops = np.frombuffer(b"".join(t + v.to_bytes(4, "little") for t, v, _, _ in seeds), np.uint8)
            cat = np.frombuffer(b"".join(s for _, _, s, _ in seeds), np.uint8)
            off = np.zeros(len(seeds) + 1, np.uint32)
            off[1:] = np.cumsum([len(s) for _, _, s, _ in seeds])
            amt = np.array([a for _, _, _, a in seeds], np.int64)
            assert L.bcp_gpu_utxo_add_coins(self.h, ops.ctypes.data, cat.ctypes.data, off.ctypes.data, amt.ctypes.data, len(seeds)) == 0

    def close(self):
        self.L.bcp_gpu_utxo_destroy(self.h)

    def size(self):
        return self.L.bcp_gpu_utxo_size(self.h)

    def apply(self, seq, blk):
        """applies + waits; returns dict(status, root, mutated, txids, spk list, amounts)"""
        tx_off, _, ins, outs = _py_scan(blk, False)
        ntx, nin, nout = len(tx_off) - 1, len(ins), len(outs)
        b = np.frombuffer(blk, np.uint8)
        a = lambda v: np.array(v if len(v) else [0], np.uint32)
        t_off, i_off, i_tx = a(tx_off), a([p for p, _, _ in ins]), a([t for _, t, _ in ins])
        o_off, o_tx, o_n = a([p for p, _, _ in outs]), a([t for _, t, _ in outs]), a([v for _, _, v in outs])
        txids = np.zeros((ntx, 32), np.uint8)
        root = np.zeros(32, np.uint8)
        mut = np.zeros(1, np.int32)
        cap = nin * 64 + 4096
        cat, off, amt = np.zeros(cap, np.uint8), np.zeros(nin + 1, np.uint32), np.zeros(max(nin, 1), np.int64)
        rc = self.L.bcp_gpu_utxo_apply_block(self.h, seq, b.ctypes.data, len(blk), t_off.ctypes.data, ntx, i_off.ctypes.data,
                                             i_tx.ctypes.data, nin, o_off.ctypes.data, o_tx.ctypes.data, o_n.ctypes.data, nout,
                                             txids.ctypes.data, root.ctypes.data, mut.ctypes.data, cat.ctypes.data, cap,
                                             off.ctypes.data, amt.ctypes.data)
        assert rc == 0, self.L.bcp_gpu_last_error()
        st = np.zeros(8, np.uint32)
        assert self.L.bcp_gpu_utxo_wait(self.h, seq, st.ctypes.data) == 0, self.L.bcp_gpu_last_error()
        spk = [cat[off[i]:off[i + 1]].tobytes() for i in range(nin)]
        return {"status": st, "root": root.tobytes(), "mutated": int(mut[0]), "txids": txids, "spk": spk, "amounts": amt[:nin]}


def model_apply(coins, blk):
    """the reference's rule on a dict; returns (missing input numbers, spk list, amounts, new dict, txids); the number of
    outputs that landed on an existing coin is left in model_apply.dups"""
    tx_off, _, ins, outs = _py_scan(blk, False)
    txids = [hashlib.sha256(hashlib.sha256(blk[tx_off[t]:tx_off[t + 1]]).digest()).digest() for t in range(len(tx_off) - 1)]
    work = dict(coins)
    missing, spk, amt = [], [], []
    by_tx, outs_by_tx = {}, {}
    for k, (pos, t, i) in enumerate(ins):
        by_tx.setdefault(t, []).append((k, pos))
    for pos, t, v in outs:
        outs_by_tx.setdefault(t, []).append((pos, v))
    res = {}
    model_apply.dups = 0
    for t in range(len(txids)):
        for k, pos in by_tx.get(t, []):
            key = blk[pos:pos + 36]
            if key in work:
                res[k] = work.pop(key)
            else:
                missing.append(k)
                res[k] = (b"", 0)
        for pos, v in outs_by_tx.get(t, []):
            sl = blk[pos + 8]
            model_apply.dups += (txids[t] + v.to_bytes(4, "little")) in work
            work[txids[t] + v.to_bytes(4, "little")] = (blk[pos + 9:pos + 9 + sl], int.from_bytes(blk[pos:pos + 8], "little"))
    for k in range(len(ins)):
        spk.append(res[k][0])
        amt.append(res[k][1])
    return missing, spk, amt, work, txids


def _lib():
    import bitcoincashplus_b200 as bcp

    bcp.Engine(0)
    return bcp.lib()


def test_apply_wait_confirm_rollback_against_the_dict_model(oracle):
    L = _lib()
    blocks, seeds, _ = synth.synthetic_chain(6, 40, inblock=(2, 3, 9, 30), equihash=False)
    coins = {t + v.to_bytes(4, "little"): (s, a) for t, v, s, a in seeds}
    T = Table(L, 12, seeds)
    assert T.size() == len(coins)
    seq = 0
    for b in blocks[:4]:
        seq += 1
        missing, spk, amt, work, txids = model_apply(coins, b)
        assert not missing
        r = T.apply(seq, b)
        assert list(r["status"][:3]) == [0, 0xFFFFFFFF, 0] and r["status"][4] == 0
        assert r["spk"] == spk and list(r["amounts"]) == amt
        assert r["root"] == b[36:68] and r["mutated"] == 0 and r["txids"].tobytes() == b"".join(txids)
        assert r["status"][3] == sum(len(s) for s in spk)
        coins = work
        assert T.size() == len(coins)
        if seq <= 2:
            assert L.bcp_gpu_utxo_confirm(T.h, seq) == 0
    # blocks 3 and 4 are still pending.  The same block again: every input is gone now (and its outputs exist already)
    missing, _, _, _, _ = model_apply(coins, blocks[3])
    r = T.apply(5, blocks[3])
    assert r["status"][0] == len(missing) == 36 and r["status"][1] == missing[0] and r["status"][2] == model_apply.dups == 37
    # rollback must go newest first
    assert L.bcp_gpu_utxo_rollback(T.h, 4) != 0 and b"rolled back first" in L.bcp_gpu_last_error()
    assert L.bcp_gpu_utxo_rollback(T.h, 5) == 0
    assert T.size() == len(coins)
    assert L.bcp_gpu_utxo_rollback(T.h, 4) == 0
    assert L.bcp_gpu_utxo_rollback(T.h, 3) == 0
    # the table is as after block 2: blocks 3.. connect again, with the same answers
    coins2 = {t + v.to_bytes(4, "little"): (s, a) for t, v, s, a in seeds}
    for b in blocks[:2]:
        coins2 = model_apply(coins2, b)[3]
    assert T.size() == len(coins2)
    for k, b in enumerate(blocks[2:]):
        missing, spk, amt, coins2, _ = model_apply(coins2, b)
        r = T.apply(6 + k, b)
        assert r["status"][0] == 0 and r["spk"] == spk and list(r["amounts"]) == amt
        assert L.bcp_gpu_utxo_confirm(T.h, 6 + k) == 0
    assert T.size() == len(coins2)
    T.close()


@pytest.mark.parametrize("kind", ["inblock_reversed", "inblock_double", "double_spend", "missing_input"])
def test_spend_rules_inside_one_block(kind):
    """An output may only be spent by a LATER transaction of its block and only once; a coin of an earlier block only
    once; an unknown outpoint never.  The missing inputs are exactly the ones the dict model refuses."""
    L = _lib()
    blocks, seeds, _ = synth.synthetic_chain(2, 20, corrupt=(1, kind), equihash=False)
    coins = {t + v.to_bytes(4, "little"): (s, a) for t, v, s, a in seeds}
    T = Table(L, 10, seeds)
    coins = model_apply(coins, blocks[0])[3]
    assert T.apply(1, blocks[0])["status"][0] == 0
    missing, spk, amt, _, _ = model_apply(coins, blocks[1])
    assert missing
    r = T.apply(2, blocks[1])
    assert r["status"][0] == len(missing), (kind, r["status"], missing)
    if kind in ("inblock_reversed", "missing_input"):  # with two spenders of one coin, WHICH of them loses is not defined
        assert r["status"][1] == missing[0]
        for i in range(len(spk)):
            if i not in missing:
                assert r["spk"][i] == spk[i] and r["amounts"][i] == amt[i]
            else:
                assert r["spk"][i] == b"" and r["amounts"][i] == 0
    assert L.bcp_gpu_utxo_rollback(T.h, 2) == 0
    assert T.size() == len(coins)
    T.close()


def test_tombstones_are_reused_once_their_block_is_final_and_long_scripts_round_trip():
    """1,024 slots, 70 blocks of 30 spends + 31 new coins each = 2,170 insertions: only possible because a spent coin's slot
    is handed to a later insertion once the spending block is confirmed — and never before (blocks applied while three
    are pending must not overwrite what a rollback would need).  One seeded coin carries a 300-byte script (the arena)."""
    L = _lib()
    blocks, seeds, _ = synth.synthetic_chain(70, 30, equihash=False)
    long_spk = bytes([0x51]) + bytes(range(256)) + bytes(43)
    seeds = list(seeds)
    seeds[7] = (seeds[7][0], seeds[7][1], long_spk, seeds[7][3])
    coins = {t + v.to_bytes(4, "little"): (s, a) for t, v, s, a in seeds}
    T = Table(L, 10, seeds)
    pending = []
    for k, b in enumerate(blocks):
        missing, spk, amt, coins, _ = model_apply(coins, b)
        r = T.apply(k + 1, b)
        assert r["status"][0] == 0 and r["status"][4] == 0, (k, r["status"])
        assert r["spk"] == spk and list(r["amounts"]) == amt
        if k == 0:
            assert r["spk"][7] == long_spk and r["status"][3] == sum(len(s) for s in spk)
        pending.append(k + 1)
        if len(pending) > 3:
            assert L.bcp_gpu_utxo_confirm(T.h, pending.pop(0)) == 0
    # undo the three pending blocks and redo them: identical answers (nothing a rollback needed was overwritten)
    for s in reversed(pending):
        assert L.bcp_gpu_utxo_rollback(T.h, s) == 0
    coins3 = {t + v.to_bytes(4, "little"): (s, a) for t, v, s, a in seeds}
    for b in blocks[:67]:
        coins3 = model_apply(coins3, b)[3]
    assert T.size() == len(coins3)
    for k in (67, 68, 69):
        missing, spk, amt, coins3, _ = model_apply(coins3, blocks[k])
        r = T.apply(100 + k, blocks[k])
        assert r["status"][0] == 0 and r["spk"] == spk and list(r["amounts"]) == amt
        assert L.bcp_gpu_utxo_confirm(T.h, 100 + k) == 0
    assert T.size() == len(coins3) == len(coins)
    T.close()


def test_table_full_and_argument_errors():
    L = _lib()
    blocks, seeds, _ = synth.synthetic_chain(1, 20, equihash=False)
    assert not L.bcp_gpu_utxo_create(0, 5, 0) and b"log2_slots" in L.bcp_gpu_last_error()
    T = Table(L, 10, seeds)
    # 2,000 more coins do not fit 1,024 slots: reported, not silently dropped
    many = [(hashlib.sha256(b"fill%d" % i).digest(), 0, b"\x51", 1) for i in range(2000)]
    ops = np.frombuffer(b"".join(t + v.to_bytes(4, "little") for t, v, _, _ in many), np.uint8)
    cat = np.frombuffer(b"".join(s for _, _, s, _ in many), np.uint8)
    off = np.arange(2001, dtype=np.uint32)
    amt = np.ones(2000, np.int64)
    assert L.bcp_gpu_utxo_add_coins(T.h, ops.ctypes.data, cat.ctypes.data, off.ctypes.data, amt.ctypes.data, 2000) != 0
    assert b"full" in L.bcp_gpu_last_error()
    st = np.zeros(8, np.uint32)
    assert L.bcp_gpu_utxo_wait(T.h, 9, st.ctypes.data) != 0 and L.bcp_gpu_utxo_confirm(T.h, 9) != 0 and L.bcp_gpu_utxo_rollback(T.h, 9) != 0
    T.close()
