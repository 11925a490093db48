"""Deterministic synthetic inputs for the hot path, shared by tests/ and bench.py.

Signature sets come from the ORACLE's RFC 6979 signer (bit-identical to the reference's
secp256k1_ecdsa_sign — pinned by tests/test_oracle_pinning.py), so this module is test/bench
infrastructure and is never imported by the product package.
"""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pyoracle as po  # noqa: E402

SEED = b"bcp-b200-v1"
P = 2**256 - 2**32 - 977
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141

CORRUPTIONS = [
    "flip_r", "flip_s", "flip_msg", "swap_pubkey", "high_s", "r_zero", "s_zero", "r_eq_n", "s_eq_n_plus_1",
    "pub_x_ge_p", "pub_x_nonresidue", "prefix_05", "hybrid_ok", "hybrid_bad_parity", "msg_plus_n",
    "uncompressed_off_curve", "prefix_00", "r_eq_n_minus_1", "s_eq_half_n", "s_eq_half_n_plus_1",
]


def prng(i, tag=b"prng"):
    return int.from_bytes(hashlib.sha256(SEED + tag + i.to_bytes(8, "little")).digest(), "little")


def _be(x):
    return np.frombuffer(x.to_bytes(32, "big"), dtype=np.uint8)


def sigset(n, nkeys=4096, corrupt_pct=1, nthreads=None):
    """n signatures (msg32 (n,32), pub65 (n,65), sig64 (n,64)) with ~corrupt_pct % corrupted entries cycling
    through CORRUPTIONS.  Returns (msg, pub, sig, kinds) where kinds[i] is -1 or the corruption index."""
    o = po.oracle()
    msg, pub, sig = o.gen_sigset(n, nkeys=nkeys, seed=SEED, nthreads=nthreads)
    kinds = np.full(n, -1, dtype=np.int32)
    if corrupt_pct <= 0:
        return msg, pub, sig, kinds
    period = max(1, 100 // corrupt_pct)
    c = 0
    for i in range(n):
        if prng(i) % period != 0:
            continue
        kind = c % len(CORRUPTIONS)
        c += 1
        kinds[i] = kind
        apply_corruption(CORRUPTIONS[kind], i, msg, pub, sig, nkeys)
    return msg, pub, sig, kinds


def apply_corruption(name, i, msg, pub, sig, nkeys):
    bit = prng(i, b"bit")
    if name == "flip_r":
        sig[i, (bit >> 3) % 32] ^= 1 << (bit & 7)
    elif name == "flip_s":
        sig[i, 32 + (bit >> 3) % 32] ^= 1 << (bit & 7)
    elif name == "flip_msg":
        msg[i, (bit >> 3) % 32] ^= 1 << (bit & 7)
    elif name == "swap_pubkey":
        j = (i + 1) % len(pub)
        if (j % nkeys) == (i % nkeys):
            j = (i + 2) % len(pub)
        pub[i] = pub[j]
    elif name == "high_s":
        s = int.from_bytes(sig[i, 32:].tobytes(), "big")
        sig[i, 32:] = _be(N - s)
    elif name == "r_zero":
        sig[i, :32] = 0
    elif name == "s_zero":
        sig[i, 32:] = 0
    elif name == "r_eq_n":
        sig[i, :32] = _be(N)
    elif name == "s_eq_n_plus_1":
        sig[i, 32:] = _be(N + 1)
    elif name == "r_eq_n_minus_1":
        sig[i, :32] = _be(N - 1)
    elif name == "s_eq_half_n":
        sig[i, 32:] = _be(N // 2)
    elif name == "s_eq_half_n_plus_1":
        sig[i, 32:] = _be(N // 2 + 1)
    elif name == "pub_x_ge_p":
        pub[i, :] = 0
        pub[i, 0] = 2
        pub[i, 1:33] = _be(P + (bit % 900))
    elif name == "pub_x_nonresidue":
        x = bit % P
        while pow((x**3 + 7) % P, (P - 1) // 2, P) == 1:
            x += 1
        pub[i, :] = 0
        pub[i, 0] = 2 + (bit & 1)
        pub[i, 1:33] = _be(x)
    elif name == "prefix_05":
        pub[i, 0] = 5
    elif name == "prefix_00":
        pub[i, 0] = 0
    elif name in ("hybrid_ok", "hybrid_bad_parity", "uncompressed_off_curve"):
        o = po.oracle()
        ok, xy = o.pubkey_parse(pub[i, : 33 if pub[i, 0] in (2, 3) else 65].tobytes())
        assert ok
        y_odd = xy[63] & 1
        pub[i, 1:] = np.frombuffer(xy, dtype=np.uint8)
        if name == "hybrid_ok":
            pub[i, 0] = 7 if y_odd else 6
        elif name == "hybrid_bad_parity":
            pub[i, 0] = 6 if y_odd else 7
        else:
            pub[i, 0] = 4
            pub[i, 64] ^= 1
    elif name == "msg_plus_n":
        m = int.from_bytes(msg[i].tobytes(), "big")
        if m + N < 2**256:
            msg[i] = _be(m + N)  # reduced silently mod n: still VALID
    else:
        raise ValueError(name)


def edge_cases():
    """Hand-built adversarial cases beyond random corruption: tiny/huge scalars, r in the r+n<p band
    handled generically by differential testing.  Returns (msg, pub, sig)."""
    o = po.oracle()
    rows = []
    for sk in [1, 2, 3, N - 1, N - 2, (N + 1) // 2, 2**128, 2**255]:
        skb = sk.to_bytes(32, "big")
        for m in [0, 1, N - 1, N, N + 1, 2**256 - 1, 2**255]:
            mb = m.to_bytes(32, "big")
            s = o.sign(skb, mb)
            for comp in (True, False):
                pk = o.pubkey_create(skb, comp)
                rows.append((mb, pk, s))
    msg = np.zeros((len(rows), 32), np.uint8)
    pub = np.zeros((len(rows), 65), np.uint8)
    sig = np.zeros((len(rows), 64), np.uint8)
    for i, (m, pk, s) in enumerate(rows):
        msg[i] = np.frombuffer(m, np.uint8)
        pub[i, : len(pk)] = np.frombuffer(pk, np.uint8)
        sig[i] = np.frombuffer(s, np.uint8)
    return msg, pub, sig


# ------------------------------------------------------------------ synthetic blocks
def _compact(n):
    if n < 253:
        return bytes([n])
    if n <= 0xFFFF:
        return b"\xfd" + n.to_bytes(2, "little")
    if n <= 0xFFFFFFFF:
        return b"\xfe" + n.to_bytes(4, "little")
    return b"\xff" + n.to_bytes(8, "little")


def seckey(j):
    d = int.from_bytes(hashlib.sha256(SEED + b"key" + j.to_bytes(8, "little")).digest(), "big") % N
    return (d or 1).to_bytes(32, "big")


def p2pkh_script(pubkey):
    return b"\x76\xa9\x14" + po.oracle().hash160(pubkey) + b"\x88\xac"


def synthetic_header(height=600000, merkle_root=bytes(32), nonce_tag=0, soln=None, n=200, k=9):
    """Serialized post-fork header (src/primitives/block.h:49-73): 140 bytes || compactsize || solution."""
    soln_len = (1 << k) * (n // (k + 1) + 1) // 8
    if soln is None:
        soln = bytes(soln_len)
    h = (4).to_bytes(4, "little") + hashlib.sha256(SEED + b"prevblk").digest() + merkle_root
    h += height.to_bytes(4, "little") + bytes(28) + (1541030400).to_bytes(4, "little") + (0x1f07ffff).to_bytes(4, "little")
    h += hashlib.sha256(SEED + b"nonce" + nonce_tag.to_bytes(8, "little")).digest()
    assert len(h) == 140
    return h + _compact(len(soln)) + soln


POSTFORK_FLAGS = (1 << 0) | (1 << 1) | (1 << 2) | (1 << 3) | (1 << 9) | (1 << 10) | (1 << 14) | (1 << 16)

# input mutations understood by synthetic_block(mutations={index: kind}) and the ScriptError the reference reports
INPUT_MUTATIONS = {
    "bad_sig": 31, "der_padding": 23, "high_s": 26, "hashtype_01": 36, "hashtype_45": 22, "hashtype_c1": 0,
    "hashtype_42": 0, "hashtype_43": 0, "hashtype_c2": 0, "hashtype_c3": 0, "wrong_pubkey": 11, "hybrid_pubkey": 28,
    "uncompressed_pubkey": 0, "empty_sig": 2, "r_34_bytes": 31, "s_zero": 31, "der_trailing": 23, "der_short": 23,
    "three_pushes": None, "pushdata1": None,  # not the P2PKH shape: fast path answers 255, the interpreter decides
}


def _der(r, s):
    def enc(v):
        b = v.to_bytes(40, "big").lstrip(b"\x00") or b"\x00"
        if b[0] & 0x80:
            b = b"\x00" + b
        return b"\x02" + bytes([len(b)]) + b
    body = enc(r) + enc(s)
    return b"\x30" + bytes([len(body)]) + body


def synthetic_block(ntx=41650, sign=True, nkeys=4096, bad_inputs=(), with_prevouts=False, mutations=None):
    """Post-fork block of `ntx` 1-in/1-out P2PKH transactions (+ coinbase).  With sign=True every input
    carries a valid SIGHASH_ALL|FORKID signature over the FORKID sighash (inputs listed in bad_inputs get a
    flipped signature bit; `mutations` maps an input index to one of INPUT_MUTATIONS); otherwise scriptSigs are
    deterministic filler of the same shape.
    Returns the block bytes, or (block, prevout_spk list, prevout_amounts) when with_prevouts."""
    o = po.oracle()
    nk = min(nkeys, max(1, ntx))
    keys = [seckey(j) for j in range(nk)]
    pubs = [o.pubkey_create(kb, True) for kb in keys]
    spks = [p2pkh_script(pb) for pb in pubs]
    amount = 10000
    cb = (2).to_bytes(4, "little") + b"\x01" + bytes(32) + b"\xff\xff\xff\xff" + b"\x04" + (600000).to_bytes(4, "little")
    cb += b"\xff\xff\xff\xff" + b"\x01" + (1250000000).to_bytes(8, "little") + bytes([25]) + spks[0] + bytes(4)
    txs = [cb]
    prev_spk, prev_amt = [], []
    bad = set(bad_inputs)
    mutations = dict(mutations or {})
    for i in bad:
        mutations.setdefault(i, "bad_sig")
    for i in range(ntx):
        j = i % nk
        prev = hashlib.sha256(hashlib.sha256(SEED + b"prev" + i.to_bytes(8, "little")).digest()).digest()
        out = (amount - 500).to_bytes(8, "little") + bytes([25]) + spks[(i + 1) % nk]

        def build(script_sig):
            return ((2).to_bytes(4, "little") + b"\x01" + prev + bytes(4) + _compact(len(script_sig)) + script_sig
                    + b"\xff\xff\xff\xff" + b"\x01" + out + bytes(4))

        pub, spk = pubs[j], spks[j]
        if sign:
            kind = mutations.get(i)
            ht = {"hashtype_c1": 0xC1, "hashtype_42": 0x42, "hashtype_43": 0x43, "hashtype_c2": 0xC2,
                  "hashtype_c3": 0xC3}.get(kind, 0x41)
            if kind == "uncompressed_pubkey":
                pub = o.pubkey_create(keys[j], False)
                spk = p2pkh_script(pub)
            elif kind == "hybrid_pubkey":
                u = o.pubkey_create(keys[j], False)
                pub = bytes([6 + (u[64] & 1)]) + u[1:]
                spk = p2pkh_script(pub)
            sh = o.sighash_forkid(spk, build(b""), 0, ht, amount)
            s64 = o.sign(keys[j], sh)
            r_, s_ = int.from_bytes(s64[:32], "big"), int.from_bytes(s64[32:], "big")
            der = _der(r_, s_)
            if kind == "bad_sig":
                der = _der(r_ ^ (1 << 100), s_)
            elif kind == "der_padding":
                body = b"\x02" + bytes([len(der[4:4 + der[3]]) + 1]) + b"\x00" + der[4:4 + der[3]] + der[4 + der[3]:]
                der = b"\x30" + bytes([len(body)]) + body
            elif kind == "high_s":
                der = _der(r_, N - s_)
            elif kind == "r_34_bytes":
                der = _der((0x7F << 264) | r_, s_)
            elif kind == "s_zero":
                der = _der(r_, 0)
            elif kind == "der_trailing":
                der = der + b"\x00"
            elif kind == "der_short":
                der = der[:-1]
            htb = {"hashtype_01": 0x01, "hashtype_45": 0x45}.get(kind, ht)
            sigpush = der + bytes([htb])
            if kind == "wrong_pubkey":
                pub = pubs[(j + 1) % nk] if nk > 1 else o.pubkey_create(seckey(10**6), True)
            if kind == "empty_sig":
                script_sig = b"\x00" + bytes([len(pub)]) + pub
            elif kind == "three_pushes":
                script_sig = b"\x01\x07" + bytes([len(sigpush)]) + sigpush + bytes([len(pub)]) + pub
            elif kind == "pushdata1":
                script_sig = b"\x4c" + bytes([len(sigpush)]) + sigpush + bytes([len(pub)]) + pub
            else:
                script_sig = bytes([len(sigpush)]) + sigpush + bytes([len(pub)]) + pub
        else:
            filler = hashlib.sha256(SEED + b"fill" + i.to_bytes(8, "little")).digest()
            script_sig = bytes([72]) + (filler * 3)[:72] + bytes([33]) + pubs[j]
        txs.append(build(script_sig))
        prev_spk.append(spk)
        prev_amt.append(amount)
    body = _compact(len(txs)) + b"".join(txs)
    root, _ = o.merkle_root(b"".join(o.sha256d(t) for t in txs))
    blk = synthetic_header(merkle_root=root) + body
    if with_prevouts:
        return blk, prev_spk, prev_amt
    return blk


def synthetic_block_multi(ntx=60, nkeys=8, seed=1):
    """Post-fork block whose transactions have 1..5 inputs and 1..4 outputs, every input signed with a hashtype drawn
    from all six FORKID types (so SIGHASH_SINGLE with and without a matching output, ANYONECANPAY, NONE are all
    present).  Returns (block, prevout_spk list, prevout_amounts)."""
    import random

    o = po.oracle()
    rnd = random.Random(seed)
    keys = [seckey(1000 + j) for j in range(nkeys)]
    pubs = [o.pubkey_create(kb, j % 3 != 2) for j, kb in enumerate(keys)]  # every third key uncompressed
    spks = [p2pkh_script(pb) for pb in pubs]
    cb = (2).to_bytes(4, "little") + b"\x01" + bytes(32) + b"\xff\xff\xff\xff" + b"\x04" + (600001).to_bytes(4, "little")
    cb += b"\xff\xff\xff\xff" + b"\x01" + (1250000000).to_bytes(8, "little") + bytes([25]) + spks[0] + bytes(4)
    txs = [cb]
    prev_spk, prev_amt = [], []
    serial = 0
    for t in range(ntx):
        nin, nout = rnd.randint(1, 5), rnd.randint(1, 4)
        ins = []
        for i in range(nin):
            j = rnd.randrange(nkeys)
            prev = hashlib.sha256(SEED + b"mprev" + serial.to_bytes(8, "little")).digest()
            serial += 1
            ins.append((j, prev, rnd.randrange(4), rnd.randrange(1000, 10**9), rnd.choice([0xFFFFFFFF, 0xFFFFFFFE, 5])))
        outs = b"".join((rnd.randrange(1, 10**8)).to_bytes(8, "little") + bytes([25]) + spks[rnd.randrange(nkeys)] for _ in range(nout))
        locktime = rnd.choice([0, 0, 500000])

        def build(sigs):
            b = (2).to_bytes(4, "little") + _compact(nin)
            for (j, prev, vout, amt_, seq), ss in zip(ins, sigs):
                b += prev + vout.to_bytes(4, "little") + _compact(len(ss)) + ss + seq.to_bytes(4, "little")
            return b + _compact(nout) + outs + locktime.to_bytes(4, "little")

        unsigned = build([b""] * nin)
        sigs = []
        for i, (j, prev, vout, amt_, seq) in enumerate(ins):
            ht = rnd.choice([0x41, 0x42, 0x43, 0xC1, 0xC2, 0xC3])
            sh = o.sighash_forkid(spks[j], unsigned, i, ht, amt_)
            s64 = o.sign(keys[j], sh)
            der = _der(int.from_bytes(s64[:32], "big"), int.from_bytes(s64[32:], "big")) + bytes([ht])
            sigs.append(bytes([len(der)]) + der + bytes([len(pubs[j])]) + pubs[j])
            prev_spk.append(spks[j])
            prev_amt.append(amt_)
        txs.append(build(sigs))
    body = _compact(len(txs)) + b"".join(txs)
    root, _ = o.merkle_root(b"".join(o.sha256d(t) for t in txs))
    return synthetic_header(merkle_root=root) + body, prev_spk, prev_amt


# ------------------------------------------------------------------ tiny Equihash (48,5) solver (regtest parameters)
def equihash_solve_48_5(input140):
    """Wagner's algorithm for the regtest parameters (N=48, K=5: 512 leaves of 6 bytes, five collision rounds on one
    byte, the last on two).  Pure Python + hashlib; used to give synthetic chain headers VALID solutions on machines
    that have neither the reference nor its solver.  Returns a list of 36-byte minimal encodings."""
    pers = b"ZcashPoW" + (48).to_bytes(4, "little") + (5).to_bytes(4, "little")
    rows = []
    for g in range(52):  # 512 indices / 10 per hash output
        d = hashlib.blake2b(input140 + g.to_bytes(4, "little"), digest_size=60, person=pers).digest()
        for j in range(10):
            i = g * 10 + j
            if i < 512:
                rows.append((d[6 * j:6 * j + 6], (i,)))
    for rnd in range(5):
        width = 2 if rnd == 4 else 1
        buckets = {}
        for h, idx in rows:
            buckets.setdefault(h[:width], []).append((h, idx))
        nxt = []
        for grp in buckets.values():
            for a in range(len(grp)):
                for b in range(a + 1, len(grp)):
                    (ha, ia), (hb, ib) = grp[a], grp[b]
                    if set(ia) & set(ib):
                        continue
                    x = bytes(p ^ q for p, q in zip(ha[width:], hb[width:]))
                    nxt.append((x, ia + ib if ia[0] < ib[0] else ib + ia))
        rows = nxt
        if not rows:
            return []
    sols = []
    for h, idx in rows:
        if len(set(idx)) == 32:
            v = 0
            for x in idx:
                v = (v << 9) | x
            sols.append(v.to_bytes(36, "big"))
    return sorted(set(sols))


STRUCTURE_KINDS = {"second_coinbase": b"bad-tx-coinbase", "no_coinbase": b"bad-cb-missing", "cb_length": b"bad-cb-length",
                   "prevout_null": b"bad-txns-prevout-null", "vin_empty": b"bad-txns-vin-empty",
                   "vout_empty": b"bad-txns-vout-empty", "vout_negative": b"bad-txns-vout-negative",
                   "vout_toolarge": b"bad-txns-vout-toolarge"}


def synthetic_chain(nblocks=6, ntx=20, nkeys=8, equihash=True, corrupt=None, inblock=()):
    """A chain of post-fork blocks with REAL UTXO dependencies: block b's transaction j spends
    output 0 of block b-1's transaction j; block 0 spends seeded "prev" coins.  With equihash=True every header
    carries a valid regtest (48,5) solution.  corrupt = (block, kind) with kind in {"bad_sig", "double_spend",
    "bad_merkle", "missing_input", "bad_equihash", "inblock_reversed", "inblock_double"} or a key of STRUCTURE_KINDS
    (value = the reference's reject reason) damages exactly one block.
    inblock = positions j (>= 1) whose transaction spends output 0 of transaction j-1 of the SAME block in every block
    (the coin it would have spent stays unspent and is handed to the next block's transaction j-1 instead);
    "inblock_reversed" makes transaction 5 spend the output of transaction 6 (created AFTER it), "inblock_double" makes
    transactions 6 and 7 both spend the output of transaction 5.
    Returns (list of block bytes, list of seed coins (txid, vout, spk, amount), per-block (spk list, amounts) of the
    coins each block's inputs were built to spend)."""
    o = po.oracle()
    keys = [seckey(5000 + j) for j in range(nkeys)]
    pubs = [o.pubkey_create(kb, True) for kb in keys]
    spks = [p2pkh_script(pb) for pb in pubs]
    amount = 10**7
    seeds = []
    prev_out = []  # (txid, vout, key index, amount) spendable by tx j of the next block
    for j in range(ntx):
        txid = hashlib.sha256(hashlib.sha256(SEED + b"chainprev" + j.to_bytes(8, "little")).digest()).digest()
        seeds.append((txid, 0, spks[j % nkeys], amount))
        prev_out.append((txid, 0, j % nkeys, amount))
    blocks, prevs = [], []
    for b in range(nblocks):
        cb = (2).to_bytes(4, "little") + b"\x01" + bytes(32) + b"\xff\xff\xff\xff" + b"\x04" + (700000 + b).to_bytes(4, "little")
        cb += b"\xff\xff\xff\xff" + b"\x01" + (1250000000).to_bytes(8, "little") + bytes([25]) + spks[0] + bytes(4)
        txs = [cb]
        new_out = []
        kind = corrupt[1] if corrupt and corrupt[0] == b else None
        order = list(range(ntx))
        if kind == "inblock_reversed":
            order[5], order[6] = 6, 5  # build transaction 6 first so that 5 can refer to it
        built = {}
        for j in order:
            txid, vout, kj, amt_ = prev_out[j]
            if kind == "double_spend" and j == 1:
                txid, vout, kj, amt_ = prev_out[0]
            if kind == "missing_input" and j == 2:
                txid = hashlib.sha256(b"nowhere").digest()
            src = None
            if j in inblock and j >= 1:
                src = j - 1
            if kind == "inblock_reversed" and j == 5:
                src = 6
            if kind == "inblock_double" and j in (6, 7):
                src = 5
            if src is not None:
                txid, vout, kj, amt_ = built[src]
            nk = (kj + 1) % nkeys
            out = (amt_ - 1000).to_bytes(8, "little") + bytes([25]) + spks[nk]

            def build(ss):
                return ((2).to_bytes(4, "little") + b"\x01" + txid + vout.to_bytes(4, "little") + _compact(len(ss)) + ss
                        + b"\xff\xff\xff\xff" + b"\x01" + out + bytes(4))
            sh = o.sighash_forkid(spks[kj], build(b""), 0, 0x41, amt_)
            s64 = o.sign(keys[kj], sh)
            r_, s_ = int.from_bytes(s64[:32], "big"), int.from_bytes(s64[32:], "big")
            if kind == "bad_sig" and j == 3:
                r_ ^= 2
            if kind and kind.startswith("bad_sig@") and j == int(kind[8:]):
                r_ ^= 2
            der = _der(r_, s_) + b"\x41"
            tx = build(bytes([len(der)]) + der + bytes([33]) + pubs[kj])
            built[j] = (o.sha256d(tx), 0, nk, amt_ - 1000)
            built[("tx", j)] = tx
            built[("prev", j)] = (spks[kj], amt_)
            if src is not None and j in inblock:
                built[("keep", src)] = prev_out[j]  # the coin this transaction did not spend goes to the next block
        for j in range(ntx):
            txs.append(built[("tx", j)])
            new_out.append(built.get(("keep", j), built[j]))
        # structure damage (CheckBlock's context-free rules, src/validation.cpp:476-570, 3255-3330); signatures do not matter:
        # the reference rejects these in CheckBlock, before any script runs
        if kind in STRUCTURE_KINDS:
            t2 = txs[3]  # transaction j = 2: version(4) nin(1) outpoint(36) ...
            sslen = t2[41]
            in_end = 42 + sslen + 4
            if kind == "second_coinbase":
                txs.insert(2, cb[:41] + b"\x04" + (800000 + b).to_bytes(4, "little") + cb[46:])
            elif kind == "no_coinbase":
                txs = txs[1:]
            elif kind == "cb_length":
                txs[0] = cb[:41] + b"\x01\x00" + cb[46:]
            elif kind == "prevout_null":
                txs[3] = t2[:4] + b"\x02" + t2[5:in_end] + bytes(32) + b"\xff\xff\xff\xff" + b"\x00" + b"\xff\xff\xff\xff" + t2[in_end:]
            elif kind == "vin_empty":
                txs[3] = t2[:4] + b"\x00" + t2[in_end:]
            elif kind == "vout_empty":
                txs[3] = t2[:in_end] + b"\x00" + t2[-4:]
            elif kind == "vout_negative":
                txs[3] = t2[:in_end + 1] + (2**64 - 1).to_bytes(8, "little") + t2[in_end + 9:]
            elif kind == "vout_toolarge":
                txs[3] = t2[:in_end + 1] + (21000000 * 10**8 + 1).to_bytes(8, "little") + t2[in_end + 9:]
        prevs.append(([built[("prev", j)][0] for j in range(ntx)], [built[("prev", j)][1] for j in range(ntx)]))
        root, _ = o.merkle_root(b"".join(o.sha256d(t) for t in txs))
        if kind == "bad_merkle":
            root = bytes([root[0] ^ 1]) + root[1:]
        nonce = 0
        while True:
            hdr = synthetic_header(height=700000 + b, merkle_root=root, nonce_tag=(b << 20) | nonce, n=48, k=5)
            if not equihash:
                break
            sols = equihash_solve_48_5(hdr[:140])
            if sols:
                soln = sols[0]
                if kind == "bad_equihash":
                    soln = bytes([soln[0] ^ 0x80]) + soln[1:]
                hdr = hdr[:140] + _compact(36) + soln
                break
            nonce += 1
        blocks.append(hdr + _compact(len(txs)) + b"".join(txs))
        prev_out = new_out
    return blocks, seeds, prevs


def synthetic_chain_fast(nblocks=100, ntx=41650, nkeys=4096, equihash=True):
    """The chain of synthetic_chain() at full size, with the transactions built by the oracle's C generator
    (orc_gen_chain_txs): ~0.2 s per 8 MB block.  Yields blocks one at a time; seeds via chain_seeds()."""
    o = po.oracle()
    prev_txid, prev_key, prev_amt = chain_seed_arrays(ntx, nkeys)
    spk0 = p2pkh_script(o.pubkey_create(seckey(5000), True))
    for b in range(nblocks):
        body = o.gen_chain_txs(SEED, 5000, nkeys, prev_txid, prev_key, prev_amt)
        cb = (2).to_bytes(4, "little") + b"\x01" + bytes(32) + b"\xff\xff\xff\xff" + b"\x04" + (700000 + b).to_bytes(4, "little")
        cb += b"\xff\xff\xff\xff" + b"\x01" + (1250000000).to_bytes(8, "little") + bytes([25]) + spk0 + bytes(4)
        leaves = o.sha256d(cb) + prev_txid.tobytes()  # prev_txid now holds this block's txids
        root, _ = o.merkle_root(leaves)
        nonce = 0
        while True:
            hdr = synthetic_header(height=700000 + b, merkle_root=root, nonce_tag=(b << 20) | nonce, n=48, k=5)
            if not equihash:
                break
            sols = equihash_solve_48_5(hdr[:140])
            if sols:
                hdr = hdr[:140] + _compact(36) + sols[0]
                break
            nonce += 1
        yield hdr + _compact(ntx + 1) + cb + body


def damage_chain_block(blk, b, input_no):
    """blk = block number b of synthetic_chain_fast.  Returns a copy whose input `input_no` carries a broken signature (one
    bit of r flipped: strict DER still holds, the signature no longer verifies) under a header that is consistent again:
    Merkle root recomputed over the changed transactions, Equihash (48,5) solved anew."""
    o = po.oracle()
    n, off = o.block_index(blk, False)
    t = input_no + 1  # transaction 0 is the coinbase, every other transaction has one input
    body = bytearray(blk)
    body[off[t] + 4 + 1 + 36 + 1 + 1 + 12] ^= 0x04  # version | nin | outpoint | script length | push | 30 len 02 lenR R[8]
    body = bytes(body)
    _, _, root, _ = o.block_txids(body, False, max_tx=n + 8)
    txs = body[177:]
    nonce = 1 << 18
    while True:
        hdr = synthetic_header(height=700000 + b, merkle_root=root, nonce_tag=(b << 20) | nonce, n=48, k=5)
        sols = equihash_solve_48_5(hdr[:140])
        if sols:
            return hdr[:140] + _compact(36) + sols[0] + txs
        nonce += 1


def chain_seed_arrays(ntx, nkeys):
    prev_txid = np.zeros((ntx, 32), dtype=np.uint8)
    for j in range(ntx):
        prev_txid[j] = np.frombuffer(hashlib.sha256(hashlib.sha256(SEED + b"chainprev" + j.to_bytes(8, "little")).digest()).digest(), np.uint8)
    prev_key = (np.arange(ntx) % nkeys).astype(np.uint32)
    prev_amt = np.full(ntx, 10**7, dtype=np.int64)
    return prev_txid, prev_key, prev_amt


def chain_seeds(ntx, nkeys):
    """(txid, vout, spk, amount) of the coins block 0 of synthetic_chain_fast spends."""
    o = po.oracle()
    spks = [p2pkh_script(o.pubkey_create(seckey(5000 + j), True)) for j in range(nkeys)]
    t, k, a = chain_seed_arrays(ntx, nkeys)
    return [(t[j].tobytes(), 0, spks[int(k[j])], int(a[j])) for j in range(ntx)]


# ------------------------------------------------------------------ header chains (row N4)
def make_header(version, prev_hash, merkle_root, height, time, bits, nonce_tag, want_pow=None, limit=(1 << 255) - 1,
                corrupt_solution=False):
    """One serialized post-fork header with a VALID Equihash (48,5) solution; the nonce is ground until
    CheckProofOfWork(hash, bits) has the wanted outcome (None: any).  Returns (header bytes, hash)."""
    from oracle import pow_oracle as PO

    for attempt in range(4000):
        nonce = hashlib.sha256(SEED + b"hn" + nonce_tag.to_bytes(8, "little") + attempt.to_bytes(4, "little")).digest()
        h140 = (version & 0xFFFFFFFF).to_bytes(4, "little") + prev_hash + merkle_root + height.to_bytes(4, "little") + bytes(28) \
            + time.to_bytes(4, "little") + bits.to_bytes(4, "little") + nonce
        for sol in equihash_solve_48_5(h140):
            hdr = h140 + _compact(len(sol)) + sol
            hh = hashlib.sha256(hashlib.sha256(hdr).digest()).digest()
            if want_pow is None or PO.check_pow(hh, bits, limit) == want_pow:
                if corrupt_solution:
                    b = bytearray(hdr)
                    b[150] ^= 0x10
                    hdr = bytes(b)
                    hh = hashlib.sha256(hashlib.sha256(hdr).digest()).digest()
                return hdr, hh
    raise RuntimeError("no nonce found")
