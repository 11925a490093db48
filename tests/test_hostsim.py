"""CPU: the DEVICE arithmetic headers (bitcoincashplus_b200/csrc/*.cuh), compiled for the host with
portable stand-ins for the PTX limb primitives (tests/hostsim), against Python integers and the oracle.
This is test infrastructure: it lets the GPU-less CI exercise the same field/scalar/group code, GLV split,
window recoding and verdict logic that the sm_100a kernels run.  The product never executes on the host."""
import ctypes as C
import os
import random
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
HS = os.path.join(ROOT, "tests", "hostsim")
u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
P = 2**256 - 2**32 - 977
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141


def build_hostsim():
    so = os.path.join(HS, "libhostsim.so")
    src = os.path.join(HS, "hostsim.cpp")
    hdrs = [os.path.join(ROOT, "bitcoincashplus_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "bitcoincashplus_b200", "csrc")) if f.endswith(".cuh")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(p) for p in hdrs + [src]):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-DBCP_GCOMB_BITS=12", "-fPIC", "-shared", "-o", so, src, "-lpthread"])
    return C.CDLL(so)


@pytest.fixture(scope="module")
def hs():
    return build_hostsim()


@pytest.fixture(scope="module")
def gcomb(hs):
    g = np.zeros(22 * 4096 * 16, dtype=np.uint32)  # the host build is compiled with -DBCP_GCOMB_BITS=12
    hs.hs_gcomb_build(g.ctypes.data_as(u32p))
    return g


def b32(x):
    return C.cast(C.c_char_p(x.to_bytes(32, "big")), u8p)


def op3(f, op, a, b):
    out = (C.c_uint8 * 32)()
    f(op, b32(a), b32(b), out)
    return int.from_bytes(bytes(out), "big")


EDGE = [0, 1, 2, P - 1, P, P + 1, 2**256 - 1, 2**256 - 2, 2**255, 2**32 + 977, 2**32 + 976, 2**256 - 2**32 - 978,
        2**224, 977, (1 << 256) - (1 << 32)]


def test_field_ops_on_weakly_reduced_inputs(hs):
    rnd = random.Random(7)
    vals = EDGE + [rnd.randrange(2**256) for _ in range(120)]
    for a in vals[:48]:
        for b in vals[:48]:
            assert op3(hs.hs_fe_op, 0, a, b) == a * b % P
            assert op3(hs.hs_fe_op, 1, a, b) == (a + b) % P
            assert op3(hs.hs_fe_op, 2, a, b) == (a - b) % P
    for a in vals:
        assert op3(hs.hs_fe_op, 3, a, 0) == a * a % P
        assert op3(hs.hs_fe_op, 4, a, 0) == (-a) % P
        if a % P:
            assert op3(hs.hs_fe_op, 5, a, 0) == pow(a, -1, P)
        assert op3(hs.hs_fe_op, 6, a, 0) == pow(a, (P + 1) // 4, P)
        assert bool(hs.hs_fe_is_zero(b32(a))) == (a % P == 0)


def test_short_cut_folds_on_crafted_operands(hs):
    """fe_add / fe_sub / fe_reduce_wide finish the folds that practically never propagate out of line.  Random inputs
    do not reach that code: these operands do (the same set the device self-test uses, tools/gen_fe_selftest.py), and
    the short-cut forms must equal both the always-full-chain forms and Python's integers."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("gen_fe_selftest", os.path.join(ROOT, "tools", "gen_fe_selftest.py"))
    g = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(g)
    for a in g.edge:
        for b in g.edge:
            for op, full, want in ((1, 7, (a + b) % P), (2, 8, (a - b) % P)):
                got, ref = op3(hs.hs_fe_op, op, a, b), op3(hs.hs_fe_op, full, a, b)
                assert got % P == want and ref % P == want and got == ref, (op, hex(a), hex(b))
            assert op3(hs.hs_fe_op, 0, a, b) % P == a * b % P
    hit = 0
    for w in g.wide:
        out1, out0 = (C.c_uint8 * 32)(), (C.c_uint8 * 32)()
        t = C.cast(C.c_char_p(w.to_bytes(64, "big")), u8p)
        hs.hs_fe_reduce(t, 1, out1)
        hs.hs_fe_reduce(t, 0, out0)
        v1, v0 = int.from_bytes(bytes(out1), "big"), int.from_bytes(bytes(out0), "big")
        assert v1 == v0 and v1 % P == w % P, hex(w)
        # does this input really need the slow path?  S = L + H*(2^32+977); T = S >> 256; low 96 bits + T*(2^32+977) carries
        S = (w % 2**256) + (w >> 256) * (2**32 + 977)
        hit += ((S % 2**96) + (S >> 256) * (2**32 + 977)) >= 2**96
    assert hit >= 8


def test_quick_forms_count_every_unfinished_fold(hs):
    """The point kernel's hot loop uses branch-free field forms that COUNT the practically-never carry instead of
    finishing it (fe.cuh "quick forms"); the kernel redoes a signature with the exact forms when the count is not zero.
    So: whenever the count is zero the value must be right, and the crafted operands must raise it where the exact
    forms take their slow path."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("gen_fe_selftest", os.path.join(ROOT, "tools", "gen_fe_selftest.py"))
    g = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(g)
    hs.hs_fe_op_q.restype = C.c_uint32
    hs.hs_fe_reduce_q.restype = C.c_uint32
    rnd = random.Random(11)
    vals = list(g.edge) + EDGE + [rnd.randrange(2**256) for _ in range(60)]
    flagged = {0: 0, 1: 0, 2: 0, 3: 0, 4: 0, 9: 0}

    def q(op, a, b):
        out = (C.c_uint8 * 32)()
        bad = hs.hs_fe_op_q(op, b32(a), b32(b), out)
        return bad, int.from_bytes(bytes(out), "big")

    for a in vals:
        for b in vals:
            for op, want in ((0, a * b % P), (1, (a + b) % P), (2, (a - b) % P)):
                bad, got = q(op, a, b)
                flagged[op] += bad != 0
                assert bad or got == want, (op, hex(a), hex(b))
        for op, want in ((3, a * a % P), (4, (-a) % P), (9, 8 * a % P)):
            bad, got = q(op, a, 0)
            flagged[op] += bad != 0
            assert bad or got == want, (op, hex(a))
    # additions and subtractions of the crafted operands do leave limb 1; -a borrows for a > p
    assert flagged[1] > 0 and flagged[2] > 0 and flagged[4] > 0, flagged
    hit = 0
    for w in g.wide:
        out = (C.c_uint8 * 32)()
        bad = hs.hs_fe_reduce_q(C.cast(C.c_char_p(w.to_bytes(64, "big")), u8p), out)
        S = (w % 2**256) + (w >> 256) * (2**32 + 977)
        needs = ((S % 2**96) + (S >> 256) * (2**32 + 977)) >= 2**96
        assert bool(bad) == needs, hex(w)
        assert bad or int.from_bytes(bytes(out), "big") == w % P
        hit += needs
    assert hit >= 8


def test_wide_products_schoolbook_and_karatsuba(hs):
    rnd = random.Random(17)
    vals = EDGE + [2**128 - 1, 2**128, 2**128 + 1, (2**128 - 1) << 128, 2**256 - 2**128] + [rnd.randrange(2**256) for _ in range(200)]
    for a in vals[:40]:
        for b in vals[:40]:
            out = (C.c_uint8 * 64)()
            hs.hs_mul_wide_karatsuba(b32(a), b32(b), out)
            assert int.from_bytes(bytes(out), "big") == a * b, (hex(a), hex(b))


def test_scalar_ops_and_reduction(hs):
    from secp_consts import lam

    rnd = random.Random(8)
    svals = [0, 1, 2, N - 1, N - 2, N // 2, N // 2 + 1, lam, N - lam, 2**128, 2**129 - 1] + [rnd.randrange(N) for _ in range(60)]
    for a in svals[:30]:
        for b in svals[:30]:
            assert op3(hs.hs_sc_op, 0, a, b) == a * b % N
            assert op3(hs.hs_sc_op, 1, a, b) == (a + b) % N
            assert op3(hs.hs_sc_op, 2, a, b) == (a - b) % N
    for a in svals:
        assert op3(hs.hs_sc_op, 4, a, 0) == (-a) % N
        if a:
            assert op3(hs.hs_sc_op, 3, a, 0) == pow(a, -1, N)
            assert op3(hs.hs_sc_op, 5, a, 0) == pow(a, -1, N)  # the division-step inversion the prep kernel uses
    # the division-step inversion on many more values, and on the shapes that stress its 30-bit limbs and its sign handling
    more = [2**k for k in range(0, 256, 7)] + [N - 2**k for k in range(0, 256, 11)] + [2**30 - 1, 2**30, 2**60 - 1, 2**240, 2**255, (N + 1) // 2,
            (N - 1) // 3, 0x3FFFFFFF << 30, int("3fffffff" * 8, 16) % N] + [rnd.randrange(1, N) for _ in range(3000)]
    for a in more:
        assert op3(hs.hs_sc_op, 5, a, 0) == pow(a, -1, N), hex(a)
    assert op3(hs.hs_sc_op, 5, 0, 0) == 0  # like the reference's scalar_inverse
    for t in [0, 1, 2**512 - 1, 2**256, 2**256 - 1, N * N, (N - 1) ** 2, 2**511] + [rnd.randrange(2**512) for _ in range(300)]:
        out = (C.c_uint8 * 32)()
        hs.hs_sc_reduce512(C.cast(C.c_char_p(t.to_bytes(64, "big")), u8p), out)
        assert int.from_bytes(bytes(out), "big") == t % N


def test_lambda_split_and_window_recoding(hs):
    from secp_consts import lam

    rnd = random.Random(9)
    mx = 0
    for k in [0, 1, N - 1, N // 2, lam, N - lam, 2**128, 2**255] + [rnd.randrange(N) for _ in range(5000)]:
        k1, k2, n1, n2 = (C.c_uint32 * 5)(), (C.c_uint32 * 5)(), C.c_int(), C.c_int()
        hs.hs_split(b32(k), k1, C.byref(n1), k2, C.byref(n2))
        a1 = sum(k1[i] << (32 * i) for i in range(5))
        a2 = sum(k2[i] << (32 * i) for i in range(5))
        assert ((-a1 if n1.value else a1) + (-a2 if n2.value else a2) * lam - k) % N == 0
        mx = max(mx, a1.bit_length(), a2.bit_length())
    assert mx <= 129
    for a in [0, 1, 7, 8, 9, 2**132 - 1, 2**131, int("8" * 33, 16), int("7" * 33, 16)] + [rnd.randrange(2**132) for _ in range(2000)]:
        dg = (C.c_int * 34)()
        arr = (C.c_uint32 * 5)(*[(a >> (32 * i)) & 0xFFFFFFFF for i in range(5)])
        hs.hs_recode(arr, dg)
        assert sum(dg[w] * 16**w for w in range(34)) == a and all(-8 <= dg[w] <= 8 for w in range(34))


def test_gcomb_entries_are_multiples_of_g(hs, gcomb):
    from secp_consts import Gx, Gy, ec_mul

    for (j, d) in [(0, 1), (0, 2), (0, 4095), (1, 1), (5, 77), (21, 15), (21, 1), (10, 2048), (13, 4095)]:
        out = (C.c_uint32 * 16)()
        hs.hs_gcomb_entry(j, d, out)
        x = sum(out[i] << (32 * i) for i in range(8))
        y = sum(out[8 + i] << (32 * i) for i in range(8))
        assert (x, y) == ec_mul(d << (12 * j), (Gx, Gy))
        assert list(out) == list(gcomb[(j * 4096 + d) * 16:(j * 4096 + d) * 16 + 16])


def _run(hs, gcomb, msg, pub, sig, flags=0):
    n = len(msg)
    ok = np.zeros(n, dtype=np.uint8)
    hs.hs_ecdsa_verify_batch(msg.ctypes.data_as(u8p), pub.ctypes.data_as(u8p), sig.ctypes.data_as(u8p), C.c_size_t(n),
                             C.c_uint32(flags), gcomb.ctypes.data_as(u32p), os.cpu_count() or 1, ok.ctypes.data_as(u8p))
    return ok


def test_verify_pipeline_matches_oracle(hs, gcomb, oracle):
    import synth

    msg, pub, sig, kinds = synth.sigset(3000, nkeys=64, corrupt_pct=20)
    got = _run(hs, gcomb, msg, pub, sig)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, os.cpu_count() or 1)
    assert (got == want).all()
    m2, p2, s2 = synth.edge_cases()
    assert (_run(hs, gcomb, m2, p2, s2) == oracle.ecdsa_verify_batch(m2, p2, s2, 1)).all()
    # every signature was walked twice, over the quick and over the exact field forms (as k_ecdsa_point does when the
    # quick walk reports an unfinished fold): a trusted quick walk never disagreed with the exact one
    hs.hs_quick_counters.restype = C.c_long
    assert hs.hs_quick_counters(1) == 0


def test_verify_pipeline_on_reference_golden(hs, gcomb):
    """tests/golden/ecdsa_ref.npz carries the REFERENCE's verdicts (both API levels)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "ecdsa_ref.npz"))
    got = _run(hs, gcomb, z["msg"], z["pub"], z["sig"])
    assert (got == z["verdict_raw"]).all()
    got2 = _run(hs, gcomb, z["msg"], z["pub"], z["sig"], flags=1)
    assert (got2 == z["verdict_cpubkey"]).all()


def test_group_law_degenerate_branches(hs, gcomb, oracle):
    """Signatures engineered so that the accumulator meets a table point it already equals (P + P -> double)
    or its negative (P + (-P) -> infinity), and R = infinity (u1*G + u2*Q = 0)."""
    import synth
    from secp_consts import Gx, Gy, ec_mul, lam

    G = (Gx, Gy)
    rows = []
    # Q = d*G.  Choose u1, u2 directly, then r, s, m from them: s = r/u2, m = u1*s.  Any r works for the
    # arithmetic path; the verdict (almost surely reject) must still agree with the oracle.
    for (d, u1, u2) in [(1, N - 5, 5), (1, 5, N - 5), (2, N - 6, 3), (3, 0, 7), (5, 1, 0x10), (7, 16, 1), (1, 1, 1),
                        (1, 0, 1), (1, 0, 2), (1, 0, 16), (1, 0, 17), (1, 0, (lam + 1) % N), (1, 0, lam), (1, 0, (2 * lam) % N),
                        (1, 0, (lam + 16) % N), (1, 0, (16 * lam + 1) % N), (1, 0, (lam - 1) % N), (9, N - 9, 1), (1, 2**12, 1),
                        (1, 4095, N - 4095), (1, 2**24 + 1, N - 2**24 - 1), (4, 8, N - 2)]:
        Q = ec_mul(d, G)
        pk = b"\x04" + Q[0].to_bytes(32, "big") + Q[1].to_bytes(32, "big")
        for r in (1, 2, Gx % N, ec_mul((u1 + u2 * d) % N, G)[0] % N if (u1 + u2 * d) % N else 1):
            if u2 == 0 or r == 0:
                continue
            s = r * pow(u2, -1, N) % N
            if s > N // 2 or s == 0:
                continue
            m = u1 * s % N
            rows.append((m.to_bytes(32, "big"), pk, r.to_bytes(32, "big") + s.to_bytes(32, "big")))
    msg = np.zeros((len(rows), 32), np.uint8)
    pub = np.zeros((len(rows), 65), np.uint8)
    sig = np.zeros((len(rows), 64), np.uint8)
    for i, (m, pk, sg) in enumerate(rows):
        msg[i] = np.frombuffer(m, np.uint8)
        pub[i] = np.frombuffer(pk, np.uint8)
        sig[i] = np.frombuffer(sg, np.uint8)
    got = _run(hs, gcomb, msg, pub, sig)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, 1)
    assert (got == want).all(), np.nonzero(got != want)[0]
    assert want.any() and not want.all()
    hs.hs_quick_counters.restype = C.c_long
    assert hs.hs_quick_counters(1) == 0  # both bracketings of the walk, over the quick and the exact field forms, agreed


def test_hash_device_functions(hs, oracle):
    rng = np.random.default_rng(5)
    for ln in list(range(0, 130)) + [191, 192, 193, 1487]:
        d = rng.bytes(ln)
        out = (C.c_uint8 * 32)()
        hs.hs_sha256d(C.cast(C.c_char_p(d), u8p), C.c_size_t(ln), out)
        assert bytes(out) == oracle.sha256d(d)
        hs.hs_sha256(C.cast(C.c_char_p(d), u8p), C.c_size_t(ln), out)
        assert bytes(out) == oracle.sha256(d)


def test_p2pkh_fast_path_matches_oracle(hs, gcomb, oracle):
    """Device code of the block fast path (script.cuh: per-tx midstates, per-input template checks, FORKID sighash,
    HASH160) + the ECDSA pipeline vs the oracle's ScriptError per input, on a block holding every mutation kind."""
    import synth

    kinds = list(synth.INPUT_MUTATIONS)
    mut = {2 + 3 * i: k for i, k in enumerate(kinds)}
    blk, spk, amt = synth.synthetic_block(80, True, nkeys=8, with_prevouts=True, mutations=mut)
    want = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, os.cpu_count() or 1)
    ntx, tx_off = oracle.block_index(blk, False)
    # input index (coinbase skipped): every synthetic tx has one input at offset 5 of the tx
    nin = ntx - 1
    in_off = (tx_off[1:ntx] + 5).astype(np.uint32)
    in_tx = np.arange(1, ntx, dtype=np.uint32)
    in_n = np.zeros(nin, dtype=np.uint32)
    cat = np.frombuffer(b"".join(spk), dtype=np.uint8).copy()
    spk_off = np.zeros(nin + 1, dtype=np.uint32)
    spk_off[1:] = np.cumsum([len(s) for s in spk])
    am = np.asarray(amt, dtype=np.int64)
    status = np.zeros(nin, np.uint8)
    msg, pub, sig = np.zeros((nin, 32), np.uint8), np.zeros((nin, 65), np.uint8), np.zeros((nin, 64), np.uint8)
    b = np.frombuffer(blk, dtype=np.uint8).copy()
    hs.hs_block_p2pkh_extract(b.ctypes.data_as(u8p), tx_off.ctypes.data_as(u32p), C.c_size_t(ntx), in_off.ctypes.data_as(u32p),
                              in_tx.ctypes.data_as(u32p), in_n.ctypes.data_as(u32p), C.c_size_t(nin), cat.ctypes.data_as(u8p),
                              spk_off.ctypes.data_as(u32p), am.ctypes.data_as(C.POINTER(C.c_int64)), status.ctypes.data_as(u8p),
                              msg.ctypes.data_as(u8p), pub.ctypes.data_as(u8p), sig.ctypes.data_as(u8p))
    ok = _run(hs, gcomb, msg, pub, sig, flags=1)
    got = np.where(status == 254, np.where(ok == 1, 0, 31), status).astype(np.uint8)
    assert (got == want).all(), [(i, mut.get(i), int(got[i]), int(want[i])) for i in np.nonzero(got != want)[0]]
    assert set(want) >= {0, 2, 11, 22, 23, 26, 28, 31, 36, 255}


def test_p2pkh_fast_path_multi_input_output(hs, gcomb, oracle):
    """Transactions with several inputs/outputs and all six FORKID hashtypes (SINGLE with and without a matching
    output, ANYONECANPAY, NONE): per-tx midstates and per-input sighash of the device code vs the oracle."""
    import synth

    blk, spk, amt = synth.synthetic_block_multi(40, seed=11)
    amt = list(amt)
    amt[3] += 1  # one wrong amount: NULLFAIL exactly there
    want = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, os.cpu_count() or 1)
    assert want[3] == 31 and (np.delete(want, 3) == 0).all()
    b = np.frombuffer(blk, dtype=np.uint8).copy()
    max_tx, max_in = 200, len(spk)
    tx_off = np.zeros(max_tx + 1, np.uint32)
    in_off, in_tx, in_n = np.zeros(max_in, np.uint32), np.zeros(max_in, np.uint32), np.zeros(max_in, np.uint32)
    ntx = C.c_long(0)
    import ctypes
    orc = ctypes.CDLL(os.path.join(ROOT, "oracle", "libbcp_oracle.so"))
    orc.orc_block_index_inputs.restype = C.c_long
    nin = orc.orc_block_index_inputs(b.ctypes.data_as(u8p), C.c_size_t(len(blk)), 0, tx_off.ctypes.data_as(u32p), C.c_size_t(max_tx),
                                     in_off.ctypes.data_as(u32p), in_tx.ctypes.data_as(u32p), in_n.ctypes.data_as(u32p),
                                     C.c_size_t(max_in), C.byref(ntx))
    assert nin == len(spk)
    cat = np.frombuffer(b"".join(spk), dtype=np.uint8).copy()
    spk_off = np.zeros(nin + 1, dtype=np.uint32)
    spk_off[1:] = np.cumsum([len(s) for s in spk])
    am = np.asarray(amt, dtype=np.int64)
    status = np.zeros(nin, np.uint8)
    msg, pub, sig = np.zeros((nin, 32), np.uint8), np.zeros((nin, 65), np.uint8), np.zeros((nin, 64), np.uint8)
    hs.hs_block_p2pkh_extract(b.ctypes.data_as(u8p), tx_off.ctypes.data_as(u32p), C.c_size_t(ntx.value), in_off.ctypes.data_as(u32p),
                              in_tx.ctypes.data_as(u32p), in_n.ctypes.data_as(u32p), C.c_size_t(nin), cat.ctypes.data_as(u8p),
                              spk_off.ctypes.data_as(u32p), am.ctypes.data_as(C.POINTER(C.c_int64)), status.ctypes.data_as(u8p),
                              msg.ctypes.data_as(u8p), pub.ctypes.data_as(u8p), sig.ctypes.data_as(u8p))
    ok = _run(hs, gcomb, msg, pub, sig, flags=1)
    got = np.where(status == 254, np.where(ok == 1, 0, 31), status).astype(np.uint8)
    assert (got == want).all(), np.nonzero(got != want)[0]
