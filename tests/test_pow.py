"""SURVEY.md §8f N4, CPU side: proof-of-work arithmetic and difficulty rules.
  * oracle/pow_oracle.py (Python integers) against the reference's own vectors (tests/golden/pow_vectors.json, extracted
    from src/test/arith_uint256_tests.cpp and src/test/pow_tests.cpp) and against the reference's arith_uint256 compiled
    in place (oracle/_ref) on random inputs;
  * the product's host code (include/compat/bcp_headers.h through libbcp_chain.so) and the device target arithmetic
    (csrc/pow.cuh through the host build of the device headers) against the same vectors and against the oracle."""
import ctypes as C
import json
import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pow_oracle as PO  # noqa: E402
from oracle import pyoracle as po  # noqa: E402
from bitcoincashplus_b200 import headers as H  # noqa: E402

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "pow_vectors.json")))
LIM_LEGACY = (1 << 224) - 1


def interesting_compacts(rng, n):
    out = [0, 0x1d00ffff, 0x207fffff, 0x1f07ffff, 0x2000ffff, 0x21000001, 0x2100ffff, 0x22000001, 0x220000ff, 0x23000001,
           0x03800000, 0x04800001, 0x01000000, 0xff123456]
    for _ in range(n):
        size = rng.choice(list(range(0, 40)) + [rng.randrange(256)])
        word = rng.choice([rng.randrange(1 << 23), rng.randrange(256), rng.randrange(1 << 16), 0, 1, 0x7fffff])
        out.append((size << 24) | (rng.randrange(2) << 23) | word)
    return out


# ---------------------------------------------------------------- golden vectors
@pytest.mark.parametrize("impl", ["oracle", "product"])
def test_compact_golden(impl):
    for v in GOLD["set_compact"]:
        t, neg, ovf = PO.set_compact(v["compact"]) if impl == "oracle" else H.compact_to_target(v["compact"])
        if "hex" in v:
            assert t == int(v["hex"], 16), hex(v["compact"])
        if "negative" in v:
            assert neg == v["negative"] and ovf == v["overflow"], hex(v["compact"])
        if "get_compact" in v and not v["get_compact_negative_arg"]:
            got = PO.get_compact(t) if impl == "oracle" else H.compact_from_target(t)
            assert got == v["get_compact"], hex(v["compact"])
    # `num = 0x80; GetCompact() == 0x02008000` (same test case of the reference)
    assert PO.get_compact(0x80) == 0x02008000 and H.compact_from_target(0x80) == 0x02008000


def _legacy_case(impl, v):
    if impl == "oracle":
        c = PO.Chain(PO.Params.main())
        c.times, c.bits, c.work = [0] * v["height"] + [v["time"]], [v["bits"]] * (v["height"] + 1), [0] * (v["height"] + 1)
        return c.calculate_next_work(v["first_time"])
    c = H.HeaderChain(H.PowParams.main())
    c.add_index(1, v["time"], v["bits"])  # CalculateNextWorkRequired reads only the tip's time and bits
    return c.calculate_next_work(v["first_time"])


@pytest.mark.parametrize("impl", ["oracle", "product"])
def test_calculate_next_work_golden(impl):
    for v in GOLD["calculate_next_work"]:
        assert _legacy_case(impl, v) == v["expect"], v


class _Both:
    """drives the oracle chain and the product chain together"""

    def __init__(self, impl):
        self.c = PO.Chain(PO.Params.main()) if impl == "oracle" else H.HeaderChain(H.PowParams.main())
        self.impl, self.t = impl, 0

    def add(self, dt_or_time, bits, absolute=False):
        self.t = dt_or_time if absolute else self.t + dt_or_time
        if self.impl == "oracle":
            self.c.add(self.t, bits)
        else:
            self.c.add_index(4, self.t, bits)

    def nxt(self):
        return self.c.cashplus_next_work(0)


@pytest.mark.parametrize("impl", ["oracle", "product"])
def test_cash_plus_difficulty_sequence_golden(impl):
    """The scripted scenario of cash_plus_difficulty_test (src/test/pow_tests.cpp:114-290), step for step."""
    cp = GOLD["cashplus_checkpoints"]
    pow_limit = LIM_LEGACY
    initial = PO.get_compact(pow_limit >> 4)
    b = _Both(impl)
    b.add(1269211443, initial, absolute=True)
    for _ in range(1, 2050):
        b.add(600, initial)
    bits = b.nxt()
    for _ in range(10):
        b.add(600, bits)
        assert b.nxt() == bits
    b.add(6000, bits)
    assert b.nxt() == bits
    b.add(2 * 600 - 6000, bits)
    assert b.nxt() == bits
    for _ in range(20):
        b.add(600, bits)
        assert b.nxt() == bits
    b.add(550, bits)
    assert b.nxt() == bits
    for _ in range(10):
        b.add(550, bits)
        nb = b.nxt()
        cur, nx = PO.set_compact(bits)[0], PO.set_compact(nb)[0]
        assert nx < cur and cur - nx < (cur >> 10)
        bits = nb
    assert bits == cp[0]
    for _ in range(20):
        b.add(10, bits)
        nb = b.nxt()
        cur, nx = PO.set_compact(bits)[0], PO.set_compact(nb)[0]
        assert nx < cur and cur - nx < (cur >> 4)
        bits = nb
    assert bits == cp[1]
    b.add(6000, bits)
    bits = b.nxt()
    assert bits == cp[2]
    for _ in range(93):
        b.add(6000, bits)
        nb = b.nxt()
        cur, nx = PO.set_compact(bits)[0], PO.set_compact(nb)[0]
        assert nx <= pow_limit and nx > cur and nx - cur < (cur >> 3)
        bits = nb
    assert bits == cp[3]
    b.add(6000, bits)
    bits = b.nxt()
    assert bits == cp[4]
    for _ in range(192):
        b.add(6000, bits)
        nb = b.nxt()
        cur, nx = PO.set_compact(bits)[0], PO.set_compact(nb)[0]
        assert nx <= pow_limit and nx > cur and nx - cur < (cur >> 3)
        bits = nb
    assert bits == cp[5]
    for _ in range(5):
        b.add(6000, bits)
        assert b.nxt() == PO.get_compact(pow_limit)


# ---------------------------------------------------------------- random: oracle vs reference arithmetic, product vs oracle
def _ref_pow():
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    L = po.ref().lib
    L.ref_compact_to_target.argtypes = [C.c_uint32, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ref_compact_from_target.argtypes = [C.c_char_p]
    L.ref_compact_from_target.restype = C.c_uint32
    L.ref_block_proof.argtypes = [C.c_uint32, C.c_char_p]
    L.ref_check_pow.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p]
    L.ref_next_work.argtypes = [C.POINTER(H.PowParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_int64, C.c_int]
    L.ref_next_work.restype = C.c_uint32
    return L


def test_oracle_matches_reference_arith_random():
    L = _ref_pow()
    rng = random.Random(11)
    for c in interesting_compacts(rng, 3000):
        out, neg, ovf = C.create_string_buffer(32), C.c_int(), C.c_int()
        L.ref_compact_to_target(c, out, C.byref(neg), C.byref(ovf))
        t = int.from_bytes(out.raw, "little")
        assert (t, bool(neg.value), bool(ovf.value)) == PO.set_compact(c), hex(c)
        assert L.ref_compact_from_target(out.raw) == PO.get_compact(t), hex(c)
        L.ref_block_proof(c, out)
        assert int.from_bytes(out.raw, "little") == PO.block_proof(c), hex(c)
    for _ in range(2000):
        t = rng.getrandbits(rng.randrange(1, 257))
        assert L.ref_compact_from_target(t.to_bytes(32, "little")) == PO.get_compact(t)
    for _ in range(3000):
        c = rng.choice(interesting_compacts(rng, 4))
        t = PO.set_compact(c)[0]
        h = rng.choice([t, t + 1, max(t - 1, 0), rng.getrandbits(256), rng.getrandbits(max(1, t.bit_length()))]) & PO.M256
        lim = rng.choice([(1 << 255) - 1, LIM_LEGACY, (1 << 236) - 1])
        hb = h.to_bytes(32, "little")
        assert bool(L.ref_check_pow(hb, c, lim.to_bytes(32, "little"))) == PO.check_pow(hb, c, lim)


def test_product_matches_oracle_arith_random():
    rng = random.Random(12)
    for c in interesting_compacts(rng, 3000):
        assert H.compact_to_target(c) == PO.set_compact(c), hex(c)
        assert H.block_proof(c) == PO.block_proof(c), hex(c)
    for _ in range(3000):
        t = rng.getrandbits(rng.randrange(1, 257))
        assert H.compact_from_target(t) == PO.get_compact(t)


def _random_chain(rng, n, params_kw):
    """a chain whose bits follow the rules most of the time (so that targets stay in range) with random block times"""
    p = PO.Params((1 << 255) - 1, LIM_LEGACY, (1 << 236) - 1, **params_kw)
    c = PO.Chain(p)
    t = 1500000000
    c.add(t, PO.get_compact(p.pow_limit_legacy >> rng.randrange(0, 12)))
    for _ in range(n - 1):
        t += rng.choice([600, 600, 550, 10, 1, 6000, 1300, -500, 0])
        t = max(t, 1)
        want = c.next_work(t) if len(c.times) > 200 else c.bits[-1]
        c.add(t, want if rng.random() < 0.95 else PO.get_compact(PO.set_compact(want)[0] >> 1 or 1))
    return p, c


PARAM_SETS = [
    dict(bcp_height=100000, premine_window=10),                                   # legacy retargeting every 2016 blocks
    dict(bcp_height=100000, premine_window=10, allow_min_difficulty=True),        # testnet min-difficulty walk-back
    dict(bcp_height=2300, premine_window=20, averaging_window=30),                # fork, premine, start limit, new rule
    dict(bcp_height=2300, premine_window=20, averaging_window=30, allow_min_difficulty=True),
    dict(bcp_height=0, premine_window=0, averaging_window=0),                     # new rule from the start
    dict(bcp_height=5, premine_window=3, no_retargeting=True),
]


def _to_struct(p):
    return H.PowParams.make(p.pow_limit, p.pow_limit_legacy, p.pow_limit_start, allow_min_difficulty=int(p.allow_min_difficulty),
                            no_retargeting=int(p.no_retargeting), bcp_height=p.bcp_height, premine_window=p.premine_window,
                            averaging_window=p.averaging_window)


@pytest.mark.parametrize("kw", PARAM_SETS)
def test_next_work_random_chains(kw):
    """GetNextWorkRequired along random chains: product == oracle at every height, and == the restated control flow
    over the reference's arith_uint256 when oracle/_ref is there."""
    rng = random.Random(hash(tuple(sorted(kw.items()))) & 0xffff)
    n = 2600
    p, c = _random_chain(rng, n, kw)
    L = po.ref().lib if po.have_ref() else None
    if L is not None:
        _ref_pow()
    st = _to_struct(p)
    hc = H.HeaderChain(st)
    times, bits = np.array(c.times, dtype=np.uint32), np.array(c.bits, dtype=np.uint32)
    probe = PO.Chain(p)
    for h in range(n):
        hc.add_index(4, c.times[h], c.bits[h])
        probe.add(c.times[h], c.bits[h])
        if h < 150 and kw.get("bcp_height", 0) + kw.get("premine_window", 0) + kw.get("averaging_window", 30) <= h + 1 \
                and not kw.get("no_retargeting"):
            continue  # the new rule needs 147 blocks of history (the reference asserts); not reachable on a real chain
        for nt in (c.times[h] + 600, c.times[h] + 1201, c.times[h] - 5):
            want = probe.next_work(nt)
            assert hc.next_work(nt) == want, (h, nt)
            if L is not None and h % 7 == 0:
                assert L.ref_next_work(C.byref(st), times.ctypes.data, bits.ctypes.data, h + 1, nt, 0) == want, (h, nt)
        assert hc.median_time_past() == probe.median_time_past()
    assert hc.chain_work() == probe.work[-1]


# ---------------------------------------------------------------- device target arithmetic (host build of pow.cuh)
def test_device_pow_check_host_build():
    import test_hostsim

    hs = test_hostsim.build_hostsim()
    hs.hs_pow_check.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    rng = random.Random(13)
    for c in interesting_compacts(rng, 4000):
        t = PO.set_compact(c)[0]
        h = rng.choice([t, (t + 1) & PO.M256, max(t - 1, 0), rng.getrandbits(256), rng.getrandbits(max(1, t.bit_length()))])
        lim = rng.choice([(1 << 255) - 1, LIM_LEGACY, (1 << 236) - 1])
        out, neg, ovf = C.create_string_buffer(32), C.c_int(), C.c_int()
        hb = h.to_bytes(32, "little")
        got = hs.hs_pow_check(hb, c, lim.to_bytes(32, "little"), out, C.byref(neg), C.byref(ovf))
        assert (int.from_bytes(out.raw, "little"), bool(neg.value), bool(ovf.value)) == PO.set_compact(c), hex(c)
        assert bool(got) == PO.check_pow(hb, c, lim), hex(c)
