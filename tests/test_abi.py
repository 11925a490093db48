"""CPU: the C-ABI library loads, exports every symbol include/bcp_gpu.h declares, and refuses to produce
verdicts without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "bcp_gpu.h")).read()
    return sorted(set(re.findall(r"BCP_API\s+[\w\s\*]+?\b(bcp_\w+)\s*\(", hdr)))


def test_header_and_python_signatures_agree():
    from bitcoincashplus_b200._lib import SIGNATURES

    assert declared_symbols() == sorted(SIGNATURES)


def test_library_exports_every_declared_symbol():
    import bitcoincashplus_b200 as bcp

    L = bcp.lib()
    for name in declared_symbols():
        assert hasattr(L, name), name


def test_no_cpu_fallback_without_device():
    import bitcoincashplus_b200 as bcp

    L = bcp.lib()
    if L.bcp_gpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    assert L.bcp_gpu_init(0) == -1  # BCP_ERR_NO_DEVICE
    assert b"no CPU fallback" in L.bcp_gpu_last_error()
    ok = np.zeros(4, dtype=np.uint8)
    z = np.zeros(4 * 65, dtype=np.uint8)
    rc = L.bcp_gpu_ecdsa_verify_batch(0, z.ctypes.data, z.ctypes.data, z.ctypes.data, 4, 0, ok.ctypes.data)
    assert rc == -3 and not ok.any()  # BCP_ERR_NOT_INIT, and no verdict written
    with pytest.raises(bcp.BcpError):
        bcp.Engine(0)


def test_product_does_not_touch_the_oracle():
    """The product package and its sources never import, link or name the oracle."""
    pkg = os.path.join(ROOT, "bitcoincashplus_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "pyoracle" not in text and "bcp_oracle" not in text and "libbcp_ref" not in text, f
    import subprocess

    out = subprocess.run(["ldd", os.path.join(pkg, "libbcp_b200.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out and "hostsim" not in out


def test_der_lax_parse_matches_reference_rules(oracle):
    """bcp_der_lax_parse (host-side) vs the oracle's restatement of src/pubkey.cpp:26-168 on hostile encodings."""
    import bitcoincashplus_b200 as bcp

    rng = np.random.default_rng(3)
    r = int.from_bytes(rng.bytes(32), "big") >> 1
    s = int.from_bytes(rng.bytes(32), "big") >> 2
    rb, sb = r.to_bytes(32, "big"), s.to_bytes(32, "big")
    N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
    cases = [
        b"\x30\x44\x02\x20" + rb + b"\x02\x20" + sb,                         # canonical
        b"\x30\x46\x02\x21\x00" + rb + b"\x02\x21\x00" + sb,                 # padded
        b"\x30\x81\x44\x02\x20" + rb + b"\x02\x20" + sb,                     # long-form sequence length
        b"\x30\x44\x02\x81\x20" + rb + b"\x02\x20" + sb,                     # long-form integer length
        b"\x30\x44\x02\x20" + rb + b"\x02\x20" + sb + b"garbage",            # trailing bytes
        b"\x30\x00\x02\x20" + rb + b"\x02\x20" + sb,                         # wrong sequence length (ignored)
        b"\x30\x45\x02\x21\x01" + rb + b"\x02\x20" + sb,                     # 33-byte R => overflow => zeros
        b"\x30\x44\x02\x20" + N.to_bytes(32, "big") + b"\x02\x20" + sb,      # r == n => overflow
        b"\x30\x44\x02\x20" + b"\xff" * 32 + b"\x02\x20" + sb,               # negative-looking R
        b"\x31\x44\x02\x20" + rb + b"\x02\x20" + sb,                         # bad tag
        b"\x30\x44\x03\x20" + rb + b"\x02\x20" + sb,                         # bad integer tag
        b"\x30\x44\x02\x20" + rb + b"\x02\x30" + sb,                         # S length runs past the end
        b"\x30\x44\x02\x88" + b"\x00" * 7 + b"\x20" + rb + b"\x02\x20" + sb,  # 8 length bytes, leading zeros
        b"\x30\x44\x02\x89" + b"\x01" * 9 + rb,                              # length-of-length too large
        b"", b"\x30", b"\x30\x44", b"\x30\x44\x02", b"\x30\x06\x02\x01\x01\x02\x01\x01", b"\x30\x04\x02\x00\x02\x00",
    ]
    for _ in range(300):
        cases.append(rng.bytes(int(rng.integers(0, 80))))
        c = bytearray(cases[0])
        c[int(rng.integers(0, len(c)))] = int(rng.integers(0, 256))
        cases.append(bytes(c))
    for der in cases:
        assert bcp.der_lax_parse(der) == oracle.der_lax_parse(der), der.hex()


def test_block_index_matches_oracle(oracle):
    import bitcoincashplus_b200 as bcp
    import synth

    blk = synth.synthetic_block(ntx=50, sign=False)
    off = bcp.block_index(blk)
    n, ooff = oracle.block_index(blk, False)
    assert n == 51 and (off == ooff).all()
    raw = open(os.path.join(ROOT, "tests", "golden", "block413567.raw"), "rb").read()
    off = bcp.block_index(raw, legacy_header=True)
    n, ooff = oracle.block_index(raw, True)
    assert n == 1557 and (off == ooff).all() and off[-1] == len(raw)
    with pytest.raises(bcp.BcpError):
        bcp.block_index(raw[:5000], legacy_header=True)


def _py_scan(blk, legacy):
    """independent Python parse of a block: per-tx offsets, output-count offsets, flat non-coinbase inputs and outputs"""
    def cs(pos):
        c = blk[pos]
        if c < 253:
            return c, pos + 1
        nb = {253: 2, 254: 4, 255: 8}[c]
        return int.from_bytes(blk[pos + 1:pos + 1 + nb], "little"), pos + 1 + nb
    pos = 80 if legacy else 140
    if not legacy:
        n, pos = cs(pos)
        pos += n
    ntx, pos = cs(pos)
    tx_off, tx_out, ins, outs = [], [], [], []
    for t in range(ntx):
        tx_off.append(pos)
        pos += 4
        nin, pos = cs(pos)
        cb = nin == 1 and blk[pos:pos + 32] == bytes(32) and blk[pos + 32:pos + 36] == b"\xff" * 4
        for i in range(nin):
            if not cb:
                ins.append((pos, t, i))
            pos += 36
            sl, pos = cs(pos)
            pos += sl + 4
        tx_out.append(pos)
        nout, pos = cs(pos)
        for v in range(nout):
            outs.append((pos, t, v))
            pos += 8
            sl, pos = cs(pos)
            pos += sl
        pos += 4
    tx_off.append(pos)
    return tx_off, tx_out, ins, outs


def test_block_index_full_lists_inputs_and_outputs():
    """bcp_block_index_full (host-side framing scan with the flat input and output lists) against an independent Python
    parse: the reference's own bench block (legacy header, 1,557 transactions of every shape) and synthetic blocks."""
    import ctypes as C

    import bitcoincashplus_b200 as bcp
    import synth

    L = bcp.lib()
    raw = open(os.path.join(ROOT, "tests", "golden", "block413567.raw"), "rb").read()
    cases = [(raw, True), (synth.synthetic_block(ntx=40, sign=False), False), (synth.synthetic_block_multi(60, seed=3)[0], False),
             (synth.synthetic_block(0, True), False)]
    for blk, legacy in cases:
        tx_off, tx_out, ins, outs = _py_scan(blk, legacy)
        mt, mi, mo = len(blk) // 60 + 2, len(blk) // 41 + 2, len(blk) // 9 + 2
        a = {k: np.zeros(n, dtype=np.uint32) for k, n in (("tx", mt + 1), ("txo", mt + 1), ("io", mi), ("it", mi), ("in", mi),
                                                          ("oo", mo), ("ot", mo), ("on", mo))}
        ntx, nout = C.c_long(0), C.c_long(0)
        keys = np.zeros(mi, dtype=np.uint64)
        nin = L.bcp_block_index_full(blk, len(blk), int(legacy), a["tx"].ctypes.data, a["txo"].ctypes.data, mt, a["io"].ctypes.data,
                                     a["it"].ctypes.data, a["in"].ctypes.data, mi, C.byref(ntx), a["oo"].ctypes.data,
                                     a["ot"].ctypes.data, a["on"].ctypes.data, mo, C.byref(nout), keys.ctypes.data)
        assert (nin, ntx.value, nout.value) == (len(ins), len(tx_off) - 1, len(outs))
        for (pos, _, _), k in zip(ins, keys[:nin]):
            want = int.from_bytes(blk[pos:pos + 8], "little") ^ ((0x9E3779B97F4A7C15 * int.from_bytes(blk[pos + 32:pos + 36], "little")) & (2**64 - 1))
            assert int(k) == want
        assert list(a["tx"][:ntx.value + 1]) == tx_off and list(a["txo"][:ntx.value]) == tx_out
        assert list(zip(a["io"][:nin], a["it"][:nin], a["in"][:nin])) == ins
        assert list(zip(a["oo"][:nout.value], a["ot"][:nout.value], a["on"][:nout.value])) == outs
        # capacity error for the outputs, and the optional arrays may be absent
        if len(outs) > 1:
            assert L.bcp_block_index_full(blk, len(blk), int(legacy), a["tx"].ctypes.data, None, mt, a["io"].ctypes.data,
                                          a["it"].ctypes.data, a["in"].ctypes.data, mi, C.byref(ntx), a["oo"].ctypes.data,
                                          a["ot"].ctypes.data, a["on"].ctypes.data, len(outs) - 1, C.byref(nout), None) == -2
        assert L.bcp_block_index_full(blk, len(blk), int(legacy), a["tx"].ctypes.data, None, mt, a["io"].ctypes.data,
                                      a["it"].ctypes.data, a["in"].ctypes.data, mi, C.byref(ntx), None, None, None, 0, None, None) == len(ins)


def test_scanners_reject_non_canonical_counts(oracle):
    """ReadCompactSize (src/serialize.h:260-300) refuses a count written in a longer form than needed; so do all three
    scan entry points — for the transaction count, an input count and a script length alike — and the oracle."""
    import ctypes as C

    import bitcoincashplus_b200 as bcp
    import synth

    L = bcp.lib()
    blk = synth.synthetic_block(ntx=5, sign=False)
    tx_off, tx_out, ins, outs = _py_scan(blk, False)
    base = tx_off[0] - 1                       # the transaction count byte (6 transactions incl. the coinbase)
    assert blk[base] == 6
    variants = {
        "ntx": blk[:base] + b"\xfd\x06\x00" + blk[base + 1:],
        "nin": blk[:tx_off[1] + 4] + b"\xfd\x01\x00" + blk[tx_off[1] + 5:],
        "script_len": blk[:ins[0][0] + 36] + b"\xfd" + blk[ins[0][0] + 36:ins[0][0] + 37] + b"\x00" + blk[ins[0][0] + 37:],
    }
    mt, mi, mo = len(blk) // 60 + 8, len(blk) // 41 + 8, len(blk) // 9 + 8
    a = [np.zeros(n, dtype=np.uint32) for n in (mt + 1, mt + 1, mi, mi, mi, mo, mo, mo)]
    ntx, nout = C.c_long(0), C.c_long(0)
    for name, bad in variants.items():
        assert L.bcp_block_index(bad, len(bad), 0, a[0].ctypes.data, mt) == -1, name
        assert L.bcp_block_index_inputs(bad, len(bad), 0, a[0].ctypes.data, mt, a[2].ctypes.data, a[3].ctypes.data,
                                        a[4].ctypes.data, mi, C.byref(ntx)) == -1, name
        assert L.bcp_block_index_full(bad, len(bad), 0, a[0].ctypes.data, a[1].ctypes.data, mt, a[2].ctypes.data, a[3].ctypes.data,
                                      a[4].ctypes.data, mi, C.byref(ntx), a[5].ctypes.data, a[6].ctypes.data, a[7].ctypes.data, mo,
                                      C.byref(nout), None) == -1, name
        with pytest.raises(Exception):
            n, _ = oracle.block_index(bad, False)
            assert n >= 0
    assert L.bcp_block_index(blk, len(blk), 0, a[0].ctypes.data, mt) == 6


def test_block_index_rules_reports_the_reference_reject_string():
    """bcp_block_index_rules: the framing scan applies CheckBlock's context-free transaction rules
    (src/validation.cpp:476-570, 3255-3330) in the same pass; every damaged block of synth.STRUCTURE_KINDS gives the
    reference's reject string and names the transaction, an undamaged block gives 0.  Host code only: runs without a GPU."""
    import ctypes as C

    import numpy as np

    import bitcoincashplus_b200 as bcp
    import synth

    L = bcp.lib()

    def rules(blk):
        b = np.frombuffer(blk, np.uint8)
        n = len(blk)
        arrs = [np.zeros(n // 9 + 8, np.uint32) for _ in range(8)]
        ntx, nout, rule, rtx = C.c_long(), C.c_long(), C.c_int(-1), C.c_long(-2)
        nin = L.bcp_block_index_rules(b.ctypes.data, n, 0, arrs[0].ctypes.data, arrs[1].ctypes.data, n // 60 + 2, arrs[2].ctypes.data,
                                      arrs[3].ctypes.data, arrs[4].ctypes.data, n // 41 + 2, C.byref(ntx), arrs[5].ctypes.data,
                                      arrs[6].ctypes.data, arrs[7].ctypes.data, n // 9 + 2, C.byref(nout), None, C.byref(rule), C.byref(rtx))
        assert nin >= 0
        return L.bcp_block_rule_name(rule.value), rtx.value

    good, _, _ = synth.synthetic_chain(2, 8, equihash=False)
    assert rules(good[1]) == (b"", -1)
    where = {"second_coinbase": 2, "no_coinbase": 0, "cb_length": 0}
    for kind, want in synth.STRUCTURE_KINDS.items():
        bad, _, _ = synth.synthetic_chain(2, 8, equihash=False, corrupt=(1, kind))
        name, tx = rules(bad[1])
        assert name == want and tx == where.get(kind, 3), (kind, name, tx)
    assert L.bcp_block_rule_name(99) == b""
