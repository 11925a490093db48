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
