"""The C++ drop-ins of include/compat/ (CCheckQueue, libsecp256k1-named API) built with g++ and linked against
libbcp_b200.so, exercised by tests/cpp/test_compat.cpp the way the node's validation.cpp / pubkey.cpp would."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "test_compat")


def build():
    srcs = [os.path.join(ROOT, "tests", "cpp", "test_compat.cpp"),
            os.path.join(ROOT, "bitcoincashplus_b200", "host", "secp256k1_compat.cpp")]
    deps = srcs + [os.path.join(ROOT, "include", "compat", f) for f in os.listdir(os.path.join(ROOT, "include", "compat"))]
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= max(os.path.getmtime(d) for d in deps):
        return
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include", "compat"),
                           "-I" + os.path.join(ROOT, "include"), "-o", EXE] + srcs +
                          ["-L" + os.path.join(ROOT, "bitcoincashplus_b200"), "-lbcp_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "bitcoincashplus_b200"), "-lpthread"])


def test_checkqueue_contract_cpu():
    """Add/Wait/IsIdle/Control RAII/nullptr queue/early abort with fake checks: no signature is recorded, so
    the GPU is never touched."""
    build()
    out = subprocess.run([EXE, "queue"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert "queue ok" in out.stdout


@pytest.mark.gpu
def test_secp256k1_api_and_recording_queue_gpu(tmp_path):
    """Reference-verdict fixture (tests/golden/ecdsa_ref.npz) through the libsecp256k1-named API and through
    CCheckQueue with recording checks (deferred batch in Wait())."""
    build()
    z = np.load(os.path.join(ROOT, "tests", "golden", "ecdsa_ref.npz"))
    n = 1500
    path = tmp_path / "triples.bin"
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", n))
        for k in ("msg", "pub", "sig"):
            f.write(np.ascontiguousarray(z[k][:n]).tobytes())
        f.write(z["verdict_raw"][:n].astype(np.uint8).tobytes())
        f.write(z["verdict_cpubkey"][:n].astype(np.uint8).tobytes())
    out = subprocess.run([EXE, "gpu", str(path)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "gpu ok" in out.stdout


def test_cointable_and_worker_team_cpu():
    """bcp::CoinTable (backward-shift deletion, growth, long scripts) against std::map over 800,000 random operations,
    and bcp::WorkerTeam; pure host code, no GPU call."""
    exe = os.path.join(ROOT, "tests", "cpp", "test_cointable")
    src = os.path.join(ROOT, "tests", "cpp", "test_cointable.cpp")
    hdr = os.path.join(ROOT, "include", "compat", "bcp_chain.h")
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-o", exe, src,
                               "-L" + os.path.join(ROOT, "bitcoincashplus_b200"), "-lbcp_b200",
                               "-Wl,-rpath," + os.path.join(ROOT, "bitcoincashplus_b200"), "-lpthread"])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "cointable ok" in out.stdout


PUBKEY_EXE = os.path.join(ROOT, "oracle", "_ref", "test_pubkey_compat")


def test_reference_pubkey_cpp_links_against_the_compat_library_alone():
    """Level B3: oracle/Makefile compiles the reference's unmodified src/pubkey.cpp with include/compat/ in front of the
    libsecp256k1 headers and links it against libbcp_secp256k1_compat.so.  Every secp256k1_* symbol the object needs must
    be exported by that library (so nothing of the stock libsecp256k1 is in the link)."""
    if not os.path.exists(PUBKEY_EXE):
        pytest.skip("oracle/_ref/test_pubkey_compat not built (needs /root/reference at build time)")
    need = subprocess.run(["nm", "-D", "--undefined-only", PUBKEY_EXE], capture_output=True, text=True).stdout
    need = {l.split()[-1] for l in need.splitlines() if "secp256k1_" in l}
    have = subprocess.run(["nm", "-D", "--defined-only", os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_secp256k1_compat.so")],
                          capture_output=True, text=True).stdout
    have = {l.split()[-1] for l in have.splitlines() if "secp256k1_" in l}
    assert {"secp256k1_ec_pubkey_tweak_add", "secp256k1_ecdsa_recover", "secp256k1_ecdsa_recoverable_signature_parse_compact",
            "secp256k1_ecdsa_verify", "secp256k1_ec_pubkey_parse"} <= need
    assert need <= have, need - have
    # the header declares what the library defines
    for hdr in ("secp256k1.h", "secp256k1_recovery.h"):
        text = open(os.path.join(ROOT, "include", "compat", hdr)).read()
        for sym in need:
            if sym in text:
                break
        else:
            raise AssertionError("%s declares none of the needed symbols" % hdr)
    for sym in need:
        assert any(sym in open(os.path.join(ROOT, "include", "compat", h)).read() for h in ("secp256k1.h", "secp256k1_recovery.h")), sym


def _der(r, s):
    def enc(v):
        b = v.to_bytes(33, "big").lstrip(b"\x00") or b"\x00"
        if b[0] & 0x80:
            b = b"\x00" + b
        return b"\x02" + bytes([len(b)]) + b
    body = enc(r) + enc(s)
    return b"\x30" + bytes([len(body)]) + body


@pytest.mark.gpu
def test_reference_pubkey_cpp_runs_on_the_gpu(tmp_path, oracle):
    """CPubKey::Verify / CheckLowS / IsFullyValid / Decompress / RecoverCompact / Derive of the reference's own pubkey.cpp,
    running over the compat library on the GPU, against reference-generated fixtures (tests/golden/ecdsa_ref.npz:
    verdict_cpubkey; tests/golden/recovery_ref.npz: CPubKey::RecoverCompact and CPubKey::Derive outputs)."""
    if not os.path.exists(PUBKEY_EXE):
        pytest.skip("oracle/_ref/test_pubkey_compat not built (needs /root/reference at build time)")
    z = np.load(os.path.join(ROOT, "tests", "golden", "ecdsa_ref.npz"))
    g = np.load(os.path.join(ROOT, "tests", "golden", "recovery_ref.npz"))
    idx = list(range(0, 1200)) + list(range(len(z["msg"]) - 112, len(z["msg"])))  # random + every hand-built edge case
    path = tmp_path / "pubkey.bin"
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(idx)))
        for i in idx:
            pub = z["pub"][i].tobytes()
            plen = 33 if pub[0] in (2, 3) else 65
            der = _der(int.from_bytes(z["sig"][i, :32].tobytes(), "big"), int.from_bytes(z["sig"][i, 32:].tobytes(), "big"))
            assert len(der) <= 80
            parse = 1 if (pub[0] in (2, 3, 4, 6, 7) and oracle.pubkey_parse(pub[:plen])[0]) else 0
            f.write(z["msg"][i].tobytes() + bytes([plen]) + pub + bytes([len(der)]) + der + bytes(80 - len(der))
                    + bytes([int(z["verdict_cpubkey"][i]), 1 if oracle.check_low_s(der) else 0, parse]))
        f.write(struct.pack("<Q", len(g["rc_len"])))
        for i in range(len(g["rc_len"])):
            f.write(g["rc_hash"][i].tobytes() + g["rc_sig"][i].tobytes() + bytes([int(g["rc_len"][i])]) + g["rc_out"][i].tobytes())
        f.write(struct.pack("<Q", len(g["dv_n"])))
        for i in range(len(g["dv_n"])):
            f.write(g["dv_pub"][i].tobytes() + g["dv_cc"][i].tobytes() + struct.pack("<I", int(g["dv_n"][i]))
                    + g["dv_out"][i].tobytes() + g["dv_ccout"][i].tobytes())
    out = subprocess.run([PUBKEY_EXE, str(path)], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "pubkey compat ok" in out.stdout
