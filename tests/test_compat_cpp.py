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
