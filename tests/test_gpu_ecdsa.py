"""-m gpu: batched ECDSA verification through the C ABI vs the oracle (bit-exact verdicts)."""
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


def test_sigset_with_every_corruption(engine, oracle):
    msg, pub, sig, kinds = synth.sigset(20000, nkeys=512, corrupt_pct=10)
    got = engine.ecdsa_verify_batch(msg, pub, sig)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, nthreads=NT)
    assert (got == want).all(), np.nonzero(got != want)[0][:20]
    assert set(kinds[kinds >= 0]) == set(range(len(synth.CORRUPTIONS)))
    # each corruption class has the verdict the reference gives (all reject but the two benign ones)
    for k, name in enumerate(synth.CORRUPTIONS):
        sel = kinds == k
        if name in ("hybrid_ok", "msg_plus_n"):
            assert got[sel].all(), name
        else:
            assert not got[sel].any(), name
    assert got[kinds < 0].all()


def test_normalize_flag_matches_cpubkey_verify(engine, oracle):
    """BCP_ECDSA_FLAG_NORMALIZE_S == CPubKey::Verify (src/pubkey.cpp:170-193) on DER-encoded signatures."""
    import bitcoincashplus_b200 as bcp

    msg, pub, sig, kinds = synth.sigset(3000, nkeys=64, corrupt_pct=20)
    want = np.zeros(len(msg), dtype=np.uint8)
    sig_parsed = np.zeros_like(sig)
    for i in range(len(msg)):
        r, s = int.from_bytes(sig[i, :32].tobytes(), "big"), int.from_bytes(sig[i, 32:].tobytes(), "big")
        # encode (r, s) as DER by hand so that out-of-range / zero components survive, like a hostile peer would
        def enc(v):
            b = v.to_bytes(33, "big").lstrip(b"\x00") or b"\x00"
            if b[0] & 0x80:
                b = b"\x00" + b
            return b"\x02" + bytes([len(b)]) + b
        body = enc(r) + enc(s)
        der = b"\x30" + bytes([len(body)]) + body
        plen = 33 if pub[i, 0] in (2, 3) else 65
        want[i] = oracle.pubkey_verify(msg[i].tobytes(), pub[i, :plen].tobytes(), der)
        ok, p64 = bcp.der_lax_parse(der)
        assert ok == 1
        sig_parsed[i] = np.frombuffer(p64, dtype=np.uint8)
    got = engine.ecdsa_verify_batch(msg, pub, sig_parsed, flags=bcp.ECDSA_FLAG_NORMALIZE_S)
    assert (got == want).all(), np.nonzero(got != want)[0][:20]
    assert got[kinds == synth.CORRUPTIONS.index("high_s")].all()


def test_edge_scalars_and_messages(engine, oracle):
    msg, pub, sig = synth.edge_cases()
    got = engine.ecdsa_verify_batch(msg, pub, sig)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, nthreads=NT)
    assert (got == want).all() and want.all()


@pytest.mark.parametrize("n", [0, 1, 2, 7, 8, 9, 127, 128, 129, 1023, 1025])
def test_ragged_sizes(engine, oracle, n):
    msg, pub, sig, _ = synth.sigset(max(n, 1), nkeys=16, corrupt_pct=25)
    msg, pub, sig = msg[:n], pub[:n], sig[:n]
    got = engine.ecdsa_verify_batch(msg, pub, sig)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, nthreads=1) if n else np.zeros(0, np.uint8)
    assert (got == want).all()


def test_chunked_large_batch_properties(engine, oracle):
    """Full-size property check (1M signatures is BASELINE config 2; here the chunking path: more than six waves, so the
    host-pointer call cuts the batch into a one-wave chunk and a second one on the other stream):
    corrupting a signature flips exactly its verdict; a sample is compared with the oracle."""
    n = 520000
    msg, pub, sig, kinds = synth.sigset(n, nkeys=4096, corrupt_pct=1)
    got = engine.ecdsa_verify_batch(msg, pub, sig)
    benign = np.isin(kinds, [synth.CORRUPTIONS.index("hybrid_ok"), synth.CORRUPTIONS.index("msg_plus_n")])
    assert (got == ((kinds < 0) | benign)).all()
    idx = np.arange(0, n, 97)
    want = oracle.ecdsa_verify_batch(msg[idx].copy(), pub[idx].copy(), sig[idx].copy(), nthreads=NT)
    assert (got[idx] == want).all()
    # idempotence
    assert (engine.ecdsa_verify_batch(msg, pub, sig) == got).all()


def test_tests_c_vectors_on_gpu(engine):
    """The reference's own self-test vectors (src/secp256k1/src/tests.c: run_ec_pubkey_parse_test :2787-3111,
    test_ecdsa_edge_cases :3973-4340; tests/golden/secp_vectors.json) through the CUDA path."""
    import json

    import bitcoincashplus_b200 as bcp

    v = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "secp_vectors.json")))
    pp = [e for e in v["pubkey_parse"] if bcp and (len(e["pub"]) // 2 == (33 if e["pub"][:2] in ("02", "03") else 65 if e["pub"][:2] in ("04", "06", "07") else -1))]
    assert len(pp) > 100
    pub = np.zeros((len(pp), 65), np.uint8)
    for i, e in enumerate(pp):
        b = bytes.fromhex(e["pub"])
        pub[i, : len(b)] = np.frombuffer(b, np.uint8)
    xy, ok = engine.pubkey_parse_batch(pub)
    for i, e in enumerate(pp):
        assert ok[i] == e["ok"], e["pub"]
        if e["ok"]:
            assert xy[i].tobytes().hex() == e["xy"]
    ed = v["ecdsa_edge"]
    msg = np.stack([np.frombuffer(bytes.fromhex(e["msg"]), np.uint8) for e in ed])
    sig = np.stack([np.frombuffer(bytes.fromhex(e["sig"]), np.uint8) for e in ed])
    pk = np.zeros((len(ed), 65), np.uint8)
    for i, e in enumerate(ed):
        b = bytes.fromhex(e["pub"])
        pk[i, : len(b)] = np.frombuffer(b, np.uint8)
    assert list(engine.ecdsa_verify_batch(msg, pk, sig)) == [e["verdict_raw"] for e in ed]
    assert list(engine.ecdsa_verify_batch(msg, pk, sig, flags=bcp.ECDSA_FLAG_NORMALIZE_S)) == [e["verdict_normalized"] for e in ed]


def test_reference_golden_verdicts_on_gpu(engine):
    """tests/golden/ecdsa_ref.npz: verdicts computed by the reference at both API levels."""
    import bitcoincashplus_b200 as bcp

    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ecdsa_ref.npz"))
    assert (engine.ecdsa_verify_batch(z["msg"], z["pub"], z["sig"]) == z["verdict_raw"]).all()
    assert (engine.ecdsa_verify_batch(z["msg"], z["pub"], z["sig"], flags=bcp.ECDSA_FLAG_NORMALIZE_S) == z["verdict_cpubkey"]).all()


def test_device_field_selftest(engine):
    """bcp_gpu_selftest: crafted operands through the short-cut and the full carry chains of the field code, on the device
    (also run by bcp_gpu_init).  0 = every pair agrees."""
    import bitcoincashplus_b200 as bcp

    assert bcp.lib().bcp_gpu_selftest(engine.device) == 0, bcp.lib().bcp_gpu_last_error()


def test_recover_and_tweak_add_against_reference_fixture_and_oracle(engine, oracle):
    """bcp_gpu_ecdsa_recover_batch / bcp_gpu_pubkey_tweak_add_batch (what CPubKey::RecoverCompact and CPubKey::Derive end
    in, src/pubkey.cpp:195-264) vs the reference-generated fixture (every recovery id of 500 signatures, damaged
    signatures and messages, the tests_impl.h edge vectors, r / s / tweak range rules, tweaks that cancel the key), then
    a larger random batch vs the oracle, which covers the multi-block launch."""
    import hashlib

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "recovery_ref.npz"))
    xy, ok = engine.ecdsa_recover_batch(g["rec_msg"], g["rec_sig"], g["rec_id"])
    assert (ok == g["rec_ok"]).all(), np.nonzero(ok != g["rec_ok"])[0][:20]
    assert (xy == g["rec_xy"]).all()
    xy, ok = engine.pubkey_tweak_add_batch(g["tw_pub"], g["tw_tweak"])
    assert (ok == g["tw_ok"]).all(), np.nonzero(ok != g["tw_ok"])[0][:20]
    assert (xy == g["tw_xy"]).all()
    # recid > 3 is refused per item, not an error of the call
    xy, ok = engine.ecdsa_recover_batch(g["rec_msg"][:4], g["rec_sig"][:4], [4, 7, 255, 0])
    assert list(ok[:3]) == [0, 0, 0] and ok[3] == g["rec_ok"][0]
    # random batch: verification set reused as recovery input (r, s of valid and corrupted signatures), all four ids
    msg, pub, sig, _ = synth.sigset(3000, nkeys=64, corrupt_pct=20)
    rid = (np.arange(3000) % 4).astype(np.uint8)
    xy, ok = engine.ecdsa_recover_batch(msg, sig, rid)
    hits = 0
    for i in range(3000):
        want = oracle.ecdsa_recover(msg[i].tobytes(), sig[i].tobytes(), int(rid[i]))
        assert (want is not None) == bool(ok[i]), i
        if want is not None:
            assert xy[i].tobytes() == want, i
            plen = 33 if pub[i, 0] in (2, 3) else 65
            p_ok, p_xy = oracle.pubkey_parse(pub[i, :plen].tobytes())
            hits += bool(p_ok and p_xy == want)
    assert hits > 300  # about a quarter of the valid signatures were asked with their true recovery id
    tweaks = np.array([np.frombuffer(hashlib.sha256(b"tw%d" % i).digest(), np.uint8) for i in range(3000)])
    xy, ok = engine.pubkey_tweak_add_batch(pub, tweaks)
    for i in range(3000):
        plen = 33 if pub[i, 0] in (2, 3) else 65
        want = oracle.pubkey_tweak_add(pub[i, :plen].tobytes(), tweaks[i].tobytes()) if pub[i, 0] in (2, 3, 4, 6, 7) else None
        assert (want is not None) == bool(ok[i]), i
        if want is not None:
            assert xy[i].tobytes() == want, i
    # empty batch
    xy, ok = engine.ecdsa_recover_batch(np.zeros((0, 32), np.uint8), np.zeros((0, 64), np.uint8), np.zeros(0, np.uint8))
    assert len(ok) == 0


def test_exact_walk_behind_the_quick_one(oracle, tmp_path):
    """k_ecdsa_point walks a signature over branch-free field forms that count the practically-never carry instead of
    finishing it, and walks it again over the exact forms when the count is not zero (about two signatures in a
    million).  BCP_POINT_FORCE_REDO=3 sends every third signature through that second walk (the variable is read once
    per process, hence the child process); the verdicts must not change."""
    import subprocess
    import sys

    msg, pub, sig, kinds = synth.sigset(30000, nkeys=256, corrupt_pct=10)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, nthreads=NT)
    np.savez(tmp_path / "in.npz", msg=msg, pub=pub, sig=sig)
    child = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "import bitcoincashplus_b200 as bcp\n"
        "z = np.load(%r)\n"
        "got = bcp.Engine(0).ecdsa_verify_batch(z['msg'], z['pub'], z['sig'])\n"
        "np.save(%r, got)\n" % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), str(tmp_path / "in.npz"), str(tmp_path / "out.npy"))
    )
    env = dict(os.environ, BCP_POINT_FORCE_REDO="3")
    subprocess.check_call([sys.executable, "-c", child], env=env)
    got = np.load(tmp_path / "out.npy")
    assert (got == want).all(), np.nonzero(got != want)[0][:20]
