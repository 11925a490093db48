"""Generates the committed fixtures under tests/golden/ from the REFERENCE itself.

Runs only in the build container (needs /root/reference and oracle/_ref/libbcp_ref.so); the outputs
travel with the repo so that the GPU box, which has neither, can check parity against them.

  block413567.raw / .json   the reference's own benchmark block (src/bench/data/block413567.raw, a real
                            mainnet block used by src/bench/checkblock.cpp) + txid digest, Merkle root and
                            block hash computed by the reference's CBlock / BlockMerkleRoot / CHashWriter
  equihash_<N>_<K>.npz      serialized headers whose solutions were FOUND by the reference's solver
                            (Equihash<N,K>::BasicSolve / OptimisedSolve, src/crypto/equihash.cpp) and
                            accepted by its IsValidSolution
  ecdsa_ref.npz             signature set with every corruption class + the reference's verdicts
                            (secp256k1_ecdsa_verify path and CPubKey::Verify path)
  recovery_ref.npz          public-key recovery and tweak-add cases (random, corrupted, and the hard-coded edge vectors of
                            src/secp256k1/src/modules/recovery/tests_impl.h:208-258) with the results of the reference's
                            secp256k1_ecdsa_recover / secp256k1_ec_pubkey_tweak_add, plus CPubKey::RecoverCompact and
                            CPubKey::Derive (src/pubkey.cpp:195-264) outputs for the B3 compat test
  sha256_kat.json           src/test/crypto_tests.cpp:282-316 SHA-256 known answers re-derived with the
                            reference's CSHA256 (and hashlib as a third opinion)
  secp_vectors.json         hard-coded vectors of src/secp256k1/src/tests.c (pubkey parse valid/invalid
                            :2787-3111, ECDSA edge cases :3973-4340) extracted by tests/golden/extract_tests_c.py
usage: python tests/golden/make_golden.py [--eh200 COUNT] [--procs P]
"""
import argparse
import hashlib
import json
import multiprocessing as mp
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"


def gen_block():
    from oracle import pyoracle as po

    r = po.ref()
    src = os.path.join(REF, "src/bench/data/block413567.raw")
    raw = open(src, "rb").read()
    shutil.copyfile(src, os.path.join(HERE, "block413567.raw"))
    n, txids, root, mut = r.block_txids(raw, True)
    meta = {"source": "src/bench/data/block413567.raw", "bytes": len(raw), "ntx": int(n), "merkle_root": root.hex(),
            "mutated": bool(mut), "txids_sha256": hashlib.sha256(txids.tobytes()).hexdigest(),
            "block_hash_raw": r.header_hash(raw[:80], legacy=True).hex()}
    json.dump(meta, open(os.path.join(HERE, "block413567.json"), "w"), indent=1)
    print("block413567:", meta)


def _solve_one(args):
    from oracle import pyoracle as po
    import synth

    n, k, tag = args
    r = po.ref()
    hdr = synth.synthetic_header(nonce_tag=tag, n=n, k=k)
    sols = r.equihash_solve(n, k, hdr[:140], optimised=False, max_solns=8)
    out = []
    for s in sols:
        h = hdr[:140] + synth._compact(len(s)) + s
        assert r.equihash_verify(n, k, h) == 1
        out.append(h)
    return out


def gen_equihash(n, k, want, procs, start_tag=0):
    name = os.path.join(HERE, "equihash_%d_%d.npz" % (n, k))
    hdrs = []
    tag = start_tag
    with mp.Pool(procs) as pool:
        while len(hdrs) < want:
            batch = [(n, k, tag + i) for i in range(procs)]
            tag += procs
            for res in pool.map(_solve_one, batch):
                hdrs.extend(res)
            print("equihash(%d,%d): %d solutions after %d nonces" % (n, k, len(hdrs), tag - start_tag), flush=True)
    arr = np.stack([np.frombuffer(h, dtype=np.uint8) for h in hdrs[:want]])
    np.savez_compressed(name, headers=arr)
    print("wrote", name, arr.shape)


def gen_ecdsa():
    from oracle import pyoracle as po
    import synth

    r = po.ref()
    msg, pub, sig, kinds = synth.sigset(4000, nkeys=64, corrupt_pct=25)
    m2, p2, s2 = synth.edge_cases()
    msg, pub, sig = np.concatenate([msg, m2]), np.concatenate([pub, p2]), np.concatenate([sig, s2])
    kinds = np.concatenate([kinds, np.full(len(m2), -2, dtype=np.int32)])
    raw = r.ecdsa_verify_batch(msg, pub, sig, 8)
    # CPubKey::Verify verdicts on a hand-rolled DER of each (r, s)
    cpk = np.zeros(len(msg), dtype=np.uint8)
    ders = []
    for i in range(len(msg)):
        rr, ss = int.from_bytes(sig[i, :32].tobytes(), "big"), int.from_bytes(sig[i, 32:].tobytes(), "big")

        def enc(v):
            b = v.to_bytes(33, "big").lstrip(b"\x00") or b"\x00"
            if b[0] & 0x80:
                b = b"\x00" + b
            return b"\x02" + bytes([len(b)]) + b
        body = enc(rr) + enc(ss)
        der = b"\x30" + bytes([len(body)]) + body
        plen = 33 if pub[i, 0] in (2, 3) else 65
        cpk[i] = r.pubkey_verify(msg[i].tobytes(), pub[i, :plen].tobytes(), der)
    np.savez_compressed(os.path.join(HERE, "ecdsa_ref.npz"), msg=msg, pub=pub, sig=sig, kinds=kinds, verdict_raw=raw,
                        verdict_cpubkey=cpk)
    print("ecdsa_ref:", len(msg), "raw valid", int(raw.sum()), "cpubkey valid", int(cpk.sum()))


def gen_recovery():
    from oracle import pyoracle as po
    import synth

    r = po.ref()
    N = synth.N
    rec = []  # (msg32, sig64, recid)
    for i in range(500):
        k = synth.seckey(20000 + i)
        m = hashlib.sha256(b"recover-msg%d" % i).digest()
        sig, rid = r.sign_recoverable(k, m)
        for rr in range(4):
            rec.append((m, sig, rr))
        if i % 5 == 0:  # damaged: a flipped bit in r, s or the message
            bit = synth.prng(i, b"recbit")
            b = bytearray(sig)
            b[(bit >> 3) % 64] ^= 1 << (bit & 7)
            rec.append((m, bytes(b), rid))
            mm = bytearray(m)
            mm[(bit >> 11) % 32] ^= 1 << ((bit >> 8) & 7)
            rec.append((bytes(mm), sig, rid))
    # edge vectors of tests_impl.h:208-258: a signature made with secret key 0 (recoverable only with recid 1) and
    # (r, s) = (4, 4), which every recid recovers
    msg_e = b"This is a very secret message..."
    sig_e = bytes.fromhex("67CB285F9CD194E840D629397AF5569662FDE446499959631 79A7DD17BD23532".replace(" ", "")
                          + "4B1B7DF34CE1F68E694FF6F11AC751DD7DD73E387EE4FC866E1BE8ECC7DD9557")
    sig_44 = (4).to_bytes(32, "big") * 2
    for rr in range(4):
        rec.append((msg_e, sig_e, rr))
        rec.append((msg_e, sig_44, rr))
    # range rules: r = 0, s = 0, r = n, s = n, r just below p - n with recid 2/3 (x = r + n < p), r = p - n (refused)
    pmn = synth.P - N
    m0 = hashlib.sha256(b"range").digest()
    for rv, sv in ((0, 5), (5, 0), (N, 5), (5, N), (pmn - 1, 7), (pmn, 7), (pmn + 1, 7), (N - 1, N - 1), (1, 1)):
        for rr in range(4):
            rec.append((m0, rv.to_bytes(32, "big") + sv.to_bytes(32, "big"), rr))
    rec_msg = np.array([np.frombuffer(a, np.uint8) for a, _, _ in rec])
    rec_sig = np.array([np.frombuffer(b, np.uint8) for _, b, _ in rec])
    rec_id = np.array([c for _, _, c in rec], dtype=np.uint8)
    rec_xy = np.zeros((len(rec), 64), np.uint8)
    rec_ok = np.zeros(len(rec), np.uint8)
    for i, (m, sg, rr) in enumerate(rec):
        xy = r.ecdsa_recover(m, sg, rr)
        if xy is not None:
            rec_ok[i] = 1
            rec_xy[i] = np.frombuffer(xy, np.uint8)
    # the two tests_impl.h expectations, asserted here so the fixture cannot silently drift from them
    base = len(rec) - 36 - 8
    assert [int(rec_ok[base + 2 * rr]) for rr in range(4)] == [0, 1, 0, 0]
    assert all(rec_ok[base + 2 * rr + 1] for rr in range(4))
    tw = []  # (pub slot 65, tweak32)
    for i in range(400):
        k = synth.seckey(30000 + i)
        d = int.from_bytes(k, "big")
        pub = r.pubkey_create(k, i % 3 != 0)
        slot = pub + bytes(65 - len(pub))
        t = hashlib.sha256(b"tweak%d" % i).digest()
        tw.append((slot, t))
        if i % 10 == 0:
            for tv in (0, 1, N - 1, N, N + 1, 2**256 - 1, (N - d) % N, (N - d + 1) % N):
                tw.append((slot, tv.to_bytes(32, "big")))
        if i % 25 == 0:  # keys that do not parse
            tw.append((bytes([5]) + slot[1:], t))
            tw.append((bytes([2]) + (synth.P).to_bytes(32, "big") + bytes(32), t))
    tw_pub = np.array([np.frombuffer(a, np.uint8) for a, _ in tw])
    tw_t = np.array([np.frombuffer(b, np.uint8) for _, b in tw])
    tw_xy = np.zeros((len(tw), 64), np.uint8)
    tw_ok = np.zeros(len(tw), np.uint8)
    for i, (slot, t) in enumerate(tw):
        plen = 33 if slot[0] in (2, 3) else 65
        xy = r.pubkey_tweak_add(slot[:plen], t)
        if xy is not None:
            tw_ok[i] = 1
            tw_xy[i] = np.frombuffer(xy, np.uint8)
    # CPubKey level (src/pubkey.cpp:195-264): RecoverCompact on 65-byte compact signatures, Derive on compressed keys
    rc_hash, rc_sig, rc_out, rc_len = [], [], [], []
    for i in range(60):
        k = synth.seckey(40000 + i)
        m = hashlib.sha256(b"compact%d" % i).digest()
        sig, rid = r.sign_recoverable(k, m)
        hdr = 27 + rid + (4 if i % 2 else 0)
        s65 = bytes([hdr]) + sig
        if i % 7 == 3:
            s65 = bytes([hdr]) + bytes([sig[0] ^ 0x40]) + sig[1:]
        out = r.cpubkey_recover_compact(m, s65)
        rc_hash.append(np.frombuffer(m, np.uint8)); rc_sig.append(np.frombuffer(s65, np.uint8))
        rc_len.append(len(out) if out else 0)
        rc_out.append(np.frombuffer((out or b"") + bytes(65 - len(out or b"")), np.uint8))
    dv_pub, dv_cc, dv_n, dv_out, dv_ccout = [], [], [], [], []
    for i in range(60):
        pub = r.pubkey_create(synth.seckey(50000 + i), True)
        cc = hashlib.sha256(b"chaincode%d" % i).digest()
        res = r.cpubkey_derive(pub, cc, i * 7919)
        assert res is not None
        dv_pub.append(np.frombuffer(pub, np.uint8)); dv_cc.append(np.frombuffer(cc, np.uint8)); dv_n.append(i * 7919)
        dv_out.append(np.frombuffer(res[0], np.uint8)); dv_ccout.append(np.frombuffer(res[1], np.uint8))
    np.savez_compressed(os.path.join(HERE, "recovery_ref.npz"), rec_msg=rec_msg, rec_sig=rec_sig, rec_id=rec_id, rec_xy=rec_xy,
                        rec_ok=rec_ok, tw_pub=tw_pub, tw_tweak=tw_t, tw_xy=tw_xy, tw_ok=tw_ok, rc_hash=np.array(rc_hash),
                        rc_sig=np.array(rc_sig), rc_out=np.array(rc_out), rc_len=np.array(rc_len, np.uint8),
                        dv_pub=np.array(dv_pub), dv_cc=np.array(dv_cc), dv_n=np.array(dv_n, np.uint32), dv_out=np.array(dv_out),
                        dv_ccout=np.array(dv_ccout))
    print("recovery_ref: recover", len(rec), "ok", int(rec_ok.sum()), "| tweak_add", len(tw), "ok", int(tw_ok.sum()),
          "| RecoverCompact", len(rc_len), "ok", sum(1 for x in rc_len if x), "| Derive", len(dv_n))


def gen_sha_kat():
    from oracle import pyoracle as po

    r = po.ref()
    msgs = [b"", b"abc", b"message digest", b"secure hash algorithm", b"SHA256 is considered to be safe",
            b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq",
            b"For this sample, this 63-byte string will be used as input data",
            b"This is exactly 64 bytes long, not counting the terminating byte", b"a" * 1000000]
    kat = []
    for m in msgs:
        d = r.sha256(m)
        assert d == hashlib.sha256(m).digest()
        kat.append({"len": len(m), "msg_hex": m.hex() if len(m) < 200 else None, "repeat_a": len(m) if len(m) >= 200 else 0,
                    "sha256": d.hex(), "sha256d": r.sha256d(m).hex(), "hash160": r.hash160(m).hex()})
    json.dump(kat, open(os.path.join(HERE, "sha256_kat.json"), "w"), indent=1)
    print("sha256_kat:", len(kat))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--eh200", type=int, default=0)
    ap.add_argument("--procs", type=int, default=6)
    ap.add_argument("--skip-small", action="store_true")
    ap.add_argument("--only-recovery", action="store_true")
    a = ap.parse_args()
    if a.only_recovery:
        gen_recovery()
        sys.exit(0)
    if not a.skip_small:
        gen_recovery()
        gen_block()
        gen_sha_kat()
        gen_ecdsa()
        gen_equihash(48, 5, 256, a.procs)
        gen_equihash(96, 5, 128, a.procs)
    if a.eh200:
        gen_equihash(200, 9, a.eh200, a.procs)
