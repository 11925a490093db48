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
    a = ap.parse_args()
    if not a.skip_small:
        gen_block()
        gen_sha_kat()
        gen_ecdsa()
        gen_equihash(48, 5, 256, a.procs)
        gen_equihash(96, 5, 128, a.procs)
    if a.eh200:
        gen_equihash(200, 9, a.eh200, a.procs)
