"""Smallest end-to-end pass over every kernel, for compute-sanitizer (memcheck): python tools/sanitize_small.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bitcoincashplus_b200 as bcp
import synth
from oracle import pyoracle as po
o = po.oracle()
eng = bcp.Engine(0)
msg, pub, sig, kinds = synth.sigset(700, nkeys=16, corrupt_pct=20)
ok = eng.ecdsa_verify_batch(msg, pub, sig)
assert (ok == o.ecdsa_verify_batch(msg, pub, sig, 4)).all()
xy, pok = eng.pubkey_parse_batch(pub)
mut = {3 + 2 * i: k for i, k in enumerate(synth.INPUT_MUTATIONS)}
blk, spk, amt = synth.synthetic_block(60, True, nkeys=8, with_prevouts=True, mutations=mut)
res = eng.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS, want_txids=True)
assert (res["script_err"] == o.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, 4)).all()
for n in (0, 1, 5, 300, 1000):
    lv = np.random.default_rng(n).integers(0, 256, size=(n, 32), dtype=np.uint8)
    assert eng.merkle_root(lv) == o.merkle_root(lv)
raw = open(os.path.join(ROOT, "tests", "golden", "block413567.raw"), "rb").read()
off = bcp.block_index(raw, legacy_header=True)
txids, root, mutd = eng.txids_merkle(raw, off)
assert root == raw[36:68]
for name, n, k in (("equihash_48_5.npz", 48, 5), ("equihash_200_9.npz", 200, 9)):
    h = np.load(os.path.join(ROOT, "tests", "golden", name))["headers"][:20]
    okh, hh = eng.equihash_verify_batch(n, k, h)
    assert okh.all()
print("sanitize_small ok")
