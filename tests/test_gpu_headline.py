"""-m gpu: the BASELINE.json configurations at their FULL sizes, every verdict compared with the CPU oracle.

The other GPU tests cover corner cases at sizes the oracle finishes instantly; these run the launch shapes the bench
measures (1,000,000 signatures = 13.2 waves of the point kernel in slices; a 41,650-input block = the sub-wave launch
with 128-thread blocks and K=4 prep; 10,000 Equihash (200,9) headers in four chunks on two streams) with corrupted
entries in every part of the index space, and compare every single output with oracle/ (a few seconds of host time
on all cores)."""
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


def test_config2_one_million_signatures_every_verdict(engine, oracle):
    """configs[1]: 1M synthetic signatures incl. 1 % corrupted (all 20 classes), host-pointer C-ABI call (chunked over two
    streams) AND the device-pointer call (one launch sequence over the slices): both equal the oracle on all 1M."""
    import torch

    n = 1000000
    msg, pub, sig, kinds = synth.sigset(n, corrupt_pct=1)
    want = oracle.ecdsa_verify_batch(msg, pub, sig, nthreads=NT)
    got = engine.ecdsa_verify_batch(msg, pub, sig)
    assert (got == want).all(), np.nonzero(got != want)[0][:20]
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        d = [torch.from_numpy(a).to(dev) for a in (msg, pub, sig)]
        d_ok = torch.full((n,), 0xEE, dtype=torch.uint8, device=dev)
        st.synchronize()
        engine.ecdsa_verify_batch_dev(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, d_ok.data_ptr(), 0, st.cuda_stream)
        st.synchronize()
        got_dev = d_ok.cpu().numpy()
    assert (got_dev == want).all(), np.nonzero(got_dev != want)[0][:20]
    bad = int((want == 0).sum())
    assert 8000 < bad < 12000 and set(kinds[kinds >= 0]) == set(range(len(synth.CORRUPTIONS)))


def _full_block_with_mutations(ntx=41650):
    kinds = list(synth.INPUT_MUTATIONS)
    rng = np.random.default_rng(41650)
    mut = {}
    # every mutation kind four times, spread over the whole index range (first and last thread blocks included) ...
    for rep in range(4):
        for i, k in enumerate(kinds):
            pos = (rep * ntx // 4 + 13 + 97 * i) % ntx
            mut[pos] = k
    mut[0] = kinds[0]
    mut[ntx - 1] = "bad_sig"
    # ... and 1 % plain bad signatures at random positions
    for pos in rng.choice(ntx, size=ntx // 100, replace=False):
        mut.setdefault(int(pos), "bad_sig")
    return synth.synthetic_block(ntx, True, with_prevouts=True, mutations=mut), mut


def test_config3_full_8mb_block_every_mutation_every_input(engine, oracle):
    """configs[2]: the 41,650-input block (the sub-wave launch shape: K = 4 prep, 128-thread point blocks) with every
    INPUT_MUTATIONS kind and > 1 % bad signatures: ScriptError of every input == the oracle's VerifyScript restatement;
    txids, Merkle root and the mutated flag too.  The same block as one entry of a window of two goes through the
    window call (256-thread point blocks over the union) with the same result."""
    (blk, spk, amt), mut = _full_block_with_mutations()
    want = oracle.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, NT)
    assert int((want != 0).sum()) >= len(spk) // 100
    res = engine.block_connect_p2pkh(blk, spk, amt, synth.POSTFORK_FLAGS, want_txids=True)
    assert (res["script_err"] == want).all(), np.nonzero(res["script_err"] != want)[0][:20]
    assert not res["all_ok"]
    for i, k in mut.items():
        exp = synth.INPUT_MUTATIONS[k]
        assert res["script_err"][i] == (255 if exp is None else exp), (i, k)
    n, txids, root, mutated = oracle.block_txids(blk, False, max_tx=len(spk) + 8)
    assert n == len(spk) + 1 and res["root"] == root == blk[36:68] and res["mutated"] == mutated
    assert (res["txids"][:n] == txids[:n]).all()
    # window of two: the damaged block and an all-valid one
    blk2, spk2, amt2 = synth.synthetic_block(3000, True, nkeys=64, with_prevouts=True)
    cat = np.frombuffer(b"".join(spk) + b"".join(spk2), dtype=np.uint8)
    lens = [len(x) for x in spk] + [len(x) for x in spk2]
    off = np.zeros(len(lens) + 1, dtype=np.uint32)
    off[1:] = np.cumsum(lens)
    am = np.array(list(amt) + list(amt2), dtype=np.int64)
    w = engine.blocks_connect_p2pkh([np.frombuffer(blk, dtype=np.uint8), np.frombuffer(blk2, dtype=np.uint8)], cat, off, am,
                                    [len(spk), len(spk2)], synth.POSTFORK_FLAGS)
    errs = np.asarray(w["script_err"])
    assert (errs[: len(spk)] == want).all() and not errs[len(spk):].any()
    assert list(w["all_ok"]) == [0, 1] and bytes(w["roots"][0]) == root


def test_config4_ten_thousand_headers_every_verdict_and_hash(engine, oracle):
    """configs[3]: 10,000 Equihash (200,9) headers — the solver-found fixture solutions tiled, half of them with one
    flipped solution bit, plus structural damage (wrong index order, duplicate indices) — every verdict and every
    header hash compared with the oracle."""
    hdrs = np.load(os.path.join(os.path.dirname(__file__), "golden", "equihash_200_9.npz"))["headers"]
    rng = np.random.default_rng(10000)
    allh = np.tile(hdrs, ((10000 + len(hdrs) - 1) // len(hdrs), 1))[:10000].copy()
    corrupt = rng.random(10000) < 0.5
    for i in np.nonzero(corrupt)[0]:
        allh[i, 143 + int(rng.integers(0, 1344))] ^= 1 << int(rng.integers(0, 8))
    # nonce / time changes on otherwise valid solutions (the base state changes: every leaf hash differs)
    for i in np.nonzero(~corrupt)[0][::7]:
        allh[i, 100 + int(rng.integers(0, 40))] ^= 0x10
    # structural damage that keeps the solution well-formed: sibling subtrees swapped at a random level (ordering rule),
    # one index duplicated (distinctness rule)
    from test_gpu_equihash import _pack, _unpack

    for c, i in enumerate(np.nonzero(~corrupt)[0][3::11]):
        idx = _unpack(allh[i, 143:].tobytes(), 21)
        if c % 2 == 0:
            w = 1 << int(rng.integers(0, 9))
            a = int(rng.integers(0, 512 // (2 * w))) * 2 * w
            idx[a:a + w], idx[a + w:a + 2 * w] = idx[a + w:a + 2 * w], idx[a:a + w]
        else:
            idx[int(rng.integers(0, 512))] = idx[int(rng.integers(0, 512))]
        allh[i, 143:] = np.frombuffer(_pack(idx, 21), dtype=np.uint8)
    got, hashes = engine.equihash_verify_batch(200, 9, allh)
    want = oracle.equihash_verify_batch(200, 9, allh, nthreads=NT)
    assert (got.astype(bool) == want.astype(bool)).all(), np.nonzero(got.astype(bool) != want.astype(bool))[0][:20]
    assert 1000 < int(want.astype(bool).sum()) < 9000
    for i in range(10000):
        assert hashes[i].tobytes() == oracle.header_hash(allh[i].tobytes()), i
