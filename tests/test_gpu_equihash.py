"""-m gpu: Equihash verification + header hashing vs the oracle on the committed solution fixtures."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    z = np.load(os.path.join(GOLD, name))
    return z["headers"]


def _corruptions(hdrs, rng, per=6):
    """Deterministic single-field corruptions of valid headers: bit flips in the solution, in the nonce,
    in nTime; swapped sibling halves; a duplicated index."""
    out = []
    L = hdrs.shape[1]
    for h in hdrs:
        for t in range(per):
            b = h.copy()
            if t == 0:
                pos = 143 + int(rng.integers(0, L - 143)) if L > 200 else 141 + int(rng.integers(0, L - 141))
                b[pos] ^= 1 << int(rng.integers(0, 8))
            elif t == 1:
                b[108 + int(rng.integers(0, 32))] ^= 1
            elif t == 2:
                b[100] ^= 1
            elif t == 3:
                s0 = L - (L - 140 - (3 if L > 400 else 1))
                half = (L - s0) // 2
                b[s0:s0 + half], b[s0 + half:] = h[s0 + half:].copy(), h[s0:s0 + half].copy()
            elif t == 4:
                b[140] ^= 1  # compact-size prefix no longer says SolutionWidth
            else:
                b[L - 1] ^= 0x80
            out.append(b)
    return np.stack(out)


@pytest.mark.parametrize("n,k,fixture", [(48, 5, "equihash_48_5.npz"), (96, 5, "equihash_96_5.npz"),
                                         (200, 9, "equihash_200_9.npz")])
def test_equihash_valid_and_corrupted(engine, oracle, n, k, fixture):
    hdrs = _load(fixture)
    ok, hashes = engine.equihash_verify_batch(n, k, hdrs)
    assert ok.all(), "fixture solutions were produced by the reference's own solver"
    for i in range(min(len(hdrs), 64)):
        assert hashes[i].tobytes() == oracle.header_hash(hdrs[i].tobytes())
    rng = np.random.default_rng(n * 100 + k)
    bad = _corruptions(hdrs[:64], rng)
    allh = np.concatenate([hdrs, bad])
    got, _ = engine.equihash_verify_batch(n, k, allh, want_hash=False)
    want = oracle.equihash_verify_batch(n, k, allh, nthreads=os.cpu_count() or 1)
    assert (got == want).all(), np.nonzero(got != want)[0][:20]
    assert want[: len(hdrs)].all() and not want[len(hdrs):].all()


def test_equihash_10k_headers(engine, oracle):
    """BASELINE config 4 shape: 10,000 (200,9) headers = the valid fixture solutions tiled with deterministic
    corruptions; checked against the oracle on a sample and by construction on the rest."""
    hdrs = _load("equihash_200_9.npz")
    rng = np.random.default_rng(4)
    reps = (10000 + len(hdrs) - 1) // len(hdrs)
    allh = np.tile(hdrs, (reps, 1))[:10000].copy()
    corrupt = rng.random(10000) < 0.5
    for i in np.nonzero(corrupt)[0]:
        allh[i, 143 + int(rng.integers(0, 1344))] ^= 1 << int(rng.integers(0, 8))
    got, hashes = engine.equihash_verify_batch(200, 9, allh)
    assert got[~corrupt].all()
    idx = np.arange(0, 10000, 37)
    want = oracle.equihash_verify_batch(200, 9, allh[idx].copy(), nthreads=os.cpu_count() or 1)
    assert (got[idx] == want).all()
    for i in idx[:50]:
        assert hashes[i].tobytes() == oracle.header_hash(allh[i].tobytes())


def _unpack(soln, bits):
    v = int.from_bytes(soln, "big")
    n = len(soln) * 8 // bits
    return [(v >> (bits * (n - 1 - i))) & ((1 << bits) - 1) for i in range(n)]


def _pack(idx, bits):
    v = 0
    for x in idx:
        v = (v << bits) | x
    return v.to_bytes(len(idx) * bits // 8, "big")


@pytest.mark.parametrize("n,k,fixture", [(48, 5, "equihash_48_5.npz"), (200, 9, "equihash_200_9.npz")])
def test_equihash_structural_corruptions(engine, oracle, n, k, fixture):
    """Index-level corruptions that keep the solution well-formed: duplicated index, swapped siblings at every
    tree level (ordering rule), index +-1 (collision rule), all-equal indices, indices = 0..2^k-1."""
    hdrs = _load(fixture)[:6]
    bits = n // (k + 1) + 1
    pre = hdrs.shape[1] - (1 << k) * bits // 8
    out = []
    for h in hdrs:
        idx = _unpack(h[pre:].tobytes(), bits)
        variants = []
        v = list(idx); v[1] = v[0]; variants.append(v)                              # duplicate
        v = list(idx); v[-1] = v[0]; variants.append(v)                             # duplicate across the root
        for lvl in range(k):                                                        # swap the first sibling pair at each level
            w = 1 << lvl
            v = list(idx); v[0:w], v[w:2 * w] = idx[w:2 * w], idx[0:w]; variants.append(v)
        v = list(idx); v[5] ^= 1; variants.append(v)                                # neighbour index: collision fails
        variants.append([idx[0]] * len(idx))
        variants.append(list(range(len(idx))))
        variants.append(sorted(idx))
        for v in variants:
            b = h.copy()
            b[pre:] = np.frombuffer(_pack(v, bits), dtype=np.uint8)
            out.append(b)
    allh = np.stack(out)
    got, _ = engine.equihash_verify_batch(n, k, allh, want_hash=False)
    want = oracle.equihash_verify_batch(n, k, allh, nthreads=os.cpu_count() or 1)
    assert (got == want).all(), np.nonzero(got != want)[0]
    assert not want.any()


def _solve_48_5_allowing_duplicates(input140, max_out=8):
    """Wagner's algorithm for (48,5) like synth.equihash_solve_48_5, but WITHOUT the "index sets must be disjoint" filter:
    returns index lists whose XOR conditions and ordering all hold and in which some index occurs twice — inputs that
    only the distinctness rule (DistinctIndices, src/crypto/equihash.cpp:742-748) rejects."""
    import hashlib

    pers = b"ZcashPoW" + (48).to_bytes(4, "little") + (5).to_bytes(4, "little")
    rows = []
    for g in range(52):
        d = hashlib.blake2b(input140 + g.to_bytes(4, "little"), digest_size=60, person=pers).digest()
        for j in range(10):
            i = g * 10 + j
            if i < 512:
                rows.append((d[6 * j:6 * j + 6], (i,)))
    for rnd in range(5):
        width = 2 if rnd == 4 else 1
        buckets = {}
        for h, idx in rows:
            buckets.setdefault(h[:width], []).append((h, idx))
        nxt = []
        for grp in buckets.values():
            for a in range(len(grp)):
                for b in range(a + 1, len(grp)):
                    (ha, ia), (hb, ib) = grp[a], grp[b]
                    if ia[0] == ib[0]:
                        continue  # the ordering rule needs different first indices
                    if rnd < 3 and set(ia) & set(ib):
                        continue  # keep the early rounds clean so that the table stays small
                    x = bytes(p ^ q for p, q in zip(ha[width:], hb[width:]))
                    nxt.append((x, ia + ib if ia[0] < ib[0] else ib + ia))
        rows = nxt
    out = []
    for h, idx in rows:
        if len(set(idx)) < 32:
            out.append(idx)
            if len(out) >= max_out:
                break
    return out


def test_equihash_duplicate_indices_are_the_only_defect(engine, oracle):
    """Solutions whose XOR chunks vanish at every level and whose sub-trees are ordered, but in which an index occurs
    twice: only the distinctness check stands between them and acceptance."""
    import synth

    found = []
    nonce = 0
    while len(found) < 12 and nonce < 400:
        hdr = synth.synthetic_header(height=700000, merkle_root=bytes(32), nonce_tag=(77 << 20) | nonce, n=48, k=5)
        for idx in _solve_48_5_allowing_duplicates(hdr[:140]):
            v = 0
            for x in idx:
                v = (v << 9) | x
            found.append(hdr[:140] + b"\x24" + v.to_bytes(36, "big"))
        nonce += 1
    assert len(found) >= 4, "no duplicate-index pseudo-solutions found"
    # and genuine solutions of the same headers' family, as the control group
    good = []
    nonce = 0
    while len(good) < 4:
        hdr = synth.synthetic_header(height=700000, merkle_root=bytes(32), nonce_tag=(78 << 20) | nonce, n=48, k=5)
        for sol in synth.equihash_solve_48_5(hdr[:140])[:1]:
            good.append(hdr[:140] + b"\x24" + sol)
        nonce += 1
    arr = np.frombuffer(b"".join(found + good), dtype=np.uint8).reshape(len(found) + len(good), 177)
    ok, _ = engine.equihash_verify_batch(48, 5, arr)
    want = oracle.equihash_verify_batch(48, 5, arr, 2)
    assert (ok.astype(bool) == (want == 1)).all()
    assert not ok[:len(found)].any() and ok[len(found):].all()
