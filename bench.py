#!/usr/bin/env python
"""bench.py — throughput of the hot path on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--sigs S] [--impl ours|reference]

Workload (BASELINE.json configs[1]): single-GPU batched secp256k1_ecdsa_verify on 1,000,000 synthetic
signatures incl. 1 % corrupted.  A "step" is one pass of the two ECDSA kernels over the whole batch.

  value   sigverifies/s with inputs resident in HBM (CUDA events around K steps on the launching stream)
  e2e     the same metric through the host-pointer C-ABI call bcp_gpu_ecdsa_verify_batch on PINNED host
          buffers: host->device copies, kernels and the verdict read-back are all inside the timed region
  roofline  bound = integer pipe: achieved = sigverifies/s x 151,228 mul32 (SURVEY.md §8d unit of the
            reference algorithm) vs the 32x32->64 multiply-add peak measured live by
            bcp_gpu_measure_int_peak (MEASURED_PEAKS.json carries no integer peak); the HBM view
            (130 B/verify vs MEASURED_PEAKS hbm_gbs) is reported alongside as `hbm`
  cpu_baseline  the reference's own libsecp256k1 (oracle/_ref, kind "reference"; falls back to the
                oracle port) on all host threads over a bounded sample of the same batch
Multi-GPU: one process per GPU under torchrun; signatures shard by index range across ranks, no
collective on the data path (verdicts are AND-ed / concatenated by the caller); weak scaling.
`--impl reference` times the reference CPU path only (rank 0), same metric/config.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MUL32_PER_VERIFY = 151228  # 64*(1348+39) + 36*(1482+253)
BYTES_PER_VERIFY = 32 + 33 + 64 + 1


def workload_config(n):
    """The `config` object, identical in both arms (`--impl ours` and `--impl reference`) so the driver can match them."""
    return {"workload": "configs[1]: batched secp256k1_ecdsa_verify, %d synthetic signatures per GPU incl. 1%% corrupted" % n,
            "sigs_per_gpu": n, "sharding": "by signature index, one process per GPU, no collective",
            "l2_policy": "inputs + scratch touched per step = %.0f MB (> 126 MB L2 at the default size)"
                         % ((n * (161 + 80 + 768 + 1)) / 1e6)}


def effective_cpus():
    """cores this process may really use: the affinity mask and the cgroup CPU quota, not just os.cpu_count()"""
    n = os.cpu_count() or 1
    try:
        n = min(n, len(os.sched_getaffinity(0)))
    except Exception:
        pass
    try:
        quota, period = open("/sys/fs/cgroup/cpu.max").read().split()
        if quota != "max":
            n = max(1, min(n, int(int(quota) / int(period))))
    except Exception:
        pass
    return n


def clocks_sampler(stop, out, gpu_index):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    while not stop.is_set():
        try:
            r = subprocess.run(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            out.append([x.strip() for x in r])
        except Exception:
            pass
        stop.wait(0.2)


def summarize_clocks(samples):
    sm = sorted(int(s[0]) for s in samples if s and s[0].isdigit())
    mx = max([int(s[1]) for s in samples if len(s) > 1 and s[1].isdigit()] or [0])
    reasons = set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for s in samples:
        for k, nm in enumerate(names):
            if len(s) > 2 + k and s[2 + k].lower().startswith("active"):
                reasons.add(nm)
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
            "samples": len(sm)}


def cpu_baseline(msg, pub, sig, budget_s=12.0):
    """Reference libsecp256k1 (or the oracle port) on all host threads over a bounded sample.  Returns the JSON object
    and the verdicts of the sample (the first `len(verdicts)` signatures of the batch), which the caller compares
    with the GPU's verdicts for the same signatures outside the timed region."""
    from oracle import pyoracle as po

    impl = po.ref() if po.have_ref() else po.oracle()
    cores = os.cpu_count() or 1
    n = len(msg)
    sample = min(n, 2000 * cores)
    impl.ecdsa_verify_batch(msg[:256 * min(cores, 8)].copy(), pub[:256 * min(cores, 8)].copy(),
                            sig[:256 * min(cores, 8)].copy(), cores)  # warm-up (context/tables)
    t0 = time.perf_counter()
    verdicts = impl.ecdsa_verify_batch(msg[:sample], pub[:sample], sig[:sample], cores)
    dt = time.perf_counter() - t0
    # grow the sample until ~budget_s of wall time is spent (bounded by the batch)
    target = min(n, int(sample * max(1.0, budget_s / max(dt, 1e-3))))
    if target > sample * 1.5:
        sample = target
        t0 = time.perf_counter()
        verdicts = impl.ecdsa_verify_batch(msg[:sample], pub[:sample], sig[:sample], cores)
        dt = time.perf_counter() - t0
    return {"value": sample / dt, "unit": "sigverifies/s", "cores": cores, "kind": impl.kind,
            "sample": "first %d of the %d-signature batch, %.1f s wall, %d threads" % (sample, n, dt, cores)}, verdicts


def traffic_from_profile(n):
    """dram__bytes_read.sum + dram__bytes_write.sum of k_ecdsa_point for one launch, from the committed ncu --set full
    capture (profiles/ecdsa_point_traffic.json, taken at the default batch size); None for other sizes."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ecdsa_point_traffic.json")))
        return t["dram_bytes_per_launch"] if t.get("sigs") == n else None
    except Exception:
        return None


def block_bench(eng, args, bcp, sync=None):
    """`sync` (multi-GPU runs): callable(seconds) -> max over ranks, entered by every rank; each rank validates its own
    copy of the block on its own GPU (weak scaling, no collective on the data path) and rank 0 adds the CPU side.
    Secondary measurement (BASELINE.json configs[2], "8 MB blocks validated/sec"): one synthetic post-fork block of
    41,650 one-input P2PKH transactions through bcp_gpu_block_connect_p2pkh (host framing scan, H2D of the raw
    block + spent outputs, txids, Merkle root, per-input script checks, batched ECDSA, verdict read-back — all
    inside the timed region), next to the reference's CPU path for the same block: CBlock deserialization + txids +
    BlockMerkleRoot (single thread, as in CheckBlock) and VerifyScript per input on all host threads
    (what CCheckQueue<CScriptCheck> fans out)."""
    import numpy as np
    import synth
    from oracle import pyoracle as po

    t0 = time.perf_counter()
    blk, spk, amt = synth.synthetic_block(args.block_txs, True, with_prevouts=True)
    gen_s = time.perf_counter() - t0
    # marshal once, as a node would hand over its buffers: raw block bytes, concatenated scriptPubKeys + offsets, amounts
    L = bcp.lib()

    def pinned(a):
        """copy a numpy array into pinned host memory (bcp_gpu_host_alloc) and view it as numpy"""
        p = L.bcp_gpu_host_alloc(max(a.nbytes, 1))
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(a.nbytes, 1),))[: a.nbytes].view(a.dtype)
        out[...] = a
        return out

    off0 = np.zeros(len(spk) + 1, dtype=np.uint32)
    off0[1:] = np.cumsum([len(x) for x in spk])
    blk_np = pinned(np.frombuffer(blk, dtype=np.uint8))
    cat, off, am = pinned(np.frombuffer(b"".join(spk), dtype=np.uint8)), pinned(off0), pinned(np.asarray(amt, dtype=np.int64))
    res = eng.block_connect_p2pkh(blk_np, (cat, off), am, synth.POSTFORK_FLAGS)  # warm-up + correctness
    assert res["all_ok"] and res["root"] == blk[36:68] and not res["mutated"]
    reps = 10
    world = 1
    if sync is not None:
        world = sync(0.0, count=True)
    t0 = time.perf_counter()
    for _ in range(reps):
        res = eng.block_connect_p2pkh(blk_np, (cat, off), am, synth.POSTFORK_FLAGS)
    gpu_s = (time.perf_counter() - t0) / reps
    if sync is not None:
        gpu_s = sync(gpu_s)
        if int(os.environ.get("RANK", "0")) != 0:
            return None
    out = {"block_bytes": len(blk), "inputs": len(spk), "gpus": world, "gpu_ms_per_block": gpu_s * 1e3,
           "gpu_blocks_per_s": world / gpu_s, "gpu_inputs_per_s": world * len(spk) / gpu_s, "gen_seconds": gen_s,
           "timed": "bcp_gpu_block_connect_p2pkh end to end (host framing scan + H2D + kernels + D2H), pinned host buffers; "
                    "one block per GPU per repetition, max over ranks"}
    cores = os.cpu_count() or 1
    if po.have_ref():
        r = po.ref()
        t0 = time.perf_counter()
        n, _, root, _ = r.block_txids(blk, False, max_tx=len(spk) + 8)
        t_hash = time.perf_counter() - t0
        t0 = time.perf_counter()
        err = r.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, cores)
        t_chk = time.perf_counter() - t0
        assert not err.any() and root == blk[36:68]
        out.update({"cpu_kind": "reference", "cpu_cores": cores, "cpu_ms_txids_merkle_1thread": t_hash * 1e3,
                    "cpu_ms_script_checks_all_threads": t_chk * 1e3, "cpu_blocks_per_s": 1.0 / (t_hash + t_chk)})
    else:
        o = po.oracle()
        t0 = time.perf_counter()
        err = o.block_script_errors(blk, False, spk, amt, synth.POSTFORK_FLAGS, cores)
        o.block_txids(blk, False, max_tx=len(spk) + 8)
        t_all = time.perf_counter() - t0
        out.update({"cpu_kind": "port", "cpu_cores": cores, "cpu_blocks_per_s": 1.0 / t_all})
    out["speedup_vs_cpu"] = out["gpu_blocks_per_s"] / out["cpu_blocks_per_s"]
    return out


def queue_dropin_bench(args):
    """Level B1 measured with the reference's own interpreter: the 8 MB block's inputs as CScriptCheck-shaped closures
    calling the reference's VerifyScript, through this project's CCheckQueue drop-in on all host threads; signatures
    recorded and verified on the GPU in Wait() ("gpu") vs the reference's CPubKey::Verify in the same queue ("cpu").
    The binary is oracle/_ref/bench_queue_ref (tests/cpp/bench_queue_ref.cpp, built where /root/reference exists)."""
    import subprocess
    import tempfile

    import numpy as np
    import synth

    exe = os.path.join(ROOT, "oracle", "_ref", "bench_queue_ref")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/bench_queue_ref not built (reference sources absent at build time)"}
    blk, spk, amt = synth.synthetic_block(args.block_txs, True, with_prevouts=True)
    off = np.zeros(len(spk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in spk])
    threads = os.cpu_count() or 1
    with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
        f.write(len(blk).to_bytes(8, "little") + blk + len(spk).to_bytes(8, "little") + off.tobytes() + b"".join(spk)
                + np.asarray(amt, dtype=np.int64).tobytes())
        path = f.name
    try:
        out = subprocess.run([exe, path, str(threads), "3"], capture_output=True, text=True, timeout=600)
    finally:
        os.unlink(path)
    if out.returncode != 0:
        return {"error": (out.stdout + out.stderr)[-300:]}
    r = json.loads(out.stdout.strip().splitlines()[-1])
    r["timed"] = ("ConnectBlock's check loop per block: PrecomputedTransactionData per tx, one closure per input through "
                  "CCheckQueueControl::Add, Wait(); best of 3 after a warm-up; host interpreter + sighash on %d threads in both "
                  "arms" % threads)
    return r


def window_bench(eng, args, bcp, sync=None, W=8):
    """Throughput form of the 8 MB block measurement (reindex / initial block download, blocks at hand): W blocks per call
    through bcp_gpu_blocks_connect_p2pkh — one pass of the per-input kernels over the union of the window's inputs.
    The W blocks are the same synthetic block in W separate pinned buffers (run time does not depend on the content)."""
    import numpy as np
    import synth

    blk, spk, amt = synth.synthetic_block(args.block_txs, True, with_prevouts=True)
    L = bcp.lib()

    def pinned(a):
        p = L.bcp_gpu_host_alloc(max(a.nbytes, 1))
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(a.nbytes, 1),))[: a.nbytes].view(a.dtype)
        out[...] = a
        return out

    n = len(spk)
    lens = np.array([len(x) for x in spk] * W, dtype=np.uint64)
    off0 = np.zeros(n * W + 1, dtype=np.uint32)
    off0[1:] = np.cumsum(lens)
    blks = [pinned(np.frombuffer(blk, dtype=np.uint8)) for _ in range(W)]
    cat, off, am = pinned(np.frombuffer(b"".join(spk) * W, dtype=np.uint8)), pinned(off0), pinned(np.asarray(list(amt) * W, dtype=np.int64))
    res = eng.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)
    assert all(res["all_ok"]) and all(bytes(r) == blk[36:68] for r in res["roots"]) and not any(res["mutated"])
    reps = 5
    if sync is not None:
        sync(0.0)  # ranks enter the timed region together
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)
    s = (time.perf_counter() - t0) / reps
    # two host threads, each with its own lane of the same GPU: one window's copies and scans run under the other's kernels
    eng2 = bcp.Engine(eng.device & 0xFF, lane=1)
    eng2.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)

    def worker(e):
        for _ in range(reps):
            r = e.blocks_connect_p2pkh(blks, cat, off, am, [n] * W, synth.POSTFORK_FLAGS)
            assert all(r["all_ok"])

    th = [threading.Thread(target=worker, args=(e,)) for e in (eng, eng2)]
    if sync is not None:
        sync(0.0)
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    s2 = (time.perf_counter() - t0) / (2 * reps)
    # four host threads on four lanes, windows of four blocks: finer interleaving of copies, scans and kernels
    W4, reps4 = 4, 10
    off4 = pinned(off0[: n * W4 + 1].copy())
    engs4 = [eng, eng2] + [bcp.Engine(eng.device & 0xFF, lane=k) for k in (2, 3)]
    args4 = (blks[:W4], cat[: int(off0[n * W4])], off4, am[: n * W4], [n] * W4, synth.POSTFORK_FLAGS)
    for e in engs4:
        assert all(e.blocks_connect_p2pkh(*args4)["all_ok"])

    def worker4(e):
        for _ in range(reps4):
            e.blocks_connect_p2pkh(*args4)

    th = [threading.Thread(target=worker4, args=(e,)) for e in engs4]
    if sync is not None:
        sync(0.0)
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    s4 = (time.perf_counter() - t0) / (4 * reps4)
    world = 1
    if sync is not None:  # every rank measured its own GPU at the same time: take the slowest
        world = sync(0.0, count=True)
        s, s2, s4 = sync(s), sync(s2), sync(s4)
        if int(os.environ.get("RANK", "0")) != 0:
            return None
    return {"window_blocks": W, "inputs_per_block": n, "gpus": world, "gpu_ms_per_window": s * 1e3, "gpu_ms_per_block": s * 1e3 / W,
            "gpu_blocks_per_s": world * W / s, "gpu_inputs_per_s": world * W * n / s,
            "two_lanes": {"gpu_ms_per_window": s2 * 1e3, "gpu_blocks_per_s": world * W / s2, "gpu_inputs_per_s": world * W * n / s2},
            "four_lanes": {"window_blocks": W4, "gpu_ms_per_window": s4 * 1e3, "gpu_blocks_per_s": world * W4 / s4,
                           "gpu_inputs_per_s": world * W4 * n / s4},
            "timed": "bcp_gpu_blocks_connect_p2pkh end to end (host framing scans on W threads + H2D + kernels + D2H), pinned "
                     "host buffers; two_lanes / four_lanes = that many host threads, each on its own lane of the same GPU"}


def equihash_bench(eng):
    """Secondary measurement (BASELINE.json configs[3]): Equihash (200,9) verification + SHA-256d header hashing over
    10,000 serialized headers (the committed solver-found solutions tiled, half of them corrupted), through the
    host-pointer C-ABI call, next to the reference's IsValidSolution on all host threads."""
    import numpy as np
    from oracle import pyoracle as po

    hdrs = np.load(os.path.join(ROOT, "tests", "golden", "equihash_200_9.npz"))["headers"]
    rng = np.random.default_rng(4)
    allh = np.tile(hdrs, ((10000 + len(hdrs) - 1) // len(hdrs), 1))[:10000].copy()
    corrupt = rng.random(10000) < 0.5
    for i in np.nonzero(corrupt)[0]:
        allh[i, 143 + int(rng.integers(0, 1344))] ^= 1 << int(rng.integers(0, 8))
    ok, hashes = eng.equihash_verify_batch(200, 9, allh)
    assert ok[~corrupt].all() and not ok[corrupt].all()
    # timed through the C ABI on pinned host buffers (headers in, verdicts + hashes out), as a node's receive path would hold them
    import bitcoincashplus_b200 as bcp
    L = bcp.lib()
    nb = allh.size
    p_in, p_ok, p_hash = L.bcp_gpu_host_alloc(nb), L.bcp_gpu_host_alloc(10000), L.bcp_gpu_host_alloc(10000 * 32)
    C.memmove(p_in, allh.ctypes.data, nb)
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        rc = L.bcp_gpu_equihash_verify_batch(eng.device, 200, 9, p_in, allh.shape[1], allh.shape[1], 10000, p_ok, p_hash)
        assert rc == 0
    gpu_s = (time.perf_counter() - t0) / reps
    ok2 = np.ctypeslib.as_array(C.cast(p_ok, C.POINTER(C.c_uint8)), shape=(10000,)).copy()
    h2 = np.ctypeslib.as_array(C.cast(p_hash, C.POINTER(C.c_uint8)), shape=(10000, 32)).copy()
    assert (ok2 == ok).all() and (h2 == hashes).all()
    for p_ in (p_in, p_ok, p_hash):
        L.bcp_gpu_host_free(p_)
    cores = os.cpu_count() or 1
    impl = po.ref() if po.have_ref() else po.oracle()
    t0 = time.perf_counter()
    want = impl.equihash_verify_batch(200, 9, allh, cores)
    cpu_s = time.perf_counter() - t0
    assert (want.astype(bool) == ok.astype(bool)).all()
    return {"headers": 10000, "header_bytes": int(allh.shape[1]), "gpu_ms": gpu_s * 1e3, "gpu_headers_per_s": 10000 / gpu_s,
            "cpu_kind": impl.kind, "cpu_cores": cores, "cpu_headers_per_s": 10000 / cpu_s,
            "speedup_vs_cpu": (10000 / gpu_s) / (10000 / cpu_s),
            "timed": "bcp_gpu_equihash_verify_batch on pinned host buffers incl. H2D of 14.9 MB (four chunks on two streams), verdicts + header hashes D2H"}


def sharded_single_caller_bench(args, devices, msg, pub, sig, expected):
    """The multi-GPU form north_star describes, measured as ONE caller sees it: one host process holds one batch in
    pinned memory, bcp::VerifyTriples (include/compat/bcp_recorder.h) splits it by signature index over the N GPUs from
    N host threads (each through bcp_gpu_ecdsa_verify_batch: H2D, kernels, D2H), the verdict bytes land in one array and
    are AND-ed on the host like `fAllOk &= fOk` (src/checkqueue.h:77).  Two batch sizes: the 1M batch of configs[1] and
    8 x that.  Runs on rank 0 over every GPU of the job after the per-rank sections are done."""
    import numpy as np

    import bitcoincashplus_b200 as bcp

    G = bcp.lib()
    L = C.CDLL(os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_chain.so"))
    L.bcp_sharded_ecdsa_verify.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                           C.c_uint32, C.c_void_p, C.POINTER(C.c_int)]
    n1 = len(msg)
    out = {"gpus": len(devices), "host_threads": len(devices),
           "timed": "bcp_sharded_ecdsa_verify = bcp::VerifyTriples: one pinned host batch split by index over the GPUs, one host "
                    "thread per GPU, H2D + kernels + D2H + the host-side AND inside the timed region; wall clock of the one caller"}
    devs = (C.c_int * len(devices))(*devices)
    for mult in (1, 8):
        n = n1 * mult
        hm, hp, hs, ho = G.bcp_gpu_host_alloc(n * 32), G.bcp_gpu_host_alloc(n * 65), G.bcp_gpu_host_alloc(n * 64), G.bcp_gpu_host_alloc(n)
        for k in range(mult):
            C.memmove(hm + k * n1 * 32, msg.ctypes.data, n1 * 32)
            C.memmove(hp + k * n1 * 65, pub.ctypes.data, n1 * 65)
            C.memmove(hs + k * n1 * 64, sig.ctypes.data, n1 * 64)
        allok = C.c_int(-1)
        reps = 5 if mult == 1 else 3
        for _ in range(2):
            assert L.bcp_sharded_ecdsa_verify(devs, len(devices), hm, hp, hs, n, 0, ho, C.byref(allok)) == 0
        t0 = time.perf_counter()
        for _ in range(reps):
            assert L.bcp_sharded_ecdsa_verify(devs, len(devices), hm, hp, hs, n, 0, ho, C.byref(allok)) == 0
        dt = (time.perf_counter() - t0) / reps
        got = np.ctypeslib.as_array(C.cast(ho, C.POINTER(C.c_uint8)), shape=(n,)).astype(bool)
        assert (got == np.tile(expected, mult)).all() and allok.value == 0
        out["batch_%dM" % (n // 1000000) if n % 1000000 == 0 else "batch_%d" % n] = {
            "signatures": n, "ms": dt * 1e3, "sigverifies_per_s": n / dt, "h2d_bytes": n * 161, "d2h_bytes": n}
        for p_ in (hm, hp, hs, ho):
            G.bcp_gpu_host_free(p_)
    return out


_CHAIN_CACHE = {}


def _chain_blocks(nb, ntx):
    """the synthetic chain, generated once per process: (blocks, seeds, seconds)"""
    import synth

    key = (nb, ntx)
    if key not in _CHAIN_CACHE:
        t0 = time.perf_counter()
        blocks = list(synth.synthetic_chain_fast(nb, ntx))
        seeds = synth.chain_seeds(ntx, 4096)
        _CHAIN_CACHE[key] = (blocks, seeds, time.perf_counter() - t0)
    return _CHAIN_CACHE[key]


def chain_bench(args, devices, impl="device"):
    """BASELINE.json configs[4]: a synthetic chain of 8 MB blocks with real UTXO dependencies (block b spends block b-1)
    and valid regtest Equihash (48,5) headers, in blk*.dat framing in one pinned buffer, connected by
      impl "device": bcp::DeviceChain (include/compat/bcp_chain_gpu.h) — coin table in HBM on GPU 0 (txids + Merkle, inserts,
                     lookups + spends as kernels in chain order), whole blocks dealt out to the GPUs for the header and
                     script checks, several blocks in flight, verdicts taken in order;
      impl "host":   bcp::ChainValidator (include/compat/bcp_chain.h) — coin map walked by host threads, each block's inputs
                     split over the GPUs, two blocks in flight.
    CPU side: the reference's deserialization + txids + Merkle + VerifyScript on all threads for one block of the chain."""
    import ctypes as C
    import synth
    from oracle import pyoracle as po

    warm = 32 if impl == "device" else 1  # at least two blocks per check thread at 8 GPUs: every lane's scratch exists before the clock starts
    nblk = args.chain_blocks if impl == "device" else min(args.chain_blocks, 20)
    ntx = args.block_txs
    G_ = len(devices)
    allb, seeds, gen_s = _chain_blocks(args.chain_blocks + 32 + 8, ntx)
    blocks = allb[: warm + nblk]
    extra = allb[warm + nblk: warm + nblk + G_]  # the next blocks of the chain: the last of them is also connected damaged
    L = C.CDLL(os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_chain.so"))
    pre = "bcp_dchain_" if impl == "device" else "bcp_chain_"
    fn = lambda name: getattr(L, pre + name)
    err = C.create_string_buffer(256)
    devs = (C.c_int * len(devices))(*devices)
    if impl == "device":
        L.bcp_dchain_new.restype = C.c_void_p
        L.bcp_dchain_new.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_char_p, C.c_size_t]
        lanes = args.chain_lanes or (3 if len(devices) == 1 else 2)
        scan_threads = max(2, min(8, effective_cpus() // 4))
        h = L.bcp_dchain_new(devs, len(devices), synth.POSTFORK_FLAGS, 48, 5, 23, lanes, scan_threads, err, 256)  # 2^23 slots of 128 B = 1 GB of HBM
        if h:
            L.bcp_dchain_reserve.argtypes = [C.c_void_p, C.c_size_t]
            L.bcp_dchain_reserve(h, max(len(b) for b in blocks) + 4096)  # every pipeline slot's pinned buffers, before the clock starts
    else:
        L.bcp_chain_new.restype = C.c_void_p
        L.bcp_chain_new.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_char_p, C.c_size_t]
        h = L.bcp_chain_new(devs, len(devices), synth.POSTFORK_FLAGS, 48, 5, err, 256)
    assert h, err.value
    fn("add_coin").argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_size_t, C.c_int64]
    for txid, vout, spk, amount in seeds:
        fn("add_coin")(h, txid, vout, spk, len(spk), amount)
    reason = C.create_string_buffer(256)
    # the blocks live in pinned buffers in blk*.dat framing (magic | size | block), as the import loop of -reindex reads them
    # (LoadExternalBlockFile, src/validation.cpp:4486-4516); the first `warm` blocks are connected untimed (allocations)
    import bitcoincashplus_b200 as bcp
    G = bcp.lib()
    magic = bytes.fromhex("e48f1d1c")
    frame = lambda bl: b"".join(magic + len(b).to_bytes(4, "little") + b for b in bl)
    bufs = []
    for part in (blocks[:warm], blocks[warm:]):
        data = frame(part)
        p_ = G.bcp_gpu_host_alloc(len(data))
        C.memmove(p_, data, len(data))
        bufs.append((p_, len(data)))
    fn("connect_file").restype = C.c_long
    fn("connect_file").argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_long)]
    fn("free").argtypes = [C.c_void_p]
    failed = C.c_long(-1)
    n_ok = fn("connect_file")(h, bufs[0][0], bufs[0][1], magic, reason, 256, C.byref(failed))
    assert n_ok == warm, (n_ok, failed.value, reason.value)
    nph = 11 if impl == "device" else 5
    fn("phase_seconds").argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    ph0 = (C.c_double * nph)()
    fn("phase_seconds")(h, ph0)
    t0 = time.perf_counter()
    n_ok = fn("connect_file")(h, bufs[1][0], bufs[1][1], magic, reason, 256, C.byref(failed))
    total = time.perf_counter() - t0
    assert n_ok == nblk, (n_ok, failed.value, reason.value)
    ph = (C.c_double * nph)()
    fn("phase_seconds")(h, ph)
    if impl == "device":
        names = ["connecting_thread_apply_calls", "connecting_thread_waiting", "check_threads_wait_for_spent_coins",
                 "check_threads_header", "check_threads_scripts", "scan_threads", "connecting_thread_confirm",
                 "latency_applied_to_table_done", "latency_applied_to_verdict", "connecting_thread_waiting_window_full",
                 "connecting_thread_waiting_for_scan"]
    else:
        names = ["header_equihash", "framing_scan", "txids_wait_after_lookups", "utxo_walk", "scripts_wait_after_inserts"]
    phases = dict(zip(names, [1e3 * (x - x0) / nblk for x, x0 in zip(ph, ph0)]))
    if impl == "host":
        wk = (C.c_double * 4)()
        L.bcp_chain_walk_seconds.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.bcp_chain_walk_seconds(h, wk)
        phases["utxo_walk_parts"] = dict(zip(["pass_a_lookups", "staging_and_launch", "pass_b_inserts", "pass_c_in_block"],
                                             [1e3 * x / (nblk + warm) for x in wk]))
    # A rejected block on the LAST device: the next G blocks of the chain, the last of them (dealt to device G-1 by the
    # device connector; input range of device G-1 in the host one) with one broken signature under a consistent header.
    # It must be refused with exactly that input, the blocks before it stay connected, and the undamaged block connects after.
    bad_input = ntx - 7
    tail = list(extra[:-1]) + [synth.damage_chain_block(extra[-1], warm + nblk + G_ - 1, bad_input)]
    data = frame(tail)
    n_ok = fn("connect_file")(h, data, len(data), magic, reason, 256, C.byref(failed))
    assert n_ok == G_ - 1 and failed.value == G_ - 1, (n_ok, failed.value, reason.value)
    assert reason.value == b"blk-bad-inputs (input %d, script error 31)" % bad_input, reason.value
    data = frame(extra[-1:])
    assert fn("connect_file")(h, data, len(data), magic, reason, 256, C.byref(failed)) == 1, reason.value
    for p_, _ in bufs:
        G.bcp_gpu_host_free(p_)
    gpus_in_use = L.bcp_dchain_gpus_in_use(h) if impl == "device" else len(devices)
    fn("free")(h)
    out = {"impl": "bcp::DeviceChain (coin table in HBM)" if impl == "device" else "bcp::ChainValidator (coin map on host threads)",
           "blocks": nblk, "inputs_per_block": ntx, "block_bytes": len(blocks[0]), "gpus": len(devices),
           "sharding": ("whole blocks dealt out to the GPUs (block b's header + script checks on GPU b mod N); coin table, txids and "
                        "Merkle roots on GPU 0 in chain order; verdicts taken in order on the host" if impl == "device" else
                        "one host chain walk; each block's inputs split by index across the GPUs, verdicts ANDed on the host"),
           "gpu_seconds_total": total, "gpu_blocks_per_s": nblk / total, "gpu_ms_per_block": 1e3 * total / nblk,
           "gpu_inputs_per_s": nblk * ntx / total,
           "ms_per_block_by_phase": phases, "gen_seconds": gen_s, "warmup_blocks": warm,
           "rejected_block_check": "block %d of the chain with a broken signature at input %d, checked by GPU %d: refused with that "
                                   "input, table rolled back, the undamaged block connected afterwards" % (warm + nblk + G_ - 1, bad_input, G_ - 1),
           "gpus_in_use": gpus_in_use,
           "peer_access_to_table_gpu": [int(G.bcp_gpu_peer_access(devices[0] & 0xFF, d & 0xFF)) for d in devices],
           "gpus_in_use_note": "GPUs without direct peer access to the coin table's GPU are left out (a staged copy per block costs more than they bring)",
           "timed": "ConnectFile over %d blocks in blk*.dat framing in one pinned buffer, wall clock of the caller" % nblk}
    if impl == "device":
        # every block crosses the host's memory twice on this path (framing scan, upload): the box's read bandwidth is a ceiling
        L.bcp_host_read_bandwidth.restype = C.c_double
        L.bcp_host_read_bandwidth.argtypes = [C.c_int, C.c_size_t]
        nthr = min(os.cpu_count() or 1, 16)
        bw = max(L.bcp_host_read_bandwidth(nthr, 256 << 20) for _ in range(2))
        out["host_threads"] = {"scan": scan_threads, "check": "%d per GPU (one less on the table's GPU when there are several)" % lanes,
                               "connecting": 1, "effective_cpus": effective_cpus(), "os_cpu_count": os.cpu_count()}
        out["host_memory"] = {"read_gbs_measured": bw, "threads": nthr,
                              "bytes_read_per_block": 2 * len(blocks[0]),
                              "chain_read_gbs": 2 * len(blocks[0]) * nblk / total / 1e9,
                              "note": "host DRAM reads the chain causes (scan + upload DMA) against what this box's cores can read"}
        out["phase_note"] = ("thread-seconds per block summed over the threads of each kind, not a critical path: the connecting thread "
                             "is the only sequential part (apply calls + waiting); check threads = %d per GPU" % lanes)
    if po.have_ref() and impl == "device":
        r = po.ref()
        cores = os.cpu_count() or 1
        spk = [s[2] for s in seeds]
        amt = [s[3] for s in seeds]
        t0 = time.perf_counter()
        n, txids, root, _ = r.block_txids(blocks[0], False, max_tx=ntx + 8)
        e = r.block_script_errors(blocks[0], False, spk, amt, synth.POSTFORK_FLAGS, cores)
        okh = r.equihash_verify(48, 5, blocks[0][:177])
        cpu_s = time.perf_counter() - t0
        assert not e.any() and okh == 1 and root == blocks[0][36:68]
        out.update({"cpu_kind": "reference", "cpu_cores": cores, "cpu_ms_per_block": cpu_s * 1e3,
                    "cpu_blocks_per_s": 1.0 / cpu_s, "speedup_vs_cpu": (nblk / total) * cpu_s,
                    "cpu_sample": "block 0 of the chain: CBlock deserialize + txids + Merkle (1 thread), VerifyScript per input (all threads), Equihash; no UTXO map work"})
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import synth

    n = args.sigs
    gen_n = min(n, 200000)  # the CPU arm only ever touches a bounded sample
    msg, pub, sig, _ = synth.sigset(gen_n, corrupt_pct=1)
    from oracle import pyoracle as po

    impl = po.ref() if po.have_ref() else po.oracle()
    cores = os.cpu_count() or 1
    sample = min(gen_n, 1500 * cores)
    for _ in range(args.warmup):
        impl.ecdsa_verify_batch(msg[:sample], pub[:sample], sig[:sample], cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        impl.ecdsa_verify_batch(msg[:sample], pub[:sample], sig[:sample], cores)
    dt = time.perf_counter() - t0
    v = sample * args.steps / dt
    line = {"metric": "ecdsa_sigverifies_per_s", "value": v, "unit": "sigverifies/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic", "impl": "reference",
            "config": workload_config(n),
            "cpu_baseline": {"value": v, "unit": "sigverifies/s", "cores": cores, "kind": impl.kind,
                             "sample": "%d signatures per step (bounded sample of the batch), %d threads" % (sample, cores)},
            "e2e": {"value": v, "unit": "sigverifies/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "sample_note": "each step verifies a bounded sample (%d signatures = 1,500 x %d threads) of the %d-signature batch, "
                           "not the whole batch; the metric is a throughput" % (sample, cores, n),
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sigs", type=int, default=1000000, help="signatures per GPU per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-block", action="store_true", help="skip the secondary 8 MB block measurement (configs[2])")
    ap.add_argument("--block-txs", type=int, default=41650)
    ap.add_argument("--chain-blocks", type=int, default=100, help="blocks of the synthetic chain reindex (configs[4]); 0 = skip")
    ap.add_argument("--chain-lanes", type=int, default=0, help="check threads per GPU of the chain section (0 = 3 on one GPU, else 2)")
    ap.add_argument("--only", default="", help="comma list of secondary sections to run (block8mb,window,queue,equihash,sharded,chain,chain_host); empty = all")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch

    import bitcoincashplus_b200 as bcp
    import synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = world > 1
    if dist:
        # The data path has no collective (signatures shard by index, verdicts are AND-ed by the caller), so there is no
        # NCCL here: torch.distributed is only the launcher's rendezvous, a barrier and the max over ranks of a few
        # timings, all on the host through gloo.
        import torch.distributed as td

        td.init_process_group("gloo")
        hostgrp = td.group.WORLD
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    eng = bcp.Engine(local)
    n = args.sigs

    # synthetic batch: every rank generates its own shard (disjoint message indices would need a rank offset;
    # the kernels' run time does not depend on it, so each rank uses the same deterministic set)
    t0 = time.perf_counter()
    msg, pub, sig, kinds = synth.sigset(n, corrupt_pct=1)
    gen_s = time.perf_counter() - t0
    expected = (kinds < 0) | np.isin(kinds, [synth.CORRUPTIONS.index("hybrid_ok"), synth.CORRUPTIONS.index("msg_plus_n")])

    # ---- device-resident leg
    d_msg = torch.from_numpy(msg).to(dev)
    d_pub = torch.from_numpy(pub).to(dev)
    d_sig = torch.from_numpy(sig).to(dev)
    d_ok = torch.zeros(n, dtype=torch.uint8, device=dev)
    eng.ecdsa_reserve(n)
    torch.cuda.synchronize(dev)
    stream = torch.cuda.Stream(dev)  # a real (non-default) stream: the C ABI launches on exactly this handle
    assert stream.cuda_stream != 0

    def step():
        eng.ecdsa_verify_batch_dev(d_msg.data_ptr(), d_pub.data_ptr(), d_sig.data_ptr(), n, d_ok.data_ptr(), 0,
                                   stream.cuda_stream)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize(dev)
    got = d_ok.cpu().numpy().astype(bool)
    assert (got == expected).all(), "verdict mismatch in bench batch: %r" % np.nonzero(got != expected)[0][:10]

    samples, stop = [], threading.Event()
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, local), daemon=True)
    th.start()
    if dist:
        td.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize(dev)
    if dist:
        td.barrier()
    ms = e0.elapsed_time(e1)

    # ---- end-to-end leg: pinned host buffers through the host-pointer C-ABI call
    L = bcp.lib()
    hm, hp, hs, ho = (L.bcp_gpu_host_alloc(n * 32), L.bcp_gpu_host_alloc(n * 65), L.bcp_gpu_host_alloc(n * 64),
                      L.bcp_gpu_host_alloc(n))
    C.memmove(hm, msg.ctypes.data, n * 32)
    C.memmove(hp, pub.ctypes.data, n * 65)
    C.memmove(hs, sig.ctypes.data, n * 64)
    for _ in range(max(1, args.warmup - 1)):
        eng.ecdsa_verify_batch_ptr(hm, hp, hs, n, ho)
    if dist:
        td.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.ecdsa_verify_batch_ptr(hm, hp, hs, n, ho)
    e2e_s = time.perf_counter() - t0
    stop.set()
    th.join(timeout=2)
    ok_host = np.ctypeslib.as_array(C.cast(ho, C.POINTER(C.c_uint8)), shape=(n,)).astype(bool)
    assert (ok_host == expected).all(), "verdict mismatch in e2e leg"

    if dist:
        t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        ms, e2e_ms = t[0].item(), t[1].item()
    else:
        e2e_ms = e2e_s * 1e3
    def host_sync(seconds, count=False):
        """barrier over the host-side group; returns the max of `seconds` over ranks (or the world size)"""
        t = torch.tensor([seconds], dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX, group=hostgrp)
        return world if count else t[0].item()

    only = set(x for x in args.only.split(",") if x)
    want = lambda name: not only or name in only
    blk_section = win_section = None
    if not args.no_block:
        if want("block8mb"):
            blk_section = block_bench(eng, args, bcp, host_sync if dist else None)
        if want("window"):
            win_section = window_bench(eng, args, bcp, host_sync if dist else None)
    if rank != 0:
        if dist:
            td.barrier(group=hostgrp)  # rank 0 is still measuring its secondary sections; leave together
            td.destroy_process_group()
        return

    total = n * world
    value = total * args.steps / (ms * 1e-3)
    e2e = total * args.steps / (e2e_ms * 1e-3)
    peak_imad = eng.measure_int_peak(1)   # carry-chained IMAD.WIDE form the field code uses
    peak_wide = eng.measure_int_peak(0)
    peak_add = eng.measure_int_peak(2)
    pipes = {"imad_wide_x": peak_imad / 1e12, "imad_wide": peak_wide / 1e12, "iadd": peak_add / 1e12,
             "imad_lo": eng.measure_int_peak(3) / 1e12, "imad_hi": eng.measure_int_peak(4) / 1e12,
             "dfma": eng.measure_int_peak(5) / 1e12, "ffma": eng.measure_int_peak(6) / 1e12}
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    per_gpu = value / world
    achieved = per_gpu * MUL32_PER_VERIFY
    line = {
        "metric": "ecdsa_sigverifies_per_s", "value": value, "unit": "sigverifies/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": workload_config(n),
        "roofline": {"bound": "int32",
                     "bound_note": "32x32->64 multiply-add pipe (IMAD.WIDE on fmaheavy): nothing on this path is HBM- or "
                                   "tensor-bound (SURVEY.md 8d); the HBM view is under 'hbm'",
                     "achieved": achieved / 1e12, "peak": peak_imad / 1e12, "unit": "Tmul32/s", "frac": achieved / peak_imad,
                     "traffic": traffic_from_profile(n),
                     "peak_source": "bcp_gpu_measure_int_peak(1): IMAD.WIDE.U32.X carry chains, measured live",
                     "pipes_tops_measured": pipes,
                     "hbm": {"achieved_gbs": per_gpu * BYTES_PER_VERIFY / 1e9, "peak_gbs": hbm_peak,
                             "frac": per_gpu * BYTES_PER_VERIFY / 1e9 / hbm_peak,
                             "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}},
        "e2e": {"value": e2e, "unit": "sigverifies/s", "h2d_bytes_per_step": n * 161 * world, "d2h_bytes_per_step": n * world},
        "gpu_launches": 2 * args.steps, "clocks": summarize_clocks(samples), "gen_seconds": gen_s,
    }
    if not args.no_cpu_baseline:
        # the CPU baseline verifies the very same signatures: its verdicts are the parity check of the timed batch
        line["cpu_baseline"], cpu_verdicts = cpu_baseline(msg, pub, sig)
        k = len(cpu_verdicts)
        bad = np.nonzero((np.asarray(cpu_verdicts) != 0) != got[:k])[0]
        assert bad.size == 0, "GPU verdicts differ from the CPU %s at %r" % (line["cpu_baseline"]["kind"], bad[:10])
        bad = np.nonzero((np.asarray(cpu_verdicts) != 0) != ok_host[:k])[0]
        assert bad.size == 0, "e2e verdicts differ from the CPU %s at %r" % (line["cpu_baseline"]["kind"], bad[:10])
        line["parity_checked"] = int(k)
        line["parity_note"] = ("verdicts of the device-resident leg and of the e2e leg compared, outside the timed regions, with the "
                               "CPU baseline's verdicts for the same %d signatures (%d rejected by both)" % (k, int((np.asarray(cpu_verdicts) == 0).sum())))
    if not args.no_block:
        # the block numbers as a fraction of the integer roofline too (same convention as `roofline`: the reference
        # algorithm's 151,228 mul32 per input against the measured IMAD.WIDE peak, per GPU)
        def roof(inputs_per_s):
            return inputs_per_s / world * MUL32_PER_VERIFY / peak_imad
        if blk_section:
            blk_section["frac_of_int_roofline"] = roof(blk_section["gpu_inputs_per_s"])
        if win_section:
            win_section["frac_of_int_roofline"] = roof(win_section["gpu_inputs_per_s"])
            for k_ in ("two_lanes", "four_lanes"):
                win_section[k_]["frac_of_int_roofline"] = roof(win_section[k_]["gpu_inputs_per_s"])
        line["block8mb"] = blk_section
        line["block8mb_window"] = win_section
        if want("queue"):
            line["checkqueue_dropin"] = queue_dropin_bench(args)
        if want("equihash"):
            line["equihash10k"] = equihash_bench(eng)
        if want("sharded"):
            line["sharded_single_caller"] = sharded_single_caller_bench(args, list(range(world)), msg, pub, sig, expected)
        if args.chain_blocks > 0 and want("chain"):
            line["chain"] = chain_bench(args, list(range(world)), "device")
        if args.chain_blocks > 0 and want("chain_host"):
            line["chain_host_table"] = chain_bench(args, list(range(world)), "host")
    print(json.dumps(line), flush=True)
    if dist:
        td.barrier(group=hostgrp)
        td.destroy_process_group()


if __name__ == "__main__":
    main()
