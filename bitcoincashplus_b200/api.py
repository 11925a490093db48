"""Python face over the C ABI (include/bcp_gpu.h).  numpy in, numpy out; device pointers for the
`_dev` variants are plain integers (e.g. torch.Tensor.data_ptr()).  Names follow the reference's
domain: signatures, pubkeys, txids, Merkle root, headers, solutions."""
import ctypes as C

import numpy as np

from ._lib import BcpError, check, lib

ECDSA_FLAG_NORMALIZE_S = 1


def _np(a, shape_last=None):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if shape_last is not None and (a.ndim != 2 or a.shape[1] != shape_last):
        raise ValueError("expected an (n, %d) uint8 array, got %r" % (shape_last, a.shape))
    return a


def device_count():
    return lib().bcp_gpu_device_count()


def der_lax_parse(der):
    """ecdsa_signature_parse_der_lax (src/pubkey.cpp:26-168): returns (ok, r||s bytes)."""
    out = (C.c_uint8 * 64)()
    buf = bytes(der)
    ok = lib().bcp_der_lax_parse(buf, len(buf), out)
    return ok, bytes(out)


def block_index(blk, legacy_header=False, max_tx=1 << 22):
    """Byte offsets of every transaction of a serialized block (host-side framing scan)."""
    blk = bytes(blk)
    off = np.zeros(max_tx + 1, dtype=np.uint32)
    n = lib().bcp_block_index(blk, len(blk), int(legacy_header), off.ctypes.data, max_tx)
    if n < 0:
        raise BcpError("malformed block (bcp_block_index returned %d)" % n)
    return off[: n + 1].copy()


class Engine:
    """One device context (bcp_gpu_init / bcp_gpu_shutdown).  `lane` (0..3) selects an independent context on the same
    GPU (BCP_DEVICE_LANE in include/bcp_gpu.h): engines on different lanes run their calls concurrently."""

    def __init__(self, device=0, lane=0):
        self.device = int(device) | (int(lane) << 8)
        check(lib().bcp_gpu_init(self.device), "bcp_gpu_init(0x%x)" % self.device)

    def close(self):
        lib().bcp_gpu_shutdown(self.device)

    # ---- ECDSA
    def ecdsa_verify_batch(self, msg32, pub65, sig64, flags=0):
        """Per-signature verdict bytes for host arrays (n,32) (n,65) (n,64)."""
        msg32, pub65, sig64 = _np(msg32, 32), _np(pub65, 65), _np(sig64, 64)
        n = msg32.shape[0]
        if pub65.shape[0] != n or sig64.shape[0] != n:
            raise ValueError("batch arrays disagree on n")
        ok = np.zeros(n, dtype=np.uint8)
        check(lib().bcp_gpu_ecdsa_verify_batch(self.device, msg32.ctypes.data, pub65.ctypes.data, sig64.ctypes.data, n,
                                               flags, ok.ctypes.data), "bcp_gpu_ecdsa_verify_batch")
        return ok

    def ecdsa_verify_batch_ptr(self, msg32_ptr, pub65_ptr, sig64_ptr, n, ok_ptr, flags=0):
        """Host-pointer form on caller-owned (pinned) buffers; used by bench.py's e2e leg."""
        check(lib().bcp_gpu_ecdsa_verify_batch(self.device, msg32_ptr, pub65_ptr, sig64_ptr, n, flags, ok_ptr),
              "bcp_gpu_ecdsa_verify_batch")

    def pubkey_parse_batch(self, pub65):
        """secp256k1_ec_pubkey_parse for (n,65) key slots -> (xy (n,64), ok (n,))."""
        pub65 = _np(pub65, 65)
        n = pub65.shape[0]
        xy = np.zeros((n, 64), dtype=np.uint8)
        ok = np.zeros(n, dtype=np.uint8)
        check(lib().bcp_gpu_pubkey_parse_batch(self.device, pub65.ctypes.data, n, xy.ctypes.data, ok.ctypes.data),
              "bcp_gpu_pubkey_parse_batch")
        return xy, ok

    def ecdsa_recover_batch(self, msg32, sig64, recid):
        """secp256k1_ecdsa_recover per item -> (xy64 (n,64), ok (n,))."""
        msg32, sig64 = _np(msg32, 32), _np(sig64, 64)
        rid = np.ascontiguousarray(recid, dtype=np.uint8)
        n = len(rid)
        xy, ok = np.zeros((n, 64), dtype=np.uint8), np.zeros(n, dtype=np.uint8)
        check(lib().bcp_gpu_ecdsa_recover_batch(self.device, msg32.ctypes.data, sig64.ctypes.data, rid.ctypes.data, n,
                                                xy.ctypes.data, ok.ctypes.data), "bcp_gpu_ecdsa_recover_batch")
        return xy, ok

    def pubkey_tweak_add_batch(self, pub65, tweak32):
        """secp256k1_ec_pubkey_tweak_add per item -> (xy64 (n,64), ok (n,))."""
        pub65, tweak32 = _np(pub65, 65), _np(tweak32, 32)
        n = len(pub65)
        xy, ok = np.zeros((n, 64), dtype=np.uint8), np.zeros(n, dtype=np.uint8)
        check(lib().bcp_gpu_pubkey_tweak_add_batch(self.device, pub65.ctypes.data, tweak32.ctypes.data, n, xy.ctypes.data,
                                                   ok.ctypes.data), "bcp_gpu_pubkey_tweak_add_batch")
        return xy, ok

    def ecdsa_reserve(self, n):
        check(lib().bcp_gpu_ecdsa_reserve(self.device, n), "bcp_gpu_ecdsa_reserve")

    def ecdsa_verify_batch_dev(self, d_msg32, d_pub65, d_sig64, n, d_ok, flags=0, stream=0):
        check(lib().bcp_gpu_ecdsa_verify_batch_dev(self.device, d_msg32, d_pub65, d_sig64, n, flags, d_ok, stream),
              "bcp_gpu_ecdsa_verify_batch_dev")

    # ---- SHA-256d / Merkle
    def txids_merkle(self, blk, tx_off, want_txids=True):
        """(txids (ntx,32), root bytes, mutated bool) of a serialized block given its tx offsets."""
        blk = np.frombuffer(bytes(blk), dtype=np.uint8) if not isinstance(blk, np.ndarray) else _np(blk)
        tx_off = np.ascontiguousarray(tx_off, dtype=np.uint32)
        ntx = len(tx_off) - 1 if len(tx_off) else 0
        txids = np.zeros((ntx, 32), dtype=np.uint8)
        root = (C.c_uint8 * 32)()
        mut = C.c_int(0)
        check(lib().bcp_gpu_sha256d_txids_merkle(self.device, blk.ctypes.data, blk.size, tx_off.ctypes.data, ntx,
                                                 txids.ctypes.data if want_txids else None, root, C.byref(mut)),
              "bcp_gpu_sha256d_txids_merkle")
        return txids, bytes(root), bool(mut.value)

    def merkle_root(self, leaves32):
        leaves32 = _np(leaves32).reshape(-1, 32)
        root = (C.c_uint8 * 32)()
        mut = C.c_int(0)
        check(lib().bcp_gpu_merkle_root(self.device, leaves32.ctypes.data if len(leaves32) else None, len(leaves32),
                                        root, C.byref(mut)), "bcp_gpu_merkle_root")
        return bytes(root), bool(mut.value)

    def sha256d_batch(self, data, off):
        data = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else _np(data)
        off = np.ascontiguousarray(off, dtype=np.uint32)
        cnt = len(off) - 1
        out = np.zeros((cnt, 32), dtype=np.uint8)
        check(lib().bcp_gpu_sha256d_batch(self.device, data.ctypes.data, data.size, off.ctypes.data, cnt,
                                          out.ctypes.data), "bcp_gpu_sha256d_batch")
        return out

    # ---- whole block
    def block_connect_p2pkh(self, blk, spks, amounts, flags, legacy_header=False, want_txids=False):
        """One call per block: returns dict(txids, root, mutated, script_err (n_inputs,), all_ok)."""
        blk = np.frombuffer(bytes(blk), dtype=np.uint8) if not isinstance(blk, np.ndarray) else _np(blk)
        if isinstance(spks, tuple):  # already marshalled: (concatenated scriptPubKeys, offsets)
            cat, off = spks
            n = len(off) - 1
        else:
            n = len(spks)
            cat = np.frombuffer(b"".join(spks) or b"\x00", dtype=np.uint8)
            off = np.zeros(n + 1, dtype=np.uint32)
            off[1:] = np.cumsum([len(s) for s in spks])
        am = np.ascontiguousarray(amounts, dtype=np.int64)
        err = np.full(n, 253, dtype=np.uint8)
        root = (C.c_uint8 * 32)()
        mut, allok = C.c_int(0), C.c_int(0)
        txids = np.zeros((blk.size // 60 + 2, 32), dtype=np.uint8) if want_txids else None
        check(lib().bcp_gpu_block_connect_p2pkh(self.device, blk.ctypes.data, blk.size, int(legacy_header), cat.ctypes.data,
                                                off.ctypes.data, am.ctypes.data if n else None, n, flags,
                                                txids.ctypes.data if want_txids else None, root, C.byref(mut),
                                                err.ctypes.data if n else None, C.byref(allok)), "bcp_gpu_block_connect_p2pkh")
        return {"txids": txids, "root": bytes(root), "mutated": bool(mut.value), "script_err": err, "all_ok": bool(allok.value)}

    def blocks_connect_p2pkh(self, blks, spk_cat, spk_off, amounts, n_inputs, flags, legacy_header=False, want_txids=False):
        """A window of blocks in one pass.  blks: list of uint8 arrays (pinned for overlap); spk_cat / spk_off / amounts:
        the concatenation over all blocks' inputs; n_inputs: per block.  Returns dict(roots (nblk,32), mutated, all_ok,
        script_err (total inputs,), ntx, txids)."""
        nb = len(blks)
        arrs = [b if isinstance(b, np.ndarray) else np.frombuffer(bytes(b), dtype=np.uint8) for b in blks]
        ptrs = (C.c_void_p * nb)(*[a.ctypes.data for a in arrs])
        lens = (C.c_size_t * nb)(*[a.size for a in arrs])
        nin = (C.c_size_t * nb)(*[int(x) for x in n_inputs])
        total = int(sum(n_inputs))
        cat, off, am = _np(spk_cat), np.ascontiguousarray(spk_off, dtype=np.uint32), np.ascontiguousarray(amounts, dtype=np.int64)
        assert off.size == total + 1 and am.size == total
        err = np.full(total, 253, dtype=np.uint8)
        roots = np.zeros((nb, 32), dtype=np.uint8)
        mut, allok = (C.c_int * nb)(), (C.c_int * nb)()
        ntx = (C.c_size_t * nb)()
        txids = np.zeros((sum(a.size // 60 + 2 for a in arrs), 32), dtype=np.uint8) if want_txids else None
        check(lib().bcp_gpu_blocks_connect_p2pkh(self.device, nb, ptrs, lens, int(legacy_header), cat.ctypes.data if total else None,
                                                 off.ctypes.data, am.ctypes.data if total else None, nin, flags,
                                                 txids.ctypes.data if want_txids else None, ntx, roots.ctypes.data, mut,
                                                 err.ctypes.data if total else None, allok), "bcp_gpu_blocks_connect_p2pkh")
        ntx = [int(x) for x in ntx]
        return {"roots": roots, "mutated": [bool(x) for x in mut], "all_ok": [bool(x) for x in allok], "script_err": err,
                "ntx": ntx, "txids": txids[:sum(ntx)] if want_txids else None}

    # ---- Equihash
    def equihash_verify_batch(self, n, k, hdrs, want_hash=True):
        """hdrs: (count, hdr_len) uint8 serialized post-fork headers -> (ok (count,), hash (count,32))."""
        hdrs = _np(hdrs)
        count, hdr_len = hdrs.shape
        ok = np.zeros(count, dtype=np.uint8)
        hashes = np.zeros((count, 32), dtype=np.uint8)
        check(lib().bcp_gpu_equihash_verify_batch(self.device, n, k, hdrs.ctypes.data, hdr_len, hdr_len, count,
                                                  ok.ctypes.data, hashes.ctypes.data if want_hash else None),
              "bcp_gpu_equihash_verify_batch")
        return ok, hashes

    def header_check_batch(self, n, k, hdrs, pow_limit, pow_limit_legacy, fork_height=0):
        """CheckBlockHeader for a batch (src/validation.cpp:3202-3220) + in-batch linkage.  pow limits: 32 bytes each,
        uint256 memory order.  -> dict(ok_solution, ok_pow, linked, hash)."""
        hdrs = _np(hdrs)
        count, hdr_len = hdrs.shape
        out = {k_: np.zeros(count, dtype=np.uint8) for k_ in ("ok_solution", "ok_pow", "linked")}
        out["hash"] = np.zeros((count, 32), dtype=np.uint8)
        lim, liml = _np(np.frombuffer(bytes(pow_limit), dtype=np.uint8)), _np(np.frombuffer(bytes(pow_limit_legacy), dtype=np.uint8))
        assert lim.size == 32 and liml.size == 32
        check(lib().bcp_gpu_header_check_batch(self.device, n, k, hdrs.ctypes.data, hdr_len, hdr_len, count, lim.ctypes.data,
                                               liml.ctypes.data, fork_height, out["ok_solution"].ctypes.data,
                                               out["ok_pow"].ctypes.data, out["linked"].ctypes.data, out["hash"].ctypes.data),
              "bcp_gpu_header_check_batch")
        return out

    # ---- measurement
    def measure_int_peak(self, which=0):
        v = C.c_double(0)
        check(lib().bcp_gpu_measure_int_peak(self.device, which, C.byref(v)), "bcp_gpu_measure_int_peak")
        return v.value
