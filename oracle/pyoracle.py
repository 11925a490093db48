"""ctypes bindings for the CPU oracle (oracle/libbcp_oracle.so) and, when present, the reference
build (oracle/_ref/libbcp_ref.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs — never by the bitcoincashplus_b200 package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libbcp_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libbcp_ref.so")

u8p = C.POINTER(C.c_uint8)


def _p(a):
    """numpy uint8 array / bytes -> uint8 pointer (keeps no reference: caller holds the object)."""
    if isinstance(a, (bytes, bytearray)):
        return C.cast(C.c_char_p(bytes(a)), u8p)
    assert a.dtype == np.uint8 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u8p)


def build(ref=True):
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


class _Lib:
    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.prefix = prefix

    def fn(self, name, restype, *argtypes):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        f.argtypes = list(argtypes)
        return f


class Impl:
    """Common Python face over either library (prefix orc_ or ref_)."""

    def __init__(self, path, prefix):
        L = _Lib(path, prefix)
        self.lib = L.lib  # the raw CDLL, for the few functions only one of the libraries has
        self.kind = "port" if prefix == "orc_" else "reference"
        sz, i, u32, i64 = C.c_size_t, C.c_int, C.c_uint32, C.c_int64
        self._sha256 = L.fn("sha256", None, u8p, sz, u8p)
        self._sha256d = L.fn("sha256d", None, u8p, sz, u8p)
        self._hash160 = L.fn("hash160", None, u8p, sz, u8p)
        self._merkle = L.fn("merkle_root", None, u8p, sz, u8p, C.POINTER(i))
        self._block_txids = L.fn("block_txids", C.c_long, u8p, sz, i, u8p, sz, u8p, C.POINTER(i))
        self._ecdsa_verify = L.fn("ecdsa_verify", i, u8p, u8p, sz, u8p)
        self._ecdsa_verify_batch = L.fn("ecdsa_verify_batch", None, u8p, u8p, u8p, sz, i, u8p)
        self._pubkey_verify = L.fn("pubkey_verify", i, u8p, u8p, sz, u8p, sz)
        self._check_low_s = L.fn("check_low_s", i, u8p, sz)
        self._pubkey_create = L.fn("pubkey_create", i, u8p, i, u8p, C.POINTER(sz))
        self._sig_to_der = L.fn("sig_to_der", i, u8p, u8p, C.POINTER(sz))
        self._equihash_verify = L.fn("equihash_verify", i, C.c_uint, C.c_uint, u8p, sz)
        self._equihash_verify_batch = L.fn("equihash_verify_batch", None, C.c_uint, C.c_uint, u8p, sz, sz, sz, i, u8p)
        self._pubkey_parse = L.fn("pubkey_parse", i, u8p, sz, u8p)
        self._recover = L.fn("ecdsa_recover", i, u8p, u8p, i, u8p)
        self._tweak_add = L.fn("pubkey_tweak_add", i, u8p, sz, u8p, u8p)
        if prefix == "orc_":
            self._sign = L.fn("sign_rfc6979", i, u8p, u8p, u8p)
            self._header_hash = L.fn("header_hash", None, u8p, sz, u8p)
            self._der_lax = L.fn("der_lax_parse", i, u8p, sz, u8p)
            self._gen_sigset = L.fn("gen_sigset", None, u8p, sz, sz, sz, i, u8p, u8p, u8p)
            self._blake2b = L.fn("blake2b", None, u8p, sz, u8p, sz, u8p)
            self._sighash = L.fn("sighash_forkid", i, u8p, sz, u8p, sz, C.c_uint, u32, i64, u8p)
            self._block_index = L.fn("block_index", C.c_long, u8p, sz, i, C.POINTER(u32), sz)
            self._check_pow = L.fn("check_pow", i, u8p, u32, u8p)
            self._ripemd160 = L.fn("ripemd160", None, u8p, sz, u8p)
            self._block_verify_p2pkh = L.fn("block_verify_p2pkh", C.c_long, u8p, sz, i, u8p, C.POINTER(u32), C.POINTER(i64),
                                            u32, i, u8p, sz)
            self._is_valid_sig_enc = L.fn("is_valid_signature_encoding", i, u8p, sz)
            self._gen_chain_txs = L.fn("gen_chain_txs", sz, u8p, sz, sz, sz, sz, u8p, C.POINTER(u32), C.POINTER(i64), i, u8p, sz)
        else:
            L.fn("init", i)()
            self._sign_rec = L.fn("sign_recoverable", i, u8p, u8p, u8p, C.POINTER(i))
            self._recover_compact = L.fn("cpubkey_recover_compact", i, u8p, u8p, u8p)
            self._derive = L.fn("cpubkey_derive", i, u8p, u8p, C.c_uint, u8p, u8p)
            self._sign = L.fn("sign", i, u8p, u8p, u8p)
            self._header_hash_ref = L.fn("header_hash", i, u8p, sz, i, u8p)
            self._solve = L.fn("equihash_solve", i, C.c_uint, C.c_uint, u8p, i, u8p, sz, i)
            self._sighash_ref = L.fn("sighash", i, u8p, sz, u8p, sz, C.c_uint, u32, i64, u32, u8p)
            self._verify_script = L.fn("verify_script", i, u8p, sz, u8p, sz, u8p, sz, C.c_uint, i64, u32, C.POINTER(i))
            self._block_verify_inputs = L.fn("block_verify_inputs", C.c_long, u8p, sz, i, u8p, C.POINTER(u32),
                                             C.POINTER(i64), u32, i, u8p, sz, u8p)

    # ---- hashing
    def _h(self, f, data, n):
        out = (C.c_uint8 * n)()
        f(_p(bytes(data)), len(data), out)
        return bytes(out)

    def sha256(self, d): return self._h(self._sha256, d, 32)
    def sha256d(self, d): return self._h(self._sha256d, d, 32)
    def hash160(self, d): return self._h(self._hash160, d, 20)
    def ripemd160(self, d): return self._h(self._ripemd160, d, 20)

    def merkle_root(self, leaves):
        """leaves: (n,32) uint8 array or bytes -> (root bytes, mutated bool)"""
        a = np.frombuffer(bytes(leaves), dtype=np.uint8) if isinstance(leaves, (bytes, bytearray)) else np.ascontiguousarray(leaves).reshape(-1)
        n = a.size // 32
        root = (C.c_uint8 * 32)()
        m = C.c_int(0)
        self._merkle(_p(a) if n else None, n, root, C.byref(m))
        return bytes(root), bool(m.value)

    def block_txids(self, blk, legacy, max_tx=1 << 20):
        txids = np.zeros((max_tx, 32), dtype=np.uint8)
        root = (C.c_uint8 * 32)()
        m = C.c_int(0)
        n = self._block_txids(_p(bytes(blk)), len(blk), int(legacy), _p(txids), max_tx, root, C.byref(m))
        if n < 0:
            return n, None, None, None
        return n, txids[:n].copy(), bytes(root), bool(m.value)

    def header_hash(self, hdr, legacy=False):
        out = (C.c_uint8 * 32)()
        if self.kind == "port":
            self._header_hash(_p(bytes(hdr)), len(hdr), out)
        else:
            assert self._header_hash_ref(_p(bytes(hdr)), len(hdr), int(legacy), out) == 1
        return bytes(out)

    # ---- ECDSA
    def ecdsa_verify(self, msg32, pub, sig64):
        return self._ecdsa_verify(_p(bytes(msg32)), _p(bytes(pub)), len(pub), _p(bytes(sig64)))

    def ecdsa_verify_batch(self, msg32, pub65, sig64, nthreads=1):
        n = msg32.shape[0]
        ok = np.zeros(n, dtype=np.uint8)
        self._ecdsa_verify_batch(_p(msg32), _p(pub65), _p(sig64), n, nthreads, _p(ok))
        return ok

    def pubkey_verify(self, hash32, pub, der):
        return self._pubkey_verify(_p(bytes(hash32)), _p(bytes(pub)), len(pub), _p(bytes(der)), len(der))

    def check_low_s(self, der): return self._check_low_s(_p(bytes(der)), len(der))

    def pubkey_create(self, seckey, compressed=True):
        out = (C.c_uint8 * 65)()
        ln = C.c_size_t(65)
        if not self._pubkey_create(_p(bytes(seckey)), int(compressed), out, C.byref(ln)):
            return None
        return bytes(out)[: ln.value]

    def sign(self, seckey, msg32):
        out = (C.c_uint8 * 64)()
        if not self._sign(_p(bytes(seckey)), _p(bytes(msg32)), out):
            return None
        return bytes(out)

    def sig_to_der(self, sig64):
        out = (C.c_uint8 * 80)()
        ln = C.c_size_t(80)
        if not self._sig_to_der(_p(bytes(sig64)), out, C.byref(ln)):
            return None
        return bytes(out)[: ln.value]

    def ecdsa_recover(self, msg32, sig64, recid):
        """secp256k1_ecdsa_recover: affine x || y (64 bytes) of the recovered key, or None."""
        out = (C.c_uint8 * 64)()
        ok = self._recover(_p(bytes(msg32)), _p(bytes(sig64)), int(recid), out)
        return bytes(out) if ok else None

    def pubkey_tweak_add(self, pub, tweak32):
        """secp256k1_ec_pubkey_tweak_add: affine x || y of pub + tweak*G, or None."""
        out = (C.c_uint8 * 64)()
        ok = self._tweak_add(_p(bytes(pub)), len(pub), _p(bytes(tweak32)), out)
        return bytes(out) if ok else None

    def sign_recoverable(self, seckey, msg32):
        """reference only: (sig64, recid) from secp256k1_ecdsa_sign_recoverable."""
        out = (C.c_uint8 * 64)()
        rid = C.c_int(-1)
        assert self._sign_rec(_p(bytes(seckey)), _p(bytes(msg32)), out, C.byref(rid)) == 1
        return bytes(out), rid.value

    def cpubkey_recover_compact(self, hash32, sig65):
        """reference only: CPubKey::RecoverCompact -> serialized key bytes or None."""
        out = (C.c_uint8 * 65)()
        n = self._recover_compact(_p(bytes(hash32)), _p(bytes(sig65)), out)
        return bytes(out)[:n] if n else None

    def cpubkey_derive(self, pub33, cc32, nchild):
        """reference only: CPubKey::Derive -> (child33, chaincode32) or None."""
        o, c = (C.c_uint8 * 33)(), (C.c_uint8 * 32)()
        ok = self._derive(_p(bytes(pub33)), _p(bytes(cc32)), int(nchild), o, c)
        return (bytes(o), bytes(c)) if ok else None

    def pubkey_parse(self, pub):
        out = (C.c_uint8 * 64)()
        ok = self._pubkey_parse(_p(bytes(pub)), len(pub), out)
        return ok, bytes(out)

    def der_lax_parse(self, der):
        out = (C.c_uint8 * 64)()
        ok = self._der_lax(_p(bytes(der)), len(der), out)
        return ok, bytes(out)

    def gen_sigset(self, n, nkeys=4096, seed=b"bcp-b200-v1", nthreads=None):
        nthreads = nthreads or os.cpu_count() or 1
        msg = np.zeros((n, 32), dtype=np.uint8)
        pub = np.zeros((n, 65), dtype=np.uint8)
        sig = np.zeros((n, 64), dtype=np.uint8)
        self._gen_sigset(_p(seed), len(seed), n, nkeys, nthreads, _p(msg), _p(pub), _p(sig))
        return msg, pub, sig

    # ---- Equihash / Blake2b
    def blake2b(self, data, person, outlen):
        out = (C.c_uint8 * outlen)()
        self._blake2b(_p(bytes(data)), len(data), _p(bytes(person)), outlen, out)
        return bytes(out)

    def equihash_verify(self, n, k, hdr):
        return self._equihash_verify(n, k, _p(bytes(hdr)), len(hdr))

    def equihash_verify_batch(self, n, k, hdrs, nthreads=1):
        cnt, stride = hdrs.shape
        ok = np.zeros(cnt, dtype=np.uint8)
        self._equihash_verify_batch(n, k, _p(hdrs), stride, stride, cnt, nthreads, _p(ok))
        return ok

    def equihash_solve(self, n, k, input140, optimised=False, max_solns=16):
        soln_len = (1 << k) * (n // (k + 1) + 1) // 8
        out = np.zeros((max_solns, soln_len), dtype=np.uint8)
        cnt = self._solve(n, k, _p(bytes(input140)), int(optimised), _p(out), soln_len, max_solns)
        return [bytes(out[i]) for i in range(max(cnt, 0))]

    # ---- sighash / script
    def sighash_forkid(self, script_code, tx, nin, hashtype, amount, flags=1 << 16):
        out = (C.c_uint8 * 32)()
        if self.kind == "port":
            ok = self._sighash(_p(bytes(script_code)), len(script_code), _p(bytes(tx)), len(tx), nin, hashtype, amount, out)
        else:
            ok = self._sighash_ref(_p(bytes(script_code)), len(script_code), _p(bytes(tx)), len(tx), nin, hashtype,
                                   amount, flags, out)
        return bytes(out) if ok else None

    def verify_script(self, script_sig, script_pubkey, tx, nin, amount, flags):
        err = C.c_int(0)
        ok = self._verify_script(_p(bytes(script_sig)), len(script_sig), _p(bytes(script_pubkey)), len(script_pubkey),
                                 _p(bytes(tx)), len(tx), nin, amount, flags, C.byref(err))
        return ok, err.value

    def block_script_errors(self, blk, legacy, spks, amounts, flags, nthreads=1):
        """Per non-coinbase input ScriptError of the block's spends (spks: list of scriptPubKey bytes, amounts: ints).
        Oracle: P2PKH fast path (255 = not that shape).  Reference: full VerifyScript."""
        n = len(spks)
        cat = b"".join(spks) or b"\x00"
        off = np.zeros(n + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(s) for s in spks])
        am = np.asarray(amounts, dtype=np.int64)
        err = np.full(n, 254, dtype=np.uint8)
        u32p, i64p = C.POINTER(C.c_uint32), C.POINTER(C.c_int64)
        if self.kind == "port":
            r = self._block_verify_p2pkh(_p(bytes(blk)), len(blk), int(legacy), _p(cat), off.ctypes.data_as(u32p),
                                         am.ctypes.data_as(i64p), flags, nthreads, _p(err), n)
        else:
            ok = np.zeros(n, dtype=np.uint8)
            r = self._block_verify_inputs(_p(bytes(blk)), len(blk), int(legacy), _p(cat), off.ctypes.data_as(u32p),
                                          am.ctypes.data_as(i64p), flags, nthreads, _p(ok), n, _p(err))
            assert r < 0 or ((ok == 1) == (err == 0)).all()
        assert r == n, (r, n)
        return err

    def gen_chain_txs(self, seed, key_base, nkeys, prev_txid, prev_key, prev_amt, nthreads=None):
        """In place: advances (prev_txid (n,32) u8, prev_key (n,) u32, prev_amt (n,) i64) by one block; returns the txs."""
        n = len(prev_key)
        out = np.zeros(n * 200, dtype=np.uint8)
        tot = self._gen_chain_txs(_p(seed), len(seed), key_base, nkeys, n, _p(prev_txid), prev_key.ctypes.data_as(C.POINTER(C.c_uint32)),
                                  prev_amt.ctypes.data_as(C.POINTER(C.c_int64)), nthreads or os.cpu_count() or 1, _p(out), out.size)
        assert tot > 0
        return out[:tot].tobytes()

    def block_index(self, blk, legacy, max_tx=1 << 20):
        off = (C.c_uint32 * (max_tx + 1))()
        n = self._block_index(_p(bytes(blk)), len(blk), int(legacy), off, max_tx)
        if n < 0:
            return n, None
        return n, np.frombuffer(off, dtype=np.uint32)[: n + 1].copy()

    def check_pow(self, hash32, nbits, pow_limit32):
        return self._check_pow(_p(bytes(hash32)), nbits, _p(bytes(pow_limit32)))


_cache = {}


def oracle():
    if "o" not in _cache:
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        _cache["o"] = Impl(ORACLE_SO, "orc_")
    return _cache["o"]


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    if "r" not in _cache:
        _cache["r"] = Impl(REF_SO, "ref_")
    return _cache["r"]
