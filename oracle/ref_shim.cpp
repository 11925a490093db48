// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Thin extern "C" glue over the REFERENCE's own CPU implementation of the hot path, compiled in
// place from /root/reference by oracle/Makefile into oracle/_ref/libbcp_ref.so.  No reference
// source is copied into this repository: this file only #includes the reference headers and
// calls the reference functions named below, so Python (ctypes) tests and bench.py's
// `--impl reference` arm can drive them.
//
// What each entry point calls (paths relative to /root/reference):
//   ref_ecdsa_verify        secp256k1_ec_pubkey_parse (src/secp256k1/src/secp256k1.c:150),
//                           secp256k1_ecdsa_signature_parse_compact (:228), secp256k1_ecdsa_verify (:293)
//   ref_pubkey_verify       CPubKey::Verify (src/pubkey.cpp:170-193) — lax DER + normalize + verify
//   ref_check_low_s         CPubKey::CheckLowS (src/pubkey.cpp:294-302)
//   ref_sign / ref_pubkey_create   secp256k1_ecdsa_sign (RFC6979 default nonce), secp256k1_ec_pubkey_create
//   ref_sha256 / ref_sha256d       CSHA256 (src/crypto/sha256.cpp:194-243), CHash256 (src/hash.h:21-43)
//   ref_hash160             CHash160 (src/hash.h:46-68)
//   ref_merkle_root         ComputeMerkleRoot (src/consensus/merkle.cpp:146-150)
//   ref_block_txids         CBlock deserialize (src/primitives/block.h), CTransaction::GetId,
//                           BlockMerkleRoot (src/consensus/merkle.cpp:175-182)
//   ref_header_hash         CBlockHeader serialization + CHashWriter (src/primitives/block.cpp:17-28)
//   ref_equihash_verify     CEquihashInput / EhInitialiseState / EhIsValidSolution
//                           (src/pow.cpp:298-323, src/crypto/equihash.cpp:724-770)
//   ref_equihash_solve      EhBasicSolve / EhOptimisedSolve (src/crypto/equihash.cpp) — fixture generation
//   ref_sighash             SignatureHash (src/script/interpreter.cpp:1354-1429)
//   ref_verify_script       VerifyScript + TransactionSignatureChecker (src/script/interpreter.cpp:1551)
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#include <functional>
#include <mutex>
#include <algorithm>

#include "secp256k1.h"
#include "secp256k1_recovery.h"
#include "pubkey.h"
#include "hash.h"
#include "uint256.h"
#include "arith_uint256.h"
#include "streams.h"
#include "version.h"
#include "crypto/sha256.h"
#include "crypto/equihash.h"
#include "consensus/merkle.h"
#include "primitives/block.h"
#include "primitives/transaction.h"
#include "script/interpreter.h"
#include "script/script.h"
#include "script/script_error.h"

// primitives/block.cpp needs chainparams; we do not build it.  Provide the two symbols the
// headers reference so the link closes (unused by the shim).
uint256 CBlockHeader::GetHash() const { return uint256(); }
std::string CBlock::ToString() const { return std::string(); }
// support/cleanse.cpp forwards to OpenSSL's OPENSSL_cleanse; the shim supplies the symbol itself so
// the oracle build has no OpenSSL dependency (zeroing only, no arithmetic).
void memory_cleanse(void *ptr, size_t len) {
    volatile unsigned char *p = (volatile unsigned char *)ptr;
    while (len--) *p++ = 0;
}

namespace {
secp256k1_context *g_ctx = nullptr;       // verify+sign context for raw libsecp calls
ECCVerifyHandle *g_handle = nullptr;      // owns pubkey.cpp's global verify context
void ensure_init() {
    static std::once_flag once;
    std::call_once(once, [] {
        g_ctx = secp256k1_context_create(SECP256K1_CONTEXT_VERIFY | SECP256K1_CONTEXT_SIGN);
        g_handle = new ECCVerifyHandle();
    });
}
template <class F>
void parallel_for(size_t n, int nthreads, F f) {
    if (nthreads <= 1) { f(0, n); return; }
    std::vector<std::thread> th;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        size_t lo = std::min(n, t * chunk), hi = std::min(n, lo + chunk);
        if (lo < hi) th.emplace_back(f, lo, hi);
    }
    for (auto &x : th) x.join();
}
}  // namespace

extern "C" {

int ref_init(void) { ensure_init(); return 0; }

// Raw libsecp path (BASELINE config 2): parse pubkey, parse compact (r||s), verify.
int ref_ecdsa_verify(const uint8_t *msg32, const uint8_t *pub, size_t publen, const uint8_t *sig64) {
    ensure_init();
    secp256k1_pubkey pk;
    secp256k1_ecdsa_signature sig;
    if (!secp256k1_ec_pubkey_parse(g_ctx, &pk, pub, publen)) return 0;
    if (!secp256k1_ecdsa_signature_parse_compact(g_ctx, &sig, sig64)) return 0;
    return secp256k1_ecdsa_verify(g_ctx, &sig, msg32, &pk);
}

// secp256k1_ec_pubkey_parse (secp256k1.c:150-163) + uncompressed re-serialization: xy64 = x || y
int ref_pubkey_parse(const uint8_t *in, size_t len, uint8_t *xy64) {
    ensure_init();
    secp256k1_pubkey pk;
    memset(xy64, 0, 64);
    if (!secp256k1_ec_pubkey_parse(g_ctx, &pk, in, len)) return 0;
    uint8_t out[65];
    size_t ol = 65;
    secp256k1_ec_pubkey_serialize(g_ctx, out, &ol, &pk, SECP256K1_EC_UNCOMPRESSED);
    memcpy(xy64, out + 1, 64);
    return 1;
}

// Batch form in the B4 layout: pubkeys at stride 65 (length decided by the prefix byte as
// CPubKey::GetLen does, src/pubkey.h:47-51; unknown prefix => length 0 => reject).
static inline size_t publen_by_prefix(uint8_t h) {
    if (h == 2 || h == 3) return 33;
    if (h == 4 || h == 6 || h == 7) return 65;
    return 0;
}
void ref_ecdsa_verify_batch(const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64, size_t n,
                            int nthreads, uint8_t *ok) {
    ensure_init();
    parallel_for(n, nthreads, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; i++) {
            size_t len = publen_by_prefix(pub65[i * 65]);
            ok[i] = len ? (uint8_t)ref_ecdsa_verify(msg32 + 32 * i, pub65 + 65 * i, len, sig64 + 64 * i) : 0;
        }
    });
}

// secp256k1_ecdsa_recoverable_signature_parse_compact + secp256k1_ecdsa_recover
// (src/secp256k1/src/modules/recovery/main_impl.h:38-58, 170-190): xy64 = x || y of the recovered key
int ref_ecdsa_recover(const uint8_t *msg32, const uint8_t *sig64, int recid, uint8_t *xy64) {
    ensure_init();
    memset(xy64, 0, 64);
    secp256k1_ecdsa_recoverable_signature rs;
    secp256k1_pubkey pk;
    if (recid < 0 || recid > 3) return 0;
    if (!secp256k1_ecdsa_recoverable_signature_parse_compact(g_ctx, &rs, sig64, recid)) return 0;
    if (!secp256k1_ecdsa_recover(g_ctx, &pk, &rs, msg32)) return 0;
    uint8_t out[65];
    size_t ol = 65;
    secp256k1_ec_pubkey_serialize(g_ctx, out, &ol, &pk, SECP256K1_EC_UNCOMPRESSED);
    memcpy(xy64, out + 1, 64);
    return 1;
}
// secp256k1_ecdsa_sign_recoverable (:123-168): compact signature + recovery id, RFC 6979 nonce (fixture generation)
int ref_sign_recoverable(const uint8_t *seckey32, const uint8_t *msg32, uint8_t *sig64, int *recid) {
    ensure_init();
    secp256k1_ecdsa_recoverable_signature rs;
    if (!secp256k1_ecdsa_sign_recoverable(g_ctx, &rs, msg32, seckey32, nullptr, nullptr)) return 0;
    secp256k1_ecdsa_recoverable_signature_serialize_compact(g_ctx, sig64, recid, &rs);
    return 1;
}
// secp256k1_ec_pubkey_tweak_add (src/secp256k1/src/secp256k1.c:477-499)
int ref_pubkey_tweak_add(const uint8_t *pub, size_t publen, const uint8_t *tweak32, uint8_t *xy64) {
    ensure_init();
    memset(xy64, 0, 64);
    secp256k1_pubkey pk;
    if (!secp256k1_ec_pubkey_parse(g_ctx, &pk, pub, publen)) return 0;
    if (!secp256k1_ec_pubkey_tweak_add(g_ctx, &pk, tweak32)) return 0;
    uint8_t out[65];
    size_t ol = 65;
    secp256k1_ec_pubkey_serialize(g_ctx, out, &ol, &pk, SECP256K1_EC_UNCOMPRESSED);
    memcpy(xy64, out + 1, 64);
    return 1;
}
// CPubKey::RecoverCompact (src/pubkey.cpp:195-218) on the 65-byte compact form: out = serialized key, returns its length or 0
int ref_cpubkey_recover_compact(const uint8_t *hash32, const uint8_t *sig65, uint8_t *out65) {
    ensure_init();
    CPubKey pk;
    uint256 h;
    memcpy(h.begin(), hash32, 32);
    if (!pk.RecoverCompact(h, std::vector<uint8_t>(sig65, sig65 + 65))) return 0;
    memcpy(out65, pk.begin(), pk.size());
    return (int)pk.size();
}
// CPubKey::Derive (src/pubkey.cpp:240-264): BIP32 public child; out33 = child key, cc_out32 = child chain code
int ref_cpubkey_derive(const uint8_t *pub33, const uint8_t *cc32, unsigned nChild, uint8_t *out33, uint8_t *cc_out32) {
    ensure_init();
    CPubKey pk(pub33, pub33 + 33), child;
    ChainCode cc, ccChild;
    memcpy(cc.begin(), cc32, 32);
    if (!pk.Derive(child, ccChild, nChild, cc)) return 0;
    memcpy(out33, child.begin(), 33);
    memcpy(cc_out32, ccChild.begin(), 32);
    return 1;
}

// CPubKey::Verify path (what the node calls): DER signature WITHOUT the hashtype byte.
int ref_pubkey_verify(const uint8_t *hash32, const uint8_t *pub, size_t publen, const uint8_t *der, size_t derlen) {
    ensure_init();
    CPubKey pk(pub, pub + publen);
    uint256 h;
    memcpy(h.begin(), hash32, 32);
    std::vector<uint8_t> sig(der, der + derlen);
    return pk.Verify(h, sig) ? 1 : 0;
}

int ref_check_low_s(const uint8_t *der, size_t derlen) {
    ensure_init();
    std::vector<uint8_t> sig(der, der + derlen);
    return CPubKey::CheckLowS(sig) ? 1 : 0;
}

int ref_pubkey_create(const uint8_t *seckey32, int compressed, uint8_t *out, size_t *outlen) {
    ensure_init();
    secp256k1_pubkey pk;
    if (!secp256k1_ec_pubkey_create(g_ctx, &pk, seckey32)) return 0;
    size_t len = compressed ? 33 : 65;
    secp256k1_ec_pubkey_serialize(g_ctx, out, &len, &pk,
                                  compressed ? SECP256K1_EC_COMPRESSED : SECP256K1_EC_UNCOMPRESSED);
    *outlen = len;
    return 1;
}

// RFC6979 deterministic signature, compact r||s (low-S as libsecp always produces).
int ref_sign(const uint8_t *seckey32, const uint8_t *msg32, uint8_t *sig64) {
    ensure_init();
    secp256k1_ecdsa_signature sig;
    if (!secp256k1_ecdsa_sign(g_ctx, &sig, msg32, seckey32, nullptr, nullptr)) return 0;
    secp256k1_ecdsa_signature_serialize_compact(g_ctx, sig64, &sig);
    return 1;
}

int ref_sig_to_der(const uint8_t *sig64, uint8_t *der, size_t *derlen) {
    ensure_init();
    secp256k1_ecdsa_signature sig;
    if (!secp256k1_ecdsa_signature_parse_compact(g_ctx, &sig, sig64)) return 0;
    return secp256k1_ecdsa_signature_serialize_der(g_ctx, der, derlen, &sig);
}

void ref_sha256(const uint8_t *data, size_t len, uint8_t *out32) {
    CSHA256().Write(data, len).Finalize(out32);
}
void ref_sha256d(const uint8_t *data, size_t len, uint8_t *out32) {
    CHash256().Write(data, len).Finalize(out32);
}
void ref_hash160(const uint8_t *data, size_t len, uint8_t *out20) {
    CHash160().Write(data, len).Finalize(out20);
}

void ref_merkle_root(const uint8_t *leaves32, size_t n, uint8_t *root32, int *mutated) {
    std::vector<uint256> v(n);
    for (size_t i = 0; i < n; i++) memcpy(v[i].begin(), leaves32 + 32 * i, 32);
    bool m = false;
    uint256 r = ComputeMerkleRoot(v, &m);
    memcpy(root32, r.begin(), 32);
    if (mutated) *mutated = m ? 1 : 0;
}

// Deserialize a block (legacy=1: 80-byte header wire format; 0: post-fork header with solution),
// write per-tx ids, the Merkle root and the mutated flag.  Returns ntx, or -1 on a parse error.
// txids may be NULL (count only).
long ref_block_txids(const uint8_t *blk, size_t len, int legacy, uint8_t *txids, size_t max_tx, uint8_t *root32,
                     int *mutated) {
    try {
        CDataStream ss((const char *)blk, (const char *)blk + len, SER_NETWORK,
                       PROTOCOL_VERSION | (legacy ? SERIALIZE_BLOCK_LEGACY : 0));
        CBlock block;
        ss >> block;
        if (txids) {
            for (size_t i = 0; i < block.vtx.size() && i < max_tx; i++)
                memcpy(txids + 32 * i, block.vtx[i]->GetId().begin(), 32);
        }
        bool m = false;
        uint256 r = BlockMerkleRoot(block, &m);
        if (root32) memcpy(root32, r.begin(), 32);
        if (mutated) *mutated = m;
        return (long)block.vtx.size();
    } catch (...) {
        return -1;
    }
}

// Header hash as CBlockHeader::GetHash(params) computes it (src/primitives/block.cpp:17-28):
// deserialize, re-serialize through CHashWriter with or without SERIALIZE_BLOCK_LEGACY.
int ref_header_hash(const uint8_t *hdr, size_t len, int legacy, uint8_t *out32) {
    try {
        int version = PROTOCOL_VERSION | (legacy ? SERIALIZE_BLOCK_LEGACY : 0);
        CDataStream ss((const char *)hdr, (const char *)hdr + len, SER_NETWORK, version);
        CBlockHeader h;
        ss >> h;
        CHashWriter w(SER_GETHASH, version);
        ::Serialize(w, h);
        uint256 r = w.GetHash();
        memcpy(out32, r.begin(), 32);
        return 1;
    } catch (...) {
        return 0;
    }
}

// CheckEquihashSolution (src/pow.cpp:298-323) on a serialized post-fork header (140 B + compactsize + solution).
// Returns 1 valid, 0 invalid, -1 parse error / unsupported parameters.
int ref_equihash_verify(unsigned n, unsigned k, const uint8_t *hdr, size_t len) {
    try {
        CDataStream in((const char *)hdr, (const char *)hdr + len, SER_NETWORK, PROTOCOL_VERSION);
        CBlockHeader h;
        in >> h;
        crypto_generichash_blake2b_state state;
        EhInitialiseState(n, k, state);
        CEquihashInput I{h};
        CDataStream ss(SER_NETWORK, PROTOCOL_VERSION);
        ss << I;
        ss << h.nNonce;
        crypto_generichash_blake2b_update(&state, (unsigned char *)&ss[0], ss.size());
        bool isValid = false;
        EhIsValidSolution(n, k, state, h.nSolution, isValid);
        return isValid ? 1 : 0;
    } catch (...) {
        return -1;
    }
}

void ref_equihash_verify_batch(unsigned n, unsigned k, const uint8_t *hdrs, size_t stride, size_t len, size_t count,
                               int nthreads, uint8_t *ok) {
    parallel_for(count, nthreads, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; i++) ok[i] = ref_equihash_verify(n, k, hdrs + i * stride, len) == 1;
    });
}

// Run the reference solver on a 140-byte Equihash input (I || nonce); copies up to max_solns
// solutions of soln_len bytes each into out; returns the number found.
int ref_equihash_solve(unsigned n, unsigned k, const uint8_t *input140, int optimised, uint8_t *out, size_t soln_len,
                       int max_solns) {
    try {
        crypto_generichash_blake2b_state state;
        EhInitialiseState(n, k, state);
        crypto_generichash_blake2b_update(&state, input140, 140);
        int found = 0;
        std::function<bool(std::vector<unsigned char>)> valid = [&](std::vector<unsigned char> soln) {
            if (soln.size() == soln_len && found < max_solns) {
                memcpy(out + (size_t)found * soln_len, soln.data(), soln_len);
                found++;
            }
            return false;  // keep enumerating
        };
        if (optimised)
            EhOptimisedSolveUncancellable(n, k, state, valid);
        else
            EhBasicSolveUncancellable(n, k, state, valid);
        return found;
    } catch (...) {
        return -1;
    }
}

// SignatureHash (src/script/interpreter.cpp:1354) over a serialized tx.
int ref_sighash(const uint8_t *script_code, size_t sc_len, const uint8_t *tx, size_t txlen, unsigned nIn,
                uint32_t hashtype, int64_t amount, uint32_t flags, uint8_t *out32) {
    try {
        CDataStream ss((const char *)tx, (const char *)tx + txlen, SER_NETWORK, PROTOCOL_VERSION);
        CTransaction t(deserialize, ss);
        CScript sc(script_code, script_code + sc_len);
        uint256 h = SignatureHash(sc, t, nIn, hashtype, Amount(amount), nullptr, flags);
        memcpy(out32, h.begin(), 32);
        return 1;
    } catch (...) {
        return 0;
    }
}

// VerifyScript with the non-caching TransactionSignatureChecker; returns 1/0, *err = ScriptError.
int ref_verify_script(const uint8_t *script_sig, size_t ss_len, const uint8_t *script_pubkey, size_t spk_len,
                      const uint8_t *tx, size_t txlen, unsigned nIn, int64_t amount, uint32_t flags, int *err) {
    ensure_init();
    try {
        CDataStream ss((const char *)tx, (const char *)tx + txlen, SER_NETWORK, PROTOCOL_VERSION);
        CTransaction t(deserialize, ss);
        CScript sig(script_sig, script_sig + ss_len), spk(script_pubkey, script_pubkey + spk_len);
        PrecomputedTransactionData txdata(t);
        ScriptError e = SCRIPT_ERR_UNKNOWN_ERROR;
        bool ok = VerifyScript(sig, spk, flags, TransactionSignatureChecker(&t, nIn, Amount(amount), txdata), &e);
        if (err) *err = (int)e;
        return ok ? 1 : 0;
    } catch (...) {
        if (err) *err = -1;
        return 0;
    }
}

// Verify every input of every tx in a serialized block of 1-in P2PKH-style spends against a
// caller-supplied prevout table (scriptPubKey bytes + amount per input, block order, coinbase
// skipped), multi-threaded; this is the per-input work CScriptCheck::operator() does
// (src/validation.cpp:1350-1359) with the non-caching checker.  ok[i] per input.
long ref_block_verify_inputs(const uint8_t *blk, size_t len, int legacy, const uint8_t *spk, const uint32_t *spk_off,
                             const int64_t *amounts, uint32_t flags, int nthreads, uint8_t *ok, size_t max_inputs,
                             uint8_t *err_out) {
    ensure_init();
    try {
        CDataStream ss((const char *)blk, (const char *)blk + len, SER_NETWORK,
                       PROTOCOL_VERSION | (legacy ? SERIALIZE_BLOCK_LEGACY : 0));
        CBlock block;
        ss >> block;
        struct Job { const CTransaction *tx; unsigned nIn; size_t idx; };
        std::vector<Job> jobs;
        std::vector<PrecomputedTransactionData> txd;
        txd.reserve(block.vtx.size());
        size_t idx = 0;
        for (size_t t = 0; t < block.vtx.size(); t++) {
            const CTransaction &tx = *block.vtx[t];
            txd.emplace_back(tx);
            if (tx.IsCoinBase()) continue;
            for (unsigned i = 0; i < tx.vin.size(); i++) jobs.push_back({&tx, i, idx++});
        }
        if (idx > max_inputs) return -2;
        // map job -> txdata index
        std::vector<size_t> txidx(jobs.size());
        {
            size_t j = 0;
            for (size_t t = 0; t < block.vtx.size(); t++) {
                if (block.vtx[t]->IsCoinBase()) continue;
                for (unsigned i = 0; i < block.vtx[t]->vin.size(); i++) txidx[j++] = t;
            }
        }
        parallel_for(jobs.size(), nthreads, [&](size_t lo, size_t hi) {
            for (size_t j = lo; j < hi; j++) {
                const Job &jb = jobs[j];
                CScript spkj(spk + spk_off[j], spk + spk_off[j + 1]);
                ScriptError e = SCRIPT_ERR_UNKNOWN_ERROR;
                ok[j] = VerifyScript(jb.tx->vin[jb.nIn].scriptSig, spkj, flags,
                                     TransactionSignatureChecker(jb.tx, jb.nIn, Amount(amounts[j]), txd[txidx[j]]), &e)
                            ? 1 : 0;
                if (err_out) err_out[j] = (uint8_t)e;
            }
        });
        return (long)jobs.size();
    } catch (...) {
        return -1;
    }
}

// ---- proof-of-work arithmetic (row N4).  src/pow.cpp itself does not compile outside the node (it includes
// util.h / validation.h, i.e. Boost), so its control flow is RESTATED here — but every 256-bit operation is the
// reference's own arith_uint256 (src/arith_uint256.cpp, compiled in place), which is what pins the product's U256.
struct ref_pow_params {
    uint8_t pow_limit[32], pow_limit_legacy[32], pow_limit_start[32];
    int32_t allow_min_difficulty, no_retargeting;
    int64_t target_spacing, target_timespan_legacy;
    int32_t bcp_height, premine_window;
    int64_t averaging_window;
    int32_t bip34_height, bip65_height, bip66_height;
    uint32_t eh_n, eh_k;
};
static arith_uint256 from_le(const uint8_t *b) {
    uint256 u;
    memcpy(u.begin(), b, 32);
    return UintToArith256(u);
}
static void to_le(const arith_uint256 &a, uint8_t *out) {
    uint256 u = ArithToUint256(a);
    memcpy(out, u.begin(), 32);
}
void ref_compact_to_target(uint32_t compact, uint8_t *out32, int *negative, int *overflow) {
    bool n, o;
    arith_uint256 t;
    t.SetCompact(compact, &n, &o);
    to_le(t, out32);
    *negative = n;
    *overflow = o;
}
uint32_t ref_compact_from_target(const uint8_t *t32) { return from_le(t32).GetCompact(); }
// GetBlockProof (src/chain.cpp:132-145)
static arith_uint256 block_proof(uint32_t nBits) {
    arith_uint256 t;
    bool n, o;
    t.SetCompact(nBits, &n, &o);
    if (n || o || t == 0) return 0;
    return (~t / (t + 1)) + 1;
}
void ref_block_proof(uint32_t nbits, uint8_t *out32) { to_le(block_proof(nbits), out32); }
// CheckProofOfWork (src/pow.cpp:141-161)
int ref_check_pow(const uint8_t *hash32, uint32_t nBits, const uint8_t *limit32) {
    bool fNegative, fOverflow;
    arith_uint256 bnTarget;
    bnTarget.SetCompact(nBits, &fNegative, &fOverflow);
    if (fNegative || bnTarget == 0 || fOverflow || bnTarget > from_le(limit32)) return 0;
    if (from_le(hash32) > bnTarget) return 0;
    return 1;
}
// GetNextWorkRequired(pindexPrev = block n-1 of the given chain, pblock->nTime = new_time) (src/pow.cpp:72-105) with
// BitcoinGetNextWorkRequired (:27-67), CalculateNextWorkRequired (:110-134), GetNextCashPlusWorkRequired (:252-295).
// which: 0 = GetNextWorkRequired, 1 = CalculateNextWorkRequired(prev, first_time = new_time), 2 = GetNextCashPlusWorkRequired
uint32_t ref_next_work(const ref_pow_params *p, const uint32_t *times, const uint32_t *bits, size_t n, int64_t new_time, int which) {
    const arith_uint256 limPost = from_le(p->pow_limit), limLegacy = from_le(p->pow_limit_legacy);
    auto limit = [&](bool postfork) { return postfork ? limPost : limLegacy; };
    if (n == 0) return limLegacy.GetCompact();
    const int prev = (int)n - 1;
    auto calc = [&](int64_t nFirstBlockTime) -> uint32_t {
        if (p->no_retargeting) return bits[prev];
        int64_t nActualTimespan = (int64_t)times[prev] - nFirstBlockTime;
        if (nActualTimespan < p->target_timespan_legacy / 4) nActualTimespan = p->target_timespan_legacy / 4;
        if (nActualTimespan > p->target_timespan_legacy * 4) nActualTimespan = p->target_timespan_legacy * 4;
        arith_uint256 bnNew;
        bnNew.SetCompact(bits[prev]);
        bnNew *= nActualTimespan;
        bnNew /= p->target_timespan_legacy;
        if (bnNew > limLegacy) bnNew = limLegacy;
        return bnNew.GetCompact();
    };
    auto cashplus = [&]() -> uint32_t {
        const bool postfork = prev >= p->bcp_height;
        if (p->allow_min_difficulty && new_time > (int64_t)times[prev] + 2 * p->target_spacing) return limit(postfork).GetCompact();
        std::vector<arith_uint256> work(n);
        for (size_t i = 0; i < n; i++) work[i] = (i ? work[i - 1] : arith_uint256(0)) + block_proof(bits[i]);
        auto suitable = [&](int h) {
            int b[3] = {h - 2, h - 1, h};
            if (times[b[0]] > times[b[2]]) std::swap(b[0], b[2]);
            if (times[b[0]] > times[b[1]]) std::swap(b[0], b[1]);
            if (times[b[1]] > times[b[2]]) std::swap(b[1], b[2]);
            return b[1];
        };
        const int last = suitable(prev), first = suitable(prev - 144);
        arith_uint256 w = work[last] - work[first];
        w *= p->target_spacing;
        int64_t nActualTimespan = int64_t(times[last]) - int64_t(times[first]);
        if (nActualTimespan > 288 * p->target_spacing) nActualTimespan = 288 * p->target_spacing;
        else if (nActualTimespan < 72 * p->target_spacing) nActualTimespan = 72 * p->target_spacing;
        w /= nActualTimespan;
        const arith_uint256 nextTarget = (-w) / w;
        if (nextTarget > limit(postfork)) return limit(postfork).GetCompact();
        return nextTarget.GetCompact();
    };
    if (which == 1) return calc(new_time);
    if (which == 2) return cashplus();
    if (p->no_retargeting) return bits[prev];
    const int nHeight = prev + 1;
    const bool postfork = nHeight >= p->bcp_height;
    if (!postfork) {
        const uint32_t nProofOfWorkLimit = limLegacy.GetCompact();
        const int64_t interval = p->target_timespan_legacy / p->target_spacing;
        if ((prev + 1) % interval != 0) {
            if (p->allow_min_difficulty) {
                if (new_time > (int64_t)times[prev] + p->target_spacing * 2) return nProofOfWorkLimit;
                int i = prev;
                while (i > 0 && i % interval != 0 && bits[i] == nProofOfWorkLimit) i--;
                return bits[i];
            }
            return bits[prev];
        }
        return calc((int64_t)times[prev - (interval - 1)]);
    }
    if (nHeight < p->bcp_height + p->premine_window) return limPost.GetCompact();
    if (nHeight < p->bcp_height + p->premine_window + p->averaging_window) return from_le(p->pow_limit_start).GetCompact();
    return cashplus();
}

}  // extern "C"
