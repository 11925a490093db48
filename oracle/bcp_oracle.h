/* oracle/bcp_oracle.h — CPU ORACLE, TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's CPU algorithms for the block-validation hot path
 * (SURVEY.md §8a).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library; the product library (libbcp_b200.so) never links,
 * loads or calls it, and fails loudly without CUDA.
 *
 * Parity pinning: tests/test_oracle_*.py check every function here against the reference's own
 * golden vectors (tests/golden/, extracted from /root/reference tests with the committed
 * scripts) and, when oracle/_ref/libbcp_ref.so is present, differentially against the reference
 * code itself.  Blake2b/Equihash: the reference holds NO vectors for them and its Blake2b lives in
 * the external libsodium 1.0.15 (depends/packages/libsodium.mk:6-9); that part is pinned by
 * RFC 7693 KATs, python hashlib.blake2b(person=...) fixtures and by solutions found by the
 * reference's own solver (oracle/_ref) — see DESIGN.md "parity pinning".
 */
#ifndef BCP_ORACLE_H
#define BCP_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- SHA-256 / SHA-256d / RIPEMD-160 (src/crypto/sha256.cpp:57-243, src/hash.h:21-68, src/crypto/ripemd160.cpp) */
void orc_sha256(const uint8_t *data, size_t len, uint8_t out32[32]);
void orc_sha256d(const uint8_t *data, size_t len, uint8_t out32[32]);
void orc_ripemd160(const uint8_t *data, size_t len, uint8_t out20[20]);
void orc_hash160(const uint8_t *data, size_t len, uint8_t out20[20]);

/* ---- Merkle root + CVE-2012-2459 flag (src/consensus/merkle.cpp:47-150) */
void orc_merkle_root(const uint8_t *leaves32, size_t n, uint8_t root32[32], int *mutated);

/* ---- Block / tx framing (src/primitives/transaction.h:223-243, src/serialize.h:260-300,
 *      src/primitives/block.h:49-73).  Fills tx_off[0..ntx] (byte offsets of each tx; tx_off[ntx]=end).
 *      Returns ntx or <0 on malformed input.  legacy=1: 80-byte header. */
long orc_block_index(const uint8_t *blk, size_t len, int legacy, uint32_t *tx_off, size_t max_tx);
long orc_block_txids(const uint8_t *blk, size_t len, int legacy, uint8_t *txids, size_t max_tx, uint8_t root32[32],
                     int *mutated);
/* header hash: SHA256d over the serialized header bytes (src/primitives/block.cpp:17-28) */
void orc_header_hash(const uint8_t *hdr, size_t len, uint8_t out32[32]);
/* CheckProofOfWork (src/pow.cpp:141-161): hash (uint256 LE bytes) <= target(nBits) and target <= powLimit */
int orc_check_pow(const uint8_t hash32[32], uint32_t nbits, const uint8_t pow_limit32[32]);

/* ---- Blake2b (RFC 7693; libsodium crypto_generichash_blake2b_* semantics) */
void orc_blake2b(const uint8_t *in, size_t inlen, const uint8_t *personal16, size_t outlen, uint8_t *out);

/* ---- Equihash (src/pow.cpp:298-323, src/crypto/equihash.cpp:35-103,176-190,724-770, equihash.tcc:10-21)
 *      hdr = serialized post-fork header: 140 B || compactsize(soln) || soln.  1 valid / 0 invalid. */
int orc_equihash_verify(unsigned n, unsigned k, const uint8_t *hdr, size_t len);
void orc_equihash_verify_batch(unsigned n, unsigned k, const uint8_t *hdrs, size_t stride, size_t len, size_t count,
                               int nthreads, uint8_t *ok);

/* ---- secp256k1 ECDSA (src/secp256k1/src/secp256k1.c:150-308, ecdsa_impl.h:201-269, eckey_impl.h:17-35,
 *      group_impl.h, field_impl.h, scalar_impl.h; src/pubkey.cpp:26-193,294-302) */
int orc_pubkey_parse(const uint8_t *in, size_t len, uint8_t xy64[64]); /* affine x||y big-endian */
int orc_sig_parse_compact(const uint8_t in64[64], uint8_t out64[64]);   /* 0 on overflow (out zeroed) */
int orc_sig_normalize(uint8_t sig64[64]);                               /* returns 1 if s was high */
int orc_der_lax_parse(const uint8_t *in, size_t len, uint8_t sig64[64]);
int orc_check_low_s(const uint8_t *der, size_t len);
int orc_ecdsa_verify(const uint8_t msg32[32], const uint8_t *pub, size_t publen, const uint8_t sig64[64]);
int orc_pubkey_verify(const uint8_t hash32[32], const uint8_t *pub, size_t publen, const uint8_t *der, size_t derlen);
/* src/secp256k1/src/modules/recovery/main_impl.h:38-58, 87-121, 170-190; src/secp256k1/src/secp256k1.c:477-499 */
int orc_ecdsa_recover(const uint8_t msg32[32], const uint8_t sig64[64], int recid, uint8_t xy64[64]);
int orc_pubkey_tweak_add(const uint8_t *pub, size_t publen, const uint8_t tweak32[32], uint8_t xy64[64]);
/* B4 batch layout: pubkeys at stride 65, length by prefix byte (src/pubkey.h:47-51) */
void orc_ecdsa_verify_batch(const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64, size_t n, int nthreads,
                            uint8_t *ok);
int orc_pubkey_create(const uint8_t seckey32[32], int compressed, uint8_t *out, size_t *outlen);
int orc_sign_rfc6979(const uint8_t seckey32[32], const uint8_t msg32[32], uint8_t sig64[64]);
int orc_sig_to_der(const uint8_t sig64[64], uint8_t *der, size_t *derlen);
/* deterministic synthetic signature set: fills msg32[n*32], pub65[n*65], sig64[n*64] */
void orc_gen_sigset(const uint8_t *seed, size_t seedlen, size_t n, size_t nkeys, int nthreads, uint8_t *msg32,
                    uint8_t *pub65, uint8_t *sig64);

/* ---- FORKID sighash (src/script/interpreter.cpp:1321-1404) for a serialized tx */
int orc_sighash_forkid(const uint8_t *script_code, size_t sc_len, const uint8_t *tx, size_t txlen, unsigned nin,
                       uint32_t hashtype, int64_t amount, uint8_t out32[32]);

/* ---- P2PKH fast path of CScriptCheck::operator() (src/validation.cpp:1350-1359 -> VerifyScript
 *      src/script/interpreter.cpp:1551-1630 on scriptSig = <sig> <pubkey>, scriptPubKey = DUP HASH160 <20> EQUALVERIFY
 *      CHECKSIG) under the post-fork flag set (P2SH|STRICTENC|DERSIG|LOW_S|NULLFAIL|FORKID [+CLTV|CSV]).
 *      Per non-coinbase input, in block order, err[i] = the ScriptError VerifyScript reports (0 = OK), or 255 when
 *      the input is not of this shape (the caller then runs the full interpreter).  spk/spk_off/amounts describe
 *      the spent outputs.  Returns the number of inputs, <0 on a malformed block. */
long orc_block_index_inputs(const uint8_t *blk, size_t len, int legacy, uint32_t *tx_off, size_t max_tx, uint32_t *in_off,
                            uint32_t *in_tx, uint32_t *in_n, size_t max_inputs, long *ntx_out);
long orc_block_verify_p2pkh(const uint8_t *blk, size_t len, int legacy, const uint8_t *spk, const uint32_t *spk_off,
                            const int64_t *amounts, uint32_t flags, int nthreads, uint8_t *err, size_t max_inputs);
int orc_is_valid_signature_encoding(const uint8_t *sig, size_t len);
/* synthetic chain generator at C speed (same shapes as tests/synth.py) */
size_t orc_gen_chain_txs(const uint8_t *seed, size_t seedlen, size_t key_base, size_t nkeys, size_t ntx, uint8_t *prev_txid,
                         uint32_t *prev_key, int64_t *prev_amt, int nthreads, uint8_t *out, size_t out_cap);

#ifdef __cplusplus
}
#endif
#endif
