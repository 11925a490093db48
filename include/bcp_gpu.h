/* bcp_gpu.h — C ABI of libbcp_b200.so: the B200 (sm_100a) block-validation hot path.
 *
 * This is the "B4" boundary of SURVEY.md §8(b): thin extern "C" entry points that take a whole
 * batch (all inputs of a block, all headers of a sync window) in plain host or device buffers.
 * The reference-side shims that call them — include/compat/secp256k1.h (same-name libsecp256k1
 * C API), include/compat/checkqueue.h (CCheckQueue drop-in) — and the cgo/ctypes-style stubs a
 * maintainer would add are shown in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns BCP_OK (0) or a negative BCP_ERR_* code; a verdict is NEVER produced on
 *     error (no CPU fallback exists: without a CUDA device every call fails with BCP_ERR_NO_DEVICE);
 *   - `device` is a CUDA ordinal; bcp_gpu_init(device) must have succeeded for it;
 *   - "_dev" variants take DEVICE pointers and a cudaStream_t (as void*), enqueue work and do not
 *     synchronise; the plain variants take HOST pointers (pinned memory from bcp_gpu_host_alloc gives
 *     full PCIe bandwidth; pageable memory works, slower), copy in, run, copy out and synchronise;
 *   - one context per device; calls on the same device are serialised by an internal mutex, calls
 *     on different devices run concurrently (shard a block's inputs across devices by index range and
 *     AND the verdicts on the host, like `fAllOk &= fOk` in src/checkqueue.h:77 — no NCCL is involved).
 */
#ifndef BCP_GPU_H
#define BCP_GPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define BCP_API __attribute__((visibility("default")))
#else
#define BCP_API
#endif

#define BCP_OK 0
#define BCP_ERR_NO_DEVICE (-1)   /* no CUDA device / driver, or device ordinal out of range */
#define BCP_ERR_CUDA (-2)        /* a CUDA runtime call failed: see bcp_gpu_last_error()     */
#define BCP_ERR_NOT_INIT (-3)    /* bcp_gpu_init(device) has not been called                 */
#define BCP_ERR_INVALID (-4)     /* bad argument (NULL buffer, size beyond a stated limit)   */

/* flag for the ECDSA entry points: normalise a high s (s := n - s) instead of rejecting it, which is
 * what CPubKey::Verify does before calling libsecp256k1 (src/pubkey.cpp:186-191).  Without the flag
 * the verdict is that of a bare secp256k1_ecdsa_verify (src/secp256k1/src/secp256k1.c:293-308). */
#define BCP_ECDSA_FLAG_NORMALIZE_S 1u

/* ---- life cycle ------------------------------------------------------------------------------- */
/* Number of usable CUDA devices (0 when there is no driver/GPU). */
BCP_API int bcp_gpu_device_count(void);
/* 1 when two GPUs can access each other's memory directly (NVLink / NVSwitch, PCIe peer-to-peer), 0 when a copy between them
 * would be staged through host memory, negative on error.  bcp::DeviceChain deals blocks only to GPUs that reach the coin
 * table's GPU this way.  No reference counterpart. */
BCP_API int bcp_gpu_peer_access(int device_a, int device_b);
/* Create the per-device context: stream, fixed-base comb table for G (generated on the device; the
 * counterpart of secp256k1_ecmult_context_build, src/secp256k1/src/ecmult_impl.h:164-194, which
 * secp256k1_context_create(SECP256K1_CONTEXT_VERIFY) runs, src/secp256k1/src/secp256k1.c:59-83).
 * Idempotent.
 * Lanes: every `device` argument of this API may carry a lane number, BCP_DEVICE_LANE(ordinal, lane) with lane in
 * 0..3.  A lane is an independent context on the same GPU — its own streams, staging buffer and scratch; the comb
 * table is shared — so host threads that each use their own lane run their calls concurrently (one's copies under
 * the other's kernels).  Calls on the SAME lane are serialised by a mutex.  Lane 0 is the plain ordinal. */
#define BCP_DEVICE_LANE(ordinal, lane) ((ordinal) | ((lane) << 8))
BCP_API int bcp_gpu_init(int device);
/* Destroy it (secp256k1_context_destroy, secp256k1.c:96-105); the plain ordinal also destroys the GPU's other lanes. */
BCP_API int bcp_gpu_shutdown(int device);
/* Device self-test of the field arithmetic's rarely taken carry paths on crafted operands (run by bcp_gpu_init as well;
 * the counterpart of ECC_InitSanityCheck, src/key.cpp:336-342, which init.cpp runs at start-up). */
BCP_API int bcp_gpu_selftest(int device);
/* Text of the last error on the calling thread. */
BCP_API const char *bcp_gpu_last_error(void);
/* Pinned host memory for staging buffers (what "pinned buffers" in the north star refers to). */
BCP_API void *bcp_gpu_host_alloc(size_t bytes);
BCP_API void bcp_gpu_host_free(void *p);

/* ---- ECDSA: replaces the per-input work of CPubKey::Verify (src/pubkey.cpp:170-193) ==
 *      secp256k1_ec_pubkey_parse (secp256k1.c:150) + secp256k1_ecdsa_signature_parse_compact (:228)
 *      [+ _normalize (:274)] + secp256k1_ecdsa_verify (:293), for n signatures at once.
 *   msg32  n x 32  sighash digests, raw bytes as uint256::begin() passes them (pubkey.cpp:191)
 *   pub65  n x 65  serialized public keys, each left-aligned in a 65-byte slot; the length is implied
 *                  by the first byte exactly as CPubKey::GetLen does (src/pubkey.h:47-51): 02/03 -> 33,
 *                  04/06/07 -> 65, anything else -> invalid key -> verdict 0
 *   sig64  n x 64  r || s big-endian (the output of the lax DER parser, see bcp_der_lax_parse)
 *   ok     n       verdict bytes, 1 = valid, 0 = invalid
 */
BCP_API int bcp_gpu_ecdsa_verify_batch(int device, const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64,
                               size_t n, uint32_t flags, uint8_t *ok);
BCP_API int bcp_gpu_ecdsa_verify_batch_dev(int device, const uint8_t *d_msg32, const uint8_t *d_pub65,
                                   const uint8_t *d_sig64, size_t n, uint32_t flags, uint8_t *d_ok, void *stream);
/* NOTE on the "_dev" forms: they use the context's scratch buffer.  Work enqueued by two "_dev" calls on the same
 * device must not overlap on the GPU (use one stream, or synchronise between calls on different streams).
 * Bytes of device scratch bcp_gpu_ecdsa_verify_batch_dev needs for n signatures (it grows the
 * context's scratch buffer on demand; calling this first with the largest n avoids a reallocation
 * inside a timed region). */
BCP_API size_t bcp_gpu_ecdsa_scratch_bytes(size_t n);
BCP_API int bcp_gpu_ecdsa_reserve(int device, size_t n);

/* secp256k1_ec_pubkey_parse (src/secp256k1/src/secp256k1.c:150-163, eckey_impl.h:17-35) for n keys in the
 * same 65-byte-slot layout: xy64[i] = affine x || y big-endian (zeros when invalid), ok[i] = 1/0. */
BCP_API int bcp_gpu_pubkey_parse_batch(int device, const uint8_t *pub65, size_t n, uint8_t *xy64, uint8_t *ok);

/* secp256k1_ecdsa_recoverable_signature_parse_compact + secp256k1_ecdsa_recover
 * (src/secp256k1/include/secp256k1_recovery.h:36,99; src/secp256k1/src/modules/recovery/main_impl.h:38-58, 87-121,
 * 170-190) for n items — what CPubKey::RecoverCompact (src/pubkey.cpp:195-218) calls: sig64[i] = r || s big-endian,
 * recid[i] in 0..3, msg32[i] = the signed hash.  xy64[i] = affine x || y of the recovered key (zeros on failure),
 * ok[i] = 1/0 (r or s out of range or zero, recid > 3, no point with that x, result at infinity => 0). */
BCP_API int bcp_gpu_ecdsa_recover_batch(int device, const uint8_t *msg32, const uint8_t *sig64, const uint8_t *recid, size_t n,
                                        uint8_t *xy64, uint8_t *ok);
/* secp256k1_ec_pubkey_tweak_add (src/secp256k1/include/secp256k1.h:544, src/secp256k1/src/secp256k1.c:477-499) for n
 * keys in the 65-byte-slot layout — what CPubKey::Derive (src/pubkey.cpp:240-264) calls: xy64[i] = pub[i] + tweak32[i]*G,
 * ok[i] = 0 when the key does not parse, the tweak is >= n or the sum is the point at infinity. */
BCP_API int bcp_gpu_pubkey_tweak_add_batch(int device, const uint8_t *pub65, const uint8_t *tweak32, size_t n, uint8_t *xy64,
                                           uint8_t *ok);

/* Host-side byte parsing that stays on the CPU by design (branchy, a few dozen bytes):
 * ecdsa_signature_parse_der_lax (src/pubkey.cpp:26-168).  Returns 1 and writes r||s (zeros when a
 * component overflows, which later verifies as invalid), or 0 on a structural failure. */
BCP_API int bcp_der_lax_parse(const uint8_t *der, size_t len, uint8_t sig64[64]);

/* ---- SHA-256d: transaction ids, Merkle root, header hashes ---------------------------------------
 * Host-side framing scan (sequential byte parsing, stays on the CPU by design): fills tx_off[0..ntx]
 * with the byte offset of every transaction of a serialized block (tx_off[ntx] = end), following
 * ReadCompactSize (src/serialize.h:260-300), UnserializeTransaction (src/primitives/transaction.h:223-234)
 * and the header layout (src/primitives/block.h:49-73; legacy_header = 1 for the 80-byte format).
 * Returns ntx, -1 on malformed input, -2 if ntx > max_tx. */
BCP_API long bcp_block_index(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, size_t max_tx);
/* txid[i] = SHA256d(blk[tx_off[i] .. tx_off[i+1])) for every transaction — CTransaction::ComputeHash
 * (src/primitives/transaction.cpp:73-75) — then BlockMerkleRoot (src/consensus/merkle.cpp:175-182):
 * root32 and the CVE-2012-2459 `mutated` flag exactly as ComputeMerkleRoot (:47-150) reports them
 * (ntx == 0 gives a zero root).  txids may be NULL. */
BCP_API int bcp_gpu_sha256d_txids_merkle(int device, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                         uint8_t *txids, uint8_t root32[32], int *mutated);
/* ComputeMerkleRoot (src/consensus/merkle.cpp:146-150) over n 32-byte leaves. */
BCP_API int bcp_gpu_merkle_root(int device, const uint8_t *leaves32, size_t n, uint8_t root32[32], int *mutated);
/* out32[i] = SHA256d(data[off[i] .. off[i+1])) — CHash256 (src/hash.h:21-43) over a batch of messages. */
BCP_API int bcp_gpu_sha256d_batch(int device, const uint8_t *data, size_t len, const uint32_t *off, size_t count,
                                  uint8_t *out32);
/* Device-pointer forms (enqueue only).  d_data must be 4-byte aligned and readable 4 bytes past its end. */
BCP_API int bcp_gpu_sha256d_batch_dev(int device, const uint8_t *d_data, const uint32_t *d_off, size_t count,
                                      uint8_t *d_out32, void *stream);
BCP_API size_t bcp_gpu_merkle_scratch_bytes(size_t n);
BCP_API int bcp_gpu_merkle_root_dev(int device, const uint8_t *d_leaves32, size_t n, uint8_t *d_root32, int *d_mutated,
                                    void *d_scratch, void *stream);

/* Allocation-free variant of bcp_gpu_merkle_root_dev.  It has no counterpart in the reference; it exists
 * for callers that cannot allocate device memory: a validation thread that must never trigger the
 * device-wide synchronisation of cudaMalloc/cudaFree while it holds the node's main lock, or an embedder
 * that runs the library inside a fixed VRAM budget.  The inner-node scratch is a static device array built
 * into the library (about 256 KiB per GPU, held until process exit) sized for at most
 * BCP_MERKLE_STATIC_MAX_LEAVES = 1,048,576 leaves; larger n returns BCP_ERR_INVALID with a message.  The call itself performs no cudaMalloc, cudaFree or other allocating call; inputs and outputs
 * (d_leaves32, d_root32[32], d_mutated[1]) are the caller's device buffers.
 *   - The first call on a device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call before starting any stream capture.
 *   - All calls on one device share the one array.  It keeps whatever the previous call left there (no
 *     part of it needs clearing: every node is written before it is read), so two calls whose GPU work
 *     overlaps — different streams, or different host threads — race on it.  Order them on one stream. */
#define BCP_MERKLE_STATIC_MAX_LEAVES (1u << 20)
BCP_API int bcp_gpu_merkle_root_dev_static(int device, const uint8_t *d_leaves32, size_t n, uint8_t *d_root32,
                                           int *d_mutated, void *stream);

/* ---- Whole-block fast path: txids + Merkle root + every input's script check for the
 *      dominant script shape, from the raw block bytes and the spent outputs ConnectBlock fetched from the UTXO
 *      view (src/validation.cpp:1460-1471).  Replaces, for P2PKH inputs, the body of CScriptCheck::operator()
 *      (src/validation.cpp:1350-1359): VerifyScript (src/script/interpreter.cpp:1551) over scriptSig `<sig> <pubkey>` and
 *      scriptPubKey `DUP HASH160 <20> EQUALVERIFY CHECKSIG`, including SignatureHash's FORKID branch (:1357-1404) and
 *      PrecomputedTransactionData (:1347-1352), under the post-fork flags of GetBlockScriptFlags (validation.cpp:1803-1852).
 *   spk / spk_off / amounts   the spent scriptPubKeys (concatenated, spk_off[n_inputs+1]) and values, one per
 *                             non-coinbase input in block order
 *   flags                     SCRIPT_VERIFY_* bits as in src/script/interpreter.h:32-121; must contain
 *                             STRICTENC|DERSIG|LOW_S|NULLFAIL|ENABLE_SIGHASH_FORKID and not ALLOW_NON_FORKID.  Of the
 *                             other bits only those that cannot change the result for this script shape are accepted
 *                             (P2SH, NULLDUMMY, SIGPUSHONLY, DISCOURAGE_UPGRADABLE_NOPS, CLEANSTACK with P2SH, CLTV,
 *                             CSV, MINIMALIF); MINIMALDATA, COMPRESSED_PUBKEYTYPE and unknown bits are refused with
 *                             BCP_ERR_INVALID rather than silently ignored
 *   script_err[i]             the ScriptError VerifyScript reports for input i (src/script/script_error.h; 0 = OK), or
 *                             255 when the input is not of this shape and must be run through the host interpreter
 *   all_ok                    1 iff every script_err is 0 (the `fAllOk` of the check queue)
 *   txids (optional), root32, mutated   as bcp_gpu_sha256d_txids_merkle */
BCP_API long bcp_block_index_inputs(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, size_t max_tx,
                                    uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs, long *ntx_out);
/* Same scan for callers that go on to read the outputs (a block connector inserting the new coins):
 *   tx_out_off[t] (optional)            offset of transaction t's output count
 *   out_off / out_tx / out_n (optional) every output, flat like the inputs: offset of its 8-byte amount, its
 *                                       transaction, its index; *nout_out = how many (max_outputs = capacity, -2 beyond)
 *   in_key (optional)                   per input, a 64-bit mix of its outpoint for hash tables:
 *                                       le64(txid[0..8)) ^ 0x9E3779B97F4A7C15 * vout */
BCP_API long bcp_block_index_full(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, uint32_t *tx_out_off,
                                  size_t max_tx, uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs,
                                  long *ntx_out, uint32_t *out_off, uint32_t *out_tx, uint32_t *out_n, size_t max_outputs,
                                  long *nout_out, uint64_t *in_key);
/* The same scan applying, in the same pass, CheckBlock's context-free transaction rules (src/validation.cpp:3255-3330 ->
 * CheckCoinbase :531-549, CheckRegularTransaction :551-570, CheckTransactionCommon :476-529) in the reference's order.
 * *rule = 0 or the first violation: 1 bad-cb-missing, 2 bad-tx-coinbase, 3 bad-txns-vin-empty, 4 bad-txns-vout-empty,
 * 5 bad-txns-oversize, 6 bad-txns-vout-negative, 7 bad-txns-vout-toolarge, 8 bad-txns-txouttotal-toolarge, 9 bad-cb-length,
 * 10 bad-txns-prevout-null (bcp_block_rule_name gives the string); *rule_tx = the transaction.  Not covered: sigop counts. */
BCP_API long bcp_block_index_rules(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, uint32_t *tx_out_off,
                                   size_t max_tx, uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs,
                                   long *ntx_out, uint32_t *out_off, uint32_t *out_tx, uint32_t *out_n, size_t max_outputs,
                                   long *nout_out, uint64_t *in_key, int *rule, long *rule_tx);
BCP_API const char *bcp_block_rule_name(int rule);
/* The script checks alone, for any subset of a block's inputs chosen by the caller (arrays describe the subset;
 * spk_off may point into a larger offsets array).  This is the unit that shards across GPUs: one call per device
 * over its index range, verdicts AND-ed on the host like `fAllOk &= fOk` (src/checkqueue.h:77).  The subset must be
 * listed in block order (in_tx non-decreasing) and every in_off must lie inside the transaction in_tx names — the
 * kernels stage the span of transactions each 128-input group touches; anything else is BCP_ERR_INVALID. */
BCP_API int bcp_gpu_block_scripts_p2pkh(int device, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                        const uint32_t *in_off, const uint32_t *in_tx, const uint32_t *in_n, size_t n_inputs,
                                        const uint8_t *spk, const uint32_t *spk_off, const int64_t *amounts, uint32_t flags,
                                        uint8_t *script_err, int *all_ok);
BCP_API int bcp_gpu_block_connect_p2pkh(int device, const uint8_t *blk, size_t len, int legacy_header, const uint8_t *spk,
                                        const uint32_t *spk_off, const int64_t *amounts, size_t n_inputs, uint32_t flags,
                                        uint8_t *txids, uint8_t root32[32], int *mutated, uint8_t *script_err, int *all_ok);

/* A WINDOW of blocks in one pass (reindex / initial block download: LoadExternalBlockFile hands over block after
 * block, src/validation.cpp:4486-4560).  Same results per block as bcp_gpu_block_connect_p2pkh; the per-input kernels
 * run once over the union of the window's inputs, so the GPU sees several waves of work instead of half of one.
 *   blks/lens     nblk raw blocks, each ending with its last transaction (as blk*.dat stores them); pinned memory
 *                 lets the copies overlap the host scans, which run one thread per block
 *   n_inputs      non-coinbase inputs of each block; spk / spk_off / amounts / script_err cover the concatenation of
 *                 all blocks' inputs in window order (spk_off has sum(n_inputs)+1 entries)
 *   txids         optional, the concatenation of every block's txids; ntx_out (optional) = transactions per block
 *   roots32, mutated, all_ok   one entry per block */
BCP_API int bcp_gpu_blocks_connect_p2pkh(int device, size_t nblk, const uint8_t *const *blks, const size_t *lens,
                                         int legacy_header, const uint8_t *spk, const uint32_t *spk_off,
                                         const int64_t *amounts, const size_t *n_inputs, uint32_t flags, uint8_t *txids,
                                         size_t *ntx_out, uint8_t *roots32, int *mutated, uint8_t *script_err, int *all_ok);

/* ---- Equihash: CheckEquihashSolution (src/pow.cpp:298-323) == Equihash<N,K>::IsValidSolution
 *      (src/crypto/equihash.cpp:724-770) for `count` serialized post-fork headers (140 bytes || compact
 *      size || solution; src/primitives/block.h:49-73) laid out at `stride` bytes, each `hdr_len` long.
 *      (n,k) in {(200,9) main/test, (48,5) regtest, (96,5), (96,3)} as src/crypto/equihash.h:196-272.
 *   ok      count verdict bytes
 *   hash32  optional: SHA-256d of each header's hdr_len bytes = CBlockHeader::GetHash
 *           (src/primitives/block.cpp:17-28), raw digest byte order (uint256 memory order) */
BCP_API int bcp_gpu_equihash_verify_batch(int device, unsigned n, unsigned k, const uint8_t *hdrs, size_t stride,
                                          size_t hdr_len, size_t count, uint8_t *ok, uint8_t *hash32);
BCP_API size_t bcp_gpu_equihash_scratch_bytes(size_t count);
BCP_API int bcp_gpu_equihash_verify_batch_dev(int device, unsigned n, unsigned k, const uint8_t *d_hdrs, size_t stride,
                                              size_t hdr_len, size_t count, uint8_t *d_ok, void *d_scratch,
                                              void *stream);

/* ---- Header pipeline (headers-first sync): for `count` serialized post-fork headers, everything
 *      CheckBlockHeader computes (src/validation.cpp:3202-3220) in one pass over one upload:
 *   ok_solution  CheckEquihashSolution (src/pow.cpp:298-323)
 *   ok_pow       CheckProofOfWork(GetHash(), nBits, IsBCPEnabled(nHeight)) (src/pow.cpp:141-161): target from the
 *                compact form (arith_uint256::SetCompact, src/arith_uint256.cpp:181-199), not negative / zero /
 *                overflowing / above the limit, hash <= target.  The limit is pow_limit32 for headers whose own
 *                nHeight field is >= fork_height, pow_limit_legacy32 otherwise (Consensus::Params::PowLimit,
 *                src/consensus/params.h:88); both as uint256 memory bytes (little-endian).
 *   linked       optional: 1 for header 0, else hashPrevBlock == hash of the header before it in the batch
 *   hash32       CBlockHeader::GetHash of each header
 * The sequential rest (difficulty, median time past, versions) is host work: include/compat/bcp_headers.h. */
BCP_API int bcp_gpu_header_check_batch(int device, unsigned n, unsigned k, const uint8_t *hdrs, size_t stride, size_t hdr_len,
                                       size_t count, const uint8_t pow_limit32[32], const uint8_t pow_limit_legacy32[32],
                                       uint32_t fork_height, uint8_t *ok_solution, uint8_t *ok_pow, uint8_t *linked,
                                       uint8_t *hash32);

/* ---- measurement helper: integer-pipe peak (SURVEY.md §8d asks the build to measure it) ----------
 * Runs a dependency-free IMAD.WIDE.U32 loop on every SM and returns 32x32->64 multiply-adds per second
 * (which == 0), the carry-chained form mad.lo.cc/madc.hi.cc used by the field code (which == 1), or
 * IADD3-class adds per second (which == 2); 3: mad.lo.u32, 4: mad.hi.u32, 5: fma.rn.f64, 6: fma.rn.f32
 * (the last four only to place the wide multiply against the other pipes). */
BCP_API int bcp_gpu_measure_int_peak(int device, int which, double *ops_per_second);

/* ---- Device-resident coin table (the device side): the working set of unspent outputs of a
 *      reindex / initial block download kept in HBM, so that what ConnectBlock does per input and per output against its
 *      coin view — "the spent coin exists and is unspent", fetch its amount and scriptPubKey, spend it, add the new coins,
 *      the BIP30 overwrite test (src/validation.cpp:1975-1990, 2030-2100; UpdateCoins :1327-1347; src/coins.cpp:71-97) —
 *      is a few kernels per block on one stream (bitcoincashplus_b200/csrc/bcp_utxo.cu).  Blocks are applied in call
 *      order; block b+1 sees the table as block b left it, whether or not b's script checks have finished, and a block
 *      whose checks fail is undone with bcp_gpu_utxo_rollback (newest first).  Sequence numbers are the caller's, start at
 *      1 and must increase; at most 32 blocks may be applied and neither confirmed nor rolled back.
 *   bcp_gpu_utxo_create     2^log2_slots slots of 128 bytes (keep the table under half full); arena_bytes for scriptPubKeys
 *                           longer than 64 bytes (never freed).  NULL on failure.
 *   bcp_gpu_utxo_add_coins  seeds coins: outpoints36[i] = txid || vout (LE), spk / spk_off / amounts as elsewhere
 *   bcp_gpu_utxo_apply_block   ASYNCHRONOUS.  Inputs: the raw block and the arrays of bcp_block_index_full (non-coinbase
 *                           inputs, all outputs).  Enqueues upload, txids + Merkle root, insertion of the outputs (an
 *                           output a transaction of the SAME block spends must belong to an earlier transaction), the
 *                           lookups + spends, and the copies of the results into the caller's HOST buffers, which must
 *                           stay valid until bcp_gpu_utxo_wait(seq) returns (pinned memory keeps the call asynchronous):
 *                           txids (optional), root32, mutated, and for the script checks amounts[nin], spk_off[nin+1],
 *                           spk_cat (spk_cap bytes: the spent scriptPubKeys back to back) — all three, or all three NULL
 *                           when the checks will run from the device copies (bcp_gpu_utxo_block_scripts)
 *   bcp_gpu_utxo_wait       blocks until block seq's results are in the host buffers.  status8: [0] inputs whose coin is
 *                           missing or already spent (bad-txns-inputs-missingorspent when non-zero), [1] the first such
 *                           input, [2] outputs that overwrite a live coin (bad-txns-BIP30), [3] bytes of spk_cat used,
 *                           [4] coins that did not fit the table, [5] long scripts that did not fit the arena, [6] spent
 *                           scripts that did not fit the script buffer
 *   bcp_gpu_utxo_confirm    block seq is final (oldest first): its undo record is dropped, coins it spent may be overwritten
 *   bcp_gpu_utxo_rollback   undoes block seq (newest first): its spent coins are unspent again, its new coins are gone
 *   bcp_gpu_utxo_size       number of unspent coins (synchronises with the table's stream) */
typedef struct bcp_utxo bcp_utxo;
BCP_API bcp_utxo *bcp_gpu_utxo_create(int device, unsigned log2_slots, size_t arena_bytes);
BCP_API void bcp_gpu_utxo_destroy(bcp_utxo *u);
BCP_API int bcp_gpu_utxo_add_coins(bcp_utxo *u, const uint8_t *outpoints36, const uint8_t *spk, const uint32_t *spk_off,
                                   const int64_t *amounts, size_t n);
BCP_API long long bcp_gpu_utxo_size(bcp_utxo *u);
/* Optional: allocate every pending-block area for blocks of up to max_block_bytes now, so that no block of a run meets a
 * device allocation (a device-wide synchronisation) inside bcp_gpu_utxo_apply_block.  No reference counterpart. */
BCP_API int bcp_gpu_utxo_reserve(bcp_utxo *u, size_t max_block_bytes);
BCP_API int bcp_gpu_utxo_apply_block(bcp_utxo *u, uint32_t seq, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                     const uint32_t *in_off, const uint32_t *in_tx, size_t nin, const uint32_t *out_off,
                                     const uint32_t *out_tx, const uint32_t *out_n, size_t nout, uint8_t *txids, uint8_t *root32,
                                     int *mutated, uint8_t *spk_cat, size_t spk_cap, uint32_t *spk_off, int64_t *amounts);
BCP_API int bcp_gpu_utxo_wait(bcp_utxo *u, uint32_t seq, uint32_t status8[8]);
/* The script checks of applied block seq (bcp_gpu_block_scripts_p2pkh's work and result) from the arrays apply_block left on
 * the table's GPU — the block, its offsets, the spent amounts and scriptPubKeys — on `device` (may carry a lane): read in
 * place on the table's own GPU, brought over by one peer-to-peer copy otherwise.  The host only supplies in_n (each input's
 * index in its transaction) and, to size the kernels' staging, its copies of tx_off and in_tx.  Waits for the block's table
 * work first and returns its status words like bcp_gpu_utxo_wait; with missing inputs no script is run. */
BCP_API int bcp_gpu_utxo_block_scripts(bcp_utxo *u, uint32_t seq, int device, const uint32_t *tx_off, const uint32_t *in_tx,
                                       const uint32_t *in_n, uint32_t flags, uint8_t *script_err, int *all_ok, uint32_t status8[8]);
/* Copies the spent-coin arrays of applied block seq to the host after the fact (synchronous): for a caller that applied
 * the block with NULL result arrays and then finds inputs it must evaluate on the host. */
BCP_API int bcp_gpu_utxo_fetch_spent(bcp_utxo *u, uint32_t seq, uint8_t *spk_cat, size_t spk_cap, uint32_t *spk_off, int64_t *amounts);
BCP_API int bcp_gpu_utxo_confirm(bcp_utxo *u, uint32_t seq);
BCP_API int bcp_gpu_utxo_rollback(bcp_utxo *u, uint32_t seq);

#ifdef __cplusplus
}
#endif
#endif
