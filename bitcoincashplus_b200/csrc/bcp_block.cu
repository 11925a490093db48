// bcp_block.cu — one call per block: txids + Merkle root + per-input P2PKH script checks with batched ECDSA.
// This is row N1 of SURVEY.md §8(f) ("GPU sighash + P2PKH template matcher") on top of the §8(a) kernels: the
// per-input work of CScriptCheck::operator() (src/validation.cpp:1350-1359) for the dominant script shape is
// moved to the device, so a block is validated from its raw bytes plus the spent outputs (scriptPubKey, amount)
// that ConnectBlock already fetched from the UTXO view (src/validation.cpp:1460-1471).
//
//   host     bcp_block_index_inputs: framing scan (tx offsets, per-input offsets; sequential byte parsing)
//   k_sha256d_msgs + k_merkle_levels   txids, Merkle root, mutated flag              (bcp_hash.cu)
//   k_tx_midstate    one thread per transaction: hashPrevouts / hashSequence / hashOutputs
//   k_p2pkh_extract  one thread per input: template match, HASH160, DER / low-S / hashtype / pubkey rules,
//                    FORKID sighash -> (sighash, pubkey, r||s) triple + status
//   k_ecdsa_prep + k_ecdsa_point       the batched verification                      (bcp_ecdsa.cu)
//   k_script_finish  status + verdict -> ScriptError per input
#include "bcp_ctx.cuh"
#include "script.cuh"
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

// Both per-transaction kernels parse transactions byte by byte.  Done straight from global memory that is one LDG.U8
// per byte and the load/store queue is the limit (ncu: lg_throttle 19 and long_scoreboard 16 stalls per issue).  The
// transactions a 128-thread block works on are CONTIGUOUS in the block buffer, so the block first copies that span into
// shared memory with coalesced 16-byte loads and parses from there; a span that does not fit (a block holding a huge
// transaction) keeps the global path — the choice is uniform for the thread block.
#define BCP_SPAN_MAX 98304u /* most shared memory a span may take (opt-in dynamic shared memory) */

// copies blk[lo16, hi) to shared memory (lo16 = lo rounded down to 16); the 16-byte loads may run up to 15 bytes past
// `hi`, which stays inside the staging allocation (the block region is followed by the offset arrays)
__device__ __forceinline__ bool stage_span(uint4 *span4, const uint8_t *blk, uint32_t lo, uint32_t hi, uint32_t cap,
                                           uint32_t &lo16) {
    lo16 = lo & ~15u;
    if (hi - lo16 > cap) return false;
    const uint4 *src = reinterpret_cast<const uint4 *>(blk + lo16);
    const uint32_t n16 = (hi - lo16 + 15u) >> 4;
    for (uint32_t k = threadIdx.x; k < n16; k += blockDim.x) span4[k] = __ldg(src + k);
    __syncthreads();
    return true;
}

// dynamic shared memory: [32 words x 128 threads of streaming-hash scratch][span of `cap` bytes]
__global__ void __launch_bounds__(128) k_tx_midstate(const uint8_t *__restrict__ blk, const uint32_t *__restrict__ tx_off,
                                                     size_t ntx, uint8_t *__restrict__ mid, uint32_t cap) {
    extern __shared__ uint4 smem4[];
    uint32_t *wbuf = reinterpret_cast<uint32_t *>(smem4);
    uint4 *span4 = smem4 + (32 * 128 * 4) / 16;
    const size_t t0 = (size_t)blockIdx.x * blockDim.x;
    const size_t t1 = t0 + blockDim.x < ntx ? t0 + blockDim.x : ntx;
    uint32_t lo16;
    const bool staged = stage_span(span4, blk, tx_off[t0], tx_off[t1], cap, lo16);
    const size_t t = t0 + threadIdx.x;
    if (t >= ntx) return;
    const uint32_t o = tx_off[t];
    const uint8_t *tx = staged ? reinterpret_cast<const uint8_t *>(span4) + (o - lo16) : blk + o;
    tx_midstate(mid + t * BCP_TXMID_BYTES, tx, tx_off[t + 1] - o, wbuf + threadIdx.x, 128);
}

__global__ void __launch_bounds__(128)
    k_p2pkh_extract(const uint8_t *__restrict__ blk, const uint32_t *__restrict__ tx_off, const uint32_t *__restrict__ in_off,
                    const uint32_t *__restrict__ in_tx, const uint32_t *__restrict__ in_n, size_t nin,
                    const uint8_t *__restrict__ spk, const uint32_t *__restrict__ spk_off, const int64_t *__restrict__ amounts,
                    const uint8_t *__restrict__ mid, uint8_t *__restrict__ status, uint8_t *__restrict__ msg32,
                    uint8_t *__restrict__ pub65, uint8_t *__restrict__ sig64, uint32_t cap) {
    extern __shared__ uint4 smem4[];
    const size_t i0 = (size_t)blockIdx.x * blockDim.x;
    const size_t i1 = i0 + blockDim.x < nin ? i0 + blockDim.x : nin;
    // inputs are listed in block order, so this thread block's inputs live in transactions in_tx[i0] .. in_tx[i1-1]
    uint32_t lo16;
    const bool staged = stage_span(smem4, blk, tx_off[in_tx[i0]], tx_off[in_tx[i1 - 1] + 1], cap, lo16);
    const size_t i = i0 + threadIdx.x;
    if (i >= nin) return;
    const uint32_t t = in_tx[i];
    const uint32_t o = tx_off[t];
    const uint8_t *tx = staged ? reinterpret_cast<const uint8_t *>(smem4) + (o - lo16) : blk + o;
    status[i] = (uint8_t)p2pkh_extract_one(tx, tx_off[t + 1] - o, in_off[i] - o, in_n[i], spk + spk_off[i],
                                           spk_off[i + 1] - spk_off[i], amounts[i], mid + (size_t)t * BCP_TXMID_BYTES,
                                           msg32 + 32 * i, pub65 + 65 * i, sig64 + 64 * i);
}

// Host side of the span staging: the largest span any thread block of the launch needs (plus the 15 bytes the start may
// be rounded down by), capped; the launch asks for exactly that much dynamic shared memory, so ordinary blocks (128
// transactions of ~200 bytes = 25 KB) keep a high occupancy and only unusual ones pay for more.
static uint32_t span_cap_tx(const uint32_t *tx_off, size_t ntx) {
    uint32_t need = 0;
    for (size_t t0 = 0; t0 < ntx; t0 += 128) {
        const size_t t1 = t0 + 128 < ntx ? t0 + 128 : ntx;
        need = std::max(need, tx_off[t1] - tx_off[t0]);
    }
    return std::min(BCP_SPAN_MAX, (need + 15u + 15u) & ~15u);
}
static uint32_t span_cap_inputs(const uint32_t *tx_off, const uint32_t *in_tx, size_t nin) {
    uint32_t need = 0;
    for (size_t i0 = 0; i0 < nin; i0 += 128) {
        const size_t i1 = i0 + 128 < nin ? i0 + 128 : nin;
        need = std::max(need, tx_off[in_tx[i1 - 1] + 1] - tx_off[in_tx[i0]]);
    }
    return std::min(BCP_SPAN_MAX, (need + 15u + 15u) & ~15u);
}
#define BCP_MID_WBUF_BYTES (32u * 128u * 4u)
// sighash midstates per transaction, then template match + sighash per input (host arrays only size the spans)
static int launch_mid_extract(bcp_ctx *c, const uint8_t *d_blk, const uint32_t *d_txoff, const uint32_t *h_txoff, size_t ntx,
                              const uint32_t *d_inoff, const uint32_t *d_intx, const uint32_t *d_inn, const uint32_t *h_intx,
                              size_t nin, const uint8_t *d_spk, const uint32_t *d_spkoff, const int64_t *d_amt, uint8_t *d_mid,
                              uint8_t *d_stat, uint8_t *d_msg, uint8_t *d_pub, uint8_t *d_sig, cudaStream_t st) {
    if (!c->block_attrs) { /* per device: allow the kernels more than the default 48 KB of dynamic shared memory */
        BCP_CUDA(cudaFuncSetAttribute(k_tx_midstate, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(BCP_SPAN_MAX + BCP_MID_WBUF_BYTES + 16)));
        BCP_CUDA(cudaFuncSetAttribute(k_p2pkh_extract, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(BCP_SPAN_MAX + 16)));
        c->block_attrs = true;
    }
    const uint32_t cap_t = span_cap_tx(h_txoff, ntx), cap_i = span_cap_inputs(h_txoff, h_intx, nin);
    k_tx_midstate<<<(unsigned)((ntx + 127) / 128), 128, BCP_MID_WBUF_BYTES + cap_t + 16, st>>>(d_blk, d_txoff, ntx, d_mid, cap_t);
    k_p2pkh_extract<<<(unsigned)((nin + 127) / 128), 128, cap_i + 16, st>>>(d_blk, d_txoff, d_inoff, d_intx, d_inn, nin, d_spk,
                                                                           d_spkoff, d_amt, d_mid, d_stat, d_msg, d_pub, d_sig, cap_i);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}

__global__ void k_script_finish(const uint8_t *__restrict__ status, const uint8_t *__restrict__ ok, size_t nin,
                                uint8_t *__restrict__ err, int *__restrict__ all_ok) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nin) return;
    uint8_t s = status[i];
    uint8_t e = s == BCP_SCRIPT_PENDING ? (ok[i] ? BCP_SCRIPT_OK : BCP_SCRIPT_ERR_SIG_NULLFAIL) : s;
    err[i] = e;
    if (e != BCP_SCRIPT_OK) atomicAnd(all_ok, 0);
}

static int rd_compact_h(const uint8_t *p, size_t len, size_t *pos, uint64_t *v) {
    if (*pos >= len) return 0;
    uint8_t ch = p[(*pos)++];
    if (ch < 253) { *v = ch; return 1; }
    int nb = ch == 253 ? 2 : ch == 254 ? 4 : 8;
    if (*pos + nb > len) return 0;
    uint64_t x = 0;
    for (int i = 0; i < nb; i++) x |= (uint64_t)p[*pos + i] << (8 * i);
    *pos += nb;
    *v = x;
    return 1;
}

#define BCP_FLAGS_POSTFORK_REQUIRED ((1u << 1) | (1u << 2) | (1u << 3) | (1u << 14) | (1u << 16))
#define BCP_FLAG_ALLOW_NON_FORKID (1u << 17)
// SCRIPT_VERIFY_* bits (src/script/interpreter.h:32-121) that cannot change the verdict or the ScriptError of a
// `<sig> <pubkey>` scriptSig (direct pushes only) against a P2PKH scriptPubKey: P2SH, NULLDUMMY, SIGPUSHONLY,
// DISCOURAGE_UPGRADABLE_NOPS, CLEANSTACK (one element is left), CLTV, CSV, MINIMALIF — next to the required ones.
// MINIMALDATA (1<<6: a one-byte direct push of 1..16 / 0x81 would be an error) and COMPRESSED_PUBKEYTYPE (1<<15:
// NONCOMPRESSED_PUBKEY for 65-byte keys) do change it and are not implemented here, so they are refused, like any
// bit this file does not know.
#define BCP_FLAGS_FASTPATH_NEUTRAL ((1u << 0) | (1u << 4) | (1u << 5) | (1u << 7) | (1u << 8) | (1u << 9) | (1u << 10) | (1u << 13))
static int check_fastpath_flags(uint32_t flags, const char *fn) {
    if ((flags & BCP_FLAGS_POSTFORK_REQUIRED) != BCP_FLAGS_POSTFORK_REQUIRED || (flags & BCP_FLAG_ALLOW_NON_FORKID)) {
        bcp_set_error("%s: the fast path implements the post-fork script flags (STRICTENC|DERSIG|LOW_S|NULLFAIL|"
                      "SIGHASH_FORKID, no ALLOW_NON_FORKID); got 0x%x", fn, flags);
        return BCP_ERR_INVALID;
    }
    const uint32_t unknown = flags & ~(BCP_FLAGS_POSTFORK_REQUIRED | BCP_FLAGS_FASTPATH_NEUTRAL);
    if (unknown) {
        bcp_set_error("%s: script flag bits 0x%x would change the result for this script shape and are not implemented by "
                      "the fast path (MINIMALDATA, COMPRESSED_PUBKEYTYPE, unknown bits)", fn, unknown);
        return BCP_ERR_INVALID;
    }
    if ((flags & (1u << 8)) && !(flags & 1u)) { /* VerifyScript asserts P2SH whenever CLEANSTACK is set (interpreter.cpp:1619) */
        bcp_set_error("%s: CLEANSTACK without P2SH is not a valid flag combination", fn);
        return BCP_ERR_INVALID;
    }
    return BCP_OK;
}

// compact size with the canonical-encoding and MAX_SIZE rules of ReadCompactSize (src/serialize.h:260-300)
static int rd_compact_canon(const uint8_t *p, size_t len, size_t *pos, uint64_t *v) {
    if (*pos >= len) return 0;
    const uint8_t ch = p[(*pos)++];
    if (ch < 253) { *v = ch; return 1; }
    const int nb = ch == 253 ? 2 : ch == 254 ? 4 : 8;
    if (*pos + nb > len) return 0;
    uint64_t x = 0;
    for (int i = 0; i < nb; i++) x |= (uint64_t)p[*pos + i] << (8 * i);
    *pos += nb;
    if ((nb == 2 && x < 253) || (nb == 4 && x < 0x10000ull) || (nb == 8 && x < 0x100000000ull)) return 0; /* non-canonical */
    if (x > 0x02000000ull) return 0;                                                                        /* MAX_SIZE */
    *v = x;
    return 1;
}

// The script checks of one block with every INPUT array already on c's device (the resident-block path of the coin table,
// bcp_utxo.cu): sighash midstates, template match + sighash, batched ECDSA, verdicts; only the ScriptError bytes travel
// back.  h_tx_off / h_in_tx are the host copies of two of the arrays (they size the shared-memory spans).  The caller
// holds c->mu and has made c's device current; work is queued on st and waited for here.
int bcp_block_scripts_resident(bcp_ctx *c, const uint8_t *d_blk, const uint32_t *d_txoff, const uint32_t *h_tx_off, size_t ntx,
                               const uint32_t *d_inoff, const uint32_t *d_intx, const uint32_t *d_inn, const uint32_t *h_in_tx,
                               size_t nin, const uint8_t *d_spk, const uint32_t *d_spkoff, const int64_t *d_amt, size_t stage_used,
                               uint8_t *script_err, int *all_ok, cudaStream_t st) {
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    size_t o = up(stage_used);
    const size_t o_flag = o; o += 256;
    const size_t o_mid = o; o += up(ntx * BCP_TXMID_BYTES + 16);
    const size_t o_stat = o; o += up(nin + 8);
    const size_t o_msg = o; o += up(nin * 32 + 32);
    const size_t o_pub = o; o += up(nin * 65 + 65);
    const size_t o_sig = o; o += up(nin * 64 + 64);
    const size_t o_ok = o; o += up(nin + 8);
    const size_t o_err = o; o += up(nin + 8);
    if (o > c->stage_bytes) { /* the caller reserved stage_used + bcp_block_scripts_resident_bytes(ntx, nin) */
        bcp_set_error("bcp_block_scripts_resident: staging area too small");
        return BCP_ERR_INVALID;
    }
    int rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(nin));
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    int one = 1;
    BCP_CUDA(cudaMemcpyAsync(d + o_flag, &one, sizeof(int), cudaMemcpyHostToDevice, st));
    rc = launch_mid_extract(c, d_blk, d_txoff, h_tx_off, ntx, d_inoff, d_intx, d_inn, h_in_tx, nin, d_spk, d_spkoff, d_amt, d + o_mid,
                            d + o_stat, d + o_msg, d + o_pub, d + o_sig, st);
    if (rc != BCP_OK) return rc;
    c->prep_throughput = true; /* this path is the chain connector's: other lanes keep the GPU busy meanwhile */
    rc = bcp_ecdsa_launch(c, d + o_msg, d + o_pub, d + o_sig, nin, BCP_ECDSA_FLAG_NORMALIZE_S, d + o_ok,
                          reinterpret_cast<uint32_t *>(c->scratch), st);
    c->prep_throughput = false;
    if (rc != BCP_OK) return rc;
    k_script_finish<<<(unsigned)((nin + 255) / 256), 256, 0, st>>>(d + o_stat, d + o_ok, nin, d + o_err, reinterpret_cast<int *>(d + o_flag));
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(script_err, d + o_err, nin, cudaMemcpyDeviceToHost, st));
    int a = 0;
    BCP_CUDA(cudaMemcpyAsync(&a, d + o_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    // this path is driven by one host thread per (GPU, lane) next to scan threads and a connecting thread — more threads
    // than cores on a small host — so the wait sleeps on a blocking-sync event instead of spinning in cudaStreamSynchronize
    BCP_CUDA(cudaEventRecord(c->ev[3], st));
    BCP_CUDA(cudaEventSynchronize(c->ev[3]));
    if (all_ok) *all_ok = a ? 1 : 0;
    return BCP_OK;
}
size_t bcp_block_scripts_resident_bytes(size_t ntx, size_t nin) {
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    return 512 + up(ntx * BCP_TXMID_BYTES + 16) + up(nin + 8) * 3 + up(nin * 32 + 32) + up(nin * 65 + 65) + up(nin * 64 + 64);
}
int bcp_block_check_flags(uint32_t flags, const char *fn) { return check_fastpath_flags(flags, fn); }

extern "C" {

// Framing scan that also lists every non-coinbase input (CTransaction::IsCoinBase: a single input with a null
// prevout, src/primitives/transaction.h): in_off = byte offset of the input's outpoint, in_tx = its transaction,
// in_n = its ordinal in that transaction.  Returns the number of inputs (-1 malformed, -2 capacity); *ntx_out = ntx.
BCP_API long bcp_block_index_inputs(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, size_t max_tx,
                                    uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs, long *ntx_out) {
    return bcp_block_index_full(blk, len, legacy_header, tx_off, nullptr, max_tx, in_off, in_tx, in_n, max_inputs, ntx_out,
                                nullptr, nullptr, nullptr, 0, nullptr, nullptr);
}

// The same scan, additionally reporting where each transaction's output list starts (the offset of its output count)
// and, flat like the inputs, every output: out_off = offset of its 8-byte amount, out_tx = its transaction, out_n = its
// index.  A block connector inserts the new coins from these without parsing anything again.  in_key[i] = a 64-bit mix
// of input i's outpoint (first 8 txid bytes, little-endian, XOR 0x9E3779B97F4A7C15 * vout) for hash-table lookups.  The
// extra arrays may be NULL (max_outputs is then ignored); more than max_outputs outputs is the capacity error -2.
BCP_API long bcp_block_index_full(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, uint32_t *tx_out_off,
                                  size_t max_tx, uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs,
                                  long *ntx_out, uint32_t *out_off, uint32_t *out_tx, uint32_t *out_n, size_t max_outputs,
                                  long *nout_out, uint64_t *in_key) {
    return bcp_block_index_rules(blk, len, legacy_header, tx_off, tx_out_off, max_tx, in_off, in_tx, in_n, max_inputs, ntx_out, out_off,
                                 out_tx, out_n, max_outputs, nout_out, in_key, nullptr, nullptr);
}

BCP_API const char *bcp_block_rule_name(int rule) {
    static const char *const names[] = {"", "bad-cb-missing", "bad-tx-coinbase", "bad-txns-vin-empty", "bad-txns-vout-empty",
                                        "bad-txns-oversize", "bad-txns-vout-negative", "bad-txns-vout-toolarge",
                                        "bad-txns-txouttotal-toolarge", "bad-cb-length", "bad-txns-prevout-null"};
    return rule >= 0 && rule <= 10 ? names[rule] : "";
}

// The same scan that, while it has each transaction's bytes in the cache anyway, also applies CheckBlock's context-free
// transaction rules (src/validation.cpp:3255-3330 calling CheckCoinbase :531-549 and CheckRegularTransaction :551-570, both
// over CheckTransactionCommon :476-529) in the reference's order: transactions in block order, and inside a transaction
// coinbase position, empty input / output list, size, output amounts, coinbase script length, null prevouts.
// *rule = 0 or the BCP_RULE_* of the FIRST violation (bcp_block_rule_name gives the reference's reject string),
// *rule_tx = that transaction.  The arrays are complete either way.
BCP_API long bcp_block_index_rules(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, uint32_t *tx_out_off,
                                   size_t max_tx, uint32_t *in_off, uint32_t *in_tx, uint32_t *in_n, size_t max_inputs,
                                   long *ntx_out, uint32_t *out_off, uint32_t *out_tx, uint32_t *out_n, size_t max_outputs,
                                   long *nout_out, uint64_t *in_key, int *rule, long *rule_tx) {
    size_t ocnt = 0, touched = 0;
    volatile unsigned touch_sink = 0; /* keeps the cache-warming loads alive */
    int first_rule = 0;
    long first_rule_tx = -1;
    static const int64_t MAX_MONEY_SAT = 21000000ll * 100000000ll; /* src/amount.h:127 */
    // One pass (this scan sits on the critical path of a block: nothing can launch before the offsets exist).
    // Same acceptance rules as bcp_block_index: canonical compact sizes, MAX_SIZE, everything inside the buffer.
    size_t pos;
    uint64_t v, ntx;
    if (legacy_header) {
        pos = 80;
    } else {
        pos = 140;
        if (len < 140 || !rd_compact_canon(blk, len, &pos, &v)) return -1;
        pos += v;
    }
    if (pos > len || !rd_compact_canon(blk, len, &pos, &ntx)) return -1;
    if (ntx > max_tx) return -2;
    size_t cnt = 0;
#define BCP_RD_COMPACT(dst)                                                                  \
    do {                                                                                     \
        if (pos < len && blk[pos] < 253) { dst = blk[pos++]; }                               \
        else if (!rd_compact_canon(blk, len, &pos, &dst)) return -1;                         \
    } while (0)
    for (uint64_t t = 0; t < ntx; t++) {
        tx_off[t] = (uint32_t)pos;
        const size_t tx_start = pos;
        uint64_t nin, nout, sl;
        // Every step of this walk depends on a byte the previous step loaded (a length), so on a block that is not in the
        // cache it runs at memory LATENCY — 0.6 ms per 8 MB block on a quiet host, 4.6 ms and 19 ms measured on busier ones
        // (other GPUs' uploads, remote memory), where prefetch hints a fixed distance ahead did not help.  The walk only
        // moves forward, so the block is pulled into the cache in 128 KB pieces by a loop of INDEPENDENT loads (one per
        // line: the core keeps a dozen misses in flight and the hardware prefetcher follows the stream) just before the
        // walk enters each piece; the dependent loads then hit the L2.
        if (pos >= touched) {
            const size_t end = touched + (128u << 10) < len ? touched + (128u << 10) : len;
            unsigned acc = 0;
            for (size_t q = touched; q < end; q += 64) acc += blk[q];
            touch_sink += acc;
            touched = end;
        }
        if (pos + 4 > len) return -1;
        pos += 4;
        BCP_RD_COMPACT(nin);
        bool coinbase = false;
        if (nin == 1 && pos + 36 <= len) {
            uint64_t first8;
            memcpy(&first8, blk + pos, 8);
            if (first8 == 0) { /* a real txid practically never starts with eight zero bytes: look closer only then */
                coinbase = true;
                for (int b = 8; b < 32; b++) coinbase = coinbase && blk[pos + b] == 0;
                for (int b = 32; b < 36; b++) coinbase = coinbase && blk[pos + b] == 0xFF;
            }
        }
        int tx_rule = 0;
        if (t == 0 && !coinbase) tx_rule = 1;
        else if (t > 0 && coinbase) tx_rule = 2;
        else if (nin == 0) tx_rule = 3;
        bool null_prevout = false;
        uint64_t cb_script_len = 0;
        for (uint64_t i = 0; i < nin; i++) {
            if (pos + 36 > len) return -1;
            if (!coinbase) {
                uint64_t h8;
                uint32_t vout;
                memcpy(&h8, blk + pos, 8);
                memcpy(&vout, blk + pos + 32, 4);
                if (h8 == 0 && vout == 0xFFFFFFFFu) { /* candidate null prevout (COutPoint::IsNull): compare the rest */
                    bool z = true;
                    for (int b = 8; b < 32; b++) z = z && blk[pos + b] == 0;
                    null_prevout = null_prevout || z;
                }
                if (in_off) {
                    if (cnt >= max_inputs) return -2;
                    in_off[cnt] = (uint32_t)pos; in_tx[cnt] = (uint32_t)t; in_n[cnt] = (uint32_t)i;
                    if (in_key) in_key[cnt] = h8 ^ (0x9E3779B97F4A7C15ull * vout); /* the 64-bit mix a hash table would use */
                    cnt++;
                }
            }
            pos += 36;
            BCP_RD_COMPACT(sl);
            if (coinbase) cb_script_len = sl;
            if (pos + sl + 4 > len) return -1;
            pos += sl + 4;
        }
        if (tx_out_off) tx_out_off[t] = (uint32_t)pos;
        BCP_RD_COMPACT(nout);
        if (!tx_rule && nout == 0) tx_rule = 4;
        int amount_rule = 0;
        int64_t total = 0;
        for (uint64_t i = 0; i < nout; i++) {
            if (pos + 8 > len) return -1;
            if (out_off) {
                if (ocnt >= max_outputs) return -2;
                out_off[ocnt] = (uint32_t)pos; out_tx[ocnt] = (uint32_t)t; out_n[ocnt] = (uint32_t)i;
                ocnt++;
            }
            if (!amount_rule) {
                int64_t val;
                memcpy(&val, blk + pos, 8);
                if (val < 0) amount_rule = 6;
                else if (val > MAX_MONEY_SAT) amount_rule = 7;
                else if ((total += val) > MAX_MONEY_SAT) amount_rule = 8;
            }
            pos += 8;
            BCP_RD_COMPACT(sl);
            if (pos + sl > len) return -1;
            pos += sl;
        }
        if (pos + 4 > len) return -1;
        pos += 4;
        if (!tx_rule && pos - tx_start > 1000000u) tx_rule = 5; /* MAX_TX_SIZE, src/consensus/consensus.h:14 */
        if (!tx_rule) tx_rule = amount_rule;
        if (!tx_rule && t == 0 && (cb_script_len < 2 || cb_script_len > 100)) tx_rule = 9;
        if (!tx_rule && null_prevout) tx_rule = 10;
        if (tx_rule && !first_rule) { first_rule = tx_rule; first_rule_tx = (long)t; }
    }
#undef BCP_RD_COMPACT
    tx_off[ntx] = (uint32_t)pos;
    if (ntx_out) *ntx_out = (long)ntx;
    if (nout_out) *nout_out = (long)ocnt;
    if (rule) *rule = ntx == 0 ? 1 : first_rule; /* no transaction at all: "first tx is not coinbase" (:3258-3261) */
    if (rule_tx) *rule_tx = first_rule_tx;
    return (long)cnt;
}

// Script checks only, for an arbitrary subset of a block's inputs (the caller did the framing scan and picked the
// inputs): this is the unit that shards across devices — each GPU gets the raw block plus its own index range of
// inputs and returns its ScriptError bytes; the host ANDs them (src/checkqueue.h:77), nothing is exchanged
// between devices.  in_off / in_tx / in_n / spk_off / amounts / script_err all describe the SUBSET (n_inputs
// entries, spk_off rebased to the subset's first scriptPubKey).
BCP_API int bcp_gpu_block_scripts_p2pkh(int device, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                        const uint32_t *in_off, const uint32_t *in_tx, const uint32_t *in_n, size_t n_inputs,
                                        const uint8_t *spk, const uint32_t *spk_off, const int64_t *amounts, uint32_t flags,
                                        uint8_t *script_err, int *all_ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (all_ok) *all_ok = 1;
    if (n_inputs == 0) return BCP_OK;
    if (!blk || !tx_off || !in_off || !in_tx || !in_n || !spk || !spk_off || !amounts || !script_err) {
        bcp_set_error("bcp_gpu_block_scripts_p2pkh: NULL buffer");
        return BCP_ERR_INVALID;
    }
    if (check_fastpath_flags(flags, "bcp_gpu_block_scripts_p2pkh") != BCP_OK) return BCP_ERR_INVALID;
    if (len > 0xFFFFFFF0ull || tx_off[ntx] > len) { bcp_set_error("bcp_gpu_block_scripts_p2pkh: offsets exceed the block"); return BCP_ERR_INVALID; }
    // The kernels stage the bytes between the first and the last transaction of each 128-input group in shared memory,
    // so the subset must be in block order and every input must lie inside the transaction it names.
    for (size_t i = 0; i < n_inputs; i++) {
        const uint32_t t = in_tx[i];
        if (t >= ntx || (i && t < in_tx[i - 1]) || in_off[i] < tx_off[t] + 5 || (uint64_t)in_off[i] + 41 > tx_off[t + 1]) {
            bcp_set_error("bcp_gpu_block_scripts_p2pkh: input %zu is out of block order or outside transaction %u", i, t);
            return BCP_ERR_INVALID;
        }
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t nin = n_inputs, spk_bytes = spk_off[nin] - spk_off[0];
    size_t o = 0;
    const size_t o_blk = o; o += up(len + 8);
    const size_t o_txoff = o; o += up((ntx + 1) * 4);
    const size_t o_inoff = o; o += up(nin * 4 + 4);
    const size_t o_intx = o; o += up(nin * 4 + 4);
    const size_t o_inn = o; o += up(nin * 4 + 4);
    const size_t o_spk = o; o += up(spk_bytes + 8);
    const size_t o_spkoff = o; o += up((nin + 1) * 4);
    const size_t o_amt = o; o += up(nin * 8 + 8);
    const size_t o_flag = o; o += 256;
    const size_t o_mid = o; o += up(ntx * BCP_TXMID_BYTES + 16);
    const size_t o_stat = o; o += up(nin + 8);
    const size_t o_msg = o; o += up(nin * 32 + 32);
    const size_t o_pub = o; o += up(nin * 65 + 65);
    const size_t o_sig = o; o += up(nin * 64 + 64);
    const size_t o_ok = o; o += up(nin + 8);
    const size_t o_err = o; o += up(nin + 8);
    int rc = bcp_ctx_reserve_stage(c, o);
    if (rc != BCP_OK) return rc;
    rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(nin));
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream;
    // spk_off may start anywhere in the caller's array: rebase on the host (tiny)
    std::vector<uint32_t> &reb = c->h_in_n;
    if (reb.size() < nin + 1) reb.resize(nin + 1);
    for (size_t i = 0; i <= nin; i++) reb[i] = spk_off[i] - spk_off[0];
    int one = 1;
    BCP_CUDA(cudaMemcpyAsync(d + o_blk, blk, len, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_txoff, tx_off, (ntx + 1) * 4, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_inoff, in_off, nin * 4, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_intx, in_tx, nin * 4, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_inn, in_n, nin * 4, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_spk, spk + spk_off[0], spk_bytes, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_spkoff, reb.data(), (nin + 1) * 4, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_amt, amounts, nin * 8, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_flag, &one, sizeof(int), cudaMemcpyHostToDevice, st));
    uint32_t *d_txoff = reinterpret_cast<uint32_t *>(d + o_txoff);
    rc = launch_mid_extract(c, d + o_blk, d_txoff, tx_off, ntx, reinterpret_cast<uint32_t *>(d + o_inoff),
                            reinterpret_cast<uint32_t *>(d + o_intx), reinterpret_cast<uint32_t *>(d + o_inn), in_tx, nin,
                            d + o_spk, reinterpret_cast<uint32_t *>(d + o_spkoff), reinterpret_cast<int64_t *>(d + o_amt),
                            d + o_mid, d + o_stat, d + o_msg, d + o_pub, d + o_sig, st);
    if (rc != BCP_OK) return rc;
    rc = bcp_ecdsa_launch(c, d + o_msg, d + o_pub, d + o_sig, nin, BCP_ECDSA_FLAG_NORMALIZE_S, d + o_ok,
                          reinterpret_cast<uint32_t *>(c->scratch), st);
    if (rc != BCP_OK) return rc;
    k_script_finish<<<(unsigned)((nin + 255) / 256), 256, 0, st>>>(d + o_stat, d + o_ok, nin, d + o_err,
                                                                   reinterpret_cast<int *>(d + o_flag));
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(script_err, d + o_err, nin, cudaMemcpyDeviceToHost, st));
    int a = 0;
    BCP_CUDA(cudaMemcpyAsync(&a, d + o_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaStreamSynchronize(st));
    if (all_ok) *all_ok = a ? 1 : 0;
    return BCP_OK;
}

BCP_API int bcp_gpu_block_connect_p2pkh(int device, const uint8_t *blk, size_t len, int legacy_header, const uint8_t *spk,
                                        const uint32_t *spk_off, const int64_t *amounts, size_t n_inputs, uint32_t flags,
                                        uint8_t *txids, uint8_t root32[32], int *mutated, uint8_t *script_err, int *all_ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (!blk || !root32 || (n_inputs && (!spk || !spk_off || !amounts || !script_err))) {
        bcp_set_error("bcp_gpu_block_connect_p2pkh: NULL buffer");
        return BCP_ERR_INVALID;
    }
    if (check_fastpath_flags(flags, "bcp_gpu_block_connect_p2pkh") != BCP_OK) return BCP_ERR_INVALID;
    if (len > 0xFFFFFFF0ull) { bcp_set_error("block too large"); return BCP_ERR_INVALID; }
    // Device layout is sized from upper bounds (a transaction is at least 60 bytes, an input at least 41) so that
    // the raw block can start its way to the device BEFORE the host framing scan runs; with a pinned `blk` the copy
    // and the scan overlap.
    const size_t max_tx = len / 60 + 2, max_in = len / 41 + 2;
    if (n_inputs > max_in) { bcp_set_error("bcp_gpu_block_connect_p2pkh: more inputs than the block can hold"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t spk_bytes = n_inputs ? spk_off[n_inputs] : 0;
    const size_t cap_in = n_inputs + 1;
    size_t o = 0;
    const size_t o_blk = o; o += up(len + 8);
    const size_t o_txoff = o; o += up((max_tx + 1) * 4);
    const size_t o_inoff = o; o += up(cap_in * 4 + 4);
    const size_t o_intx = o; o += up(cap_in * 4 + 4);
    const size_t o_inn = o; o += up(cap_in * 4 + 4);
    const size_t o_spk = o; o += up(spk_bytes + 8);
    const size_t o_spkoff = o; o += up((cap_in + 1) * 4);
    const size_t o_amt = o; o += up(cap_in * 8 + 8);
    const size_t o_txid = o; o += up(max_tx * 32 + 32);
    const size_t o_mtmp = o; o += up(bcp_merkle_tmp_words(max_tx) * 4);
    const size_t o_root = o; o += 256; /* root[32], mutated(int), all_ok(int) */
    const size_t o_mid = o; o += up(max_tx * BCP_TXMID_BYTES + 16);
    const size_t o_stat = o; o += up(cap_in + 8);
    const size_t o_msg = o; o += up(cap_in * 32 + 32);
    const size_t o_pub = o; o += up(cap_in * 65 + 65);
    const size_t o_sig = o; o += up(cap_in * 64 + 64);
    const size_t o_ok = o; o += up(cap_in + 8);
    const size_t o_err = o; o += up(cap_in + 8);
    int rc = bcp_ctx_reserve_stage(c, o);
    if (rc != BCP_OK) return rc;
    rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(n_inputs));
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream;
    BCP_CUDA(cudaMemcpyAsync(d + o_blk, blk, len, cudaMemcpyHostToDevice, st));
    // host: framing scan (overlaps the copy above when blk is pinned)
    std::vector<uint32_t> &tx_off = c->h_tx_off, &in_off = c->h_in_off, &in_tx = c->h_in_tx, &in_n = c->h_in_n;
    if (tx_off.size() < max_tx + 1) tx_off.resize(max_tx + 1);
    if (in_off.size() < cap_in) { in_off.resize(cap_in); in_tx.resize(cap_in); in_n.resize(cap_in); }
    long ntx = 0;
    long nin = bcp_block_index_inputs(blk, len, legacy_header, tx_off.data(), max_tx, in_off.data(), in_tx.data(), in_n.data(),
                                      cap_in, &ntx);
    if (nin < 0 || (size_t)nin != n_inputs) {
        cudaStreamSynchronize(st);
        if (nin == -1) bcp_set_error("bcp_gpu_block_connect_p2pkh: malformed block");
        else bcp_set_error("bcp_gpu_block_connect_p2pkh: the block's non-coinbase inputs (%ld) do not match the %zu spent "
                           "outputs the caller described", nin, n_inputs);
        return BCP_ERR_INVALID;
    }
    BCP_CUDA(cudaMemcpyAsync(d + o_txoff, tx_off.data(), (ntx + 1) * 4, cudaMemcpyHostToDevice, st));
    uint32_t *d_txoff = reinterpret_cast<uint32_t *>(d + o_txoff);
    // txids + Merkle on the second stream, beside the script checks (nothing below needs them)
    cudaStream_t st2 = c->copy_stream;
    BCP_CUDA(cudaEventRecord(c->ev[0], st));
    BCP_CUDA(cudaStreamWaitEvent(st2, c->ev[0], 0));
    rc = bcp_sha256d_offsets_dev(c, d + o_blk, d_txoff, ntx, d + o_txid, st2);
    if (rc != BCP_OK) return rc;
    int *d_flags = reinterpret_cast<int *>(d + o_root + 32);
    rc = bcp_merkle_dev(c, d + o_txid, ntx, d + o_root, d_flags, reinterpret_cast<uint32_t *>(d + o_mtmp), st2);
    if (rc != BCP_OK) return rc;
    if (txids) BCP_CUDA(cudaMemcpyAsync(txids, d + o_txid, (size_t)ntx * 32, cudaMemcpyDeviceToHost, st2));
    BCP_CUDA(cudaEventRecord(c->ev[1], st2));
    int one = 1;
    BCP_CUDA(cudaMemcpyAsync(d_flags + 1, &one, sizeof(int), cudaMemcpyHostToDevice, st));
    if (nin) {
        BCP_CUDA(cudaMemcpyAsync(d + o_inoff, in_off.data(), nin * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_intx, in_tx.data(), nin * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_inn, in_n.data(), nin * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_spk, spk, spk_bytes, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_spkoff, spk_off, (nin + 1) * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_amt, amounts, nin * 8, cudaMemcpyHostToDevice, st));
        rc = launch_mid_extract(c, d + o_blk, d_txoff, tx_off.data(), ntx, reinterpret_cast<uint32_t *>(d + o_inoff),
                                reinterpret_cast<uint32_t *>(d + o_intx), reinterpret_cast<uint32_t *>(d + o_inn), in_tx.data(), nin,
                                d + o_spk, reinterpret_cast<uint32_t *>(d + o_spkoff), reinterpret_cast<int64_t *>(d + o_amt),
                                d + o_mid, d + o_stat, d + o_msg, d + o_pub, d + o_sig, st);
        if (rc != BCP_OK) return rc;
        rc = bcp_ecdsa_launch(c, d + o_msg, d + o_pub, d + o_sig, nin, BCP_ECDSA_FLAG_NORMALIZE_S, d + o_ok,
                              reinterpret_cast<uint32_t *>(c->scratch), st);
        if (rc != BCP_OK) return rc;
        k_script_finish<<<(unsigned)((nin + 255) / 256), 256, 0, st>>>(d + o_stat, d + o_ok, nin, d + o_err, d_flags + 1);
        BCP_CUDA(cudaGetLastError());
        BCP_CUDA(cudaMemcpyAsync(script_err, d + o_err, nin, cudaMemcpyDeviceToHost, st));
    }
    BCP_CUDA(cudaStreamWaitEvent(st, c->ev[1], 0));
    uint8_t res[40];
    BCP_CUDA(cudaMemcpyAsync(res, d + o_root, 40, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaStreamSynchronize(st));
    memcpy(root32, res, 32);
    int m, a;
    memcpy(&m, res + 32, 4);
    memcpy(&a, res + 36, 4);
    if (mutated) *mutated = m ? 1 : 0;
    if (all_ok) *all_ok = a ? 1 : 0;
    return BCP_OK;
}

// A window of blocks in one pass (reindex / initial block download, where the next blocks are already at hand).
// Only the transaction bytes of each block go to the device, laid back to back with no gap, so that ONE offsets
// array describes every transaction of the window and the per-transaction and per-input kernels (txids, sighash
// midstates, template match + sighash, ECDSA prep + point, verdict) each run once over the union; only the Merkle
// reduction is per block, on the second stream, next to the ECDSA kernels.  A single 8 MB block fills barely half
// of one wave of the point kernel; a window fills several.
static long first_tx_offset(const uint8_t *blk, size_t len, int legacy_header) {
    size_t pos;
    uint64_t v;
    if (legacy_header) {
        pos = 80;
    } else {
        pos = 140;
        if (len < 140 || !rd_compact_h(blk, len, &pos, &v)) return -1;
        pos += v;
    }
    if (pos > len || !rd_compact_h(blk, len, &pos, &v)) return -1;
    return pos <= len ? (long)pos : -1;
}

BCP_API int bcp_gpu_blocks_connect_p2pkh(int device, size_t nblk, const uint8_t *const *blks, const size_t *lens,
                                         int legacy_header, const uint8_t *spk, const uint32_t *spk_off,
                                         const int64_t *amounts, const size_t *n_inputs, uint32_t flags, uint8_t *txids,
                                         size_t *ntx_out, uint8_t *roots32, int *mutated, uint8_t *script_err, int *all_ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (nblk == 0) return BCP_OK;
    if (!blks || !lens || !n_inputs || !roots32) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: NULL buffer"); return BCP_ERR_INVALID; }
    if (check_fastpath_flags(flags, "bcp_gpu_blocks_connect_p2pkh") != BCP_OK) return BCP_ERR_INVALID;
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    // body[b] = where block b's transactions start in its host buffer; base[b] = where they start in the device buffer
    std::vector<size_t> body(nblk), base(nblk + 1, 0), in_base(nblk + 1, 0);
    size_t max_tx_all = 0, max_tx_one = 0;
    for (size_t b = 0; b < nblk; b++) {
        if (!blks[b]) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: NULL block"); return BCP_ERR_INVALID; }
        const long f = first_tx_offset(blks[b], lens[b], legacy_header);
        if (f < 0) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: block %zu is malformed", b); return BCP_ERR_INVALID; }
        body[b] = (size_t)f;
        base[b + 1] = base[b] + (lens[b] - body[b]);
        in_base[b + 1] = in_base[b] + n_inputs[b];
        if (n_inputs[b] > lens[b] / 41 + 2) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: more inputs than block %zu can hold", b); return BCP_ERR_INVALID; }
        max_tx_all += lens[b] / 60 + 2;
        max_tx_one = std::max(max_tx_one, lens[b] / 60 + 2);
    }
    const size_t total_in = in_base[nblk];
    if (base[nblk] > 0xFFFFFFF0ull) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: window larger than 4 GiB"); return BCP_ERR_INVALID; }
    if (total_in && (!spk || !spk_off || !amounts || !script_err)) { bcp_set_error("bcp_gpu_blocks_connect_p2pkh: NULL buffer"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    const size_t spk_bytes = total_in ? spk_off[total_in] : 0, cap_in = total_in + 1;
    size_t o = 0;
    const size_t o_blk = o; o += up(base[nblk] + 64);
    const size_t o_txoff = o; o += up((max_tx_all + 1) * 4);
    const size_t o_inoff = o; o += up(cap_in * 4 + 4);
    const size_t o_intx = o; o += up(cap_in * 4 + 4);
    const size_t o_inn = o; o += up(cap_in * 4 + 4);
    const size_t o_spk = o; o += up(spk_bytes + 8);
    const size_t o_spkoff = o; o += up((cap_in + 1) * 4);
    const size_t o_amt = o; o += up(cap_in * 8 + 8);
    const size_t o_txid = o; o += up(max_tx_all * 32 + 32);
    const size_t o_mtmp = o; o += up(bcp_merkle_tmp_words(max_tx_one) * 4);
    const size_t o_root = o; o += up(nblk * 64); /* per block: root[32], mutated(int), pad */
    const size_t o_flag = o; o += 256;
    const size_t o_mid = o; o += up(max_tx_all * BCP_TXMID_BYTES + 16);
    const size_t o_stat = o; o += up(cap_in + 8);
    const size_t o_msg = o; o += up(cap_in * 32 + 32);
    const size_t o_pub = o; o += up(cap_in * 65 + 65);
    const size_t o_sig = o; o += up(cap_in * 64 + 64);
    const size_t o_ok = o; o += up(cap_in + 8);
    const size_t o_err = o; o += up(cap_in + 8);
    int rc = bcp_ctx_reserve_stage(c, o);
    if (rc != BCP_OK) return rc;
    rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(total_in));
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream, st2 = c->copy_stream;
    for (size_t b = 0; b < nblk; b++)
        BCP_CUDA(cudaMemcpyAsync(d + o_blk + base[b], blks[b] + body[b], lens[b] - body[b], cudaMemcpyHostToDevice, st));
    // host: one framing scan per block, on as many threads as there are blocks (up to the core count)
    struct scan_t { std::vector<uint32_t> tx_off, in_off, in_tx, in_n; long ntx = 0, nin = 0; };
    std::vector<scan_t> sc(nblk);
    {
        const unsigned T = (unsigned)std::min<size_t>(nblk, std::max(1u, std::thread::hardware_concurrency()));
        std::vector<std::thread> th;
        for (unsigned w = 0; w < T; w++)
            th.emplace_back([&, w] {
                for (size_t b = w; b < nblk; b += T) {
                    scan_t &x = sc[b];
                    const size_t mt = lens[b] / 60 + 2, mi = n_inputs[b] + 1;
                    x.tx_off.resize(mt + 1); x.in_off.resize(mi); x.in_tx.resize(mi); x.in_n.resize(mi);
                    x.nin = bcp_block_index_inputs(blks[b], lens[b], legacy_header, x.tx_off.data(), mt, x.in_off.data(),
                                                   x.in_tx.data(), x.in_n.data(), mi, &x.ntx);
                }
            });
        for (auto &t : th) t.join();
    }
    // one index space for the window: byte offsets rebased to the device buffer, transaction numbers to the window
    std::vector<size_t> t_base(nblk + 1, 0);
    for (size_t b = 0; b < nblk; b++) {
        if (sc[b].nin < 0 || (size_t)sc[b].nin != n_inputs[b]) {
            cudaStreamSynchronize(st);
            if (sc[b].nin == -1) bcp_set_error("bcp_gpu_blocks_connect_p2pkh: block %zu is malformed", b);
            else bcp_set_error("bcp_gpu_blocks_connect_p2pkh: block %zu: its non-coinbase inputs (%ld) do not match the %zu spent "
                               "outputs the caller described", b, sc[b].nin, n_inputs[b]);
            return BCP_ERR_INVALID;
        }
        if (sc[b].tx_off[sc[b].ntx] != lens[b]) { // the gap-free layout needs each block to end with its last transaction
            cudaStreamSynchronize(st);
            bcp_set_error("bcp_gpu_blocks_connect_p2pkh: block %zu has %zu bytes after its last transaction (use the "
                          "single-block call for such a buffer)", b, lens[b] - (size_t)sc[b].tx_off[sc[b].ntx]);
            return BCP_ERR_INVALID;
        }
        t_base[b + 1] = t_base[b] + (size_t)sc[b].ntx;
    }
    const size_t total_tx = t_base[nblk];
    std::vector<uint32_t> &tx_off = c->h_tx_off, &in_off = c->h_in_off, &in_tx = c->h_in_tx, &in_n = c->h_in_n;
    if (tx_off.size() < total_tx + 1) tx_off.resize(total_tx + 1);
    if (in_off.size() < cap_in) { in_off.resize(cap_in); in_tx.resize(cap_in); in_n.resize(cap_in); }
    for (size_t b = 0; b < nblk; b++) {
        const scan_t &x = sc[b];
        const uint32_t shift = (uint32_t)base[b] - (uint32_t)body[b], tb = (uint32_t)t_base[b]; // modulo 2^32 on purpose
        for (long t = 0; t < x.ntx; t++) tx_off[t_base[b] + t] = x.tx_off[t] + shift;
        for (long i = 0; i < x.nin; i++) {
            in_off[in_base[b] + i] = x.in_off[i] + shift;
            in_tx[in_base[b] + i] = x.in_tx[i] + tb;
            in_n[in_base[b] + i] = x.in_n[i];
        }
    }
    tx_off[total_tx] = (uint32_t)base[nblk]; // a block's transactions end where the next block's begin
    BCP_CUDA(cudaMemcpyAsync(d + o_txoff, tx_off.data(), (total_tx + 1) * 4, cudaMemcpyHostToDevice, st));
    uint32_t *d_txoff = reinterpret_cast<uint32_t *>(d + o_txoff);
    rc = bcp_sha256d_offsets_dev(c, d + o_blk, d_txoff, total_tx, d + o_txid, st);
    if (rc != BCP_OK) return rc;
    // Merkle roots on the second stream, beside the script checks
    BCP_CUDA(cudaEventRecord(c->ev[0], st));
    BCP_CUDA(cudaStreamWaitEvent(st2, c->ev[0], 0));
    for (size_t b = 0; b < nblk; b++) {
        rc = bcp_merkle_dev(c, d + o_txid + t_base[b] * 32, (size_t)sc[b].ntx, d + o_root + b * 64,
                            reinterpret_cast<int *>(d + o_root + b * 64 + 32), reinterpret_cast<uint32_t *>(d + o_mtmp), st2);
        if (rc != BCP_OK) return rc;
    }
    std::vector<uint8_t> res(nblk * 64);
    BCP_CUDA(cudaMemcpyAsync(res.data(), d + o_root, nblk * 64, cudaMemcpyDeviceToHost, st2));
    if (txids) BCP_CUDA(cudaMemcpyAsync(txids, d + o_txid, total_tx * 32, cudaMemcpyDeviceToHost, st2));
    BCP_CUDA(cudaEventRecord(c->ev[1], st2));
    if (total_in) {
        int one = 1;
        int *d_flag = reinterpret_cast<int *>(d + o_flag);
        BCP_CUDA(cudaMemcpyAsync(d_flag, &one, sizeof(int), cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_inoff, in_off.data(), total_in * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_intx, in_tx.data(), total_in * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_inn, in_n.data(), total_in * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_spk, spk, spk_bytes, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_spkoff, spk_off, (total_in + 1) * 4, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_amt, amounts, total_in * 8, cudaMemcpyHostToDevice, st));
        rc = launch_mid_extract(c, d + o_blk, d_txoff, tx_off.data(), total_tx, reinterpret_cast<uint32_t *>(d + o_inoff),
                                reinterpret_cast<uint32_t *>(d + o_intx), reinterpret_cast<uint32_t *>(d + o_inn), in_tx.data(), total_in,
                                d + o_spk, reinterpret_cast<uint32_t *>(d + o_spkoff), reinterpret_cast<int64_t *>(d + o_amt),
                                d + o_mid, d + o_stat, d + o_msg, d + o_pub, d + o_sig, st);
        if (rc != BCP_OK) return rc;
        rc = bcp_ecdsa_launch(c, d + o_msg, d + o_pub, d + o_sig, total_in, BCP_ECDSA_FLAG_NORMALIZE_S, d + o_ok,
                              reinterpret_cast<uint32_t *>(c->scratch), st);
        if (rc != BCP_OK) return rc;
        k_script_finish<<<(unsigned)((total_in + 255) / 256), 256, 0, st>>>(d + o_stat, d + o_ok, total_in, d + o_err, d_flag);
        BCP_CUDA(cudaGetLastError());
        BCP_CUDA(cudaMemcpyAsync(script_err, d + o_err, total_in, cudaMemcpyDeviceToHost, st));
    }
    BCP_CUDA(cudaStreamWaitEvent(st, c->ev[1], 0));
    BCP_CUDA(cudaStreamSynchronize(st));
    for (size_t b = 0; b < nblk; b++) {
        memcpy(roots32 + b * 32, &res[b * 64], 32);
        int m;
        memcpy(&m, &res[b * 64 + 32], 4);
        if (mutated) mutated[b] = m ? 1 : 0;
        if (ntx_out) ntx_out[b] = (size_t)sc[b].ntx;
        if (all_ok) {
            int a = 1;
            for (size_t i = in_base[b]; i < in_base[b + 1]; i++) a &= script_err[i] == 0;
            all_ok[b] = a;
        }
    }
    return BCP_OK;
}

} // extern "C"
