// bcp_utxo.cu — the unspent-output table of a block connector, resident in HBM, and the kernels that apply a block to
// it.  C-ABI in include/bcp_gpu.h ("Device-resident coin table").
//
// What it replaces: the in-memory coin map a connector walks for every block — per input "the spent coin must exist and
// be unspent" with the lookup of its amount and scriptPubKey (CCoinsViewCache::AccessCoin / HaveInputs / SpendCoin as
// used by ConnectBlock, src/validation.cpp:2030-2100, and CheckTxInputs :1363-1420), per output AddCoin (UpdateCoins
// :1327-1347, src/coins.cpp:71-97), and the BIP30 overwrite test (src/validation.cpp:1975-1990).  The node's LevelDB
// chainstate itself stays out of scope; this is the working set of a reindex / initial block
// download, sized by the caller, in device memory, so that the sequential part of connecting a chain — lookups, spends
// and inserts of ~41,650 coins per 8 MB block — is three kernels on ONE stream instead of a walk on ONE host thread.
// Stream order gives exactly the reference's order: block b+1 sees the table as block b left it.
//
// Table: 2^k slots of 128 bytes, open addressing with linear probing from a hash of the outpoint.  A spent coin stays
// in place as a tombstone (so probe runs of other keys are never cut) and is reused by a later insert once the block
// that spent it is final; every slot remembers which block created and which block spent it, so a rejected block — or a
// run of unconfirmed blocks on top of it — is undone by two small kernels from the slot numbers recorded when it was
// applied.
//
// Per block (`bcp_gpu_utxo_apply_block`, all on the table's stream, block b's upload on a second stream under block
// b-1's kernels):
//   k_sha256d_msgs + k_merkle_levels   txids, Merkle root, mutated flag                (bcp_hash.cu)
//   k_utxo_insert   one thread per output: the new coins (tagged block b, transaction t)
//   k_utxo_spend    one thread per input: find the coin, refuse one created by the same or a later transaction of this
//                   block, claim it with one atomicCAS (a second spender of the same coin loses: double spend)
//   k_utxo_offsets  exclusive scan of the spent scripts' lengths
//   k_utxo_gather   the spent scriptPubKeys back to back + amounts: the arrays bcp_gpu_block_scripts_p2pkh takes
#include "bcp_ctx.cuh"
#include <cstring>
#include <new>

#define UTXO_EMPTY 0u
#define UTXO_BUSY 1u
#define UTXO_LIVE 2u
#define UTXO_SPENT 3u
#define UTXO_NONE 0xFFFFFFFFu
#define UTXO_INLINE 64u
#define UTXO_RING 32 /* blocks that may be applied and not yet confirmed or rolled back */

// status words of one applied block (device -> host); [6] = spent scripts that did not fit the script buffer
enum { ST_MISSING = 0, ST_FIRST_MISSING = 1, ST_DUPLICATE = 2, ST_SPK_BYTES = 3, ST_TABLE_FULL = 4, ST_ARENA_FULL = 5, ST_SPK_OVERFLOW = 6, ST_WORDS = 8 };

struct __align__(16) utxo_slot {
    uint32_t state;
    uint32_t created_seq; /* block that created the coin (0: seeded by the caller) */
    uint32_t created_tx;  /* its transaction number inside that block */
    uint32_t spent_seq;   /* block that spent it (state SPENT); 0 = reusable at once */
    uint32_t key[9];      /* outpoint: txid (32 bytes, memory order) || vout, as little-endian words */
    uint32_t spk_len;
    int64_t amount;
    uint8_t spk[UTXO_INLINE]; /* the script, or for a longer one the 64-bit offset of its bytes in the arena */
};
static_assert(sizeof(utxo_slot) == 128, "slot is 128 bytes");

struct utxo_dev { /* what the kernels need, by value */
    utxo_slot *slots;
    uint32_t mask;
    uint8_t *arena;
    unsigned long long arena_cap;
    unsigned long long *arena_next;
    long long *live; /* number of live coins */
};

__device__ __forceinline__ uint32_t utxo_home(const uint32_t *key, uint32_t mask) {
    const uint64_t h = ((uint64_t)key[1] << 32 | key[0]) ^ (0x9E3779B97F4A7C15ull * key[8]);
    return (uint32_t)(h ^ (h >> 29)) & mask;
}
__device__ __forceinline__ bool utxo_key_eq(const utxo_slot *s, const uint32_t *key) {
    bool eq = true;
#pragma unroll
    for (int k = 0; k < 9; k++) eq = eq && s->key[k] == key[k];
    return eq;
}
__device__ __forceinline__ uint32_t ld_state(const utxo_slot *s) { return *reinterpret_cast<const volatile uint32_t *>(&s->state); }

// Puts one coin into the table.  Returns the slot number, UTXO_NONE when the table is full; status[ST_DUPLICATE] counts the
// coins put on top of a LIVE coin with the same outpoint that an EARLIER block (or the caller) created — the BIP30
// situation; the new coin is stored all the same and the block is expected to be rolled back.
__device__ uint32_t utxo_put(const utxo_dev &T, const uint32_t *key, const uint8_t *spk, uint32_t len, int64_t amount, uint32_t seq,
                             uint32_t tx, uint32_t confirmed, uint32_t *status) {
    uint32_t idx = utxo_home(key, T.mask);
    uint32_t reuse = UTXO_NONE;
    for (uint32_t probes = 0; probes <= T.mask; probes++, idx = (idx + 1) & T.mask) {
        utxo_slot *s = T.slots + idx;
        const uint32_t st = ld_state(s);
        if (st == UTXO_LIVE) {
            // BIP30 looks at the table as it was BEFORE this block: a twin made by an identical transaction of the same block
            // does not count (that transaction fails on its spent inputs instead)
            if (s->created_seq != seq && utxo_key_eq(s, key)) atomicAdd(&status[ST_DUPLICATE], 1u);
            continue;
        }
        if (st == UTXO_SPENT && reuse == UTXO_NONE && s->spent_seq <= confirmed) reuse = idx; /* a final tombstone */
        if (st != UTXO_EMPTY) continue;
        // end of the probe run: take the remembered tombstone, else this empty slot
        uint32_t take = reuse != UTXO_NONE ? reuse : idx;
        utxo_slot *d = T.slots + take;
        const uint32_t expect = reuse != UTXO_NONE ? UTXO_SPENT : UTXO_EMPTY;
        if (atomicCAS(&d->state, expect, UTXO_BUSY) != expect) { /* another output of this launch was faster */
            reuse = UTXO_NONE;              /* forget the tombstone, if that is what was lost ...                       */
            idx = (idx + T.mask) & T.mask;  /* ... and look at the current slot again (the loop steps forward by one) */
            continue;
        }
        d->created_seq = seq;
        d->created_tx = tx;
        d->spent_seq = 0;
#pragma unroll
        for (int k = 0; k < 9; k++) d->key[k] = key[k];
        d->spk_len = len;
        d->amount = amount;
        if (len <= UTXO_INLINE) {
            for (uint32_t k = 0; k < len; k++) d->spk[k] = spk[k];
        } else {
            const unsigned long long need = (len + 7ull) & ~7ull;
            const unsigned long long at = atomicAdd(T.arena_next, need);
            if (at + need > T.arena_cap) {
                atomicAdd(&status[ST_ARENA_FULL], 1u);
                d->spk_len = 0;
            } else {
                for (uint32_t k = 0; k < len; k++) T.arena[at + k] = spk[k];
            }
            *reinterpret_cast<unsigned long long *>(d->spk) = at;
        }
        __threadfence();
        *reinterpret_cast<volatile uint32_t *>(&d->state) = UTXO_LIVE;
        return take;
    }
    atomicAdd(&status[ST_TABLE_FULL], 1u); /* every slot looked at, none free (tombstones of pending blocks do not count) */
    return UTXO_NONE;
}

__device__ __forceinline__ uint64_t utxo_compact(const uint8_t *p, uint32_t &pos) {
    const uint8_t c = p[pos++];
    if (c < 253) return c;
    const int nb = c == 253 ? 2 : c == 254 ? 4 : 8;
    uint64_t v = 0;
    for (int i = 0; i < nb; i++) v |= (uint64_t)p[pos + i] << (8 * i);
    pos += nb;
    return v;
}

// seeding: coins given by the caller (created_seq 0)
__global__ void __launch_bounds__(128) k_utxo_seed(utxo_dev T, const uint8_t *__restrict__ outpoints36, const uint8_t *__restrict__ spk,
                                                   const uint32_t *__restrict__ spk_off, const int64_t *__restrict__ amounts, size_t n,
                                                   uint32_t *__restrict__ status) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t key[9];
    const uint8_t *p = outpoints36 + 36 * i;
#pragma unroll
    for (int k = 0; k < 9; k++) key[k] = p[4 * k] | (uint32_t)p[4 * k + 1] << 8 | (uint32_t)p[4 * k + 2] << 16 | (uint32_t)p[4 * k + 3] << 24;
    const uint32_t r = utxo_put(T, key, spk + spk_off[i], spk_off[i + 1] - spk_off[i], amounts[i], 0, 0, 0xFFFFFFFFu, status);
    if (r != UTXO_NONE) atomicAdd(reinterpret_cast<unsigned long long *>(T.live), 1ull);
}

// one thread per output of the block; rec_out[o] = its slot (UTXO_NONE: unspendable output, not stored —
// CScript::IsUnspendable, src/script/script.h: OP_RETURN first or longer than MAX_SCRIPT_SIZE 10,000)
__global__ void __launch_bounds__(128)
    k_utxo_insert(utxo_dev T, uint32_t seq, uint32_t confirmed, const uint8_t *__restrict__ blk, const uint8_t *__restrict__ txids,
                  const uint32_t *__restrict__ out_off, const uint32_t *__restrict__ out_tx, const uint32_t *__restrict__ out_n, size_t nout,
                  uint32_t *__restrict__ rec_out, uint32_t *__restrict__ status) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= nout) return;
    const uint32_t t = out_tx[o];
    uint32_t key[9];
    const uint32_t *id = reinterpret_cast<const uint32_t *>(txids + 32 * (size_t)t); /* digests are 32-byte aligned */
#pragma unroll
    for (int k = 0; k < 8; k++) key[k] = id[k];
    key[8] = out_n[o];
    uint32_t pos = out_off[o];
    uint64_t amount_u = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) amount_u |= (uint64_t)blk[pos + k] << (8 * k);
    const int64_t amount = (int64_t)amount_u;
    pos += 8;
    const uint32_t len = (uint32_t)utxo_compact(blk, pos);
    uint32_t r = UTXO_NONE;
    if (!((len > 0 && blk[pos] == 0x6a) || len > 10000u)) {
        r = utxo_put(T, key, blk + pos, len, amount, seq, t, confirmed, status);
        if (r != UTXO_NONE) atomicAdd(reinterpret_cast<unsigned long long *>(T.live), 1ull);
    }
    rec_out[o] = r;
}

// one thread per input of the block; rec_in[i] = the slot it spent (UTXO_NONE: no such coin), lens[i] / amounts[i] = what
// the script check needs.  A coin created by this very block counts only when its transaction comes BEFORE the spender
// (UpdateCoins applies each transaction before the next one is looked at, src/validation.cpp:2095).
__global__ void __launch_bounds__(128)
    k_utxo_spend(utxo_dev T, uint32_t seq, const uint8_t *__restrict__ blk, const uint32_t *__restrict__ in_off,
                 const uint32_t *__restrict__ in_tx, size_t nin, uint32_t *__restrict__ rec_in, int64_t *__restrict__ amounts,
                 uint32_t *__restrict__ lens, uint32_t *__restrict__ status) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nin) return;
    uint32_t key[9];
    const uint8_t *p = blk + in_off[i];
#pragma unroll
    for (int k = 0; k < 9; k++) key[k] = p[4 * k] | (uint32_t)p[4 * k + 1] << 8 | (uint32_t)p[4 * k + 2] << 16 | (uint32_t)p[4 * k + 3] << 24;
    uint32_t idx = utxo_home(key, T.mask), found = UTXO_NONE;
    for (uint32_t probes = 0; probes <= T.mask; probes++, idx = (idx + 1) & T.mask) {
        utxo_slot *s = T.slots + idx;
        const uint32_t st = ld_state(s);
        if (st == UTXO_EMPTY) break;
        if (st != UTXO_LIVE || !utxo_key_eq(s, key)) continue;
        if (s->created_seq == seq && s->created_tx >= in_tx[i]) break; /* created by the same or a later transaction */
        if (atomicCAS(&s->state, UTXO_LIVE, UTXO_SPENT) == UTXO_LIVE) { /* the only spender that gets it */
            s->spent_seq = seq;
            found = idx;
        }
        break;
    }
    rec_in[i] = found;
    if (found != UTXO_NONE) {
        amounts[i] = T.slots[found].amount;
        lens[i] = T.slots[found].spk_len;
        atomicAdd(reinterpret_cast<unsigned long long *>(T.live), (unsigned long long)-1ll);
    } else {
        amounts[i] = 0;
        lens[i] = 0;
        atomicAdd(&status[ST_MISSING], 1u);
        atomicMax(&status[ST_FIRST_MISSING], ~(uint32_t)i); /* the smallest index wins; decoded by the host side */
    }
}

// spk_off[0..n] = exclusive prefix sums of lens.  One block walks the array in tiles of 1024 threads x 8 items (a block of
// 41,650 inputs is six tiles); n is a block's input count.
#define UTXO_SCAN_ITEMS 8
__global__ void __launch_bounds__(1024) k_utxo_offsets(const uint32_t *__restrict__ lens, size_t n, uint32_t *__restrict__ spk_off,
                                                       uint32_t *__restrict__ status) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (size_t base = 0; base < n; base += (size_t)blockDim.x * UTXO_SCAN_ITEMS) {
        const size_t i0 = base + (size_t)threadIdx.x * UTXO_SCAN_ITEMS;
        uint32_t v[UTXO_SCAN_ITEMS], mine = 0;
#pragma unroll
        for (int k = 0; k < UTXO_SCAN_ITEMS; k++) {
            v[k] = i0 + k < n ? lens[i0 + k] : 0u;
            mine += v[k];
        }
        uint32_t x = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d);
            if (lane >= (unsigned)d) x += y;
        }
        if (lane == 31) warp_sum[w] = x;
        __syncthreads();
        if (w == 0) {
            uint32_t s = warp_sum[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, s, d);
                if (lane >= (unsigned)d) s += y;
            }
            warp_sum[lane] = s; /* inclusive over warps */
        }
        __syncthreads();
        uint32_t run = carry + (w ? warp_sum[w - 1] : 0u) + x - mine;
#pragma unroll
        for (int k = 0; k < UTXO_SCAN_ITEMS; k++) {
            if (i0 + k < n) spk_off[i0 + k] = run;
            run += v[k];
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry += warp_sum[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        spk_off[n] = carry;
        status[ST_SPK_BYTES] = carry;
    }
}

__global__ void __launch_bounds__(128)
    k_utxo_gather(utxo_dev T, const uint32_t *__restrict__ rec_in, const uint32_t *__restrict__ spk_off, size_t nin,
                  uint8_t *__restrict__ spk_cat, size_t cap, uint32_t *__restrict__ status) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nin) return;
    const uint32_t r = rec_in[i];
    if (r == UTXO_NONE) return;
    const utxo_slot *s = T.slots + r;
    const uint32_t len = s->spk_len;
    if ((size_t)spk_off[i] + len > cap) { /* the buffer the caller sized cannot hold this block's spent scripts */
        atomicAdd(&status[ST_SPK_OVERFLOW], 1u);
        return;
    }
    const uint8_t *src = len <= UTXO_INLINE ? s->spk : T.arena + *reinterpret_cast<const unsigned long long *>(s->spk);
    uint8_t *dst = spk_cat + spk_off[i];
    for (uint32_t k = 0; k < len; k++) dst[k] = src[k];
}

// undo of one applied block, newest first: its spent coins are live again, its new coins become reusable tombstones
// (a coin this block created AND spent is touched by its output's thread only; coins of the block that later in-flight
// blocks spent are LIVE again by now, because those blocks are rolled back first)
__global__ void __launch_bounds__(128) k_utxo_undo(utxo_dev T, uint32_t seq, const uint32_t *__restrict__ rec_in, size_t nin,
                                                   const uint32_t *__restrict__ rec_out, size_t nout) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nin) {
        const uint32_t r = rec_in[i];
        if (r != UTXO_NONE && T.slots[r].created_seq != seq) {
            T.slots[r].state = UTXO_LIVE;
            atomicAdd(reinterpret_cast<unsigned long long *>(T.live), 1ull);
        }
    }
    if (i < nout) {
        const uint32_t r = rec_out[i];
        if (r != UTXO_NONE) {
            if (T.slots[r].state == UTXO_LIVE) atomicAdd(reinterpret_cast<unsigned long long *>(T.live), (unsigned long long)-1ll);
            T.slots[r].spent_seq = 0;
            T.slots[r].state = UTXO_SPENT;
        }
    }
}

// Run when a block becomes final, over the coins it spent.  A tombstone only exists so that the probe runs of OTHER coins
// are not cut: coin j (home slot h, stored at j) needs every slot from h to j to be non-empty.  So the tombstone at i may be
// emptied iff no coin stored after it in the same run has i on its path, i.e. iff every later entry of the run that may
// still have to be found — live coins, and the spent coins of blocks that are not final yet (a rollback makes them live
// again) — has its home behind i.  Each thread decides that for one tombstone by walking to the end of the run.  The
// decisions are independent of one another (emptying a slot that is on nobody's path changes nobody's path), so a whole
// block's tombstones go in one launch.  In a table kept under half full runs are a few slots long and almost every spent
// coin is freed this way; what stays is reused by insertions that walk over it (utxo_put).
__global__ void __launch_bounds__(128) k_utxo_reclaim(utxo_dev T, uint32_t seq, const uint32_t *__restrict__ rec_in, size_t nin) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nin) return;
    const uint32_t i = rec_in[k];
    if (i == UTXO_NONE) return;
    utxo_slot *s = T.slots + i;
    if (ld_state(s) != UTXO_SPENT || s->spent_seq > seq) return; /* seq = the newest final block */
    uint32_t j = (i + 1) & T.mask;
    for (uint32_t steps = 0; steps < 512; steps++, j = (j + 1) & T.mask) {
        const utxo_slot *e = T.slots + j;
        const uint32_t st = ld_state(e);
        if (st == UTXO_EMPTY) { /* end of the run: nobody needs slot i */
            *reinterpret_cast<volatile uint32_t *>(&s->state) = UTXO_EMPTY;
            return;
        }
        if (st == UTXO_SPENT && e->spent_seq <= seq) continue; /* a final tombstone never has to be found again */
        const uint32_t h = utxo_home(e->key, T.mask);
        if (((i - h) & T.mask) <= ((j - h) & T.mask)) return; /* slot i lies between that coin's home and the coin */
    }
}

struct ring_slot {
    uint8_t *d = nullptr; /* one allocation carved below */
    size_t cap = 0;
    uint32_t seq = 0;
    bool used = false;
    size_t nin = 0, nout = 0, ntx = 0, len = 0;
    size_t o_rec_in = 0, o_rec_out = 0;
    size_t o_txoff = 0, o_inoff = 0, o_intx = 0, o_amt = 0, o_spkoff = 0, in_area = 0; /* script-check inputs: one contiguous area */
    cudaEvent_t uploaded = nullptr, done = nullptr, reclaimed = nullptr;
    bool reclaim_pending = false;
    // host destinations of the results
    uint8_t *h_root32 = nullptr;
    int *h_mutated = nullptr;
    cudaEvent_t hashed = nullptr, merkled = nullptr;
    uint32_t *h_status = nullptr;
    uint8_t *h_spk_cat = nullptr;
    size_t spk_cap = 0, spk_copied = 0, o_spk = 0;
    uint32_t h_status_buf[ST_WORDS];
};

struct bcp_utxo {
    int device = 0;
    bcp_ctx *ctx = nullptr;
    utxo_dev T{};
    unsigned log2_slots = 0;
    cudaStream_t st = nullptr, up = nullptr, aux = nullptr; /* table kernels | uploads | Merkle reductions */
    uint32_t confirmed = 0;
    ring_slot ring[UTXO_RING];
    uint32_t *d_seed_status = nullptr;
    uint32_t *h_pinned_status = nullptr; /* UTXO_RING x 32 words, pinned: per pending block status[8] | root[32 bytes] | mutated */
    bool peer_tried[BCP_MAX_DEVICES] = {false}; /* peer access between the table's GPU and that GPU has been asked for */
    std::mutex mu;
};

static size_t up256(size_t x) { return (x + 255) / 256 * 256; }

extern "C" {

BCP_API bcp_utxo *bcp_gpu_utxo_create(int device, unsigned log2_slots, size_t arena_bytes) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return nullptr;
    if (log2_slots < 10 || log2_slots > 30) {
        bcp_set_error("bcp_gpu_utxo_create: log2_slots %u out of range (10..30)", log2_slots);
        return nullptr;
    }
    if (cudaSetDevice(c->device) != cudaSuccess) return nullptr;
    bcp_utxo *u = new (std::nothrow) bcp_utxo();
    if (!u) return nullptr;
    u->device = c->device;
    u->ctx = c;
    u->log2_slots = log2_slots;
    const size_t nslots = (size_t)1 << log2_slots;
    cudaError_t e = cudaMalloc(&u->T.slots, nslots * sizeof(utxo_slot));
    if (e == cudaSuccess) e = cudaMemset(u->T.slots, 0, nslots * sizeof(utxo_slot));
    if (arena_bytes < 4096) arena_bytes = 4096;
    if (e == cudaSuccess) e = cudaMalloc(&u->T.arena, arena_bytes);
    uint8_t *ctr = nullptr;
    if (e == cudaSuccess) e = cudaMalloc(&ctr, 256);
    if (e == cudaSuccess) e = cudaMemset(ctr, 0, 256);
    // The table's kernels are small (tens of microseconds per block) and everything downstream waits for them, while the
    // same GPU usually also runs script checks whose point kernel fills the SMs for a millisecond at a time: the table's
    // streams get the highest priority, so its thread blocks are placed first whenever an SM has room.
    int prLow = 0, prHigh = 0;
    if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prLow, &prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&u->st, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&u->up, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&u->aux, cudaStreamNonBlocking, prHigh);
    if (e == cudaSuccess) e = cudaHostAlloc(&u->h_pinned_status, UTXO_RING * 32 * sizeof(uint32_t), cudaHostAllocPortable);
    for (int k = 0; k < UTXO_RING && e == cudaSuccess; k++) {
        e = cudaEventCreateWithFlags(&u->ring[k].uploaded, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&u->ring[k].done, cudaEventDisableTiming | cudaEventBlockingSync); /* waited for by host threads: sleep, do not spin */
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&u->ring[k].reclaimed, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&u->ring[k].hashed, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&u->ring[k].merkled, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        bcp_cuda_fail(e, "bcp_gpu_utxo_create");
        delete u;
        return nullptr;
    }
    u->T.mask = (uint32_t)(nslots - 1);
    u->T.arena_cap = arena_bytes;
    u->T.arena_next = reinterpret_cast<unsigned long long *>(ctr);
    u->T.live = reinterpret_cast<long long *>(ctr + 8);
    u->d_seed_status = reinterpret_cast<uint32_t *>(ctr + 64);
    return u;
}

BCP_API void bcp_gpu_utxo_destroy(bcp_utxo *u) {
    if (!u) return;
    cudaSetDevice(u->device);
    if (u->st) cudaStreamSynchronize(u->st);
    if (u->up) cudaStreamSynchronize(u->up);
    if (u->aux) cudaStreamSynchronize(u->aux);
    for (int k = 0; k < UTXO_RING; k++) {
        if (u->ring[k].d) cudaFree(u->ring[k].d);
        if (u->ring[k].uploaded) cudaEventDestroy(u->ring[k].uploaded);
        if (u->ring[k].done) cudaEventDestroy(u->ring[k].done);
        if (u->ring[k].reclaimed) cudaEventDestroy(u->ring[k].reclaimed);
        if (u->ring[k].hashed) cudaEventDestroy(u->ring[k].hashed);
        if (u->ring[k].merkled) cudaEventDestroy(u->ring[k].merkled);
    }
    if (u->T.slots) cudaFree(u->T.slots);
    if (u->T.arena) cudaFree(u->T.arena);
    if (u->T.arena_next) cudaFree(u->T.arena_next);
    if (u->h_pinned_status) cudaFreeHost(u->h_pinned_status);
    if (u->st) cudaStreamDestroy(u->st);
    if (u->up) cudaStreamDestroy(u->up);
    if (u->aux) cudaStreamDestroy(u->aux);
    delete u;
}

BCP_API int bcp_gpu_utxo_add_coins(bcp_utxo *u, const uint8_t *outpoints36, const uint8_t *spk, const uint32_t *spk_off,
                                   const int64_t *amounts, size_t n) {
    if (!u) return BCP_ERR_INVALID;
    if (n == 0) return BCP_OK;
    if (!outpoints36 || !spk || !spk_off || !amounts) { bcp_set_error("bcp_gpu_utxo_add_coins: NULL buffer"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(u->mu);
    BCP_CUDA(cudaSetDevice(u->device));
    const size_t spk_bytes = spk_off[n];
    uint8_t *d = nullptr;
    const size_t o_op = 0, o_spk = up256(n * 36), o_off = o_spk + up256(spk_bytes + 8), o_amt = o_off + up256((n + 1) * 4);
    BCP_CUDA(cudaMalloc(&d, o_amt + up256(n * 8)));
    cudaError_t e = cudaMemcpyAsync(d + o_op, outpoints36, n * 36, cudaMemcpyHostToDevice, u->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_spk, spk, spk_bytes, cudaMemcpyHostToDevice, u->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_off, spk_off, (n + 1) * 4, cudaMemcpyHostToDevice, u->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + o_amt, amounts, n * 8, cudaMemcpyHostToDevice, u->st);
    if (e == cudaSuccess) e = cudaMemsetAsync(u->d_seed_status, 0, ST_WORDS * 4, u->st);
    uint32_t hs[ST_WORDS] = {0};
    if (e == cudaSuccess) {
        k_utxo_seed<<<(unsigned)((n + 127) / 128), 128, 0, u->st>>>(u->T, d + o_op, d + o_spk, reinterpret_cast<uint32_t *>(d + o_off),
                                                                   reinterpret_cast<int64_t *>(d + o_amt), n, u->d_seed_status);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(hs, u->d_seed_status, sizeof hs, cudaMemcpyDeviceToHost, u->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(u->st);
    cudaFree(d);
    if (e != cudaSuccess) return bcp_cuda_fail(e, "bcp_gpu_utxo_add_coins");
    if (hs[ST_TABLE_FULL] || hs[ST_ARENA_FULL]) {
        bcp_set_error("bcp_gpu_utxo_add_coins: the coin table is full (%u coins did not fit, %u long scripts did not fit the arena)",
                      hs[ST_TABLE_FULL], hs[ST_ARENA_FULL]);
        return BCP_ERR_INVALID;
    }
    return BCP_OK;
}

BCP_API long long bcp_gpu_utxo_size(bcp_utxo *u) {
    if (!u) return -1;
    std::lock_guard<std::mutex> lk(u->mu);
    if (cudaSetDevice(u->device) != cudaSuccess) return -1;
    long long v = -1;
    if (cudaMemcpyAsync(&v, u->T.live, 8, cudaMemcpyDeviceToHost, u->st) != cudaSuccess) return -1;
    if (cudaStreamSynchronize(u->st) != cudaSuccess) return -1;
    return v;
}

// Allocate every pending-block area for blocks of up to max_block_bytes now (counts bounded as for P2PKH-shaped
// transactions: the layout below with ntx <= len/60, nin <= len/41, nout <= len/34), so that no block of a run meets
// cudaMalloc — a device-wide synchronisation — on its way through bcp_gpu_utxo_apply_block.  Optional: areas still grow on demand.
BCP_API int bcp_gpu_utxo_reserve(bcp_utxo *u, size_t max_block_bytes) {
    if (!u) return BCP_ERR_INVALID;
    const size_t len = max_block_bytes, ntx = len / 60 + 2, nin = len / 41 + 2, nout = len / 34 + 2;
    size_t o = up256(len + 16) + up256((ntx + 1) * 4) + 2 * up256(nin * 4 + 4) + 3 * up256(nout * 4 + 4);
    o += up256(nin * 8 + 8) + up256((nin + 1) * 4) + up256(nin * 80 + 65536 + 16);
    o += up256(ntx * 32 + 32) + up256(bcp_merkle_tmp_words(ntx) * 4) + 256;
    o += 2 * up256(nin * 4 + 4) + up256(nout * 4 + 4);
    std::lock_guard<std::mutex> lk(u->mu);
    BCP_CUDA(cudaSetDevice(u->device));
    for (int k = 0; k < UTXO_RING; k++) {
        ring_slot &r = u->ring[k];
        if (r.used || r.cap >= o) continue;
        if (r.d) BCP_CUDA(cudaFree(r.d));
        r.d = nullptr;
        r.cap = 0;
        BCP_CUDA(cudaMalloc(&r.d, o));
        r.cap = o;
    }
    return BCP_OK;
}

BCP_API int bcp_gpu_utxo_apply_block(bcp_utxo *u, uint32_t seq, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                     const uint32_t *in_off, const uint32_t *in_tx, size_t nin, const uint32_t *out_off,
                                     const uint32_t *out_tx, const uint32_t *out_n, size_t nout, uint8_t *txids, uint8_t *root32,
                                     int *mutated, uint8_t *spk_cat, size_t spk_cap, uint32_t *spk_off, int64_t *amounts) {
    if (!u) return BCP_ERR_INVALID;
    if (seq == 0 || !blk || !tx_off || ntx == 0 || !root32 || !mutated || (nin && (!in_off || !in_tx)) ||
        (nout && (!out_off || !out_tx || !out_n)) || ((spk_cat || spk_off || amounts) && !(spk_cat && spk_off && amounts))) {
        bcp_set_error("bcp_gpu_utxo_apply_block: NULL buffer, no transaction or sequence number 0 (numbers start at 1)");
        return BCP_ERR_INVALID;
    }
    if (len > 0xFFFFFFF0ull || tx_off[ntx] > len) { bcp_set_error("bcp_gpu_utxo_apply_block: offsets exceed the block"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(u->mu);
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (r.used) {
        bcp_set_error("bcp_gpu_utxo_apply_block: block %u is still pending in the slot block %u needs (at most %d blocks may be "
                      "applied and neither confirmed nor rolled back)", r.seq, seq, UTXO_RING);
        return BCP_ERR_INVALID;
    }
    BCP_CUDA(cudaSetDevice(u->device));
    // device layout of this block's slot
    // The arrays the script checks need come first and back to back (block, offsets, inputs, outputs, spent amounts /
    // script offsets / scripts): bcp_gpu_utxo_block_scripts reads them in place or moves them to another GPU in one peer
    // copy.  The six offset arrays are adjacent, each padded to 256 bytes: a caller that keeps them laid out the same way
    // in host memory gets them uploaded with one copy.
    size_t o = 0;
    const size_t o_blk = o; o += up256(len + 16);
    const size_t o_txoff = o; o += up256((ntx + 1) * 4);
    const size_t o_inoff = o; o += up256(nin * 4 + 4);
    const size_t o_intx = o; o += up256(nin * 4 + 4);
    const size_t o_outoff = o; o += up256(nout * 4 + 4);
    const size_t o_outtx = o; o += up256(nout * 4 + 4);
    const size_t o_outn = o; o += up256(nout * 4 + 4);
    const size_t arrays_bytes = o - o_txoff;
    const size_t o_amt = o; o += up256(nin * 8 + 8);
    const size_t o_spkoff = o; o += up256((nin + 1) * 4);
    const size_t dev_spk_cap = spk_cat ? spk_cap : nin * 80 + 65536; /* without a host buffer: room for P2PKH-sized scripts and more */
    const size_t o_spk = o; o += up256(dev_spk_cap + 16);
    const size_t in_area = o;
    const size_t o_txid = o; o += up256(ntx * 32 + 32);
    const size_t o_mtmp = o; o += up256(bcp_merkle_tmp_words(ntx) * 4);
    const size_t o_stat = o; o += 256; /* status[8 words] | root[32 bytes] | mutated(int): one copy brings all three home */
    const size_t o_root = o_stat + ST_WORDS * 4;
    const size_t o_recin = o; o += up256(nin * 4 + 4);
    const size_t o_recout = o; o += up256(nout * 4 + 4);
    const size_t o_lens = o; o += up256(nin * 4 + 4);
    if (o > r.cap) { /* grow-only; happens while the first blocks of a run pass through */
        BCP_CUDA(cudaStreamSynchronize(u->st));
        BCP_CUDA(cudaStreamSynchronize(u->up));
        BCP_CUDA(cudaStreamSynchronize(u->aux));
        if (r.d) BCP_CUDA(cudaFree(r.d));
        r.d = nullptr;
        r.cap = 0;
        BCP_CUDA(cudaMalloc(&r.d, o + o / 8));
        r.cap = o + o / 8;
    }
    uint8_t *d = r.d;
    // upload on the second stream (it runs under the kernels of the block before); the slot's previous user is done except,
    // possibly, for the clean-up kernel its confirmation queued, which still reads its spent-slot list
    if (r.reclaim_pending) {
        BCP_CUDA(cudaStreamWaitEvent(u->up, r.reclaimed, 0));
        r.reclaim_pending = false;
    }
    BCP_CUDA(cudaMemcpyAsync(d + o_blk, blk, len, cudaMemcpyHostToDevice, u->up));
    const uint8_t *hb = reinterpret_cast<const uint8_t *>(tx_off);
    const bool packed = nin && nout && reinterpret_cast<const uint8_t *>(in_off) == hb + (o_inoff - o_txoff) &&
                        reinterpret_cast<const uint8_t *>(in_tx) == hb + (o_intx - o_txoff) &&
                        reinterpret_cast<const uint8_t *>(out_off) == hb + (o_outoff - o_txoff) &&
                        reinterpret_cast<const uint8_t *>(out_tx) == hb + (o_outtx - o_txoff) &&
                        reinterpret_cast<const uint8_t *>(out_n) == hb + (o_outn - o_txoff);
    if (packed) { /* the caller's arrays are laid out like the device's: one copy (every runtime call costs the connecting thread) */
        BCP_CUDA(cudaMemcpyAsync(d + o_txoff, tx_off, arrays_bytes, cudaMemcpyHostToDevice, u->up));
    } else {
        BCP_CUDA(cudaMemcpyAsync(d + o_txoff, tx_off, (ntx + 1) * 4, cudaMemcpyHostToDevice, u->up));
        if (nin) {
            BCP_CUDA(cudaMemcpyAsync(d + o_inoff, in_off, nin * 4, cudaMemcpyHostToDevice, u->up));
            BCP_CUDA(cudaMemcpyAsync(d + o_intx, in_tx, nin * 4, cudaMemcpyHostToDevice, u->up));
        }
        if (nout) {
            BCP_CUDA(cudaMemcpyAsync(d + o_outoff, out_off, nout * 4, cudaMemcpyHostToDevice, u->up));
            BCP_CUDA(cudaMemcpyAsync(d + o_outtx, out_tx, nout * 4, cudaMemcpyHostToDevice, u->up));
            BCP_CUDA(cudaMemcpyAsync(d + o_outn, out_n, nout * 4, cudaMemcpyHostToDevice, u->up));
        }
    }
    BCP_CUDA(cudaEventRecord(r.uploaded, u->up));
    cudaStream_t st = u->st;
    BCP_CUDA(cudaStreamWaitEvent(st, r.uploaded, 0));
    uint32_t *d_stat = reinterpret_cast<uint32_t *>(d + o_stat);
    BCP_CUDA(cudaMemsetAsync(d_stat, 0, ST_WORDS * 4, st)); /* [ST_FIRST_MISSING] holds ~index under atomicMax: 0 = none */
    int rc = bcp_sha256d_offsets_dev(u->ctx, d + o_blk, reinterpret_cast<uint32_t *>(d + o_txoff), ntx, d + o_txid, st);
    if (rc != BCP_OK) return rc;
    // the Merkle reduction needs the txids only: it runs on the third stream while the table kernels go on (in its
    // eight-levels-per-launch form: off the critical path its latency does not matter, the connecting thread's API calls do)
    BCP_CUDA(cudaEventRecord(r.hashed, st));
    BCP_CUDA(cudaStreamWaitEvent(u->aux, r.hashed, 0));
    rc = bcp_merkle_dev_few_launches(u->ctx, d + o_txid, ntx, d + o_root, reinterpret_cast<int *>(d + o_root + 32),
                                     reinterpret_cast<uint32_t *>(d + o_mtmp), u->aux);
    if (rc != BCP_OK) return rc;
    BCP_CUDA(cudaEventRecord(r.merkled, u->aux));
    if (nout)
        k_utxo_insert<<<(unsigned)((nout + 127) / 128), 128, 0, st>>>(u->T, seq, u->confirmed, d + o_blk, d + o_txid,
                                                                     reinterpret_cast<uint32_t *>(d + o_outoff), reinterpret_cast<uint32_t *>(d + o_outtx),
                                                                     reinterpret_cast<uint32_t *>(d + o_outn), nout,
                                                                     reinterpret_cast<uint32_t *>(d + o_recout), d_stat);
    if (nin) {
        k_utxo_spend<<<(unsigned)((nin + 127) / 128), 128, 0, st>>>(u->T, seq, d + o_blk, reinterpret_cast<uint32_t *>(d + o_inoff),
                                                                   reinterpret_cast<uint32_t *>(d + o_intx), nin,
                                                                   reinterpret_cast<uint32_t *>(d + o_recin), reinterpret_cast<int64_t *>(d + o_amt),
                                                                   reinterpret_cast<uint32_t *>(d + o_lens), d_stat);
        k_utxo_offsets<<<1, 1024, 0, st>>>(reinterpret_cast<uint32_t *>(d + o_lens), nin, reinterpret_cast<uint32_t *>(d + o_spkoff), d_stat);
        k_utxo_gather<<<(unsigned)((nin + 127) / 128), 128, 0, st>>>(u->T, reinterpret_cast<uint32_t *>(d + o_recin),
                                                                    reinterpret_cast<uint32_t *>(d + o_spkoff), nin, d + o_spk, dev_spk_cap, d_stat);
    }
    BCP_CUDA(cudaGetLastError());
    // results to the host: the fixed-size ones now; of the scripts as much as P2PKH-sized coins would need (the rest, if the
    // block spent longer scripts, is fetched by bcp_gpu_utxo_wait once the total is known)
    uint32_t *hst = u->h_pinned_status + (seq % UTXO_RING) * 32;
    BCP_CUDA(cudaStreamWaitEvent(st, r.merkled, 0)); /* the Merkle root ran beside the table kernels on the third stream */
    BCP_CUDA(cudaMemcpyAsync(hst, d_stat, ST_WORDS * 4 + 32 + 4, cudaMemcpyDeviceToHost, st));
    if (txids) BCP_CUDA(cudaMemcpyAsync(txids, d + o_txid, ntx * 32, cudaMemcpyDeviceToHost, st));
    size_t first = 0;
    if (!spk_cat) spk_cap = 0; /* the spent-coin arrays stay on the device (bcp_gpu_utxo_block_scripts reads them there) */
    if (nin && spk_cat) {
        BCP_CUDA(cudaMemcpyAsync(amounts, d + o_amt, nin * 8, cudaMemcpyDeviceToHost, st));
        BCP_CUDA(cudaMemcpyAsync(spk_off, d + o_spkoff, (nin + 1) * 4, cudaMemcpyDeviceToHost, st));
        first = nin * 25 + 1024 < spk_cap ? nin * 25 + 1024 : spk_cap;
        BCP_CUDA(cudaMemcpyAsync(spk_cat, d + o_spk, first, cudaMemcpyDeviceToHost, st));
    }
    BCP_CUDA(cudaEventRecord(r.done, st));
    r.used = true;
    r.seq = seq;
    r.nin = nin; r.nout = nout; r.ntx = ntx; r.len = len;
    r.o_rec_in = o_recin; r.o_rec_out = o_recout;
    r.o_txoff = o_txoff; r.o_inoff = o_inoff; r.o_intx = o_intx; r.o_amt = o_amt; r.o_spkoff = o_spkoff; r.in_area = in_area;
    r.h_status = hst;
    r.h_root32 = root32;
    r.h_mutated = mutated;
    r.h_spk_cat = spk_cat; r.spk_cap = spk_cap; r.spk_copied = first; r.o_spk = o_spk;
    return BCP_OK;
}

BCP_API int bcp_gpu_utxo_wait(bcp_utxo *u, uint32_t seq, uint32_t status8[8]) {
    if (!u || !status8) return BCP_ERR_INVALID;
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (!r.used || r.seq != seq) { bcp_set_error("bcp_gpu_utxo_wait: block %u is not pending", seq); return BCP_ERR_INVALID; }
    BCP_CUDA(cudaSetDevice(u->device));
    BCP_CUDA(cudaEventSynchronize(r.done));
    memcpy(status8, r.h_status, ST_WORDS * 4);
    status8[ST_FIRST_MISSING] = ~status8[ST_FIRST_MISSING]; /* 0xFFFFFFFF when nothing is missing */
    memcpy(r.h_root32, r.h_status + ST_WORDS, 32);
    memcpy(r.h_mutated, r.h_status + ST_WORDS + 8, sizeof(int));
    const size_t total = status8[ST_SPK_BYTES];
    if (!r.h_spk_cat) return BCP_OK; /* the spent-coin arrays were left on the device */
    if (total > r.spk_cap) {
        bcp_set_error("bcp_gpu_utxo_wait: the spent scripts of block %u need %zu bytes, the caller's buffer holds %zu", seq, total, r.spk_cap);
        return BCP_ERR_INVALID;
    }
    if (total > r.spk_copied) { /* rare: longer scripts than the first copy allowed for; slots and scripts are still in place */
        std::lock_guard<std::mutex> lk(u->mu);
        BCP_CUDA(cudaMemcpy(r.h_spk_cat + r.spk_copied, r.d + r.o_spk + r.spk_copied, total - r.spk_copied, cudaMemcpyDeviceToHost));
        r.spk_copied = total;
    }
    return BCP_OK;
}

// The script checks of applied block `seq` straight from the arrays bcp_gpu_utxo_apply_block left on the table's GPU: the
// raw block, its offsets, and the spent (amount, scriptPubKey) of every input.  On the table's own GPU they are read in
// place; on another GPU one peer-to-peer copy over NVLink brings them over — either way the block does not cross the
// host's memory a second time and the spent-coin arrays never visit the host at all.
BCP_API int bcp_gpu_utxo_block_scripts(bcp_utxo *u, uint32_t seq, int device, const uint32_t *tx_off, const uint32_t *in_tx,
                                       const uint32_t *in_n, uint32_t flags, uint8_t *script_err, int *all_ok, uint32_t status8[8]) {
    if (!u || !status8) return BCP_ERR_INVALID;
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (!r.used || r.seq != seq) { bcp_set_error("bcp_gpu_utxo_block_scripts: block %u is not pending", seq); return BCP_ERR_INVALID; }
    if (all_ok) *all_ok = 1;
    if (r.nin && (!tx_off || !in_tx || !in_n || !script_err)) { bcp_set_error("bcp_gpu_utxo_block_scripts: NULL buffer"); return BCP_ERR_INVALID; }
    if (bcp_block_check_flags(flags, "bcp_gpu_utxo_block_scripts") != BCP_OK) return BCP_ERR_INVALID;
    BCP_CUDA(cudaSetDevice(u->device));
    BCP_CUDA(cudaEventSynchronize(r.done));
    memcpy(status8, r.h_status, ST_WORDS * 4);
    status8[ST_FIRST_MISSING] = ~status8[ST_FIRST_MISSING];
    memcpy(r.h_root32, r.h_status + ST_WORDS, 32);
    memcpy(r.h_mutated, r.h_status + ST_WORDS + 8, sizeof(int));
    if (status8[ST_SPK_OVERFLOW]) {
        bcp_set_error("bcp_gpu_utxo_block_scripts: the spent scripts of block %u (%u bytes) do not fit the device buffer", seq, status8[ST_SPK_BYTES]);
        return BCP_ERR_INVALID;
    }
    if (r.nin == 0 || status8[ST_MISSING]) return BCP_OK; /* nothing to check / the block is refused anyway */
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const size_t nin = r.nin, ntx = r.ntx;
    const size_t in_bytes = r.o_spk + ((size_t)status8[ST_SPK_BYTES] + 255) / 256 * 256; /* the part of the input area in use */
    const size_t o_inn = (in_bytes + 255) / 256 * 256, used = o_inn + (nin * 4 + 4 + 255) / 256 * 256;
    int rc = bcp_ctx_reserve_stage(c, used + bcp_block_scripts_resident_bytes(ntx, nin) + 1024);
    if (rc != BCP_OK) return rc;
    const uint8_t *base = r.d;
    if (c->device != u->device) { /* bring the input area over: one device-to-device copy, NVLink when peer access is on */
        if (!u->peer_tried[c->device]) { /* once per pair, both directions; "already enabled" and "not supported" are fine */
            std::lock_guard<std::mutex> lk2(u->mu);
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c->device, u->device) == cudaSuccess && can) cudaDeviceEnablePeerAccess(u->device, 0);
            cudaSetDevice(u->device);
            if (cudaDeviceCanAccessPeer(&can, u->device, c->device) == cudaSuccess && can) cudaDeviceEnablePeerAccess(c->device, 0);
            cudaGetLastError();
            BCP_CUDA(cudaSetDevice(c->device));
            u->peer_tried[c->device] = true;
        }
        BCP_CUDA(cudaMemcpyPeerAsync(c->stage, c->device, r.d, u->device, in_bytes, st));
        base = c->stage;
    }
    uint8_t *d_inn = c->stage + o_inn; /* the one array the table did not need */
    BCP_CUDA(cudaMemcpyAsync(d_inn, in_n, nin * 4, cudaMemcpyHostToDevice, st));
    return bcp_block_scripts_resident(c, base, reinterpret_cast<const uint32_t *>(base + r.o_txoff), tx_off, ntx,
                                      reinterpret_cast<const uint32_t *>(base + r.o_inoff), reinterpret_cast<const uint32_t *>(base + r.o_intx),
                                      reinterpret_cast<const uint32_t *>(d_inn), in_tx, nin, base + r.o_spk,
                                      reinterpret_cast<const uint32_t *>(base + r.o_spkoff), reinterpret_cast<const int64_t *>(base + r.o_amt),
                                      used, script_err, all_ok, st);
}

// The spent-coin arrays of applied block seq, for a caller that left them on the device at apply time and needs them on
// the host after all (inputs the fast path hands back to the host interpreter).  Synchronous.
BCP_API int bcp_gpu_utxo_fetch_spent(bcp_utxo *u, uint32_t seq, uint8_t *spk_cat, size_t spk_cap, uint32_t *spk_off, int64_t *amounts) {
    if (!u || !spk_cat || !spk_off || !amounts) return BCP_ERR_INVALID;
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (!r.used || r.seq != seq) { bcp_set_error("bcp_gpu_utxo_fetch_spent: block %u is not pending", seq); return BCP_ERR_INVALID; }
    BCP_CUDA(cudaSetDevice(u->device));
    BCP_CUDA(cudaEventSynchronize(r.done));
    const size_t total = r.h_status[ST_SPK_BYTES];
    if (total > spk_cap || r.h_status[ST_SPK_OVERFLOW]) {
        bcp_set_error("bcp_gpu_utxo_fetch_spent: the spent scripts of block %u need %zu bytes, the buffer holds %zu", seq, total, spk_cap);
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(u->mu);
    BCP_CUDA(cudaMemcpy(amounts, r.d + r.o_amt, r.nin * 8, cudaMemcpyDeviceToHost));
    BCP_CUDA(cudaMemcpy(spk_off, r.d + r.o_spkoff, (r.nin + 1) * 4, cudaMemcpyDeviceToHost));
    if (total) BCP_CUDA(cudaMemcpy(spk_cat, r.d + r.o_spk, total, cudaMemcpyDeviceToHost));
    return BCP_OK;
}

BCP_API int bcp_gpu_utxo_confirm(bcp_utxo *u, uint32_t seq) {
    if (!u) return BCP_ERR_INVALID;
    std::lock_guard<std::mutex> lk(u->mu);
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (!r.used || r.seq != seq) { bcp_set_error("bcp_gpu_utxo_confirm: block %u is not pending", seq); return BCP_ERR_INVALID; }
    r.used = false;
    if (seq > u->confirmed) u->confirmed = seq; /* tombstones left by blocks up to here may be reused from now on */
    if (r.nin) {
        BCP_CUDA(cudaSetDevice(u->device));
        k_utxo_reclaim<<<(unsigned)((r.nin + 127) / 128), 128, 0, u->st>>>(u->T, u->confirmed, reinterpret_cast<uint32_t *>(r.d + r.o_rec_in), r.nin);
        BCP_CUDA(cudaGetLastError());
        BCP_CUDA(cudaEventRecord(r.reclaimed, u->st));
        r.reclaim_pending = true;
    }
    return BCP_OK;
}

BCP_API int bcp_gpu_utxo_rollback(bcp_utxo *u, uint32_t seq) {
    if (!u) return BCP_ERR_INVALID;
    std::lock_guard<std::mutex> lk(u->mu);
    ring_slot &r = u->ring[seq % UTXO_RING];
    if (!r.used || r.seq != seq) { bcp_set_error("bcp_gpu_utxo_rollback: block %u is not pending", seq); return BCP_ERR_INVALID; }
    for (int k = 0; k < UTXO_RING; k++)
        if (u->ring[k].used && u->ring[k].seq > seq) {
            bcp_set_error("bcp_gpu_utxo_rollback: block %u was applied after block %u and must be rolled back first", u->ring[k].seq, seq);
            return BCP_ERR_INVALID;
        }
    BCP_CUDA(cudaSetDevice(u->device));
    const size_t n = r.nin > r.nout ? r.nin : r.nout;
    if (n)
        k_utxo_undo<<<(unsigned)((n + 127) / 128), 128, 0, u->st>>>(u->T, seq, reinterpret_cast<uint32_t *>(r.d + r.o_rec_in), r.nin,
                                                                   reinterpret_cast<uint32_t *>(r.d + r.o_rec_out), r.nout);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaStreamSynchronize(u->st));
    r.used = false;
    return BCP_OK;
}

} // extern "C"
