// include/compat/bcp_chain.h — host-side block connector over the C ABI: the part of ConnectBlock
// (src/validation.cpp:1868-2188) and of the import loop (LoadExternalBlockFile, src/validation.cpp:4461-4560) that
// feeds the data-parallel hot path, with an in-memory UTXO map standing in for CCoinsViewCache (the node's
// LevelDB-backed chainstate is out of scope).  It exists to drive BASELINE.json configs[4]
// ("100-block synthetic chain reindex") end to end and is row N2 of SURVEY.md §8(f) on the host side.
//
// Per block, in the reference's order:
//   header:  CheckEquihashSolution + header hash (src/validation.cpp:3202-3220, 1139-1149)     -> GPU (bcp_gpu_equihash_verify_batch)
//   CheckBlock: Merkle root vs header, then the `mutated` flag (src/validation.cpp:3236-3252)  -> GPU (txids + Merkle)
//   for each tx, for each input: the spent coin must exist and be unspent — including coins created earlier in the same
//            block (UpdateCoins after every tx, src/validation.cpp:2095)                        -> host, hash map
//   script checks (CScriptCheck, src/validation.cpp:1350-1359) for every input                 -> GPUs, sharded by input index range,
//            one bcp_gpu_block_scripts_p2pkh call per device, verdicts AND-ed here (src/checkqueue.h:77)
//   commit: spend inputs, add outputs
// Context-free structure rules of CheckBlock / CheckCoinbase / CheckRegularTransaction (src/validation.cpp:476-570,
// 3255-3330) that need nothing but the framing are applied by the scan: bad-cb-missing, bad-cb-length, bad-tx-coinbase,
// bad-txns-vin-empty, bad-txns-vout-empty, bad-txns-oversize, bad-txns-vout-negative / -toolarge /
// -txouttotal-toolarge, bad-txns-prevout-null; and ConnectBlock's BIP30 rule (bad-txns-BIP30).
// Not done (outside the hot path): fees/maturity (Consensus::CheckTxInputs), sigop limits, block size limit, BIP68
// locks, difficulty, undo data.  Inputs that are not of the P2PKH shape are reported (reason "script-not-in-fast-path"):
// a node would run them through its interpreter.
#ifndef BCP_COMPAT_CHAIN_H
#define BCP_COMPAT_CHAIN_H

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../bcp_gpu.h"
#include "bcp_other_scripts.h"

namespace bcp {

struct OutPoint {
    uint8_t b[36];
    bool operator==(const OutPoint &o) const { return memcmp(b, o.b, 36) == 0; }
};
struct OutPointHash {
    size_t operator()(const OutPoint &o) const {
        uint64_t h;
        uint32_t v;
        memcpy(&h, o.b, 8); // txids are SHA-256d outputs: any 8 bytes are a fine hash
        memcpy(&v, o.b + 32, 4);
        return (size_t)(h ^ (0x9E3779B97F4A7C15ull * v));
    }
};
// Open-addressing UTXO table: linear probing, deletion by backward shift (no tombstones, so the table never needs
// a clean-up rehash and probe sequences stay short however many blocks have been connected), coins stored inline in
// 128-byte slots = two adjacent cache lines that prefetch() pulls in together.  scriptPubKeys up to 64 bytes live
// in the slot (P2PKH is 25), longer ones in a side vector.  One table serves lookups, in-block creations and
// spends; a per-block journal undoes everything when the block is rejected, so no second map is needed.
struct alignas(128) CoinSlot {
    // first cache line: everything a probe looks at
    uint8_t state;      // 0 empty, 1 live
    uint8_t pad[3];
    uint32_t spk_len;
    OutPoint key;       // bytes 8..43
    uint32_t stamp;     // which block created the coin (0: seeded) — tells an in-block twin from a coin that was there before
    int64_t amount;     // 48
    std::vector<uint8_t> *big; // 56; only for scripts longer than the inline space
    // second cache line
    uint8_t spk[64];
};
static_assert(sizeof(CoinSlot) == 128, "CoinSlot is two cache lines");
class CoinTable {
public:
    explicit CoinTable(size_t cap_pow2 = 1 << 17) : slots(cap_pow2), mask(cap_pow2 - 1) {
        for (auto &s : slots) { s.state = 0; s.big = nullptr; }
    }
    ~CoinTable() { for (auto &s : slots) delete s.big; }
    CoinTable(const CoinTable &) = delete;
    CoinTable &operator=(const CoinTable &) = delete;
    size_t size() const { return live; }
    // the table is far larger than the CPU caches and every probe is a random access: callers hint the slots of the
    // next few outpoints while they work on the current one
    void prefetch(const OutPoint &k) const { prefetch_hash(OutPointHash()(k)); }
    void prefetch_hash(size_t h) const {
        const char *p = reinterpret_cast<const char *>(&slots[h & mask]);
        __builtin_prefetch(p);
        __builtin_prefetch(p + 64);
        __builtin_prefetch(p + 128); // probe line of the next slot: erase() and colliding inserts look at it
    }
    CoinSlot *find(const OutPoint &k) {
        size_t i = OutPointHash()(k) & mask;
        for (;;) {
            CoinSlot &s = slots[i];
            if (s.state == 0) return nullptr;
            if (s.key == k) return &s;
            i = (i + 1) & mask;
        }
    }
    CoinSlot *insert(const OutPoint &k, const uint8_t *spk, size_t len, int64_t amount, uint32_t stamp = 0) {
        if ((live + 1) * 2 > slots.size()) grow();
        size_t i = OutPointHash()(k) & mask;
        for (;;) {
            CoinSlot &s = slots[i];
            if (s.state == 0) break;
            if (s.key == k) { fill(s, spk, len, amount); return &s; } // overwrite (BIP30-style duplicates)
            i = (i + 1) & mask;
        }
        CoinSlot &d = slots[i];
        d.key = k;
        d.state = 1;
        d.big = nullptr;
        d.stamp = stamp;
        fill(d, spk, len, amount);
        live++;
        return &d;
    }
    // Removes the coin and closes the gap: every following entry of the probe run moves back one hole if its home
    // position allows it, so find() can keep stopping at the first empty slot.  Invalidates CoinSlot pointers.
    void erase(CoinSlot *s) {
        delete s->big;
        size_t i = (size_t)(s - slots.data()), j = i;
        for (;;) {
            j = (j + 1) & mask;
            if (slots[j].state == 0) break;
            const size_t h = OutPointHash()(slots[j].key) & mask;
            if (i <= j ? (i < h && h <= j) : (i < h || h <= j)) continue; // home lies after the hole: must stay
            slots[i] = slots[j];
            i = j;
        }
        slots[i].state = 0;
        slots[i].big = nullptr;
        live--;
    }
    static const uint8_t *script(const CoinSlot &s) { return s.big ? s.big->data() : s.spk; }

private:
    void fill(CoinSlot &s, const uint8_t *spk, size_t len, int64_t amount) {
        s.amount = amount;
        s.spk_len = (uint32_t)len;
        if (len <= sizeof s.spk) {
            memcpy(s.spk, spk, len);
            delete s.big;
            s.big = nullptr;
        } else {
            if (!s.big) s.big = new std::vector<uint8_t>();
            s.big->assign(spk, spk + len);
        }
    }
    void grow() {
        std::vector<CoinSlot> old;
        old.swap(slots);
        slots.resize(old.size() * 2);
        mask = slots.size() - 1;
        for (auto &s : slots) { s.state = 0; s.big = nullptr; }
        live = 0;
        for (auto &o : old)
            if (o.state == 1) {
                CoinSlot *n = insert(o.key, script(o), o.spk_len, o.amount, o.stamp);
                (void)n;
                delete o.big;
                o.big = nullptr;
            }
    }
    std::vector<CoinSlot> slots;
    size_t mask, live = 0;
};

// A fixed team of host threads for the two table passes of a block: Run(f) calls f(worker) on every worker (the
// caller is worker 0) and returns when all are done.
class WorkerTeam {
public:
    explicit WorkerTeam(int n) : n_(n < 1 ? 1 : n) {
        for (int w = 1; w < n_; w++) th_.emplace_back([this, w] { Loop(w); });
    }
    ~WorkerTeam() {
        {
            std::lock_guard<std::mutex> l(mu_);
            quit_ = true;
            gen_++;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    int size() const { return n_; }
    void Run(const std::function<void(int)> &f) {
        if (n_ == 1) { f(0); return; }
        {
            std::lock_guard<std::mutex> l(mu_);
            job_ = &f;
            pending_ = n_ - 1;
            gen_++;
        }
        cv_.notify_all();
        f(0);
        std::unique_lock<std::mutex> l(mu_);
        done_.wait(l, [this] { return pending_ == 0; });
        job_ = nullptr;
    }

private:
    void Loop(int w) {
        unsigned long seen = 0;
        for (;;) {
            const std::function<void(int)> *f;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return gen_ != seen; });
                seen = gen_;
                if (quit_) return;
                f = job_;
            }
            (*f)(w);
            {
                std::lock_guard<std::mutex> l(mu_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const std::function<void(int)> *job_ = nullptr;
    int pending_ = 0;
    unsigned long gen_ = 0;
    bool quit_ = false;
};

// Growable array WITHOUT value initialisation: the framing arrays are sized for the worst case a block of that length
// could hold (an output needs 9 bytes: 900 k entries for 8 MB) while a real block uses a few percent of that; a
// std::vector would zero-fill, and so page in, all of it.
template <class T> struct RawBuf {
    T *p = nullptr;
    size_t cap = 0;
    RawBuf() {}
    RawBuf(const RawBuf &) = delete;
    RawBuf &operator=(const RawBuf &) = delete;
    ~RawBuf() { delete[] p; }
    void need(size_t n) {
        if (cap >= n) return;
        delete[] p;
        p = new T[n];
        cap = n;
    }
    T *data() { return p; }
    const T *data() const { return p; }
    T &operator[](size_t i) { return p[i]; }
    const T &operator[](size_t i) const { return p[i]; }
    void swap(RawBuf &o) { std::swap(p, o.p); std::swap(cap, o.cap); }
};

class ChainValidator {
public:
    // walk_threads: host threads for the UTXO passes (rounded down to a power of two, at most 16).  The table is split
    // into that many shards by the top bits of the outpoint hash; during a pass every shard is touched by exactly one
    // thread, so the passes need no locks.
    ChainValidator(const std::vector<int> &devices, uint32_t script_flags, unsigned eh_n, unsigned eh_k, int walk_threads = 8)
        : devs(devices), flags(script_flags), ehN(eh_n), ehK(eh_k), shardBits(0) {
        while ((2 << shardBits) <= walk_threads && shardBits < 4) shardBits++;
        for (int i = 0; i < (1 << shardBits); i++) shards.emplace_back(new CoinTable());
        team.reset(new WorkerTeam(1 << shardBits));
    }
    ~ChainValidator() {
        for (Scan &a : ahead)
            if (a.th.joinable()) a.th.join();
        if (getenv("BCP_CHAIN_DEBUG")) {
            double f = 0, d = 0;
            for (Scan &a : ahead) { f += a.frameSecs; d += a.deviceSecs; }
            for (Work &w : work) { f += w.cur.frameSecs; d += w.cur.deviceSecs; }
            fprintf(stderr, "[bcp_chain] scan threads: framing %.1f ms, device pre-checks %.1f ms in total\n", f * 1e3, d * 1e3);
        }
        for (Work &w : work)
            for (auto &t : w.th)
                if (t.joinable()) t.join();
    }

    bool Init(std::string &err) {
        if (devs.empty()) { err = "no device"; return false; }
        // Lanes (independent contexts on one GPU): the two blocks ConnectFile keeps in flight send their script checks
        // through lanes 0 and 1 of each device, so that block b+1's copies and extraction run under block b's ECDSA
        // kernels; header checks and txids of the blocks being scanned ahead go through lane 2 of the first device.
        // A `devices` entry may itself carry a lane base in its upper bits (BCP_DEVICE_LANE(ordinal, base)): that entry then
        // uses lanes base and base+1 of the GPU, so one GPU can stand in for two devices (tests on a one-GPU box).
        for (int d : devs)
            for (int lane = 0; lane < 2; lane++)
                if (bcp_gpu_init(BCP_DEVICE_LANE(d & 0xFF, (d >> 8) + lane)) != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        devAux = BCP_DEVICE_LANE(devs[0] & 0xFF, 2);
        if (bcp_gpu_init(devAux) != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        return true;
    }
    void AddCoin(const uint8_t txid[32], uint32_t vout, const uint8_t *spk, size_t len, int64_t amount) {
        OutPoint op;
        memcpy(op.b, txid, 32);
        memcpy(op.b + 32, &vout, 4);
        ShardOf(op).insert(op, spk, len, amount);
    }
    size_t UtxoSize() const {
        size_t n = 0;
        for (const auto &t : shards) n += t->size();
        return n;
    }
    size_t InputsChecked() const { return nInputs; }
    //! Who evaluates the inputs the device fast path does not (bcp_other_scripts.h).  Unset: such a block is refused with
    //! "script-not-in-fast-path".  Called on the connecting thread, once per block at most.
    void SetOtherScripts(OtherScripts f) { other = std::move(f); }

    // seconds since construction: [0] header (Equihash + hash), [1] framing scan on the connecting thread (0 when the
    // scan ran ahead), [2] waiting for the txids after the lookups, [3] the UTXO passes, [4] waiting for script checks
    double phase[5] = {0, 0, 0, 0, 0};
    double walk[4] = {0, 0, 0, 0}; // inside [3]: pass A, staging for the script call, pass B, pass C

    // Returns true and updates the UTXO set, or false with `reason` (the reference's reject strings where one exists).
    // `next` / `next_len` (optional): the block that will be connected after this one; its framing scan then runs on a
    // helper thread while this block is being connected.  Those bytes must stay valid until that block is connected.
    bool ConnectBlock(const uint8_t *blk, size_t len, std::string &reason, const uint8_t *next = nullptr, size_t next_len = 0) {
        Work &w = work[0];
        const Hint h(next, next_len);
        if (!Begin(w, blk, len, reason, &h, next ? 1 : 0)) return false;
        if (!Collect(w, reason)) { Rollback(w); return false; }
        Commit(w);
        return true;
    }
    //! Optional: start the framing scan of the block that will be connected next on a helper thread, before the first
    //! ConnectBlock of a run (afterwards the `next` argument of ConnectBlock keeps the scans one block ahead).
    void ScanAhead(const uint8_t *blk, size_t len) {
        for (Scan &a : ahead)
            if (a.blk == blk && a.len == len && (a.th.joinable() || a.valid)) return; // already on its way
        Scan &a = ahead[(nextSlot++) % NAHEAD];
        if (a.th.joinable()) a.th.join();
        a.valid = false;
        a.blk = blk;
        a.len = len;
        const int dev = devAux, sb = shardBits;
        const unsigned n_ = ehN, k_ = ehK;
        Scan *pa = &a;
        a.th = std::thread([pa, blk, len, sb, dev, n_, k_] {
            pa->Frame(blk, len, sb);
            pa->Device(dev, n_, k_, sb);
        });
    }
    const uint8_t *LastBlockHash() const { return lastHash; }
    const std::vector<uint8_t> &LastScriptErrors() const { return work[lastWork].errs; }

    // blk*.dat framing (src/validation.cpp:4486-4516): [magic 4][size u32 LE][block] ... ; returns blocks connected,
    // stops at the first failure (reason set, failed_index = its ordinal).  Two blocks are in flight: while the devices
    // run block b's script checks, the host already scans, looks up and applies block b+1 on top of b's (still
    // unconfirmed) result; if b turns out invalid, b+1 is undone first, then b.
    long ConnectFile(const uint8_t *data, size_t len, const uint8_t magic[4], std::string &reason, long &failed_index) {
        size_t pos = 0;
        long n = 0;
        failed_index = -1;
        Work *prev = nullptr;
        while (pos + 8 <= len) {
            if (memcmp(data + pos, magic, 4) != 0) { pos++; continue; } // resynchronise on the magic like :4489-4498
            uint32_t sz;
            memcpy(&sz, data + pos + 4, 4);
            if (pos + 8 + sz > len) break;
            // the framing scans and device pre-checks of the next few blocks run on helper threads meanwhile
            Hint hints[NAHEAD];
            size_t nh = 0, nx = pos + 8 + (size_t)sz;
            while (nh < NAHEAD && nx + 8 <= len && memcmp(data + nx, magic, 4) == 0) {
                uint32_t nsz;
                memcpy(&nsz, data + nx + 4, 4);
                if (nx + 8 + nsz > len) break;
                hints[nh++] = Hint(data + nx + 8, nsz);
                nx += 8 + (size_t)nsz;
            }
            Work &w = work[n & 1];
            std::string r;
            const bool begun = Begin(w, data + pos + 8, sz, r, hints, nh); // on failure this block is already undone
            if (prev) {
                std::string rp;
                if (!Collect(*prev, rp)) { // the block before is invalid: this one goes first, then that one
                    if (begun) { Abandon(w); Rollback(w); }
                    Rollback(*prev);
                    reason = rp;
                    failed_index = n - 1;
                    return n - 1;
                }
                Commit(*prev);
                prev = nullptr;
            }
            if (!begun) { reason = r; failed_index = n; return n; }
            prev = &w;
            pos += 8 + (size_t)sz;
            n++;
        }
        if (prev) {
            if (!Collect(*prev, reason)) { Rollback(*prev); failed_index = n - 1; return n - 1; }
            Commit(*prev);
        }
        return n;
    }

private:
    struct Scan { // result of the framing scan of one block
        RawBuf<uint32_t> tx_off, tx_out_off, in_off, in_tx, in_n, out_off, out_tx, out_n;
        std::vector<uint32_t> by_shard, shard_start; // input numbers grouped by table shard (block order inside a group)
        RawBuf<uint64_t> in_key;                     // OutPointHash of every input's outpoint, computed by the scan
        std::vector<uint32_t> fillTmp;
        std::vector<uint32_t> out_by_shard, out_shard_start; // output numbers grouped by shard (needs the txids: Device())
        RawBuf<uint64_t> out_key;
        long ntx = 0, nin = -1, nout = 0;
        const uint8_t *blk = nullptr;
        size_t len = 0;
        bool valid = false, deviceDone = false;
        std::thread th;
        // device pre-checks, which need nothing but the block's own bytes: header (Equihash + hash), txids + Merkle root
        std::vector<uint8_t> txids;
        uint8_t root[32], hash[32];
        int mutated = 0, rcTx = BCP_OK, rcEq = BCP_OK;
        uint8_t eqOk = 1;
        std::string errTx, errEq;
        double frameSecs = 0, deviceSecs = 0; // time this slot's thread spent in each part (diagnostics)
        const char *structErr = nullptr;      // first structure rule of CheckBlock the block breaks (reference reject string)
        void Device(int dev, unsigned ehN, unsigned ehK, int shardBits) {
            const auto td0 = std::chrono::steady_clock::now();
            const size_t soln = ehN ? ((size_t)1 << ehK) * (ehN / (ehK + 1) + 1) / 8 : 0;
            const size_t hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
            rcEq = rcTx = BCP_OK;
            eqOk = 1;
            if (ehN && len >= hdr_len) {
                rcEq = bcp_gpu_equihash_verify_batch(dev, ehN, ehK, blk, hdr_len, hdr_len, 1, &eqOk, hash);
                if (rcEq != BCP_OK) errEq = bcp_gpu_last_error();
            }
            if (nin >= 0 && ntx >= 1) {
                txids.resize((size_t)ntx * 32);
                rcTx = bcp_gpu_sha256d_txids_merkle(dev, blk, len, tx_off.data(), (size_t)ntx, txids.data(), root, &mutated);
                if (rcTx != BCP_OK) errTx = bcp_gpu_last_error();
            }
            // the new coins' hash keys need the txids; with them the outputs are grouped by shard here, off the
            // connecting thread, exactly like the inputs were by Frame()
            const size_t S = (size_t)1 << shardBits;
            out_shard_start.assign(S + 1, 0);
            if (rcTx == BCP_OK && nin >= 0 && nout > 0) {
                out_key.need((size_t)nout);
                out_by_shard.resize((size_t)nout);
                const int sh = 64 - shardBits;
                for (long o = 0; o < nout; o++) {
                    uint64_t h8;
                    memcpy(&h8, &txids[(size_t)out_tx[(size_t)o] * 32], 8);
                    const uint64_t k = h8 ^ (0x9E3779B97F4A7C15ull * out_n[(size_t)o]);
                    out_key[(size_t)o] = k;
                    out_shard_start[(shardBits ? (size_t)(k >> sh) : 0) + 1]++;
                }
                for (size_t k = 0; k < S; k++) out_shard_start[k + 1] += out_shard_start[k];
                fillTmp.assign(out_shard_start.begin(), out_shard_start.end() - 1);
                for (long o = 0; o < nout; o++) out_by_shard[fillTmp[shardBits ? (size_t)(out_key[(size_t)o] >> sh) : 0]++] = (uint32_t)o;
            }
            deviceDone = true;
            deviceSecs += std::chrono::duration<double>(std::chrono::steady_clock::now() - td0).count();
        }
        void Frame(const uint8_t *b, size_t l, int shardBits) {
            const auto tf0 = std::chrono::steady_clock::now();
            deviceDone = false;
            const size_t max_tx = l / 60 + 2, max_in = l / 41 + 2, max_out = l / 9 + 2;
            tx_off.need(max_tx + 1); tx_out_off.need(max_tx + 1);
            in_off.need(max_in); in_tx.need(max_in); in_n.need(max_in); in_key.need(max_in);
            out_off.need(max_out); out_tx.need(max_out); out_n.need(max_out);
            ntx = nout = 0;
            // the scan applies CheckBlock's context-free transaction rules while it has the bytes in the cache
            int rule = 0;
            long ruleTx = -1;
            nin = bcp_block_index_rules(b, l, 0, tx_off.data(), tx_out_off.data(), max_tx, in_off.data(), in_tx.data(), in_n.data(),
                                        max_in, &ntx, out_off.data(), out_tx.data(), out_n.data(), max_out, &nout, in_key.data(), &rule,
                                        &ruleTx);
            structErr = (nin >= 0 && rule) ? bcp_block_rule_name(rule) : nullptr;
            // counting sort of the inputs by shard, here (on the scan-ahead thread) so that each worker of pass A only
            // ever touches its own inputs
            const size_t S = (size_t)1 << shardBits;
            shard_start.assign(S + 1, 0);
            if (nin > 0) {
                by_shard.resize((size_t)nin);
                const int sh = 64 - shardBits;
                if (shardBits)
                    for (long i = 0; i < nin; i++) shard_start[(size_t)(in_key[(size_t)i] >> sh) + 1]++;
                else
                    shard_start[1] = (uint32_t)nin;
                for (size_t k = 0; k < S; k++) shard_start[k + 1] += shard_start[k];
                fillTmp.assign(shard_start.begin(), shard_start.end() - 1);
                for (long i = 0; i < nin; i++) by_shard[fillTmp[shardBits ? (size_t)(in_key[(size_t)i] >> sh) : 0]++] = (uint32_t)i;
            }
            blk = b;
            len = l;
            valid = true;
            frameSecs += std::chrono::duration<double>(std::chrono::steady_clock::now() - tf0).count();
        }
    };
    static void SwapScan(Scan &a, Scan &b) {
        a.tx_off.swap(b.tx_off); a.tx_out_off.swap(b.tx_out_off); a.in_off.swap(b.in_off); a.in_tx.swap(b.in_tx); a.in_n.swap(b.in_n);
        a.out_off.swap(b.out_off); a.out_tx.swap(b.out_tx); a.out_n.swap(b.out_n);
        a.by_shard.swap(b.by_shard); a.shard_start.swap(b.shard_start); a.in_key.swap(b.in_key);
        a.out_by_shard.swap(b.out_by_shard); a.out_shard_start.swap(b.out_shard_start); a.out_key.swap(b.out_key);
        a.txids.swap(b.txids); a.errTx.swap(b.errTx); a.errEq.swap(b.errEq);
        std::swap(a.root, b.root); std::swap(a.hash, b.hash); std::swap(a.mutated, b.mutated); std::swap(a.rcTx, b.rcTx);
        std::swap(a.rcEq, b.rcEq); std::swap(a.eqOk, b.eqOk); std::swap(a.deviceDone, b.deviceDone);
        std::swap(a.ntx, b.ntx); std::swap(a.nin, b.nin); std::swap(a.nout, b.nout); std::swap(a.blk, b.blk); std::swap(a.len, b.len);
        std::swap(a.valid, b.valid); std::swap(a.structErr, b.structErr);
        std::swap(a.frameSecs, b.frameSecs); std::swap(a.deviceSecs, b.deviceSecs);
    }

    // what pass A / C found for one input; 128 bytes so that two threads never write the same cache line
    enum : uint8_t { ST_MISS = 0, ST_TAKEN = 1, ST_BIG = 2, ST_INBLOCK = 3, ST_INBLOCK_BIG = 4 };
    static const size_t STAGE_SPK = 64;
    struct alignas(128) Staged {
        int64_t amount;
        uint32_t len, big;
        uint8_t state;
        uint8_t spk[STAGE_SPK];
    };
    // Everything that belongs to one block on its way through: its framing, txids, the staging records of its inputs
    // (also the undo log), the arrays of the script-check call and the helper threads running that call.
    struct Work {
        Scan cur;
        const uint8_t *blk = nullptr;
        size_t len = 0;
        std::vector<uint8_t> spk_cat, errs;
        std::vector<uint32_t> spk_off;
        std::vector<int64_t> amounts;
        std::vector<Staged> stage;
        std::vector<std::vector<uint8_t>> bigStore;
        std::vector<uint8_t> dupOut;           // outputs NOT inserted: 1 = a coin with that outpoint was there before the block
                                               // (BIP30), 2 = an identical transaction earlier in this block created it
        uint32_t stamp = 0;                    // this block's mark on the coins it creates
        size_t spentUpTo = 0, createdUpTo = 0; // how far the block has been applied to the table (for Rollback)
        std::vector<std::thread> th;
        std::vector<int> rc;
        std::vector<std::string> msg;
        uint8_t hash[32];
    };

    int ShardIndex(const OutPoint &op) const { return shardBits ? (int)(OutPointHash()(op) >> (64 - shardBits)) : 0; }
    CoinTable &ShardOf(const OutPoint &op) { return *shards[(size_t)ShardIndex(op)]; }
    static double Secs(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double>(b - a).count();
    }

    // Header, framing, txids + Merkle, the table passes, and the launch of the script checks.  true: the block is
    // applied to the table and its script checks are running (Collect gives the verdict); false: rejected, nothing of it
    // is left in the table.
    typedef std::pair<const uint8_t *, size_t> Hint;
    bool Begin(Work &w, const uint8_t *blk, size_t len, std::string &reason, const Hint *hints, size_t nhints) {
        auto now = [] { return std::chrono::steady_clock::now(); };
        w.blk = blk;
        w.len = len;
        w.spentUpTo = w.createdUpTo = 0;
        auto t0 = now();
        const size_t soln = ehN ? ((size_t)1 << ehK) * (ehN / (ehK + 1) + 1) / 8 : 0;
        const size_t hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
        if (ehN && len < hdr_len) { reason = "bad-blk-length"; return false; }
        // Framing and the device pre-checks (header, txids + Merkle) come from a scan-ahead slot when one holds this very
        // block; otherwise the framing is done now and the device part runs on a helper thread under pass A.
        Scan *slot = nullptr;
        for (Scan &a : ahead)
            if (a.blk == blk && a.len == len && (a.th.joinable() || a.valid)) slot = &a;
        std::thread thDev;
        if (slot) {
            if (slot->th.joinable()) slot->th.join();
            SwapScan(w.cur, *slot);
            slot->valid = false;
            slot->blk = nullptr;
        } else {
            w.cur.Frame(blk, len, shardBits);
            Scan *pc = &w.cur;
            const int dev = devAux;
            const unsigned n_ = ehN, k_ = ehK;
            const int sb_ = shardBits;
            thDev = std::thread([pc, dev, n_, k_, sb_] { pc->Device(dev, n_, k_, sb_); });
        }
        for (size_t k = 0; k < nhints; k++) ScanAhead(hints[k].first, hints[k].second);
        const Scan &cur = w.cur;
        const long ntx = cur.ntx, nin = cur.nin, nout = cur.nout;
        auto t2 = now();
        phase[1] += Secs(t0, t2);
        if (nin < 0 || ntx < 1) {
            if (thDev.joinable()) thDev.join();
            if (cur.rcEq == BCP_OK && !cur.eqOk) reason = "invalid-solution"; // the header is looked at before the body
            else reason = "bad-blk-framing";
            return false;
        }
        // Pass A, all shards in parallel: every input whose coin is in the table as it was BEFORE this block is staged
        // (amount + scriptPubKey for the script checks) and removed.  What is not found is either created earlier in
        // this block or missing; pass C settles that.
        if (w.stage.size() < (size_t)nin) w.stage.resize((size_t)nin);
        w.bigStore.clear();
        w.spentUpTo = (size_t)nin;
        team->Run([&](int s_) { PassA(w, s_); });
        walk[0] += Secs(t2, now());
        // what pass A left open, counted by the team over contiguous slices of the staging records
        size_t unresolved = 0, nBig = 0;
        {
            const int P = team->size();
            std::vector<size_t> miss((size_t)P, 0), big((size_t)P, 0);
            team->Run([&](int s_) {
                const size_t lo = (size_t)nin * (size_t)s_ / (size_t)P, hi = (size_t)nin * ((size_t)s_ + 1) / (size_t)P;
                size_t m = 0, b = 0;
                for (size_t i = lo; i < hi; i++) {
                    m += w.stage[i].state == ST_MISS;
                    b += w.stage[i].state == ST_BIG;
                }
                miss[(size_t)s_] = m;
                big[(size_t)s_] = b;
            });
            for (int k = 0; k < P; k++) { unresolved += miss[(size_t)k]; nBig += big[(size_t)k]; }
        }
        for (size_t i = 0; nBig && i < (size_t)nin; i++) { // rare: scripts too long for the staging slot, done on one thread
            Staged &st = w.stage[i];
            if (st.state != ST_BIG) continue;
            OutPoint op;
            memcpy(op.b, blk + cur.in_off[i], 36);
            CoinTable &T = ShardOf(op);
            CoinSlot *c = T.find(op);
            if (!c) { // spent a moment ago by another input of this block
                st.state = ST_MISS;
                unresolved++;
                continue;
            }
            st.amount = c->amount;
            st.len = c->spk_len;
            st.big = (uint32_t)w.bigStore.size();
            w.bigStore.emplace_back(CoinTable::script(*c), CoinTable::script(*c) + c->spk_len);
            T.erase(c);
        }
        auto t3 = now();
        if (thDev.joinable()) thDev.join();
        auto t3b = now();
        phase[3] += Secs(t2, t3);
        phase[2] += Secs(t3, t3b); // device pre-checks the lookups did not hide
        if (cur.rcEq != BCP_OK) { Rollback(w); reason = "device-error: " + cur.errEq; return false; }
        if (!cur.eqOk) { Rollback(w); reason = "invalid-solution"; return false; } // CheckEquihashSolution (src/pow.cpp:298-323)
        if (cur.rcTx != BCP_OK) { Rollback(w); reason = "device-error: " + cur.errTx; return false; }
        if (memcmp(cur.root, blk + 36, 32) != 0) { Rollback(w); reason = "bad-txnmrklroot"; return false; }
        if (cur.mutated) { Rollback(w); reason = "bad-txns-duplicate"; return false; }
        if (cur.structErr) { Rollback(w); reason = cur.structErr; return false; } // after the Merkle checks, like CheckBlock
        memcpy(w.hash, cur.hash, 32);
        // Script checks, sharded by input index range over the devices.  When pass A resolved every input they start
        // now, from helper threads, and run while the team inserts the block's outputs (pass B).
        w.errs.assign((size_t)nin, 0);
        const size_t G = devs.size(), per = ((size_t)nin + G - 1) / G;
        w.rc.assign(G, BCP_OK);
        w.msg.assign(G, std::string());
        w.th.clear();
        auto launchScripts = [&] {
            BuildStaging(w);
            for (size_t g = 0; g < G && nin > 0; g++) {
                const size_t lo = std::min((size_t)nin, g * per), hi = std::min((size_t)nin, lo + per);
                if (lo >= hi) continue;
                Work *pw = &w;
                const int dev = BCP_DEVICE_LANE(devs[g] & 0xFF, (devs[g] >> 8) + (int)(&w - work)); /* work[0] -> lane 0, work[1] -> lane 1 */
                const uint32_t fl = flags;
                w.th.emplace_back([pw, g, lo, hi, dev, fl] {
                    int all = 0;
                    const Scan &c = pw->cur;
                    pw->rc[g] = bcp_gpu_block_scripts_p2pkh(dev, pw->blk, pw->len, c.tx_off.data(), (size_t)c.ntx, &c.in_off[lo],
                                                            &c.in_tx[lo], &c.in_n[lo], hi - lo, pw->spk_cat.data(), &pw->spk_off[lo],
                                                            &pw->amounts[lo], fl, &pw->errs[lo], &all);
                    if (pw->rc[g] != BCP_OK) pw->msg[g] = bcp_gpu_last_error();
                });
            }
        };
        const bool overlapped = unresolved == 0;
        if (overlapped) launchScripts();
        auto tb0 = now();
        walk[1] += Secs(t3b, tb0);
        // Pass B, all shards in parallel: the block's outputs become coins.  An output whose outpoint is still unspent in
        // the table is the BIP30 case (src/validation.cpp:1975-1990: no transaction may overwrite an unspent one) — the
        // reference tests it before it looks at any input.
        w.dupOut.assign((size_t)nout, 0);
        w.stamp = ++blockStamp;
        w.createdUpTo = (size_t)nout;
        team->Run([&](int s_) { PassB(w, s_); });
        auto tb1 = now();
        walk[2] += Secs(tb0, tb1);
        for (size_t o = 0; o < (size_t)nout; o++)
            if (w.dupOut[o] == 1) {
                if (overlapped) Abandon(w);
                Rollback(w);
                reason = "bad-txns-BIP30";
                return false;
            }
        // Pass C, only when pass A left inputs open: they may spend coins created EARLIER in this block (the creating
        // transaction must come first, src/validation.cpp:2095 applies each transaction before the next is looked at) and
        // each such coin only once.
        bool missing = false;
        if (unresolved) {
            // first 8 bytes of a txid -> transaction number.  A multimap: two transactions of a block may share those
            // 8 bytes (2^32 work to craft), and every candidate is compared on all 32 bytes below.
            std::unordered_multimap<uint64_t, uint32_t> txIndex;
            txIndex.reserve((size_t)ntx * 2);
            for (long t = 0; t < ntx; t++) txIndex.emplace(TxKey(&cur.txids[(size_t)t * 32]), (uint32_t)t);
            for (size_t i = 0; i < (size_t)nin && !missing; i++) {
                Staged &st = w.stage[i];
                if (st.state != ST_MISS) continue;
                OutPoint op;
                memcpy(op.b, blk + cur.in_off[i], 36);
                // the creating transaction: the EARLIEST one of this block with that txid (identical transactions have
                // identical outputs, and UpdateCoins applies each as it is met)
                bool created_before = false;
                auto range = txIndex.equal_range(TxKey(op.b));
                for (auto it = range.first; it != range.second; ++it)
                    if (it->second < cur.in_tx[i] && memcmp(&cur.txids[(size_t)it->second * 32], op.b, 32) == 0) created_before = true;
                CoinTable &T = ShardOf(op);
                CoinSlot *c = T.find(op);
                if (!c || !created_before) {
                    missing = true;
                    break;
                }
                st.amount = c->amount;
                st.len = c->spk_len;
                if (c->spk_len <= STAGE_SPK) {
                    memcpy(st.spk, CoinTable::script(*c), c->spk_len);
                    st.state = ST_INBLOCK;
                } else {
                    st.big = (uint32_t)w.bigStore.size();
                    w.bigStore.emplace_back(CoinTable::script(*c), CoinTable::script(*c) + c->spk_len);
                    st.state = ST_INBLOCK_BIG;
                }
                T.erase(c);
            }
        }
        auto t4 = now();
        walk[3] += Secs(tb1, t4);
        phase[3] += Secs(t3b, t4);
        if (missing) { // only reachable when the scripts were not started
            Rollback(w);
            reason = "bad-txns-inputs-missingorspent";
            return false;
        }
        if (!overlapped) launchScripts();
        return true;
    }
    // Waits for the block's script checks; false + reason when they reject it (the table is NOT touched here).
    bool Collect(Work &w, std::string &reason) {
        auto t0 = std::chrono::steady_clock::now();
        for (auto &t : w.th) t.join();
        w.th.clear();
        phase[4] += Secs(t0, std::chrono::steady_clock::now());
        lastWork = (int)(&w - work);
        for (size_t g = 0; g < w.rc.size(); g++)
            if (w.rc[g] != BCP_OK) { reason = "device-error: " + w.msg[g]; return false; }
        std::vector<OtherInput> others; // inputs that are not of the fast path's shape: decided after the ones the device decided
        for (size_t k = 0; k < (size_t)w.cur.nin; k++) {
            if (w.errs[k] == 255) {
                others.push_back(OtherInput{(uint32_t)k, w.cur.in_tx[k], w.cur.in_n[k], &w.spk_cat[w.spk_off[k]], w.spk_off[k + 1] - w.spk_off[k],
                                            w.amounts[k]});
                continue;
            }
            if (w.errs[k] != 0) { // ConnectBlock: "blk-bad-inputs" (src/validation.cpp:2124-2128)
                reason = "blk-bad-inputs (input " + std::to_string(k) + ", script error " + std::to_string(w.errs[k]) + ")";
                return false;
            }
        }
        if (others.empty()) return true;
        if (!other) { reason = "script-not-in-fast-path (input " + std::to_string(others[0].input_no) + ")"; return false; }
        std::string why;
        if (other(w.blk, w.len, w.cur.tx_off.data(), (size_t)w.cur.ntx, others.data(), others.size(), flags, why)) return true;
        reason = why.empty() ? "blk-bad-inputs" : why;
        return false;
    }
    void Abandon(Work &w) { // a block whose verdict no longer matters: just let its device calls finish
        for (auto &t : w.th) t.join();
        w.th.clear();
    }
    void Commit(Work &w) { // the table already holds the new state
        nInputs += (size_t)w.cur.nin;
        memcpy(lastHash, w.hash, 32);
        w.spentUpTo = w.createdUpTo = 0;
    }

    // One worker's share of pass A: the inputs whose outpoint hashes to its shard (grouped by the scan), in block order,
    // with the table slots of the next few prefetched while the current one is handled.
    void PassA(Work &wk, int w) {
        CoinTable &T = *shards[(size_t)w];
        const Scan &cur = wk.cur;
        const uint8_t *blk = wk.blk;
        const size_t RING = 12;
        size_t ring[RING], head = 0, count = 0;
        auto take = [&](size_t i) {
            OutPoint op;
            memcpy(op.b, blk + cur.in_off[i], 36);
            Staged &s = wk.stage[i];
            CoinSlot *c = T.find(op);
            if (!c) { s.state = ST_MISS; return; }
            if (c->spk_len > STAGE_SPK) { s.state = ST_BIG; return; } // left in the table for the connecting thread
            s.amount = c->amount;
            s.len = c->spk_len;
            memcpy(s.spk, CoinTable::script(*c), c->spk_len);
            s.state = ST_TAKEN;
            T.erase(c);
        };
        for (size_t q = cur.shard_start[(size_t)w]; q < cur.shard_start[(size_t)w + 1]; q++) {
            const size_t i = cur.by_shard[q];
            T.prefetch_hash((size_t)cur.in_key[i]);      /* the table slot, from the key the scan computed ...      */
            __builtin_prefetch(blk + cur.in_off[i]);     /* ... the outpoint bytes, last touched by the scan thread */
            __builtin_prefetch(blk + cur.in_off[i] + 35);
            __builtin_prefetch(&wk.stage[i], 1);         /* ... and the staging record this input will be written to */
            __builtin_prefetch(reinterpret_cast<const char *>(&wk.stage[i]) + 64, 1);
            if (count == RING) {
                take(ring[head]);
                ring[head] = i;
                head = (head + 1) % RING;
            } else {
                ring[(head + count++) % RING] = i;
            }
        }
        for (size_t k = 0; k < count; k++) take(ring[(head + k) % RING]);
    }
    // One worker's share of pass B: the outputs whose outpoint hashes to its shard.
    void PassB(Work &wk, int w) {
        CoinTable &T = *shards[(size_t)w];
        const Scan &cur = wk.cur;
        const uint8_t *blk = wk.blk;
        const std::vector<uint8_t> &txids = cur.txids;
        const size_t nout = (size_t)cur.nout;
        const size_t RING = 8;
        size_t ring[RING], head = 0, count = 0;
        auto put = [&](size_t o) {
            OutPoint op;
            memcpy(op.b, &txids[(size_t)cur.out_tx[o] * 32], 32);
            memcpy(op.b + 32, &cur.out_n[o], 4);
            size_t pos = cur.out_off[o];
            int64_t amt;
            memcpy(&amt, blk + pos, 8);
            pos += 8;
            const uint64_t sl = ReadCompact(blk, pos);
            // The reference's BIP30 test looks at the view as it was BEFORE the block; a twin created by an identical
            // transaction of this same block is not its business (that transaction then fails on its spent inputs).
            if (CoinSlot *have = T.find(op)) { wk.dupOut[o] = have->stamp == wk.stamp ? 2 : 1; return; }
            T.insert(op, blk + pos, sl, amt, wk.stamp);
        };
        (void)nout;
        for (size_t q = cur.out_shard_start[(size_t)w]; q < cur.out_shard_start[(size_t)w + 1]; q++) {
            const size_t o = cur.out_by_shard[q];
            T.prefetch_hash((size_t)cur.out_key[o]);
            __builtin_prefetch(blk + cur.out_off[o]); /* amount + script of the output, read when it is inserted */
            if (count == RING) {
                put(ring[head]);
                ring[head] = o;
                head = (head + 1) % RING;
            } else {
                ring[(head + count++) % RING] = o;
            }
        }
        for (size_t k = 0; k < count; k++) put(ring[(head + k) % RING]);
    }
    // staging records -> the concatenated scriptPubKeys / offsets / amounts the script-check call takes
    void BuildStaging(Work &w) { // two rounds of the team over contiguous slices: lengths, then the copies
        const size_t nin = (size_t)w.cur.nin;
        const int P = team->size();
        w.spk_off.resize(nin + 1);
        w.amounts.resize(nin);
        std::vector<size_t> sum((size_t)P + 1, 0);
        team->Run([&](int s_) {
            const size_t lo = nin * (size_t)s_ / (size_t)P, hi = nin * ((size_t)s_ + 1) / (size_t)P;
            size_t t = 0;
            for (size_t i = lo; i < hi; i++) t += w.stage[i].len;
            sum[(size_t)s_ + 1] = t;
        });
        for (int k = 0; k < P; k++) sum[(size_t)k + 1] += sum[(size_t)k];
        w.spk_cat.resize(sum[(size_t)P] + 1);
        team->Run([&](int s_) {
            const size_t lo = nin * (size_t)s_ / (size_t)P, hi = nin * ((size_t)s_ + 1) / (size_t)P;
            uint32_t o = (uint32_t)sum[(size_t)s_];
            for (size_t i = lo; i < hi; i++) {
                const Staged &s = w.stage[i];
                w.spk_off[i] = o;
                w.amounts[i] = s.amount;
                const uint8_t *src = (s.state == ST_BIG || s.state == ST_INBLOCK_BIG) ? w.bigStore[s.big].data() : s.spk;
                memcpy(&w.spk_cat[o], src, s.len);
                o += s.len;
            }
        });
        w.spk_off[nin] = (uint32_t)sum[(size_t)P];
    }
    // Key of the in-block transaction index.  BCP_CHAIN_TEST_TXKEY_BITS (tests only) narrows it so that collisions,
    // which take 2^32 hashes to find for the real 64-bit key, can be exercised.
    static uint64_t TxKey(const uint8_t *txid) {
        uint64_t k;
        memcpy(&k, txid, 8);
#ifdef BCP_CHAIN_TEST_TXKEY_BITS
        k &= (1ull << BCP_CHAIN_TEST_TXKEY_BITS) - 1;
#endif
        return k;
    }
    static uint64_t ReadCompact(const uint8_t *p, size_t &pos) { // already validated by the framing scan
        uint8_t c = p[pos++];
        if (c < 253) return c;
        int nb = c == 253 ? 2 : c == 254 ? 4 : 8;
        uint64_t v = 0;
        for (int i = 0; i < nb; i++) v |= (uint64_t)p[pos + i] << (8 * i);
        pos += nb;
        return v;
    }
    // Undo a (partially) applied block (one thread; a rejected block is the rare case): drop the coins it created, then
    // put back the coins it spent that existed before it — from the staging records, which still hold them.  Blocks are
    // undone in reverse order of application.
    void Rollback(Work &w) {
        const Scan &cur = w.cur;
        for (size_t o = 0; o < w.createdUpTo; o++) {
            if (o < w.dupOut.size() && w.dupOut[o]) continue; // was there before this block and is not this block's to remove
            OutPoint op;
            memcpy(op.b, &cur.txids[(size_t)cur.out_tx[o] * 32], 32);
            memcpy(op.b + 32, &cur.out_n[o], 4);
            CoinTable &T = ShardOf(op);
            CoinSlot *c = T.find(op);
            if (c) T.erase(c);
        }
        for (size_t i = 0; i < w.spentUpTo; i++) {
            const Staged &s = w.stage[i];
            if (s.state != ST_TAKEN && s.state != ST_BIG) continue; // misses, and coins this block itself had created
            OutPoint op;
            memcpy(op.b, w.blk + cur.in_off[i], 36);
            const uint8_t *src = s.state == ST_BIG ? w.bigStore[s.big].data() : s.spk;
            ShardOf(op).insert(op, src, s.len, s.amount);
        }
        w.spentUpTo = w.createdUpTo = 0;
    }

    std::vector<int> devs;
    int devAux = 0;
    uint32_t flags;
    unsigned ehN, ehK;
    int shardBits;
    std::vector<std::unique_ptr<CoinTable>> shards;
    std::unique_ptr<WorkerTeam> team;
    static const unsigned NAHEAD = 4; // blocks being scanned / pre-checked ahead of the one being connected
    Scan ahead[NAHEAD];
    unsigned nextSlot = 0;
    Work work[2];
    OtherScripts other;
    uint32_t blockStamp = 0;
    int lastWork = 0;
    uint8_t lastHash[32] = {0};
    size_t nInputs = 0;
};

} // namespace bcp
#endif
