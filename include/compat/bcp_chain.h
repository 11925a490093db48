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
// Not done (outside the hot path): amounts/fees/maturity (Consensus::CheckTxInputs), sigop limits, BIP68 locks,
// difficulty, undo data.  Inputs that are not of the P2PKH shape are reported (reason "script-not-in-fast-path"):
// a node would run them through its interpreter.
#ifndef BCP_COMPAT_CHAIN_H
#define BCP_COMPAT_CHAIN_H

#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../bcp_gpu.h"

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
    uint32_t pad2;
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
    void prefetch(const OutPoint &k) const {
        const char *p = reinterpret_cast<const char *>(&slots[OutPointHash()(k) & mask]);
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
    CoinSlot *insert(const OutPoint &k, const uint8_t *spk, size_t len, int64_t amount) {
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
                CoinSlot *n = insert(o.key, script(o), o.spk_len, o.amount);
                (void)n;
                delete o.big;
                o.big = nullptr;
            }
    }
    std::vector<CoinSlot> slots;
    size_t mask, live = 0;
};

class ChainValidator {
public:
    ChainValidator(const std::vector<int> &devices, uint32_t script_flags, unsigned eh_n, unsigned eh_k)
        : devs(devices), flags(script_flags), ehN(eh_n), ehK(eh_k) {}

    bool Init(std::string &err) {
        if (devs.empty()) { err = "no device"; return false; }
        for (int d : devs)
            if (bcp_gpu_init(d) != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        return true;
    }
    void AddCoin(const uint8_t txid[32], uint32_t vout, const uint8_t *spk, size_t len, int64_t amount) {
        OutPoint op;
        memcpy(op.b, txid, 32);
        memcpy(op.b + 32, &vout, 4);
        utxo.insert(op, spk, len, amount);
    }
    size_t UtxoSize() const { return utxo.size(); }
    size_t InputsChecked() const { return nInputs; }

    // Returns true and updates the UTXO set, or false with `reason` (the reference's reject strings where one exists).
    // seconds spent per phase since construction: header (Equihash + hash), framing scan, txids + Merkle, UTXO walk,
    // script checks (GPU calls incl. copies)
    double phase[5] = {0, 0, 0, 0, 0};

    bool ConnectBlock(const uint8_t *blk, size_t len, std::string &reason) {
        const int dev0 = devs[0];
        auto now = [] { return std::chrono::steady_clock::now(); };
        auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
            return std::chrono::duration<double>(b - a).count();
        };
        auto t0 = now();
        // header: Equihash + hash (ehN == 0 skips it, for chains built without solutions)
        const size_t soln = ehN ? ((size_t)1 << ehK) * (ehN / (ehK + 1) + 1) / 8 : 0;
        const size_t hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
        if (ehN) {
            if (len < hdr_len) { reason = "bad-blk-length"; return false; }
            uint8_t ok = 0, hash[32];
            if (bcp_gpu_equihash_verify_batch(dev0, ehN, ehK, blk, hdr_len, hdr_len, 1, &ok, hash) != BCP_OK) {
                reason = std::string("device-error: ") + bcp_gpu_last_error();
                return false;
            }
            if (!ok) { reason = "invalid-solution"; return false; } // CheckEquihashSolution (src/pow.cpp:298-323)
            memcpy(lastHash, hash, 32);
        }
        auto t1 = now();
        phase[0] += secs(t0, t1);
        // framing
        const size_t max_tx = len / 60 + 2, max_in = len / 41 + 2;
        tx_off.resize(max_tx + 1);
        in_off.resize(max_in); in_tx.resize(max_in); in_n.resize(max_in);
        long ntx = 0;
        long nin = bcp_block_index_inputs(blk, len, 0, tx_off.data(), max_tx, in_off.data(), in_tx.data(), in_n.data(), max_in, &ntx);
        if (nin < 0 || ntx < 1) { reason = "bad-blk-framing"; return false; }
        auto t2 = now();
        phase[1] += secs(t1, t2);
        // txids + Merkle root (CheckBlock, src/validation.cpp:3236-3252) on the device, from a helper thread, while
        // this thread looks up the spent coins: the lookups need the previous outputs, not this block's txids
        txids.resize((size_t)ntx * 32);
        uint8_t root[32];
        int mutated = 0, rcTx = BCP_OK;
        std::string errTx;
        std::thread thTx([&] {
            rcTx = bcp_gpu_sha256d_txids_merkle(dev0, blk, len, tx_off.data(), (size_t)ntx, txids.data(), root, &mutated);
            if (rcTx != BCP_OK) errTx = bcp_gpu_last_error();
        });
        // Spent coins.  Everything is applied to the one table and journalled so that a rejected block can be rolled
        // back.  Pass A takes the inputs in block order for as long as they are found in the table as it was before
        // this block (the normal case: all of them); the first miss — an output created earlier in this block, or a
        // missing coin — ends it at input `m`.
        spk_cat.clear();
        spk_off.assign(1, 0);
        amounts.clear();
        journalSpent.clear();
        journalCreated.clear();
        const size_t AHEAD = 16;
        size_t m = 0;
        for (; m < (size_t)nin; m++) {
            if (m + AHEAD < (size_t)nin) {
                OutPoint nx;
                memcpy(nx.b, blk + in_off[m + AHEAD], 36);
                utxo.prefetch(nx);
            }
            if (!Spend(blk + in_off[m])) break;
        }
        auto t3 = now();
        thTx.join();
        auto t3b = now();
        phase[3] += secs(t2, t3);
        phase[2] += secs(t3, t3b); // what the lookups did not hide
        if (rcTx != BCP_OK) { Rollback(); reason = "device-error: " + errTx; return false; }
        if (memcmp(root, blk + 36, 32) != 0) { Rollback(); reason = "bad-txnmrklroot"; return false; }
        if (mutated) { Rollback(); reason = "bad-txns-duplicate"; return false; }
        // Script checks, sharded by input index range over the devices.  When pass A resolved every input they start
        // now, from helper threads, and run while this thread inserts the block's outputs (pass B).
        errs.assign((size_t)nin, 0);
        const size_t G = devs.size(), per = ((size_t)nin + G - 1) / G;
        std::vector<int> rc(G, BCP_OK);
        std::vector<std::string> msg(G);
        std::vector<std::thread> th;
        auto launchScripts = [&] {
            for (size_t g = 0; g < G && nin > 0; g++) {
                const size_t lo = std::min((size_t)nin, g * per), hi = std::min((size_t)nin, lo + per);
                if (lo >= hi) continue;
                th.emplace_back([&, g, lo, hi] {
                    int all = 0;
                    rc[g] = bcp_gpu_block_scripts_p2pkh(devs[g], blk, len, tx_off.data(), (size_t)ntx, &in_off[lo], &in_tx[lo],
                                                        &in_n[lo], hi - lo, spk_cat.data(), &spk_off[lo], &amounts[lo], flags,
                                                        &errs[lo], &all);
                    if (rc[g] != BCP_OK) msg[g] = bcp_gpu_last_error();
                });
            }
        };
        const bool overlapped = m == (size_t)nin;
        if (overlapped) launchScripts();
        // Pass B, transaction by transaction: the inputs pass A did not reach (they may spend outputs created earlier in
        // this block, and only those), then the transaction's own outputs.
        size_t i = m;
        bool missing = false;
        for (long t = 0; t < ntx && !missing; t++) {
            if ((size_t)t + AHEAD < (size_t)ntx) { // where transaction t+AHEAD's first output will land
                OutPoint nx;
                memcpy(nx.b, &txids[((size_t)t + AHEAD) * 32], 32);
                memset(nx.b + 32, 0, 4);
                utxo.prefetch(nx);
            }
            for (; i < (size_t)nin && in_tx[i] == (uint32_t)t; i++)
                if (!Spend(blk + in_off[i])) { missing = true; break; }
            if (!missing) AddOutputs(blk, (uint32_t)t);
        }
        auto t4 = now();
        phase[3] += secs(t3b, t4);
        if (missing) { // only reachable when the scripts were not started
            Rollback();
            reason = "bad-txns-inputs-missingorspent";
            return false;
        }
        if (!overlapped) launchScripts();
        for (auto &t : th) t.join();
        phase[4] += secs(t4, now()); // what pass B did not hide (everything when the passes could not overlap)
        for (size_t g = 0; g < G; g++)
            if (rc[g] != BCP_OK) { Rollback(); reason = "device-error: " + msg[g]; return false; }
        for (size_t k = 0; k < (size_t)nin; k++) {
            if (errs[k] == 255) { Rollback(); reason = "script-not-in-fast-path (input " + std::to_string(k) + ")"; return false; }
            if (errs[k] != 0) { // ConnectBlock: "blk-bad-inputs" (src/validation.cpp:2124-2128)
                Rollback();
                reason = "blk-bad-inputs (input " + std::to_string(k) + ", script error " + std::to_string(errs[k]) + ")";
                return false;
            }
        }
        // commit: the table already holds the new state
        nInputs += (size_t)nin;
        journalSpent.clear();
        journalCreated.clear();
        return true;
    }
    const uint8_t *LastBlockHash() const { return lastHash; }
    const std::vector<uint8_t> &LastScriptErrors() const { return errs; }

    // blk*.dat framing (src/validation.cpp:4486-4516): [magic 4][size u32 LE][block] ... ; returns blocks connected,
    // stops at the first failure (reason set, failed_index = its ordinal).
    long ConnectFile(const uint8_t *data, size_t len, const uint8_t magic[4], std::string &reason, long &failed_index) {
        size_t pos = 0;
        long n = 0;
        failed_index = -1;
        while (pos + 8 <= len) {
            if (memcmp(data + pos, magic, 4) != 0) { pos++; continue; } // resynchronise on the magic like :4489-4498
            uint32_t sz;
            memcpy(&sz, data + pos + 4, 4);
            if (pos + 8 + sz > len) break;
            if (!ConnectBlock(data + pos + 8, sz, reason)) { failed_index = n; return n; }
            pos += 8 + (size_t)sz;
            n++;
        }
        return n;
    }

private:
    // Looks up the coin an input spends, appends its scriptPubKey / amount to the staging arrays of the script checks,
    // removes it from the table and journals it.  false: no such coin (already spent, or never there).
    bool Spend(const uint8_t *outpoint36) {
        OutPoint op;
        memcpy(op.b, outpoint36, 36);
        CoinSlot *c = utxo.find(op);
        if (!c) return false;
        const uint8_t *sp = CoinTable::script(*c);
        spk_cat.insert(spk_cat.end(), sp, sp + c->spk_len);
        spk_off.push_back((uint32_t)spk_cat.size());
        amounts.push_back(c->amount);
        utxo.erase(c);
        journalSpent.push_back(op);
        return true;
    }
    void AddOutputs(const uint8_t *blk, uint32_t t) {
        // skip inputs to reach the outputs of transaction t
        size_t pos = tx_off[t] + 4;
        uint64_t nin = ReadCompact(blk, pos);
        for (uint64_t k = 0; k < nin; k++) {
            pos += 36;
            uint64_t sl = ReadCompact(blk, pos);
            pos += sl + 4;
        }
        uint64_t nout = ReadCompact(blk, pos);
        for (uint64_t v = 0; v < nout; v++) {
            int64_t amt;
            memcpy(&amt, blk + pos, 8);
            pos += 8;
            uint64_t sl = ReadCompact(blk, pos);
            OutPoint op;
            memcpy(op.b, &txids[(size_t)t * 32], 32);
            uint32_t vv = (uint32_t)v;
            memcpy(op.b + 32, &vv, 4);
            utxo.insert(op, blk + pos, sl, amt);
            journalCreated.push_back(op);
            pos += sl;
        }
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
    // Undo a partially applied block: drop what it created, then put back what it spent.  The spent coins are
    // re-inserted from the block's own staging arrays (outpoint journal + the scriptPubKey / amount arrays that were
    // built for the GPU call), because a coin created by this block may have reused a spent coin's slot.
    void Rollback() {
        for (const OutPoint &op : journalCreated) {
            CoinSlot *c = utxo.find(op);
            if (c) utxo.erase(c);
        }
        for (size_t i = 0; i < journalSpent.size(); i++)
            utxo.insert(journalSpent[i], &spk_cat[spk_off[i]], spk_off[i + 1] - spk_off[i], amounts[i]);
        journalSpent.clear();
        journalCreated.clear();
    }

    std::vector<int> devs;
    uint32_t flags;
    unsigned ehN, ehK;
    CoinTable utxo;
    std::vector<OutPoint> journalSpent, journalCreated;
    std::vector<uint32_t> tx_off, in_off, in_tx, in_n, spk_off;
    std::vector<uint8_t> txids, spk_cat, errs;
    std::vector<int64_t> amounts;
    uint8_t lastHash[32] = {0};
    size_t nInputs = 0;
};

} // namespace bcp
#endif
