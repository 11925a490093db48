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

#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
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
struct Coin {
    std::vector<uint8_t> spk;
    int64_t amount;
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
        utxo[op] = Coin{std::vector<uint8_t>(spk, spk + len), amount};
    }
    size_t UtxoSize() const { return utxo.size(); }
    size_t InputsChecked() const { return nInputs; }

    // Returns true and updates the UTXO set, or false with `reason` (the reference's reject strings where one exists).
    bool ConnectBlock(const uint8_t *blk, size_t len, std::string &reason) {
        const int dev0 = devs[0];
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
        // framing
        const size_t max_tx = len / 60 + 2, max_in = len / 41 + 2;
        tx_off.resize(max_tx + 1);
        in_off.resize(max_in); in_tx.resize(max_in); in_n.resize(max_in);
        long ntx = 0;
        long nin = bcp_block_index_inputs(blk, len, 0, tx_off.data(), max_tx, in_off.data(), in_tx.data(), in_n.data(), max_in, &ntx);
        if (nin < 0 || ntx < 1) { reason = "bad-blk-framing"; return false; }
        // txids + Merkle root (CheckBlock, src/validation.cpp:3236-3252)
        txids.resize((size_t)ntx * 32);
        uint8_t root[32];
        int mutated = 0;
        if (bcp_gpu_sha256d_txids_merkle(dev0, blk, len, tx_off.data(), (size_t)ntx, txids.data(), root, &mutated) != BCP_OK) {
            reason = std::string("device-error: ") + bcp_gpu_last_error();
            return false;
        }
        if (memcmp(root, blk + 36, 32) != 0) { reason = "bad-txnmrklroot"; return false; }
        if (mutated) { reason = "bad-txns-duplicate"; return false; }
        // spent coins, transaction by transaction, seeing outputs created earlier in this block
        std::unordered_map<OutPoint, Coin, OutPointHash> created;
        std::vector<OutPoint> spent;
        spk_cat.clear();
        spk_off.assign(1, 0);
        amounts.clear();
        spent.reserve(nin);
        size_t i = 0;
        for (long t = 0; t < ntx; t++) {
            for (; i < (size_t)nin && in_tx[i] == (uint32_t)t; i++) {
                OutPoint op;
                memcpy(op.b, blk + in_off[i], 36);
                const Coin *c = nullptr;
                auto it = created.find(op);
                if (it != created.end()) c = &it->second;
                else {
                    auto iu = utxo.find(op);
                    if (iu != utxo.end()) c = &iu->second;
                }
                bool again = false;
                for (const OutPoint &s : spentThisTx) again = again || s == op;
                if (!c || again || spentInBlock.count(op)) { Reset(); reason = "bad-txns-inputs-missingorspent"; return false; }
                spk_cat.insert(spk_cat.end(), c->spk.begin(), c->spk.end());
                spk_off.push_back((uint32_t)spk_cat.size());
                amounts.push_back(c->amount);
                spentThisTx.push_back(op);
            }
            for (const OutPoint &s : spentThisTx) { spentInBlock.emplace(s, true); spent.push_back(s); }
            spentThisTx.clear();
            AddOutputs(blk, (uint32_t)t, created);
        }
        // script checks, sharded by input index range over the devices
        errs.assign((size_t)nin, 0);
        const size_t G = devs.size(), per = ((size_t)nin + G - 1) / G;
        std::vector<int> rc(G, BCP_OK);
        std::vector<std::string> msg(G);
        std::vector<std::thread> th;
        for (size_t g = 0; g < G && nin > 0; g++) {
            const size_t lo = std::min((size_t)nin, g * per), hi = std::min((size_t)nin, lo + per);
            if (lo >= hi) continue;
            auto work = [&, g, lo, hi] {
                int all = 0;
                rc[g] = bcp_gpu_block_scripts_p2pkh(devs[g], blk, len, tx_off.data(), (size_t)ntx, &in_off[lo], &in_tx[lo], &in_n[lo],
                                                    hi - lo, spk_cat.data(), &spk_off[lo], &amounts[lo], flags, &errs[lo], &all);
                if (rc[g] != BCP_OK) msg[g] = bcp_gpu_last_error();
            };
            if (G == 1) work();
            else th.emplace_back(work);
        }
        for (auto &t : th) t.join();
        for (size_t g = 0; g < G; g++)
            if (rc[g] != BCP_OK) { Reset(); reason = "device-error: " + msg[g]; return false; }
        for (size_t k = 0; k < (size_t)nin; k++) {
            if (errs[k] == 255) { Reset(); reason = "script-not-in-fast-path (input " + std::to_string(k) + ")"; return false; }
            if (errs[k] != 0) { // ConnectBlock: "blk-bad-inputs" (src/validation.cpp:2124-2128)
                Reset();
                reason = "blk-bad-inputs (input " + std::to_string(k) + ", script error " + std::to_string(errs[k]) + ")";
                return false;
            }
        }
        // commit
        for (const OutPoint &s : spent) utxo.erase(s);
        for (auto &kv : created)
            if (!spentInBlock.count(kv.first)) utxo[kv.first] = std::move(kv.second);
        nInputs += (size_t)nin;
        Reset();
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
    void AddOutputs(const uint8_t *blk, uint32_t t, std::unordered_map<OutPoint, Coin, OutPointHash> &created) {
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
            created[op] = Coin{std::vector<uint8_t>(blk + pos, blk + pos + sl), amt};
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
    void Reset() {
        spentInBlock.clear();
        spentThisTx.clear();
    }

    std::vector<int> devs;
    uint32_t flags;
    unsigned ehN, ehK;
    std::unordered_map<OutPoint, Coin, OutPointHash> utxo;
    std::unordered_map<OutPoint, bool, OutPointHash> spentInBlock;
    std::vector<OutPoint> spentThisTx;
    std::vector<uint32_t> tx_off, in_off, in_tx, in_n, spk_off;
    std::vector<uint8_t> txids, spk_cat, errs;
    std::vector<int64_t> amounts;
    uint8_t lastHash[32] = {0};
    size_t nInputs = 0;
};

} // namespace bcp
#endif
