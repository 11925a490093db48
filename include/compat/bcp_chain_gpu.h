// include/compat/bcp_chain_gpu.h — block connector for a chain of blocks (reindex / initial block download,
// LoadExternalBlockFile src/validation.cpp:4461-4560 feeding ConnectBlock :1868-2188) whose coin table lives in HBM
// (bcp_gpu_utxo_*, bitcoincashplus_b200/csrc/bcp_utxo.cu) and whose blocks are spread over the GPUs of the box.
// BASELINE.json configs[4]: "100-block synthetic chain reindex, signature batches sharded across 2/4/8 B200".
//
// bcp::ChainValidator (bcp_chain.h) walks an in-memory coin map on host threads and splits every block's inputs over the
// devices; its connecting thread (~1 ms of table passes per 8 MB block) is the ceiling whatever the number of GPUs.  Here
// nothing that is proportional to a block's size runs on the connecting thread:
//
//   scan threads     framing scan of the next blocks (bcp_block_index_full) + CheckBlock's context-free structure rules
//   connecting thread  (the caller) per block: ONE asynchronous call, bcp_gpu_utxo_apply_block — the table's stream does
//                    txids + Merkle root, inserts the outputs, looks up and spends the inputs in block order and returns the
//                    spent (amount, scriptPubKey) arrays; stream order is the chain's order, so block b+1 is applied on top
//                    of block b while b's script checks are still running
//   check threads    one per (GPU, lane); block b belongs to thread b mod W: waits for b's table work, verifies the
//                    header (Equihash + hash) and runs the script checks of the WHOLE block on its GPU
//                    (bcp_gpu_utxo_block_scripts: from the copy of the block and the spent-coin arrays that are already in
//                    device memory — in place on the table's GPU, one NVLink copy away on the others; the host's memory
//                    is crossed once per block, by the upload).  Blocks, not inputs, are what is dealt out to the devices: a block is
//                    less than one wave of the point kernel on a B200, so splitting it further would not shorten it
//   connecting thread  takes the verdicts in order: confirm, or — first failure — let the blocks applied after it finish,
//                    roll them back newest first, then the failed block; the table is then exactly as before that block
//
// Reject reasons and their order are those of bcp::ChainValidator (the reference's strings).
#ifndef BCP_COMPAT_CHAIN_GPU_H
#define BCP_COMPAT_CHAIN_GPU_H

#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../bcp_gpu.h"
#include "bcp_other_scripts.h"

namespace bcp {

class DeviceChain {
public:
    // devices: GPU ordinals; the coin table lives on devices[0].  lanes: check threads per GPU (1..3).
    DeviceChain(const std::vector<int> &devices, uint32_t script_flags, unsigned eh_n, unsigned eh_k, unsigned log2_slots = 22,
                int lanes = 2, int scan_threads = 4)
        : devs(devices), flags(script_flags), ehN(eh_n), ehK(eh_k), log2Slots(log2_slots), lanes_(lanes < 1 ? 1 : lanes > 3 ? 3 : lanes),
          nScan(scan_threads < 1 ? 1 : scan_threads) {}
    ~DeviceChain() {
        Stop();
        for (Job &j : jobs) j.Free();
        if (utxo) bcp_gpu_utxo_destroy(utxo);
    }
    DeviceChain(const DeviceChain &) = delete;
    DeviceChain &operator=(const DeviceChain &) = delete;

    bool Init(std::string &err) {
        if (devs.empty()) { err = "no device"; return false; }
        // A block's input area reaches its checking GPU by ONE device-to-device copy from the table's GPU.  Without direct peer
        // access that copy is staged through host memory by the runtime (measured on a 2-GPU box without it: 8.7 ms per block
        // instead of 2, and every other runtime call of the process stalls behind it: 340 blocks/s instead of 1,200).  Such a GPU
        // is left out; BCP_CHAIN_NO_PEER=1 treats every other GPU that way (test hook).
        {
            const char *np = getenv("BCP_CHAIN_NO_PEER");
            std::vector<int> usable(1, devs[0]);
            for (size_t g = 1; g < devs.size(); g++) {
                const bool same = (devs[g] & 0xFF) == (devs[0] & 0xFF);
                if (same || (!(np && atoi(np)) && bcp_gpu_peer_access(devs[0], devs[g]) == 1)) usable.push_back(devs[g]);
                else skipped.push_back(devs[g] & 0xFF);
            }
            devs.swap(usable);
            // alone on its GPU the table shares it with all the check threads there are: give it the full count
        }
        for (int d : devs)
            for (int l = 0; l < lanes_; l++)
                if (bcp_gpu_init(BCP_DEVICE_LANE(d & 0xFF, l)) != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        for (int d : devs) // lane 3: the header checks, made by the scan threads as soon as a block's bytes are at hand
            if (bcp_gpu_init(BCP_DEVICE_LANE(d & 0xFF, 3)) != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        utxo = bcp_gpu_utxo_create(devs[0] & 0xFF, log2Slots, (size_t)64 << 20);
        if (!utxo) { err = bcp_gpu_last_error(); return false; }
        // The check threads, in the order blocks are dealt to them: lane by lane across the GPUs, so that consecutive
        // blocks go to different GPUs.  With more than one GPU the table's GPU gets one check thread less: its SMs must
        // have room for the table's kernels, which every other block waits for (a sub-wave point kernel holds the
        // registers of the SMs it runs on for a millisecond; two of them at once fill the chip).
        workers.clear();
        for (int l = 0; l < lanes_; l++)
            for (size_t g = 0; g < devs.size(); g++)
                if (!(devs.size() > 1 && g == 0 && l == lanes_ - 1 && lanes_ > 1)) workers.push_back(BCP_DEVICE_LANE(devs[g] & 0xFF, l));
        const size_t W = workers.size();
        inflightMax = 2 * W + 2 < 30 ? 2 * W + 2 : 30; /* blocks applied and awaiting their verdict: two per check thread (one
                                                          being checked, one waiting), and the table keeps undo records of at most 32 */
        jobs.resize(inflightMax + (size_t)nScan + 2);
        return true;
    }
    //! Optional: allocate the pinned per-block buffers of every pipeline slot now, for blocks of up to max_block_bytes, so
    //! that no allocation (a pinned allocation mapped into every GPU takes milliseconds and stalls the runtime) happens
    //! while blocks are being connected.  Without it the buffers grow on first use.
    void Reserve(size_t max_block_bytes) {
        const size_t l = max_block_bytes, max_tx = l / 60 + 2, max_in = l / 41 + 2, max_out = l / 34 + 2;
        auto up = [](size_t x) { return (x + 255) / 256 * 256; };
        for (Job &j : jobs) {
            j.tx_off.need(max_tx + 1); j.tx_out_off.need(max_tx + 1);
            j.in_off.need(max_in); j.in_tx.need(max_in); j.in_n.need(max_in);
            j.out_off.need(max_out); j.out_tx.need(max_out); j.out_n.need(max_out);
            j.errs.need(max_in + 1);
            j.small.need(64);
            j.packed.need(up((max_tx + 1) * 4) + 2 * up(max_in * 4 + 4) + 3 * up(max_out * 4 + 4));
        }
        bcp_gpu_utxo_reserve(utxo, max_block_bytes); /* the table's pending-block areas on the device, likewise */
    }
    void AddCoin(const uint8_t txid[32], uint32_t vout, const uint8_t *spk, size_t len, int64_t amount) {
        seedOp.insert(seedOp.end(), txid, txid + 32);
        for (int k = 0; k < 4; k++) seedOp.push_back((uint8_t)(vout >> (8 * k)));
        seedSpk.insert(seedSpk.end(), spk, spk + len);
        seedOff.push_back((uint32_t)seedSpk.size());
        seedAmt.push_back(amount);
    }
    long long UtxoSize() {
        std::string e;
        if (!FlushSeeds(e)) return -1;
        return bcp_gpu_utxo_size(utxo);
    }
    size_t InputsChecked() const { return nInputs; }
    const uint8_t *LastBlockHash() const { return lastHash; }
    //! Who evaluates the inputs the device fast path does not (bcp_other_scripts.h).  Unset: such a block is refused with
    //! "script-not-in-fast-path".  Called from check threads: at most one call per block, several blocks concurrently.
    void SetOtherScripts(OtherScripts f) { other = std::move(f); }

    // seconds, summed over blocks: [0] connecting thread inside bcp_gpu_utxo_apply_block, [1] connecting thread waiting
    // (for a scan, a free slot or a verdict), [2] check threads waiting for the spent-coin arrays, [3] check threads in
    // the header check, [4] check threads in the script checks, [5] scan threads, [6] connecting thread confirming,
    // [7] from "applied" to "table work done" (the table stream's queue + work), [8] from "applied" to "verdict taken"
    // (how long a block occupies one of the in-flight slots)
    // [9] of [1]: waiting with the in-flight window full (for a verdict), [10] of [1]: waiting for the next block's scan
    double phase[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<int> skipped; //!< GPUs of the constructor's list that were left out (no direct peer access to the table's GPU)
    size_t GpusInUse() const { return devs.size(); }

    bool ConnectBlock(const uint8_t *blk, size_t len, std::string &reason) {
        std::vector<std::pair<const uint8_t *, size_t>> one{{blk, len}};
        long failed = -1;
        return Run(one, reason, failed) == 1;
    }
    // blk*.dat framing (src/validation.cpp:4486-4516): [magic 4][size u32 LE][block] ...; returns the number of blocks
    // connected, stops at the first failure (reason set, failed_index = its ordinal, table as before that block).
    long ConnectFile(const uint8_t *data, size_t len, const uint8_t magic[4], std::string &reason, long &failed_index) {
        std::vector<std::pair<const uint8_t *, size_t>> blocks;
        size_t pos = 0;
        while (pos + 8 <= len) {
            if (memcmp(data + pos, magic, 4) != 0) { pos++; continue; } // resynchronise on the magic like :4489-4498
            uint32_t sz;
            memcpy(&sz, data + pos + 4, 4);
            if (pos + 8 + sz > len) break;
            blocks.emplace_back(data + pos + 8, (size_t)sz);
            pos += 8 + (size_t)sz;
        }
        return Run(blocks, reason, failed_index);
    }

private:
    typedef std::chrono::steady_clock Clock;
    static double Secs(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double>(b - a).count(); }

    template <class T> struct Pinned { // host arrays the device copies from / into: pinned, grow-only, never value-initialised
        T *p = nullptr;
        size_t cap = 0;
        void need(size_t n) {
            if (cap >= n) return;
            if (p) bcp_gpu_host_free(p);
            p = static_cast<T *>(bcp_gpu_host_alloc((n + n / 4 + 64) * sizeof(T)));
            cap = p ? n + n / 4 + 64 : 0;
        }
        void release() {
            if (p) bcp_gpu_host_free(p);
            p = nullptr;
            cap = 0;
        }
    };
    enum State { FREE = 0, SCANNING, SCANNED, APPLIED, CHECKING, DONE };
    struct Job {
        State state = FREE;
        long index = -1; // ordinal of the block in this run
        const uint8_t *blk = nullptr;
        size_t len = 0;
        Pinned<uint32_t> tx_off, tx_out_off, in_off, in_tx, in_n, out_off, out_tx, out_n, spk_off;
        Pinned<int64_t> amounts;
        Pinned<uint8_t> spk_cat, errs, small; // small: root[32] | mutated(int) at 32
        Pinned<uint8_t> packed;               // the six offset arrays back to back, each padded to 256 bytes (one upload)
        uint32_t *p_tx_off = nullptr, *p_in_off = nullptr, *p_in_tx = nullptr, *p_out_off = nullptr, *p_out_tx = nullptr, *p_out_n = nullptr;
        long ntx = 0, nin = -1, nout = 0;
        const char *structErr = nullptr;
        uint32_t seq = 0;
        int rcEq = BCP_OK;     // header check (Equihash + hash), done by the scan thread
        uint8_t eqOk = 1;
        std::string eqErr;
        Clock::time_point tApplied;
        bool ok = false;
        std::string reason;
        uint8_t hash[32];
        void Free() {
            tx_off.release(); tx_out_off.release(); in_off.release(); in_tx.release(); in_n.release(); out_off.release();
            out_tx.release(); out_n.release(); spk_off.release(); amounts.release(); spk_cat.release(); errs.release(); small.release();
            packed.release();
        }
    };

    bool FlushSeeds(std::string &err) {
        if (seedAmt.empty()) return true;
        std::vector<uint32_t> off(1, 0);
        off.insert(off.end(), seedOff.begin(), seedOff.end());
        if (seedSpk.empty()) seedSpk.push_back(0);
        const int rc = bcp_gpu_utxo_add_coins(utxo, seedOp.data(), seedSpk.data(), off.data(), seedAmt.data(), seedAmt.size());
        seedOp.clear(); seedSpk.clear(); seedOff.clear(); seedAmt.clear();
        if (rc != BCP_OK) { err = bcp_gpu_last_error(); return false; }
        return true;
    }

    // framing scan + structure rules of one block into its job, and its header check (scan thread `who`)
    void Scan(Job &j, size_t who) {
        const uint8_t *b = j.blk;
        const size_t l = j.len;
        // CheckBlockHeader's part that needs nothing but the header (src/validation.cpp:3202-3220): Equihash + block hash,
        // one small launch on lane 3 of one of the GPUs — here, not on a check thread, whose time is the GPU's
        j.rcEq = BCP_OK;
        j.eqOk = 1;
        if (ehN) {
            const size_t soln = ((size_t)1 << ehK) * (ehN / (ehK + 1) + 1) / 8;
            const size_t hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
            if (l >= hdr_len) {
                j.rcEq = bcp_gpu_equihash_verify_batch(BCP_DEVICE_LANE(devs[who % devs.size()] & 0xFF, 3), ehN, ehK, b, hdr_len, hdr_len, 1, &j.eqOk, j.hash);
                if (j.rcEq != BCP_OK) j.eqErr = bcp_gpu_last_error();
            }
        }
        size_t max_tx = l / 60 + 2, max_in = l / 41 + 2, max_out = l / 34 + 2; /* outputs: first guess, P2PKH-sized */
        int rule = 0; // the scan applies CheckBlock's context-free transaction rules while it has the bytes in the cache
        long ruleTx = -1;
        for (int attempt = 0; attempt < 2; attempt++) {
            j.tx_off.need(max_tx + 1); j.tx_out_off.need(max_tx + 1);
            j.in_off.need(max_in); j.in_tx.need(max_in); j.in_n.need(max_in);
            j.out_off.need(max_out); j.out_tx.need(max_out); j.out_n.need(max_out);
            j.ntx = j.nout = 0;
            rule = 0;
            j.nin = bcp_block_index_rules(b, l, 0, j.tx_off.p, j.tx_out_off.p, max_tx, j.in_off.p, j.in_tx.p, j.in_n.p, max_in, &j.ntx,
                                          j.out_off.p, j.out_tx.p, j.out_n.p, max_out, &j.nout, nullptr, &rule, &ruleTx);
            if (j.nin != -2) break;
            max_out = l / 9 + 2; /* the worst case a block of this size can hold */
        }
        j.structErr = (j.nin >= 0 && rule) ? bcp_block_rule_name(rule) : nullptr;
        if (j.nin >= 0) {
            j.errs.need((size_t)j.nin + 1);
            // the arrays bcp_gpu_utxo_apply_block uploads, laid out as it lays them out on the device: it then needs ONE
            // copy for all six (the connecting thread pays for every runtime call)
            auto up = [](size_t x) { return (x + 255) / 256 * 256; };
            const size_t ntx = (size_t)j.ntx, nin = (size_t)j.nin, nout = (size_t)j.nout;
            const size_t s_tx = up((ntx + 1) * 4), s_in = up(nin * 4 + 4), s_out = up(nout * 4 + 4);
            j.packed.need(s_tx + 2 * s_in + 3 * s_out);
            uint8_t *q = j.packed.p;
            j.p_tx_off = reinterpret_cast<uint32_t *>(q); memcpy(q, j.tx_off.p, (ntx + 1) * 4); q += s_tx;
            j.p_in_off = reinterpret_cast<uint32_t *>(q); memcpy(q, j.in_off.p, nin * 4); q += s_in;
            j.p_in_tx = reinterpret_cast<uint32_t *>(q); memcpy(q, j.in_tx.p, nin * 4); q += s_in;
            j.p_out_off = reinterpret_cast<uint32_t *>(q); memcpy(q, j.out_off.p, nout * 4); q += s_out;
            j.p_out_tx = reinterpret_cast<uint32_t *>(q); memcpy(q, j.out_tx.p, nout * 4); q += s_out;
            j.p_out_n = reinterpret_cast<uint32_t *>(q); memcpy(q, j.out_n.p, nout * 4);
        }
        j.small.need(64);
    }
    // everything a check thread does for one applied block
    void Check(Job &j, int dev) {
        auto t0 = Clock::now();
        uint32_t st[8] = {0};
        int rcw = bcp_gpu_utxo_wait(utxo, j.seq, st);
        auto t1 = Clock::now();
        const double tableLatency = Secs(j.tApplied, t1);
        j.ok = false;
        // header first, like CheckBlockHeader before anything of the body (src/validation.cpp:3202-3220); the scan thread
        // has already run it
        const uint8_t eqOk = j.eqOk;
        const int rcEq = j.rcEq;
        if (rcEq != BCP_OK) j.reason = std::string("device-error: ") + j.eqErr;
        auto t2 = Clock::now();
        double tScripts = 0;
        if (rcEq != BCP_OK) {
        } else if (!eqOk) {
            j.reason = "invalid-solution"; // CheckEquihashSolution (src/pow.cpp:298-323)
        } else if (rcw != BCP_OK) {
            j.reason = std::string("device-error: ") + bcp_gpu_last_error();
        } else if (memcmp(j.small.p, j.blk + 36, 32) != 0) {
            j.reason = "bad-txnmrklroot";
        } else if (*reinterpret_cast<const int *>(j.small.p + 32)) {
            j.reason = "bad-txns-duplicate";
        } else if (j.structErr) {
            j.reason = j.structErr;
        } else if (st[4] || st[5]) {
            j.reason = "device-error: coin table full";
        } else if (st[2]) {
            j.reason = "bad-txns-BIP30"; // src/validation.cpp:1975-1990
        } else if (st[0]) {
            j.reason = "bad-txns-inputs-missingorspent";
        } else {
            int all = 1;
            int rc = BCP_OK;
            if (j.nin > 0) // from the block and the spent-coin arrays that are already on the table's GPU
                rc = bcp_gpu_utxo_block_scripts(utxo, j.seq, dev, j.tx_off.p, j.in_tx.p, j.in_n.p, flags, j.errs.p, &all, st);
            tScripts = Secs(t2, Clock::now());
            if (rc != BCP_OK) {
                j.reason = std::string("device-error: ") + bcp_gpu_last_error();
            } else {
                j.ok = true;
                long firstOther = -1, nOther = 0;
                for (long k = 0; k < j.nin; k++) {
                    const uint8_t e = j.errs.p[k];
                    if (e == 0) continue;
                    if (e == 255) { // not the fast path's shape: decided below, after the inputs the device did decide
                        if (firstOther < 0) firstOther = k;
                        nOther++;
                        continue;
                    }
                    j.ok = false;
                    j.reason = "blk-bad-inputs (input " + std::to_string(k) + ", script error " + std::to_string(e) + ")"; // :2124-2128
                    break;
                }
                if (j.ok && nOther) j.ok = CheckOthers(j, firstOther, nOther);
            }
        }
        std::lock_guard<std::mutex> l(mu);
        phase[2] += Secs(t0, t1);
        phase[3] += Secs(t1, t2);
        phase[4] += tScripts;
        phase[7] += tableLatency;
    }

    // the inputs the fast path answered 255 for, through the caller's interpreter (bcp_other_scripts.h)
    bool CheckOthers(Job &j, long firstOther, long nOther) {
        if (!other) {
            j.reason = "script-not-in-fast-path (input " + std::to_string(firstOther) + ")";
            return false;
        }
        // the spent scripts and amounts were left on the device: fetch them now (rare path)
        std::vector<uint8_t> spk((size_t)j.nin * 80 + 65536);
        std::vector<uint32_t> off((size_t)j.nin + 1);
        std::vector<int64_t> amt((size_t)j.nin);
        if (bcp_gpu_utxo_fetch_spent(utxo, j.seq, spk.data(), spk.size(), off.data(), amt.data()) != BCP_OK) {
            j.reason = std::string("device-error: ") + bcp_gpu_last_error();
            return false;
        }
        std::vector<OtherInput> in;
        in.reserve((size_t)nOther);
        for (long k = firstOther; k < j.nin; k++)
            if (j.errs.p[k] == 255)
                in.push_back(OtherInput{(uint32_t)k, j.in_tx.p[k], j.in_n.p[k], spk.data() + off[(size_t)k], off[(size_t)k + 1] - off[(size_t)k], amt[(size_t)k]});
        std::string why;
        if (other(j.blk, j.len, j.tx_off.p, (size_t)j.ntx, in.data(), in.size(), flags, why)) return true;
        j.reason = why.empty() ? "blk-bad-inputs" : why;
        return false;
    }

    void Stop() {
        {
            std::lock_guard<std::mutex> l(mu);
            quit = true;
        }
        cv.notify_all();
        for (auto &t : threads) t.join();
        threads.clear();
        quit = false;
    }

    long Run(const std::vector<std::pair<const uint8_t *, size_t>> &blocks, std::string &reason, long &failed_index) {
        failed_index = -1;
        if (!FlushSeeds(reason)) return 0;
        const long n = (long)blocks.size();
        if (n == 0) return 0;
        const size_t R = jobs.size(), W = workers.size();
        for (Job &j : jobs) { j.state = FREE; j.index = -1; }
        long nextScan = 0;     // next block a scan thread takes
        long nextApply = 0;    // next block the connecting thread applies
        long nextCommit = 0;   // next verdict the connecting thread takes
        long stopAt = n;       // blocks at or beyond this ordinal are not started (set on the first failure)
        auto jobOf = [&](long b) -> Job & { return jobs[(size_t)b % R]; };
        // scan threads: block b may be scanned once its ring slot is free (its previous user was committed)
        for (int s = 0; s < nScan; s++)
            threads.emplace_back([&, s] {
                for (;;) {
                    long b;
                    {
                        std::unique_lock<std::mutex> l(mu);
                        cv.wait(l, [&] { return quit || (nextScan < stopAt && nextScan < nextCommit + (long)R && jobOf(nextScan).state == FREE); });
                        if (quit) return;
                        b = nextScan++;
                        Job &j = jobOf(b);
                        j.state = SCANNING;
                        j.index = b;
                        j.blk = blocks[(size_t)b].first;
                        j.len = blocks[(size_t)b].second;
                    }
                    auto t0 = Clock::now();
                    Scan(jobOf(b), (size_t)s);
                    const double dt = Secs(t0, Clock::now());
                    {
                        std::lock_guard<std::mutex> l(mu);
                        jobOf(b).state = SCANNED;
                        phase[5] += dt;
                    }
                    cv.notify_all();
                }
            });
        // check threads: thread w owns one (GPU, lane) pair and takes the next applied block whenever it is free (the
        // threads are not equally fast — the table's GPU is busier, a peer copy costs the others a little — and verdicts
        // are taken in order, so a fixed assignment would make every thread wait for the slowest one)
        long nextCheck = 0;
        for (size_t w = 0; w < W; w++)
            threads.emplace_back([&, w] {
                const int dev = workers[w];
                for (;;) {
                    long b;
                    {
                        std::unique_lock<std::mutex> l(mu);
                        cv.wait(l, [&] { return quit || (nextCheck < nextApply && jobOf(nextCheck).index == nextCheck && jobOf(nextCheck).state == APPLIED); });
                        if (quit) return;
                        b = nextCheck++;
                        jobOf(b).state = CHECKING;
                    }
                    Check(jobOf(b), dev);
                    {
                        std::lock_guard<std::mutex> l(mu);
                        jobOf(b).state = DONE;
                    }
                    cv.notify_all();
                }
            });
        // the connecting thread
        long connected = 0;
        std::string devErr;
        for (;;) {
            long b = -1;
            bool commit = false;
            auto tw0 = Clock::now();
            {
                std::unique_lock<std::mutex> l(mu);
                // what the thread is short of when it starts to wait (for the report): a scanned block, or room in flight
                const bool windowFull = nextApply - nextCommit >= (long)inflightMax;
                const bool scanMissing = !windowFull && nextApply < stopAt && !(jobOf(nextApply).index == nextApply && jobOf(nextApply).state == SCANNED);
                cv.wait(l, [&] {
                    if (nextCommit < nextApply && jobOf(nextCommit).state == DONE) return true;                      // a verdict is in
                    if (nextApply < stopAt && nextApply - nextCommit < (long)inflightMax && jobOf(nextApply).index == nextApply &&
                        jobOf(nextApply).state == SCANNED) return true;                                               // a block can go
                    return nextCommit >= stopAt && nextCommit >= nextApply;                                           // nothing left
                });
                const double waited = Secs(tw0, Clock::now());
                phase[1] += waited;
                if (windowFull) phase[9] += waited;
                else if (scanMissing) phase[10] += waited;
                if (nextCommit < nextApply && jobOf(nextCommit).state == DONE) { commit = true; b = nextCommit; }
                else if (nextApply < stopAt && nextApply - nextCommit < (long)inflightMax && jobOf(nextApply).index == nextApply &&
                         jobOf(nextApply).state == SCANNED) b = nextApply;
                else break;
            }
            Job &j = jobOf(b);
            if (!commit) {
                j.seq = ++seqCounter;
                if (j.nin < 0 || j.ntx < 1) { // framing broken: nothing to apply; the header still decides the reason
                    const size_t soln = ehN ? ((size_t)1 << ehK) * (ehN / (ehK + 1) + 1) / 8 : 0;
                    const size_t hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
                    j.ok = false;
                    j.reason = (ehN && j.len < hdr_len) ? "bad-blk-length" : "bad-blk-framing";
                    if (ehN && j.len >= hdr_len && j.rcEq == BCP_OK && !j.eqOk) j.reason = "invalid-solution";
                    j.seq = 0; // never applied
                    --seqCounter;
                    std::lock_guard<std::mutex> l(mu);
                    j.state = DONE;
                    nextApply++;
                    if (nextCheck == b) nextCheck++; // nothing to check: the check threads pass over it
                    stopAt = b + 1 < stopAt ? b + 1 : stopAt;
                    cv.notify_all();
                    continue;
                }
                auto ta0 = Clock::now();
                const int rc = bcp_gpu_utxo_apply_block(utxo, j.seq, j.blk, j.len, j.p_tx_off, (size_t)j.ntx, j.p_in_off, j.p_in_tx, (size_t)j.nin,
                                                        j.p_out_off, j.p_out_tx, j.p_out_n, (size_t)j.nout, nullptr, j.small.p,
                                                        reinterpret_cast<int *>(j.small.p + 32), nullptr, 0, nullptr, nullptr);
                j.tApplied = Clock::now();
                std::lock_guard<std::mutex> l(mu);
                phase[0] += Secs(ta0, j.tApplied);
                nextApply++;
                if (rc != BCP_OK) { // could not even be enqueued: treat like a failed block that was never applied
                    j.ok = false;
                    j.reason = std::string("device-error: ") + bcp_gpu_last_error();
                    j.seq = 0;
                    --seqCounter;
                    j.state = DONE;
                    if (nextCheck == b) nextCheck++;
                    stopAt = b + 1 < stopAt ? b + 1 : stopAt;
                } else {
                    j.state = APPLIED;
                }
                cv.notify_all();
                continue;
            }
            // a verdict, in chain order
            if (j.ok) {
                auto tc0 = Clock::now();
                bcp_gpu_utxo_confirm(utxo, j.seq);
                nInputs += (size_t)j.nin;
                memcpy(lastHash, j.hash, 32);
                connected++;
                std::lock_guard<std::mutex> l(mu);
                phase[6] += Secs(tc0, Clock::now());
                phase[8] += Secs(j.tApplied, tc0);
                j.state = FREE;
                j.index = -1;
                nextCommit++;
                cv.notify_all();
                continue;
            }
            // First failure: no further block is started; the ones applied after it finish their checks (their verdicts no
            // longer matter), then everything is undone newest first.
            reason = j.reason;
            failed_index = b;
            {
                std::unique_lock<std::mutex> l(mu);
                stopAt = b + 1 < stopAt ? b + 1 : stopAt;
                cv.wait(l, [&] {
                    for (long k = b + 1; k < nextApply; k++)
                        if (jobOf(k).state != DONE) return false;
                    return true;
                });
            }
            for (long k = nextApply - 1; k >= b; k--)
                if (jobOf(k).seq) bcp_gpu_utxo_rollback(utxo, jobOf(k).seq);
            break;
        }
        Stop();
        return connected;
    }

    std::vector<int> devs;
    std::vector<int> workers; // (GPU, lane) of each check thread
    uint32_t flags;
    unsigned ehN, ehK, log2Slots;
    int lanes_, nScan;
    bcp_utxo *utxo = nullptr;
    OtherScripts other;
    size_t inflightMax = 4;
    std::vector<Job> jobs;
    std::vector<std::thread> threads;
    std::mutex mu;
    std::condition_variable cv;
    bool quit = false;
    uint32_t seqCounter = 0;
    std::vector<uint8_t> seedOp, seedSpk;
    std::vector<uint32_t> seedOff;
    std::vector<int64_t> seedAmt;
    uint8_t lastHash[32] = {0};
    size_t nInputs = 0;
};

} // namespace bcp
#endif
