// tests/cpp/bench_queue_ref.cpp — TEST / BENCH INFRASTRUCTURE ONLY (never part of the product).
//
// Level B1 of the boundary measured with the reference's own code on the host side: every input of a serialized block
// becomes a CScriptCheck-shaped closure that calls the REFERENCE's VerifyScript (src/script/interpreter.cpp, compiled in
// place into oracle/_ref/libbcp_ref.so) with PrecomputedTransactionData, and the closures go through THIS project's
// CCheckQueue drop-in (include/compat/checkqueue.h) on all host threads.
//   arm "gpu": the checker's VerifySignature records into bcp::Recorder; Wait() verifies the batch on the GPU
//   arm "cpu": the reference's own TransactionSignatureChecker::VerifySignature (CPubKey::Verify -> libsecp256k1)
// Same queue, same threads, same interpreter; only where the signature arithmetic runs differs.  Prints one JSON line.
//
// usage: bench_queue_ref <file> <threads> <reps>      file = u64 len | block | u64 n | u32 spk_off[n+1] | spk | i64 amount[n]
// Built by oracle/Makefile (target ref).  Needs /root/reference to BUILD, not to run.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <thread>
#include <vector>

#include "primitives/block.h"
#include "primitives/transaction.h"
#include "pubkey.h"
#include "script/interpreter.h"
#include "streams.h"
#include "version.h"

#include "compat/checkqueue.h" // this project's drop-in

extern "C" int ref_init(void);

static bool g_record = true;

class Checker : public TransactionSignatureChecker {
public:
    using TransactionSignatureChecker::TransactionSignatureChecker;

protected:
    bool VerifySignature(const std::vector<uint8_t> &vchSig, const CPubKey &pubkey, const uint256 &sighash) const override {
        if (!g_record) return TransactionSignatureChecker::VerifySignature(vchSig, pubkey, sighash);
        return bcp::Recorder::current()->record(sighash.begin(), pubkey.begin(), pubkey.size(), vchSig.data(), vchSig.size());
    }
};

struct ScriptCheck { // the shape of CScriptCheck (src/validation.h:499-537)
    CScript scriptPubKey;
    Amount amount;
    const CTransaction *ptxTo = nullptr;
    const PrecomputedTransactionData *txdata = nullptr;
    unsigned nIn = 0;
    uint32_t nFlags = 0;
    ScriptCheck() : amount(0) {}
    ScriptCheck(const CScript &spk, Amount a, const CTransaction *tx, const PrecomputedTransactionData *td, unsigned n, uint32_t f)
        : scriptPubKey(spk), amount(a), ptxTo(tx), txdata(td), nIn(n), nFlags(f) {}
    bool operator()() {
        ScriptError e;
        return VerifyScript(ptxTo->vin[nIn].scriptSig, scriptPubKey, nFlags, Checker(ptxTo, nIn, amount, *txdata), &e);
    }
    void swap(ScriptCheck &o) {
        scriptPubKey.swap(o.scriptPubKey);
        std::swap(amount, o.amount);
        std::swap(ptxTo, o.ptxTo);
        std::swap(txdata, o.txdata);
        std::swap(nIn, o.nIn);
        std::swap(nFlags, o.nFlags);
    }
};

int main(int argc, char **argv) {
    if (argc < 4) return 2;
    ref_init();
    std::ifstream f(argv[1], std::ios::binary);
    uint64_t len = 0, n = 0;
    f.read((char *)&len, 8);
    std::vector<char> blk(len);
    f.read(blk.data(), len);
    f.read((char *)&n, 8);
    std::vector<uint32_t> off(n + 1);
    f.read((char *)off.data(), (n + 1) * 4);
    std::vector<uint8_t> spk(off[n]);
    f.read((char *)spk.data(), off[n]);
    std::vector<int64_t> amt(n);
    f.read((char *)amt.data(), n * 8);
    if (!f) { fprintf(stderr, "short input file\n"); return 2; }
    const int threads = atoi(argv[2]), reps = atoi(argv[3]);
    const uint32_t flags = SCRIPT_VERIFY_P2SH | SCRIPT_VERIFY_STRICTENC | SCRIPT_VERIFY_DERSIG | SCRIPT_VERIFY_LOW_S |
                           SCRIPT_VERIFY_NULLFAIL | SCRIPT_ENABLE_SIGHASH_FORKID | SCRIPT_VERIFY_CHECKLOCKTIMEVERIFY |
                           SCRIPT_VERIFY_CHECKSEQUENCEVERIFY;
    CBlock block;
    {
        CDataStream ss(blk.data(), blk.data() + blk.size(), SER_NETWORK, PROTOCOL_VERSION);
        ss >> block;
    }
    if (bcp_gpu_init(0) != BCP_OK) { fprintf(stderr, "no CUDA device: %s\n", bcp_gpu_last_error()); return 2; }

    CCheckQueue<ScriptCheck> queue(128); // the node's batch size (src/validation.cpp: scriptcheckqueue(128))
    std::vector<std::thread> workers;
    for (int i = 0; i < threads - 1; i++) workers.emplace_back([&] { queue.Thread(); });

    double produce_s[2] = {0, 0};
    auto run = [&](bool record, double &best, bool &ok) {
        g_record = record;
        best = 1e30;
        ok = true;
        for (int r = 0; r < reps + 1; r++) { // first pass is the warm-up
            auto t0 = std::chrono::steady_clock::now();
            {
                // ConnectBlock's loop (src/validation.cpp:2063-2131): per transaction precompute the sighash midstates,
                // queue one check per input
                std::vector<PrecomputedTransactionData> txd;
                txd.reserve(block.vtx.size());
                CCheckQueueControl<ScriptCheck> control(&queue);
                size_t j = 0;
                for (const auto &tx : block.vtx) {
                    txd.emplace_back(*tx);
                    if (tx->IsCoinBase()) continue;
                    std::vector<ScriptCheck> v;
                    v.reserve(tx->vin.size());
                    for (unsigned i = 0; i < tx->vin.size(); i++, j++)
                        v.emplace_back(CScript(spk.begin() + off[j], spk.begin() + off[j + 1]), Amount(amt[j]), tx.get(), &txd.back(), i, flags);
                    control.Add(v);
                }
                if (r > 0) produce_s[record ? 0 : 1] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                ok = control.Wait() && ok;
            }
            const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (r > 0 && s < best) best = s;
        }
    };
    double gpu_s, cpu_s;
    bool gpu_ok, cpu_ok;
    run(true, gpu_s, gpu_ok);
    run(false, cpu_s, cpu_ok);
    queue.Quit();
    for (auto &t : workers) t.join();
    printf("{\"inputs\": %llu, \"threads\": %d, \"gpu_ms\": %.3f, \"cpu_ms\": %.3f, \"gpu_ok\": %s, \"cpu_ok\": %s, "
           "\"gpu_inputs_per_s\": %.0f, \"cpu_inputs_per_s\": %.0f, \"speedup\": %.2f, \"gpu_producer_loop_ms\": %.3f, "
           "\"cpu_producer_loop_ms\": %.3f}\n",
           (unsigned long long)n, threads, gpu_s * 1e3, cpu_s * 1e3, gpu_ok ? "true" : "false", cpu_ok ? "true" : "false", n / gpu_s,
           n / cpu_s, cpu_s / gpu_s, produce_s[0] * 1e3, produce_s[1] * 1e3);
    return gpu_ok == cpu_ok ? 0 : 1;
}
