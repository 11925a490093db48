// bcp_chain.cpp — C face over bcp::ChainValidator (include/compat/bcp_chain.h) so that the tests and bench.py can
// drive the host-side block connector through ctypes.  Built into bitcoincashplus_b200/libbcp_chain.so by
// csrc/Makefile (plain g++, links libbcp_b200.so).
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/compat/bcp_chain.h"
#include "../../include/compat/bcp_chain_gpu.h"
#include "../../include/compat/bcp_recorder.h"

#define CHAIN_API extern "C" __attribute__((visibility("default")))

CHAIN_API void *bcp_chain_new_threads(const int *devices, int ndev, uint32_t script_flags, unsigned eh_n, unsigned eh_k,
                                      int walk_threads, char *err, size_t errlen) {
    std::vector<int> d(devices, devices + ndev);
    auto *v = new bcp::ChainValidator(d, script_flags, eh_n, eh_k, walk_threads);
    std::string e;
    if (!v->Init(e)) {
        if (err && errlen) snprintf(err, errlen, "%s", e.c_str());
        delete v;
        return nullptr;
    }
    return v;
}
CHAIN_API void *bcp_chain_new(const int *devices, int ndev, uint32_t script_flags, unsigned eh_n, unsigned eh_k, char *err,
                              size_t errlen) {
    std::vector<int> d(devices, devices + ndev);
    auto *v = new bcp::ChainValidator(d, script_flags, eh_n, eh_k);
    std::string e;
    if (!v->Init(e)) {
        if (err && errlen) snprintf(err, errlen, "%s", e.c_str());
        delete v;
        return nullptr;
    }
    return v;
}
CHAIN_API void bcp_chain_free(void *h) { delete static_cast<bcp::ChainValidator *>(h); }
CHAIN_API void bcp_chain_add_coin(void *h, const uint8_t *txid32, uint32_t vout, const uint8_t *spk, size_t len, int64_t amount) {
    static_cast<bcp::ChainValidator *>(h)->AddCoin(txid32, vout, spk, len, amount);
}
CHAIN_API size_t bcp_chain_utxo_size(void *h) { return static_cast<bcp::ChainValidator *>(h)->UtxoSize(); }
CHAIN_API size_t bcp_chain_inputs_checked(void *h) { return static_cast<bcp::ChainValidator *>(h)->InputsChecked(); }
CHAIN_API int bcp_chain_connect_block_hint(void *h, const uint8_t *blk, size_t len, const uint8_t *next, size_t next_len, char *reason,
                                           size_t reasonlen) {
    std::string r;
    bool ok = static_cast<bcp::ChainValidator *>(h)->ConnectBlock(blk, len, r, next, next_len);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    return ok ? 1 : 0;
}
CHAIN_API int bcp_chain_connect_block(void *h, const uint8_t *blk, size_t len, char *reason, size_t reasonlen) {
    std::string r;
    bool ok = static_cast<bcp::ChainValidator *>(h)->ConnectBlock(blk, len, r);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    return ok ? 1 : 0;
}
CHAIN_API long bcp_chain_connect_file(void *h, const uint8_t *data, size_t len, const uint8_t *magic4, char *reason, size_t reasonlen,
                                      long *failed_index) {
    std::string r;
    long fi = -1;
    long n = static_cast<bcp::ChainValidator *>(h)->ConnectFile(data, len, magic4, r, fi);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    if (failed_index) *failed_index = fi;
    return n;
}
CHAIN_API void bcp_chain_last_hash(void *h, uint8_t *out32) { memcpy(out32, static_cast<bcp::ChainValidator *>(h)->LastBlockHash(), 32); }
CHAIN_API void bcp_chain_phase_seconds(void *h, double *out5) {
    for (int i = 0; i < 5; i++) out5[i] = static_cast<bcp::ChainValidator *>(h)->phase[i];
}

CHAIN_API void bcp_chain_walk_seconds(void *h, double out4[4]) {
    auto *v = static_cast<bcp::ChainValidator *>(h);
    for (int i = 0; i < 4; i++) out4[i] = v->walk[i];
}

// One caller, one batch, every GPU: bcp::VerifyTriples (include/compat/bcp_recorder.h) splits the batch by index range over
// `devices`, one host thread per device, and the per-signature verdicts land in one array; *all_ok is their AND, the
// `fAllOk &= fOk` of the check queue (src/checkqueue.h:77).  Returns 0, or -1 on a device failure (never a verdict then).
CHAIN_API int bcp_sharded_ecdsa_verify(const int *devices, int ndev, const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64,
                                       size_t n, uint32_t flags, uint8_t *ok, int *all_ok) {
    if (!bcp::SetDevices(std::vector<int>(devices, devices + ndev))) return -1;
    if (!bcp::VerifyTriples(msg32, pub65, sig64, n, ok, flags)) return -1;
    if (all_ok) {
        int a = 1;
        for (size_t i = 0; i < n; i++) a &= ok[i] != 0;
        *all_ok = a;
    }
    return 0;
}

// ---- bcp::DeviceChain (include/compat/bcp_chain_gpu.h): the connector whose coin table lives on the device ----
CHAIN_API void *bcp_dchain_new(const int *devices, int ndev, uint32_t script_flags, unsigned eh_n, unsigned eh_k, unsigned log2_slots,
                               int lanes, int scan_threads, char *err, size_t errlen) {
    std::vector<int> d(devices, devices + ndev);
    auto *v = new bcp::DeviceChain(d, script_flags, eh_n, eh_k, log2_slots, lanes, scan_threads);
    std::string e;
    if (!v->Init(e)) {
        if (err && errlen) snprintf(err, errlen, "%s", e.c_str());
        delete v;
        return nullptr;
    }
    return v;
}
CHAIN_API void bcp_dchain_reserve(void *h, size_t max_block_bytes) { static_cast<bcp::DeviceChain *>(h)->Reserve(max_block_bytes); }
CHAIN_API void bcp_dchain_free(void *h) { delete static_cast<bcp::DeviceChain *>(h); }
// GPUs the connector deals blocks to: the constructor's list minus those without direct peer access to the table's GPU
CHAIN_API int bcp_dchain_gpus_in_use(void *h) { return (int)static_cast<bcp::DeviceChain *>(h)->GpusInUse(); }
CHAIN_API void bcp_dchain_add_coin(void *h, const uint8_t *txid32, uint32_t vout, const uint8_t *spk, size_t len, int64_t amount) {
    static_cast<bcp::DeviceChain *>(h)->AddCoin(txid32, vout, spk, len, amount);
}
CHAIN_API size_t bcp_dchain_utxo_size(void *h) { return (size_t)static_cast<bcp::DeviceChain *>(h)->UtxoSize(); }
CHAIN_API size_t bcp_dchain_inputs_checked(void *h) { return static_cast<bcp::DeviceChain *>(h)->InputsChecked(); }
CHAIN_API int bcp_dchain_connect_block(void *h, const uint8_t *blk, size_t len, char *reason, size_t reasonlen) {
    std::string r;
    bool ok = static_cast<bcp::DeviceChain *>(h)->ConnectBlock(blk, len, r);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    return ok ? 1 : 0;
}
CHAIN_API long bcp_dchain_connect_file(void *h, const uint8_t *data, size_t len, const uint8_t *magic4, char *reason, size_t reasonlen,
                                       long *failed_index) {
    std::string r;
    long fi = -1;
    long n = static_cast<bcp::DeviceChain *>(h)->ConnectFile(data, len, magic4, r, fi);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    if (failed_index) *failed_index = fi;
    return n;
}
CHAIN_API void bcp_dchain_last_hash(void *h, uint8_t *out32) { memcpy(out32, static_cast<bcp::DeviceChain *>(h)->LastBlockHash(), 32); }
CHAIN_API void bcp_dchain_phase_seconds(void *h, double *out11) {
    for (int i = 0; i < 11; i++) out11[i] = static_cast<bcp::DeviceChain *>(h)->phase[i];
}

// Measurement helper for bench.py (not part of the validation path): how fast `threads` host threads can READ memory that
// is not in their caches — each sums its own buffer of `bytes_per_thread` bytes once, after writing it.  The chain connectors
// move every block through host memory (framing scan, upload), so this is the box's ceiling for them; returns GB/s.
CHAIN_API double bcp_host_read_bandwidth(int threads, size_t bytes_per_thread) {
    if (threads < 1) threads = 1;
    std::vector<std::vector<uint64_t>> bufs((size_t)threads);
    for (auto &b : bufs) b.assign(bytes_per_thread / 8, 0x0101010101010101ull);
    std::vector<uint64_t> sums((size_t)threads, 0);
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([&, t] {
            uint64_t s = 0;
            const uint64_t *p = bufs[(size_t)t].data();
            const size_t n = bufs[(size_t)t].size();
            for (size_t i = 0; i < n; i += 8) s += p[i]; /* one word per 64-byte line */
            sums[(size_t)t] = s;
        });
    for (auto &x : th) x.join();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    uint64_t keep = 0;
    for (uint64_t v : sums) keep += v;
    return keep == 0 ? 0.0 : (double)threads * (double)bytes_per_thread / dt / 1e9;
}
