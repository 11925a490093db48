// include/compat/bcp_recorder.h — level B2 glue: the recording signature checker behind the CCheckQueue
// drop-in.  The reference's plug point is the protected virtual
//   TransactionSignatureChecker::VerifySignature(vchSig /*DER, hashtype stripped*/, pubkey, sighash)
// (src/script/interpreter.h:160-162), overridden by CachingTransactionSignatureChecker
// (src/script/sigcache.cpp:83-99).  With this project the override asks bcp::Recorder::current(): when a
// recorder is installed (inside CCheckQueue worker/master loops) the triple is recorded and `true` is
// returned; the verdict is settled in CCheckQueue::Wait().  The recorder never touches the signature cache, so
// an optimistic `true` can not poison it.
#ifndef BCP_COMPAT_RECORDER_H
#define BCP_COMPAT_RECORDER_H

#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

#include "../bcp_gpu.h"

namespace bcp {

class Recorder {
public:
    std::vector<uint8_t> msg32, pub65, sig64; // n x 32, n x 65, n x 64 (the B4 batch layout)
    size_t size() const { return msg32.size() / 32; }

    //! CPubKey::Verify up to the point where arithmetic starts (src/pubkey.cpp:170-185): invalid key
    //! encoding, empty signature or a structurally broken DER fail immediately; otherwise the triple is
    //! queued for the device and the call reports success.
    bool record(const uint8_t *sighash32, const uint8_t *pubkey, size_t publen, const uint8_t *der, size_t derlen) {
        if (publen == 0) return false;
        const uint8_t h = pubkey[0];
        const size_t want = (h == 2 || h == 3) ? 33 : (h == 4 || h == 6 || h == 7) ? 65 : 0; // src/pubkey.h:47-51
        if (want == 0 || want != publen) return false;
        if (derlen == 0) return false;
        uint8_t rs[64];
        if (!bcp_der_lax_parse(der, derlen, rs)) return false;
        const size_t n = size();
        msg32.resize((n + 1) * 32);
        pub65.resize((n + 1) * 65, 0);
        sig64.resize((n + 1) * 64);
        memcpy(&msg32[n * 32], sighash32, 32);
        memcpy(&pub65[n * 65], pubkey, publen);
        memcpy(&sig64[n * 64], rs, 64);
        return true;
    }

    static Recorder *&slot() {
        static thread_local Recorder *cur = nullptr;
        return cur;
    }
    static Recorder *current() { return slot(); }

    //! Installs a recorder for the calling thread for the lifetime of the scope.
    struct Scope {
        Recorder *prev;
        explicit Scope(Recorder *r) : prev(slot()) { slot() = r; }
        ~Scope() { slot() = prev; }
    };
};

//! Devices the deferred batches are spread over.  Default: device 0 only; SetDevices() widens it.
inline std::vector<int> &Devices() {
    static std::vector<int> d{0};
    return d;
}
inline bool SetDevices(const std::vector<int> &devs) {
    for (int d : devs)
        if (bcp_gpu_init(d) != BCP_OK) return false;
    Devices() = devs;
    return !devs.empty();
}

//! Concatenates the per-thread batches, shards them by index range over Devices(), verifies, ANDs.
inline bool VerifyRecorded(const std::vector<std::unique_ptr<Recorder>> &recs) {
    size_t n = 0;
    for (const auto &r : recs) n += r->size();
    if (n == 0) return true; // nothing reached a signature check: the GPU is not touched at all
    std::vector<uint8_t> msg(n * 32), pub(n * 65), sig(n * 64), ok(n, 0);
    size_t o = 0;
    for (const auto &r : recs) {
        const size_t k = r->size();
        if (!k) continue;
        memcpy(&msg[o * 32], r->msg32.data(), k * 32);
        memcpy(&pub[o * 65], r->pub65.data(), k * 65);
        memcpy(&sig[o * 64], r->sig64.data(), k * 64);
        o += k;
    }
    const std::vector<int> devs = Devices();
    const size_t G = devs.size();
    const size_t per = (n + G - 1) / G;
    std::vector<int> rc(G, BCP_OK);
    std::vector<std::thread> th;
    for (size_t g = 0; g < G; g++) {
        const size_t lo = std::min(n, g * per), hi = std::min(n, lo + per);
        if (lo >= hi) continue;
        auto work = [&, g, lo, hi] {
            rc[g] = bcp_gpu_init(devs[g]);
            if (rc[g] == BCP_OK)
                rc[g] = bcp_gpu_ecdsa_verify_batch(devs[g], &msg[lo * 32], &pub[lo * 65], &sig[lo * 64], hi - lo,
                                                   BCP_ECDSA_FLAG_NORMALIZE_S, &ok[lo]);
        };
        if (G == 1) work();
        else th.emplace_back(work);
    }
    for (auto &t : th) t.join();
    for (int r : rc)
        if (r != BCP_OK) return false; // device failure: never a silent accept
    for (uint8_t v : ok)
        if (!v) return false;
    return true;
}

} // namespace bcp
#endif
