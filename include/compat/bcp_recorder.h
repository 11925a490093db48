// include/compat/bcp_recorder.h — level B2 glue: the recording signature checker behind the CCheckQueue
// drop-in.  The reference's plug point is the protected virtual
//   TransactionSignatureChecker::VerifySignature(vchSig /*DER, hashtype stripped*/, pubkey, sighash)
// (src/script/interpreter.h:160-162), overridden by CachingTransactionSignatureChecker
// (src/script/sigcache.cpp:83-99).  With this project the override asks bcp::Recorder::current(): when a
// recorder is installed (inside CCheckQueue worker/master loops) the triple is recorded and `true` is
// returned; the verdict is settled in CCheckQueue::Wait().  The recorder never touches the signature cache, so
// an optimistic `true` can not poison it; an integration that wants the cache filled passes the
// cache entry as the triple's cookie and inserts it from bcp::VerifiedHook(), i.e. only after a device said VALID.
#ifndef BCP_COMPAT_RECORDER_H
#define BCP_COMPAT_RECORDER_H

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../bcp_gpu.h"

namespace bcp {

//! Verdicts the devices have already returned for this block, keyed by the exact bytes that were verified
//! (sighash ‖ pubkey padded to 65 ‖ r ‖ s).  Used when a check is replayed (see checkqueue.h, "replay").
typedef std::unordered_map<std::string, bool> KnownVerdicts;

class Recorder {
public:
    std::vector<uint8_t> msg32, pub65, sig64; // n x 32, n x 65, n x 64 (the B4 batch layout)
    std::vector<uint8_t> cookie32, hasCookie; // optional caller tag per triple (e.g. the signature-cache entry)
    const KnownVerdicts *known = nullptr;     // replay mode: answer from here when the triple is known
    size_t size() const { return msg32.size() / 32; }

    static std::string key(const uint8_t *m, const uint8_t *p, const uint8_t *s) {
        std::string k(161, '\0');
        memcpy(&k[0], m, 32);
        memcpy(&k[32], p, 65);
        memcpy(&k[97], s, 64);
        return k;
    }
    std::string keyAt(size_t i) const { return key(&msg32[i * 32], &pub65[i * 65], &sig64[i * 64]); }

    //! CPubKey::Verify up to the point where arithmetic starts (src/pubkey.cpp:170-185): invalid key
    //! encoding, empty signature or a structurally broken DER fail immediately; otherwise the triple is
    //! queued for the device and the call reports success — unless a verdict for exactly this triple is
    //! already known (replay), in which case that verdict is the answer and nothing is queued.
    //! `cookie` (32 bytes, optional) is handed to the VerifiedHook once the device has confirmed the triple.
    bool record(const uint8_t *sighash32, const uint8_t *pubkey, size_t publen, const uint8_t *der, size_t derlen,
                const uint8_t *cookie = nullptr) {
        if (publen == 0) return false;
        const uint8_t h = pubkey[0];
        const size_t want = (h == 2 || h == 3) ? 33 : (h == 4 || h == 6 || h == 7) ? 65 : 0; // src/pubkey.h:47-51
        if (want == 0 || want != publen) return false;
        if (derlen == 0) return false;
        uint8_t rs[64];
        if (!bcp_der_lax_parse(der, derlen, rs)) return false;
        uint8_t p65[65] = {0};
        memcpy(p65, pubkey, publen);
        if (known) {
            auto it = known->find(key(sighash32, p65, rs));
            if (it != known->end()) return it->second;
        }
        const size_t n = size();
        msg32.resize((n + 1) * 32);
        pub65.resize((n + 1) * 65);
        sig64.resize((n + 1) * 64);
        cookie32.resize((n + 1) * 32, 0);
        hasCookie.push_back(cookie ? 1 : 0);
        memcpy(&msg32[n * 32], sighash32, 32);
        memcpy(&pub65[n * 65], p65, 65);
        memcpy(&sig64[n * 64], rs, 64);
        if (cookie) memcpy(&cookie32[n * 32], cookie, 32);
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

//! Called (from the thread inside CCheckQueue::Wait) once per triple that carried a cookie and that a device
//! confirmed VALID — the place to do what CachingTransactionSignatureChecker does after a successful Verify
//! (signatureCache.Set(entry), src/script/sigcache.cpp:95-97).  Never called for a triple that was only assumed.
inline std::function<void(const uint8_t *cookie32)> &VerifiedHook() {
    static std::function<void(const uint8_t *)> h;
    return h;
}

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

//! Verifies n triples, sharded by index range over Devices(); ok[i] = verdict.  false = device failure.
inline bool VerifyTriples(const uint8_t *msg, const uint8_t *pub, const uint8_t *sig, size_t n, uint8_t *ok,
                          uint32_t flags = BCP_ECDSA_FLAG_NORMALIZE_S) {
    if (n == 0) return true;
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
                rc[g] = bcp_gpu_ecdsa_verify_batch(devs[g], msg + lo * 32, pub + lo * 65, sig + lo * 64, hi - lo, flags, ok + lo);
        };
        if (G == 1) work();
        else th.emplace_back(work);
    }
    for (auto &t : th) t.join();
    for (int r : rc)
        if (r != BCP_OK) return false; // device failure: never a silent accept
    return true;
}

//! Verdicts for everything the given recorders hold, in recorder order (ok.size() == total triples); fires the
//! VerifiedHook for confirmed triples that carried a cookie.  false = device failure.
inline bool VerifyRecorders(const std::vector<Recorder *> &recs, std::vector<uint8_t> &ok) {
    size_t n = 0;
    for (const Recorder *r : recs) n += r->size();
    ok.assign(n, 0);
    if (n == 0) return true; // nothing reached a signature check: the GPU is not touched at all
    std::vector<uint8_t> msg(n * 32), pub(n * 65), sig(n * 64);
    size_t o = 0;
    for (const Recorder *r : recs) {
        const size_t k = r->size();
        if (!k) continue;
        memcpy(&msg[o * 32], r->msg32.data(), k * 32);
        memcpy(&pub[o * 65], r->pub65.data(), k * 65);
        memcpy(&sig[o * 64], r->sig64.data(), k * 64);
        o += k;
    }
    if (!VerifyTriples(msg.data(), pub.data(), sig.data(), n, ok.data())) return false;
    if (VerifiedHook()) {
        o = 0;
        for (const Recorder *r : recs)
            for (size_t i = 0; i < r->size(); i++, o++)
                if (ok[o] && r->hasCookie[i]) VerifiedHook()(&r->cookie32[i * 32]);
    }
    return true;
}

//! All-or-nothing form: true iff every recorded triple verifies.
inline bool VerifyRecorded(const std::vector<std::unique_ptr<Recorder>> &recs) {
    std::vector<Recorder *> p;
    for (const auto &r : recs) p.push_back(r.get());
    std::vector<uint8_t> ok;
    if (!VerifyRecorders(p, ok)) return false;
    for (uint8_t v : ok)
        if (!v) return false;
    return true;
}

} // namespace bcp
#endif
