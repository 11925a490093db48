// tests/cpp/test_replay_ref.cpp — TEST INFRASTRUCTURE ONLY (never part of the product).
//
// The integration the drop-in is designed for, executed for real: the REFERENCE's own script interpreter
// (src/script/interpreter.cpp, compiled in place into oracle/_ref/libbcp_ref.so) is driven through THIS
// project's CCheckQueue (include/compat/checkqueue.h) with a signature checker whose VerifySignature override
// (the reference's plug point, src/script/interpreter.h:160-162) records into bcp::Recorder.  Every verdict is
// compared with the reference's plain CPU checker on the same script.
//
// What it covers: m-of-n OP_CHECKMULTISIG signed by arbitrary key subsets (the interpreter
// walks the keys, so assumed-true comparisons are refuted and the check is replayed), wrong signature order, bad
// signatures with and without SCRIPT_VERIFY_NULLFAIL, P2SH-wrapped multisig, scripts that branch on a failed
// OP_CHECKSIG (NOT / IF-ELSE), P2PKH, and the VerifiedHook contract (only device-confirmed triples reach it).
//
// Built by oracle/Makefile (target ref) into oracle/_ref/test_replay_ref; run on the GPU box by
// tests/test_gpu_replay.py.  Needs /root/reference to BUILD, not to run.
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <random>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "crypto/sha256.h"
#include "hash.h"
#include "primitives/transaction.h"
#include "pubkey.h"
#include "script/interpreter.h"
#include "script/script.h"
#include "streams.h"
#include "uint256.h"
#include "version.h"

#include "compat/checkqueue.h" // this project's drop-in (include/compat/), not the reference's src/checkqueue.h

extern "C" {
int ref_init(void);
int ref_pubkey_create(const uint8_t *seckey32, int compressed, uint8_t *out, size_t *outlen);
int ref_sign(const uint8_t *seckey32, const uint8_t *msg32, uint8_t *sig64);
int ref_sig_to_der(const uint8_t *sig64, uint8_t *der, size_t *derlen);
}

static std::mutex g_mu;
static std::set<std::string> g_confirmed;            // cookies the VerifiedHook received
static std::vector<std::pair<std::string, std::string>> g_seen; // (cookie, triple bytes) of everything recorded

static std::string cookie_of(const uint256 &sighash, const CPubKey &pk, const std::vector<uint8_t> &sig) {
    // same ingredients as CSignatureCache::ComputeEntry (src/script/sigcache.cpp:36-44), without the salt
    uint8_t out[32];
    CSHA256().Write(sighash.begin(), 32).Write(pk.begin(), pk.size()).Write(sig.data(), sig.size()).Finalize(out);
    return std::string((const char *)out, 32);
}

class RecordingChecker : public TransactionSignatureChecker {
public:
    using TransactionSignatureChecker::TransactionSignatureChecker;

protected:
    bool VerifySignature(const std::vector<uint8_t> &vchSig, const CPubKey &pubkey, const uint256 &sighash) const override {
        bcp::Recorder *r = bcp::Recorder::current();
        if (!r) {
            fprintf(stderr, "signature check outside a recording scope\n");
            abort();
        }
        const std::string ck = cookie_of(sighash, pubkey, vchSig);
        {
            std::lock_guard<std::mutex> l(g_mu);
            std::string t((const char *)sighash.begin(), 32);
            t.append((const char *)pubkey.begin(), pubkey.size());
            t.push_back('|');
            t.append((const char *)vchSig.data(), vchSig.size());
            g_seen.emplace_back(ck, t);
        }
        return r->record(sighash.begin(), pubkey.begin(), pubkey.size(), vchSig.data(), vchSig.size(), (const uint8_t *)ck.data());
    }
};

// the shape of the reference's CScriptCheck (src/validation.h:499-537)
struct ScriptCheck {
    CScript scriptPubKey;
    Amount amount;
    const CTransaction *ptxTo = nullptr;
    unsigned nIn = 0;
    uint32_t nFlags = 0;
    ScriptError error = SCRIPT_ERR_UNKNOWN_ERROR;
    ScriptCheck() : amount(0) {}
    ScriptCheck(const CScript &spk, Amount a, const CTransaction *tx, unsigned n, uint32_t f)
        : scriptPubKey(spk), amount(a), ptxTo(tx), nIn(n), nFlags(f) {}
    bool operator()() {
        return VerifyScript(ptxTo->vin[nIn].scriptSig, scriptPubKey, nFlags, RecordingChecker(ptxTo, nIn, amount), &error);
    }
    void swap(ScriptCheck &o) {
        scriptPubKey.swap(o.scriptPubKey);
        std::swap(amount, o.amount);
        std::swap(ptxTo, o.ptxTo);
        std::swap(nIn, o.nIn);
        std::swap(nFlags, o.nFlags);
        std::swap(error, o.error);
    }
};

struct Key {
    uint8_t sk[32];
    std::vector<uint8_t> pub;
};
static Key make_key(uint32_t seed, uint32_t i, bool compressed) {
    Key k;
    uint8_t in[8];
    memcpy(in, &seed, 4);
    memcpy(in + 4, &i, 4);
    CSHA256().Write(in, 8).Finalize(k.sk);
    uint8_t out[65];
    size_t ol = 65;
    if (!ref_pubkey_create(k.sk, compressed ? 1 : 0, out, &ol)) abort();
    k.pub.assign(out, out + ol);
    return k;
}
static std::vector<uint8_t> sign(const Key &k, const uint256 &h, uint8_t hashtype) {
    uint8_t s64[64], der[80];
    size_t dl = sizeof der;
    if (!ref_sign(k.sk, h.begin(), s64) || !ref_sig_to_der(s64, der, &dl)) abort();
    std::vector<uint8_t> v(der, der + dl);
    v.push_back(hashtype);
    return v;
}

struct Case {
    std::string name;
    CTransactionRef tx;
    CScript spk;
    Amount amount{0};
    uint32_t flags = 0;
    bool truth = false;
};

static const uint32_t BASE_FLAGS = SCRIPT_VERIFY_P2SH | SCRIPT_VERIFY_STRICTENC | SCRIPT_VERIFY_DERSIG | SCRIPT_VERIFY_LOW_S |
                                   SCRIPT_ENABLE_SIGHASH_FORKID;
static const uint8_t HT = SIGHASH_ALL | SIGHASH_FORKID;

// builds the spending transaction of one case; `build_sig` receives the sighash function and returns scriptSig
template <class F> static Case finish(const std::string &name, const CScript &spk, const CScript &scriptCode, uint32_t flags,
                                      uint32_t seed, F build_sig) {
    Case c;
    c.name = name;
    c.spk = spk;
    c.flags = flags;
    c.amount = Amount(int64_t(100000 + seed));
    CMutableTransaction m;
    m.nVersion = 2;
    m.vin.resize(1);
    uint8_t prev[32];
    CSHA256().Write((const uint8_t *)&seed, 4).Finalize(prev);
    memcpy(m.vin[0].prevout.hash.begin(), prev, 32);
    m.vin[0].prevout.n = seed % 3;
    m.vin[0].nSequence = 0xffffffff;
    m.vout.resize(1);
    m.vout[0].nValue = Amount(int64_t(90000));
    m.vout[0].scriptPubKey = CScript() << OP_1;
    const CTransaction pre(m);
    const uint256 h = SignatureHash(scriptCode, pre, 0, HT, c.amount, nullptr, flags);
    m.vin[0].scriptSig = build_sig(h);
    c.tx = MakeTransactionRef(CTransaction(m));
    ScriptError e;
    c.truth = VerifyScript(c.tx->vin[0].scriptSig, c.spk, c.flags, TransactionSignatureChecker(c.tx.get(), 0, c.amount), &e);
    return c;
}

static CScript multisig_script(int m, const std::vector<Key> &keys) {
    CScript s;
    s << CScript::EncodeOP_N(m);
    for (const Key &k : keys) s << k.pub;
    s << CScript::EncodeOP_N((int)keys.size()) << OP_CHECKMULTISIG;
    return s;
}

static void make_cases(std::vector<Case> &out) {
    std::mt19937 rng(20260101);
    uint32_t seed = 1;
    // --- multisig: every signer subset of small m-of-n, in order
    for (int n = 1; n <= 4; n++)
        for (int m = 1; m <= n; m++)
            for (unsigned mask = 0; mask < (1u << n); mask++) {
                if (__builtin_popcount(mask) != m) continue;
                for (int nullfail = 0; nullfail < 2; nullfail++) {
                    std::vector<Key> keys;
                    for (int i = 0; i < n; i++) keys.push_back(make_key(seed, i, (seed + i) % 3 != 0));
                    const CScript spk = multisig_script(m, keys);
                    const uint32_t fl = BASE_FLAGS | (nullfail ? SCRIPT_VERIFY_NULLFAIL | SCRIPT_VERIFY_NULLDUMMY : 0);
                    out.push_back(finish("multisig " + std::to_string(m) + "of" + std::to_string(n) + " mask " + std::to_string(mask),
                                         spk, spk, fl, seed, [&](const uint256 &h) {
                                             CScript s;
                                             s << OP_0;
                                             for (int i = 0; i < n; i++)
                                                 if (mask >> i & 1) s << sign(keys[i], h, HT);
                                             return s;
                                         }));
                    seed++;
                }
            }
    // --- larger multisig, random subsets, random defects
    for (int rep = 0; rep < 120; rep++) {
        const int n = 2 + rng() % 14, m = 1 + rng() % n;
        std::vector<Key> keys;
        for (int i = 0; i < n; i++) keys.push_back(make_key(seed, i, rng() % 4 != 0));
        std::vector<int> idx(n);
        for (int i = 0; i < n; i++) idx[i] = i;
        std::shuffle(idx.begin(), idx.end(), rng);
        std::vector<int> signers(idx.begin(), idx.begin() + m);
        const int defect = rng() % 6; // 0,1: none; 2: wrong order; 3: one corrupted sig; 4: one sig by a stranger; 5: a sig missing
        if (defect != 2) std::sort(signers.begin(), signers.end());
        const bool nullfail = rng() % 2;
        const bool p2sh = rng() % 3 == 0;
        const CScript redeem = multisig_script(m, keys);
        CScript spk = redeem;
        if (p2sh) {
            const uint160 id = Hash160(redeem.begin(), redeem.end());
            spk = CScript() << OP_HASH160 << std::vector<uint8_t>(id.begin(), id.end()) << OP_EQUAL;
        }
        const uint32_t fl = BASE_FLAGS | (nullfail ? SCRIPT_VERIFY_NULLFAIL : 0);
        const Key stranger = make_key(seed, 999, true);
        out.push_back(finish("random multisig " + std::to_string(m) + "of" + std::to_string(n) + " defect " + std::to_string(defect) +
                                 (p2sh ? " p2sh" : ""),
                             spk, redeem, fl, seed, [&](const uint256 &h) {
                                 CScript s;
                                 s << OP_0;
                                 const int victim = rng() % m;
                                 for (int j = 0; j < m; j++) {
                                     if (defect == 5 && j == victim) continue;
                                     std::vector<uint8_t> sg = sign(defect == 4 && j == victim ? stranger : keys[signers[j]], h, HT);
                                     if (defect == 3 && j == victim) sg[8 + rng() % 20] ^= 1 << (rng() % 8);
                                     s << sg;
                                 }
                                 if (p2sh) s << std::vector<uint8_t>(redeem.begin(), redeem.end());
                                 return s;
                             }));
        seed++;
    }
    // --- scripts that use the RESULT of a signature check (no NULLFAIL): CHECKSIG NOT, IF/ELSE on a CHECKSIG
    for (int rep = 0; rep < 40; rep++) {
        const Key a = make_key(seed, 0, true), b = make_key(seed, 1, rep & 1);
        const bool good_a = rng() % 2, good_b = rng() % 2;
        const int form = rep % 3;
        CScript spk;
        if (form == 0) spk = CScript() << a.pub << OP_CHECKSIG << OP_NOT;
        else if (form == 1) spk = CScript() << a.pub << OP_CHECKSIG << OP_IF << OP_1 << OP_ELSE << b.pub << OP_CHECKSIG << OP_ENDIF;
        else spk = CScript() << a.pub << OP_CHECKSIG << OP_SWAP << b.pub << OP_CHECKSIG << OP_BOOLOR;
        const uint32_t fl = BASE_FLAGS | (rep % 5 == 4 ? SCRIPT_VERIFY_NULLFAIL : 0);
        out.push_back(finish("branching form " + std::to_string(form), spk, spk, fl, seed, [&](const uint256 &h) {
            std::vector<uint8_t> sa = sign(a, h, HT), sb = sign(b, h, HT);
            if (!good_a) sa[10] ^= 4;
            if (!good_b) sb[12] ^= 2;
            CScript s;
            if (form == 0) s << sa;
            else s << sb << sa; // sa is consumed first
            return s;
        }));
        seed++;
    }
    // --- P2PKH, good and bad
    for (int rep = 0; rep < 40; rep++) {
        const Key a = make_key(seed, 0, rep % 4 != 0);
        const uint160 id = Hash160(a.pub.begin(), a.pub.end());
        const CScript spk = CScript() << OP_DUP << OP_HASH160 << std::vector<uint8_t>(id.begin(), id.end()) << OP_EQUALVERIFY << OP_CHECKSIG;
        const bool bad = rep % 5 == 0;
        out.push_back(finish("p2pkh", spk, spk, BASE_FLAGS | SCRIPT_VERIFY_NULLFAIL, seed, [&](const uint256 &h) {
            std::vector<uint8_t> sa = sign(a, h, HT);
            if (bad) sa[9] ^= 1;
            return CScript() << sa << a.pub;
        }));
        seed++;
    }
}

// ---- the reference's own script corpus (src/test/data/script_tests.json as binary scripts, tests/golden/script_tests.json.gz
// turned into a flat file by tests/test_gpu_replay.py): [u64 n] n x { u32 flags, i64 amount, u8 ok, u32 len + scriptSig,
// u32 len + scriptPubKey, u32 len + spending transaction }.  Each vector is evaluated like DoTest (src/test/script_tests.cpp:
// 153-170) — but through this project's CCheckQueue with the recording checker, signatures on the GPU.
struct Vector {
    uint32_t flags = 0;
    Amount amount{0};
    bool ok = false;
    CScript sig, spk;
    CTransactionRef tx;
};
static bool load_vectors(const char *path, std::vector<Vector> &out) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    std::vector<uint8_t> buf;
    uint8_t tmp[65536];
    size_t k;
    while ((k = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + k);
    fclose(f);
    size_t pos = 0;
    auto rd = [&](void *dst, size_t n) { memcpy(dst, &buf[pos], n); pos += n; };
    uint64_t n;
    rd(&n, 8);
    for (uint64_t i = 0; i < n; i++) {
        Vector v;
        int64_t amt;
        uint8_t ok;
        uint32_t len;
        rd(&v.flags, 4); rd(&amt, 8); rd(&ok, 1);
        v.amount = Amount(amt);
        v.ok = ok != 0;
        rd(&len, 4); v.sig = CScript(&buf[pos], &buf[pos] + len); pos += len;
        rd(&len, 4); v.spk = CScript(&buf[pos], &buf[pos] + len); pos += len;
        rd(&len, 4);
        CDataStream ss((const char *)&buf[pos], (const char *)&buf[pos] + len, SER_NETWORK, PROTOCOL_VERSION);
        pos += len;
        v.tx = MakeTransactionRef(CTransaction(deserialize, ss));
        out.push_back(std::move(v));
    }
    return pos == buf.size();
}
static int run_vectors(const char *path, bool gpu) {
    std::vector<Vector> vs;
    if (!load_vectors(path, vs)) {
        fprintf(stderr, "cannot read %s\n", path);
        return 2;
    }
    int failures = 0;
    size_t n_ok = 0;
    for (const Vector &v : vs) { // the fixture's expectation against the reference's plain CPU checker
        ScriptError e;
        const bool cpu = VerifyScript(v.sig, v.spk, v.flags, TransactionSignatureChecker(v.tx.get(), 0, v.amount), &e);
        if (cpu != v.ok) { printf("FIXTURE MISMATCH: expected %d, reference CPU %d\n", (int)v.ok, (int)cpu); failures++; }
        n_ok += v.ok;
    }
    printf("%zu vectors, %zu expected to verify\n", vs.size(), n_ok);
    if (!gpu) return failures ? 1 : 0;
    if (bcp_gpu_init(0) != BCP_OK) {
        fprintf(stderr, "no CUDA device: %s\n", bcp_gpu_last_error());
        return 2;
    }
    CCheckQueue<ScriptCheck> queue(16);
    std::vector<std::thread> workers;
    for (int i = 0; i < 3; i++) workers.emplace_back([&] { queue.Thread(); });
    g_seen.clear();
    // every vector alone
    for (size_t i = 0; i < vs.size(); i++) {
        const Vector &v = vs[i];
        CCheckQueueControl<ScriptCheck> control(&queue);
        std::vector<ScriptCheck> c;
        c.emplace_back(v.spk, v.amount, v.tx.get(), 0, v.flags);
        control.Add(c);
        const bool got = control.Wait();
        if (got != v.ok) { printf("MISMATCH vector %zu: queue %d, expected %d\n", i, (int)got, (int)v.ok); failures++; }
    }
    const size_t comparisons = g_seen.size();
    // all verifying vectors as one block
    {
        CCheckQueueControl<ScriptCheck> control(&queue);
        std::vector<ScriptCheck> c;
        for (const Vector &v : vs)
            if (v.ok) c.emplace_back(v.spk, v.amount, v.tx.get(), 0, v.flags);
        control.Add(c);
        if (!control.Wait()) { printf("MISMATCH: block of all verifying vectors rejected\n"); failures++; }
    }
    // that block plus one failing vector that reaches a signature comparison, a few times
    size_t k = 0;
    for (const Vector &bad : vs) {
        if (bad.ok) continue;
        const std::vector<uint8_t> raw(bad.spk.begin(), bad.spk.end()), raws(bad.sig.begin(), bad.sig.end());
        bool has_sigop = false;
        for (uint8_t b : raw) has_sigop = has_sigop || b == 0xac || b == 0xad || b == 0xae || b == 0xaf;
        for (uint8_t b : raws) has_sigop = has_sigop || b == 0xac || b == 0xae;
        if (!has_sigop || (k++ % 5) != 0) continue;
        CCheckQueueControl<ScriptCheck> control(&queue);
        std::vector<ScriptCheck> c;
        for (const Vector &v : vs)
            if (v.ok) c.emplace_back(v.spk, v.amount, v.tx.get(), 0, v.flags);
        c.emplace(c.begin() + c.size() / 2, bad.spk, bad.amount, bad.tx.get(), 0, bad.flags);
        control.Add(c);
        if (control.Wait()) { printf("MISMATCH: block with a failing vector accepted\n"); failures++; }
    }
    queue.Quit();
    for (auto &t : workers) t.join();
    printf("%zu signature comparisons went to the device while the vectors ran one by one\n", comparisons);
    if (failures) {
        printf("vectors FAILED: %d mismatches\n", failures);
        return 1;
    }
    printf("vectors ok\n");
    return 0;
}

int main(int argc, char **argv) {
    ref_init();
    if (argc > 2 && std::string(argv[1]) == "vectors-cpu") return run_vectors(argv[2], false);
    if (argc > 2 && std::string(argv[1]) == "vectors") return run_vectors(argv[2], true);
    if (argc > 1 && std::string(argv[1]) == "cases") { // CPU-only: the case set and the reference's verdicts on it
        std::vector<Case> cases;
        make_cases(cases);
        size_t n_true = 0;
        for (const Case &c : cases) n_true += c.truth;
        printf("cases %zu valid %zu\n", cases.size(), n_true);
        return 0;
    }
    if (bcp_gpu_init(0) != BCP_OK) {
        fprintf(stderr, "no CUDA device: %s\n", bcp_gpu_last_error());
        return 2;
    }
    bcp::VerifiedHook() = [](const uint8_t *c) {
        std::lock_guard<std::mutex> l(g_mu);
        g_confirmed.insert(std::string((const char *)c, 32));
    };
    std::vector<Case> cases;
    make_cases(cases);
    size_t n_true = 0;
    for (const Case &c : cases) n_true += c.truth;
    printf("%zu cases, %zu valid by the reference's CPU checker\n", cases.size(), n_true);

    CCheckQueue<ScriptCheck> queue(16);
    std::vector<std::thread> workers;
    for (int i = 0; i < 3; i++) workers.emplace_back([&] { queue.Thread(); });
    int failures = 0;
    // 1. every case alone: per-input verdict
    for (const Case &c : cases) {
        CCheckQueueControl<ScriptCheck> control(&queue);
        std::vector<ScriptCheck> v;
        v.emplace_back(c.spk, c.amount, c.tx.get(), 0, c.flags);
        control.Add(v);
        const bool got = control.Wait();
        if (got != c.truth) {
            printf("MISMATCH %s: queue %d, reference %d\n", c.name.c_str(), (int)got, (int)c.truth);
            failures++;
        }
    }
    // 2. blocks: all valid cases together must pass; adding any single invalid case must fail
    {
        CCheckQueueControl<ScriptCheck> control(&queue);
        std::vector<ScriptCheck> v;
        for (const Case &c : cases)
            if (c.truth) v.emplace_back(c.spk, c.amount, c.tx.get(), 0, c.flags);
        control.Add(v);
        if (!control.Wait()) {
            printf("MISMATCH: block of all valid cases rejected\n");
            failures++;
        }
    }
    size_t k = 0;
    for (const Case &bad : cases) {
        if (bad.truth || (k++ % 7) != 0) continue;
        CCheckQueueControl<ScriptCheck> control(&queue);
        for (int part = 0; part < 2; part++) { // two Add calls, the invalid one in the middle
            std::vector<ScriptCheck> v;
            size_t j = 0;
            for (const Case &c : cases)
                if (c.truth && (j++ % 2) == (size_t)part) v.emplace_back(c.spk, c.amount, c.tx.get(), 0, c.flags);
            if (part == 0) v.emplace_back(bad.spk, bad.amount, bad.tx.get(), 0, bad.flags);
            control.Add(v);
        }
        if (control.Wait()) {
            printf("MISMATCH: block with invalid case '%s' accepted\n", bad.name.c_str());
            failures++;
        }
    }
    queue.Quit();
    for (auto &t : workers) t.join();

    // 3. VerifiedHook: exactly the triples that verify on the CPU may have been confirmed
    size_t hook_true = 0;
    for (const auto &s : g_seen) {
        const std::string &t = s.second;
        // pubkey length from its header byte
        const uint8_t h0 = (uint8_t)t[32];
        const size_t pl = (h0 == 2 || h0 == 3) ? 33 : 65;
        CPubKey pk((const uint8_t *)&t[32], (const uint8_t *)&t[32] + pl);
        std::vector<uint8_t> sig(t.begin() + 32 + pl + 1, t.end());
        uint256 hh;
        memcpy(hh.begin(), t.data(), 32);
        const bool cpu = pk.Verify(hh, sig);
        const bool hooked = g_confirmed.count(s.first) != 0;
        hook_true += hooked;
        if (hooked && !cpu) {
            printf("MISMATCH: VerifiedHook fired for a triple the reference rejects\n");
            failures++;
        }
        // (a triple that verifies but was never confirmed is fine: its block was abandoned before it was sent)
    }
    printf("%zu signature comparisons recorded, %zu confirmed through the hook\n", g_seen.size(), hook_true);
    if (failures) {
        printf("replay FAILED: %d mismatches\n", failures);
        return 1;
    }
    printf("replay ok\n");
    return 0;
}
