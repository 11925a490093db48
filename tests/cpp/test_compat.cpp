// tests/cpp/test_compat.cpp — exercises the C++ drop-ins (include/compat/) the way the node would.
//   test_compat queue            CPU only: CCheckQueue / CCheckQueueControl contract with fake checks
//   test_compat gpu <file>       needs a GPU: signature triples from <file> through (a) the libsecp256k1-named
//                                C API and (b) CCheckQueue with recording checks; verdicts compared with the
//                                expected bytes stored in the file
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <thread>
#include <vector>

#include "checkqueue.h"
#include "secp256k1.h"

#define CHECK(x)                                                         \
    do {                                                                 \
        if (!(x)) {                                                      \
            fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #x); \
            exit(1);                                                     \
        }                                                                \
    } while (0)

static std::atomic<int> g_executed{0};
struct FakeCheck {
    int result = 1;
    FakeCheck() {}
    explicit FakeCheck(int r) : result(r) {}
    bool operator()() {
        g_executed++;
        return result != 0;
    }
    void swap(FakeCheck &o) { std::swap(result, o.result); }
};

static int test_queue() {
    CCheckQueue<FakeCheck> q(128);
    std::vector<std::thread> workers;
    for (int i = 0; i < 3; i++) workers.emplace_back([&] { q.Thread(); });
    // give the workers a moment to register; IsIdle must then hold
    for (int spin = 0; spin < 1000 && !q.IsIdle(); spin++) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    CHECK(q.IsIdle());
    for (int block = 0; block < 20; block++) {
        g_executed = 0;
        const int n = 1 + (block * 977) % 5000;
        const bool bad = block % 3 == 1;
        {
            CCheckQueueControl<FakeCheck> control(&q);
            int added = 0;
            while (added < n) {
                std::vector<FakeCheck> v;
                int k = std::min(n - added, 1 + (added * 31 + block) % 300);
                for (int i = 0; i < k; i++) v.emplace_back((bad && added + i == n / 2) ? 0 : 1);
                control.Add(v);
                CHECK((int)v.size() == k); // elements are moved out, the vector keeps its size (:147-149)
                added += k;
            }
            bool ok = control.Wait();
            CHECK(ok == !bad);
            if (!bad) CHECK(g_executed == n);   // every check ran exactly once
            else CHECK(g_executed <= n);        // early abort may skip the rest (:121-126)
        }
        CHECK(q.IsIdle());
    }
    { // destructor waits
        g_executed = 0;
        {
            CCheckQueueControl<FakeCheck> control(&q);
            std::vector<FakeCheck> v(1000, FakeCheck(1));
            control.Add(v);
        }
        CHECK(g_executed == 1000);
        CHECK(q.IsIdle());
    }
    { // nullptr queue: checks are skipped, Wait is true
        CCheckQueueControl<FakeCheck> control(nullptr);
        std::vector<FakeCheck> v(10, FakeCheck(0));
        control.Add(v);
        CHECK(control.Wait());
    }
    { // a queue without workers: the master does everything
        CCheckQueue<FakeCheck> solo(16);
        CCheckQueueControl<FakeCheck> control(&solo);
        std::vector<FakeCheck> v(100, FakeCheck(1));
        g_executed = 0;
        control.Add(v);
        CHECK(control.Wait());
        CHECK(g_executed == 100);
    }
    q.Quit();
    for (auto &t : workers) t.join();
    printf("queue ok\n");
    return 0;
}

// One recorded signature check: what CachingTransactionSignatureChecker::VerifySignature becomes
struct SigCheck {
    const uint8_t *msg = nullptr, *pub = nullptr, *der = nullptr;
    size_t publen = 0, derlen = 0;
    SigCheck() {}
    bool operator()() {
        bcp::Recorder *r = bcp::Recorder::current();
        if (!r) return false; // this test only runs inside the queue
        return r->record(msg, pub, publen, der, derlen);
    }
    void swap(SigCheck &o) {
        std::swap(msg, o.msg); std::swap(pub, o.pub); std::swap(der, o.der);
        std::swap(publen, o.publen); std::swap(derlen, o.derlen);
    }
};

static std::vector<uint8_t> der_of(const uint8_t *rs) { // minimal DER of (r, s)
    auto enc = [](const uint8_t *v, std::vector<uint8_t> &out) {
        int i = 0;
        while (i < 31 && v[i] == 0) i++;
        std::vector<uint8_t> b(v + i, v + 32);
        if (b[0] & 0x80) b.insert(b.begin(), 0);
        out.push_back(0x02); out.push_back((uint8_t)b.size());
        out.insert(out.end(), b.begin(), b.end());
    };
    std::vector<uint8_t> body;
    enc(rs, body); enc(rs + 32, body);
    std::vector<uint8_t> d{0x30, (uint8_t)body.size()};
    d.insert(d.end(), body.begin(), body.end());
    return d;
}

static int test_gpu(const char *path) {
    // file: u64 n | msg n*32 | pub n*65 | sig n*64 | verdict_raw n | verdict_cpubkey n
    std::ifstream f(path, std::ios::binary);
    CHECK(f.good());
    uint64_t n = 0;
    f.read((char *)&n, 8);
    std::vector<uint8_t> msg(n * 32), pub(n * 65), sig(n * 64), vraw(n), vcpk(n);
    f.read((char *)msg.data(), msg.size()); f.read((char *)pub.data(), pub.size()); f.read((char *)sig.data(), sig.size());
    f.read((char *)vraw.data(), n); f.read((char *)vcpk.data(), n);
    CHECK(f.good());

    // (a) the libsecp256k1-named API, as CPubKey::Verify uses it (src/pubkey.cpp:170-193)
    secp256k1_context *ctx = secp256k1_context_create(SECP256K1_CONTEXT_VERIFY);
    CHECK(ctx != nullptr);
    std::vector<secp256k1_pubkey> pks;
    std::vector<secp256k1_ecdsa_signature> sigs;
    std::vector<uint8_t> msgs, want;
    for (uint64_t i = 0; i < n && i < 400; i++) {
        const uint8_t *p = &pub[i * 65];
        size_t len = (p[0] == 2 || p[0] == 3) ? 33 : (p[0] == 4 || p[0] == 6 || p[0] == 7) ? 65 : 0;
        secp256k1_pubkey pk;
        secp256k1_ecdsa_signature s;
        int pok = len ? secp256k1_ec_pubkey_parse(ctx, &pk, p, len) : 0;
        int sok = secp256k1_ecdsa_signature_parse_compact(ctx, &s, &sig[i * 64]);
        if (!pok || !sok) { CHECK(vraw[i] == 0); continue; }
        if (i < 40) { // single-call path
            CHECK(secp256k1_ecdsa_verify(ctx, &s, &msg[i * 32], &pk) == vraw[i]);
            secp256k1_ecdsa_signature sn;
            secp256k1_ecdsa_signature_normalize(ctx, &sn, &s);
            CHECK(secp256k1_ecdsa_verify(ctx, &sn, &msg[i * 32], &pk) == vcpk[i]);
            unsigned char ser[65];
            size_t sl = 65;
            CHECK(secp256k1_ec_pubkey_serialize(ctx, ser, &sl, &pk, len == 33 ? SECP256K1_EC_COMPRESSED : SECP256K1_EC_UNCOMPRESSED));
            if (p[0] == 2 || p[0] == 3 || p[0] == 4) CHECK(sl == len && memcmp(ser, p, len) == 0);
        }
        pks.push_back(pk); sigs.push_back(s);
        msgs.insert(msgs.end(), &msg[i * 32], &msg[i * 32] + 32);
        want.push_back(vraw[i]);
    }
    std::vector<uint8_t> got(want.size());
    CHECK(secp256k1_ecdsa_verify_batch(ctx, sigs.data(), msgs.data(), pks.data(), want.size(), got.data()) == 1);
    CHECK(got == want);
    secp256k1_context_destroy(ctx);

    // (b) CCheckQueue with recording checks: all-valid subset => true; with one invalid => false
    std::vector<std::vector<uint8_t>> ders(n);
    for (uint64_t i = 0; i < n; i++) ders[i] = der_of(&sig[i * 64]);
    CCheckQueue<SigCheck> q(128);
    std::vector<std::thread> workers;
    for (int i = 0; i < 3; i++) workers.emplace_back([&] { q.Thread(); });
    for (int round = 0; round < 3; round++) {
        CCheckQueueControl<SigCheck> control(&q);
        std::vector<SigCheck> v;
        bool expect = true;
        int poisoned = 0;
        for (uint64_t i = 0; i < n; i++) {
            const uint8_t *p = &pub[i * 65];
            size_t len = (p[0] == 2 || p[0] == 3) ? 33 : (p[0] == 4 || p[0] == 6 || p[0] == 7) ? 65 : 0;
            bool take = vcpk[i] == 1;
            if (round == 1 && !vcpk[i] && len && !poisoned) { take = true; poisoned = 1; } // exactly one bad signature
            if (round == 2) take = true;                                                      // everything, good and bad
            if (!take) continue;
            SigCheck c;
            c.msg = &msg[i * 32]; c.pub = p; c.publen = len; c.der = ders[i].data(); c.derlen = ders[i].size();
            v.push_back(c);
            if (!vcpk[i]) expect = false;
            if (v.size() == 257) { control.Add(v); v.clear(); }
        }
        control.Add(v);
        bool ok = control.Wait();
        CHECK(ok == expect);
        CHECK(round == 0 ? ok : !ok);
    }
    q.Quit();
    for (auto &t : workers) t.join();
    printf("gpu ok: %llu triples\n", (unsigned long long)n);
    return 0;
}

int main(int argc, char **argv) {
    if (argc >= 2 && !strcmp(argv[1], "queue")) return test_queue();
    if (argc >= 3 && !strcmp(argv[1], "gpu")) return test_gpu(argv[2]);
    fprintf(stderr, "usage: test_compat queue | gpu <file>\n");
    return 2;
}
