// tests/cpp/test_pubkey_compat.cpp — level B3 proof: the reference's UNMODIFIED src/pubkey.cpp, compiled with
// include/compat/ in front of the libsecp256k1 headers and linked against libbcp_secp256k1_compat.so (no stock
// libsecp256k1 in the link), run through its own public interface: CPubKey::Verify, RecoverCompact, Derive, Decompress,
// IsFullyValid, CheckLowS (src/pubkey.cpp:170-302) and ECCVerifyHandle (:304-323).  Every curve operation inside lands
// on the GPU.  Expected results come from the fixture file written by tests/test_compat_cpp.py (reference-generated
// golden vectors).  Built by oracle/Makefile where /root/reference exists; the binary travels with oracle/_ref/.
//
//   file layout (little-endian u64 counts):
//     nV | nV x { hash32, publen u8, pub65, derlen u8, der80, expect u8 (Verify), lows u8 (CheckLowS), parse u8 (IsFullyValid) }
//     nR | nR x { hash32, sig65, outlen u8, out65 }                       (RecoverCompact; outlen 0 = must fail)
//     nD | nD x { pub33, cc32, nchild u32, child33, ccchild32 }           (Derive)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <vector>

#include "pubkey.h"
#include "uint256.h"

#define CHECK(x)                                                           \
    do {                                                                   \
        if (!(x)) {                                                        \
            fprintf(stderr, "FAILED %s:%d: %s (item %zu)\n", __FILE__, __LINE__, #x, item); \
            exit(1);                                                       \
        }                                                                  \
    } while (0)

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    std::vector<unsigned char> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    size_t pos = 0, item = 0;
    auto u64 = [&] { uint64_t v; memcpy(&v, &buf[pos], 8); pos += 8; return v; };
    ECCVerifyHandle handle; // secp256k1_context_create(SECP256K1_CONTEXT_VERIFY): brings the device up
    const uint64_t nV = u64();
    size_t accepted = 0, decompressed = 0;
    for (item = 0; item < nV; item++) {
        uint256 h;
        memcpy(h.begin(), &buf[pos], 32); pos += 32;
        const size_t publen = buf[pos++];
        const unsigned char *pub = &buf[pos]; pos += 65;
        const size_t derlen = buf[pos++];
        std::vector<uint8_t> der(&buf[pos], &buf[pos] + derlen); pos += 80;
        const int expect = buf[pos++], lows = buf[pos++], parse = buf[pos++];
        CPubKey pk(pub, pub + publen);
        CHECK(pk.Verify(h, der) == (expect != 0));
        accepted += expect != 0;
        CHECK(CPubKey::CheckLowS(der) == (lows != 0));
        CHECK(pk.IsFullyValid() == (parse != 0));
        if (parse && publen == 33) { // Decompress: parse + uncompressed re-serialisation; the result must verify the same
            CPubKey full = pk;
            CHECK(full.Decompress());
            CHECK(full.size() == 65 && full[0] == 4 && memcmp(&full[1], pub + 1, 32) == 0);
            CHECK(full.IsFullyValid());
            if (item % 16 == 0) CHECK(full.Verify(h, der) == (expect != 0));
            decompressed++;
        }
    }
    const uint64_t nR = u64();
    size_t recovered = 0;
    for (item = 0; item < nR; item++) {
        uint256 h;
        memcpy(h.begin(), &buf[pos], 32); pos += 32;
        std::vector<uint8_t> sig(&buf[pos], &buf[pos] + 65); pos += 65;
        const size_t outlen = buf[pos++];
        const unsigned char *out = &buf[pos]; pos += 65;
        CPubKey pk;
        const bool ok = pk.RecoverCompact(h, sig);
        CHECK(ok == (outlen != 0));
        if (ok) {
            CHECK(pk.size() == outlen && memcmp(pk.begin(), out, outlen) == 0);
            recovered++;
        }
    }
    const uint64_t nD = u64();
    for (item = 0; item < nD; item++) {
        const unsigned char *pub = &buf[pos]; pos += 33;
        ChainCode cc, ccChild;
        memcpy(cc.begin(), &buf[pos], 32); pos += 32;
        uint32_t nchild;
        memcpy(&nchild, &buf[pos], 4); pos += 4;
        const unsigned char *child = &buf[pos]; pos += 33;
        const unsigned char *ccexp = &buf[pos]; pos += 32;
        CPubKey pk(pub, pub + 33), out;
        CHECK(pk.Derive(out, ccChild, nchild, cc));
        CHECK(out.size() == 33 && memcmp(out.begin(), child, 33) == 0);
        CHECK(memcmp(ccChild.begin(), ccexp, 32) == 0);
    }
    CHECK(pos == buf.size());
    printf("pubkey compat ok: %llu Verify (%zu accepted, %zu decompressed), %llu RecoverCompact (%zu recovered), %llu Derive\n",
           (unsigned long long)nV, accepted, decompressed, (unsigned long long)nR, recovered, (unsigned long long)nD);
    return 0;
}
