// include/compat/bcp_headers.h — SURVEY.md §8f row N4: the header-chain pipeline of headers-first sync.
//
// What the node does per header in AcceptBlockHeader (src/validation.cpp:3515-3590):
//   CheckBlockHeader            :3202-3220   Equihash solution, proof of work            -> GPU, whole batch at once
//   prev lookup                 :3548-3559   "bad-prevblk"                                -> GPU (in-batch linkage) + tip
//   ContextualCheckBlockHeader  :3362-3411   difficulty, height, median time past, version -> host, sequential
// The batch part is one call of bcp_gpu_header_check_batch; the sequential part below needs the chain so far
// (times, bits, accumulated work) and is restated here with its own 256-bit arithmetic:
//   GetNextWorkRequired              src/pow.cpp:72-105
//   BitcoinGetNextWorkRequired       src/pow.cpp:27-67     (pre-fork heights)
//   CalculateNextWorkRequired        src/pow.cpp:110-134
//   GetNextCashPlusWorkRequired      src/pow.cpp:252-295   (ComputeTarget :168-200, GetSuitableBlock :206-237)
//   GetBlockProof                    src/chain.cpp:132-145
//   CBlockIndex::GetMedianTimePast   src/chain.h:312-325
//   arith_uint256 Set/GetCompact     src/arith_uint256.cpp:181-232
// The chain is linear (one vector indexed by height): forks and reorganisation are the node's business.
#ifndef BCP_COMPAT_HEADERS_H
#define BCP_COMPAT_HEADERS_H

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../bcp_gpu.h"

namespace bcp {

// 256-bit unsigned integer on 8 little-endian 32-bit limbs; only what the difficulty code needs.
struct U256 {
    uint32_t w[8];
    U256() { memset(w, 0, sizeof w); }
    explicit U256(uint64_t v) {
        memset(w, 0, sizeof w);
        w[0] = (uint32_t)v;
        w[1] = (uint32_t)(v >> 32);
    }
    static U256 FromBytesLE(const uint8_t b[32]) { // uint256 memory order
        U256 r;
        for (int i = 0; i < 8; i++) r.w[i] = (uint32_t)b[4 * i] | (uint32_t)b[4 * i + 1] << 8 | (uint32_t)b[4 * i + 2] << 16 | (uint32_t)b[4 * i + 3] << 24;
        return r;
    }
    void ToBytesLE(uint8_t b[32]) const {
        for (int i = 0; i < 8; i++)
            for (int j = 0; j < 4; j++) b[4 * i + j] = (uint8_t)(w[i] >> (8 * j));
    }
    static U256 FromHex(const char *hex) { // big-endian text, like uint256S
        U256 r;
        size_t n = strlen(hex);
        for (size_t i = 0; i < n && i < 64; i++) {
            char ch = hex[n - 1 - i];
            uint32_t v = ch >= '0' && ch <= '9' ? ch - '0' : ch >= 'a' && ch <= 'f' ? ch - 'a' + 10 : ch >= 'A' && ch <= 'F' ? ch - 'A' + 10 : 0;
            r.w[i / 8] |= v << (4 * (i % 8));
        }
        return r;
    }
    int bits() const {
        for (int i = 7; i >= 0; i--)
            if (w[i]) return 32 * i + 32 - __builtin_clz(w[i]);
        return 0;
    }
    bool isZero() const { return bits() == 0; }
    int cmp(const U256 &o) const {
        for (int i = 7; i >= 0; i--) {
            if (w[i] < o.w[i]) return -1;
            if (w[i] > o.w[i]) return 1;
        }
        return 0;
    }
    bool operator>(const U256 &o) const { return cmp(o) > 0; }
    bool operator>=(const U256 &o) const { return cmp(o) >= 0; }
    bool operator==(const U256 &o) const { return cmp(o) == 0; }
    U256 operator+(const U256 &o) const {
        U256 r;
        uint64_t c = 0;
        for (int i = 0; i < 8; i++) {
            c += (uint64_t)w[i] + o.w[i];
            r.w[i] = (uint32_t)c;
            c >>= 32;
        }
        return r;
    }
    U256 operator-(const U256 &o) const {
        U256 r;
        int64_t c = 0;
        for (int i = 0; i < 8; i++) {
            c += (int64_t)w[i] - o.w[i];
            r.w[i] = (uint32_t)c;
            c >>= 32;
        }
        return r;
    }
    U256 operator~() const {
        U256 r;
        for (int i = 0; i < 8; i++) r.w[i] = ~w[i];
        return r;
    }
    U256 neg() const { return (~*this) + U256(1); } // two's complement, base_uint::operator-
    U256 shl(unsigned s) const {
        U256 r;
        const unsigned k = s / 32, b = s % 32;
        for (unsigned i = 0; i < 8; i++) {
            if (b && i + k + 1 < 8) r.w[i + k + 1] |= w[i] >> (32 - b);
            if (i + k < 8) r.w[i + k] |= w[i] << b;
        }
        return r;
    }
    U256 shr(unsigned s) const {
        U256 r;
        const unsigned k = s / 32, b = s % 32;
        for (unsigned i = 0; i < 8; i++) {
            if (b && i >= k + 1) r.w[i - k - 1] |= w[i] << (32 - b);
            if (i >= k) r.w[i - k] |= w[i] >> b;
        }
        return r;
    }
    // times a 32-bit factor, truncated to 256 bits: base_uint::operator*=(uint32_t) (arith_uint256.cpp:50-59).  That
    // is the overload `bn *= <int64_t>` selects in the reference (an integral conversion beats the user-defined one
    // to base_uint), so the callers below cast their int64 factors to uint32_t exactly like the reference does.
    U256 mul32(uint32_t f) const {
        U256 r;
        uint64_t c = 0;
        for (int i = 0; i < 8; i++) {
            c += (uint64_t)w[i] * f;
            r.w[i] = (uint32_t)c;
            c >>= 32;
        }
        return r;
    }
    // floor(this / d); d != 0.  Schoolbook shift-subtract like arith_uint256.cpp:76-104 (the quotient is what is
    // specified, not the method).
    U256 div(const U256 &d) const {
        U256 q, num = *this;
        const int nb = num.bits(), db = d.bits();
        if (db == 0 || db > nb) return q;
        int shift = nb - db;
        U256 dv = d.shl((unsigned)shift);
        for (; shift >= 0; shift--) {
            if (num >= dv) {
                num = num - dv;
                q.w[shift / 32] |= 1u << (shift & 31);
            }
            dv = dv.shr(1);
        }
        return q;
    }
    uint64_t low64() const { return (uint64_t)w[0] | (uint64_t)w[1] << 32; }

    // arith_uint256::SetCompact (src/arith_uint256.cpp:181-199)
    static U256 FromCompact(uint32_t c, bool *negative = nullptr, bool *overflow = nullptr) {
        const int size = (int)(c >> 24);
        uint32_t word = c & 0x007fffffu;
        U256 r;
        if (size <= 3) {
            word >>= 8 * (3 - size);
            r = U256(word);
        } else {
            r = U256(word).shl(8u * (unsigned)(size - 3));
        }
        if (negative) *negative = word != 0 && (c & 0x00800000u) != 0;
        if (overflow) *overflow = word != 0 && (size > 34 || (word > 0xffu && size > 33) || (word > 0xffffu && size > 32));
        return r;
    }
    // arith_uint256::GetCompact (src/arith_uint256.cpp:201-223), fNegative = false
    uint32_t GetCompact() const {
        int size = (bits() + 7) / 8;
        uint32_t c;
        if (size <= 3) c = (uint32_t)(low64() << 8 * (3 - size));
        else c = (uint32_t)shr(8u * (unsigned)(size - 3)).low64();
        if (c & 0x00800000u) {
            c >>= 8;
            size++;
        }
        return c | (uint32_t)size << 24;
    }
};

// GetBlockProof (src/chain.cpp:132-145)
inline U256 BlockProof(uint32_t nBits) {
    bool neg, ovf;
    const U256 t = U256::FromCompact(nBits, &neg, &ovf);
    if (neg || ovf || t.isZero()) return U256();
    return (~t).div(t + U256(1)) + U256(1);
}

// The fields of Consensus::Params (src/consensus/params.h:40-93) and CChainParams the header rules read.
struct PowParams {
    U256 powLimit, powLimitLegacy, powLimitStart;
    bool fPowAllowMinDifficultyBlocks = false, fPowNoRetargeting = false;
    int64_t nPowTargetSpacing = 600, nPowTargetTimespanLegacy = 14 * 24 * 60 * 60;
    int BCPHeight = 0, BCPPremineWindow = 0;
    int64_t nPowAveragingWindow = 30;
    int BIP34Height = 0, BIP65Height = 0, BIP66Height = 0;
    unsigned ehN = 200, ehK = 9;
    int64_t DifficultyAdjustmentInterval() const { return nPowTargetTimespanLegacy / nPowTargetSpacing; }
    const U256 &PowLimit(bool postfork) const { return postfork ? powLimit : powLimitLegacy; }

    static PowParams Main() { // src/chainparams.cpp:98-160
        PowParams p;
        p.BIP34Height = 227931; p.BIP65Height = 388381; p.BIP66Height = 363725;
        p.BCPHeight = 509696; p.BCPPremineWindow = 16000;
        p.powLimit = U256::FromHex("7fffffffffffffffffffffffffffffffffffffffffffffffffffffffffffffff");
        p.powLimitStart = U256::FromHex("00000fffffffffffffffffffffffffffffffffffffffffffffffffffffffffff");
        p.powLimitLegacy = U256::FromHex("00000000ffffffffffffffffffffffffffffffffffffffffffffffffffffffff");
        return p;
    }
    static PowParams Regtest() { // src/chainparams.cpp:335-392
        PowParams p;
        p.BIP34Height = 100000000; p.BIP65Height = 1351; p.BIP66Height = 1251;
        p.BCPHeight = 3000; p.BCPPremineWindow = 20;
        p.powLimit = p.powLimitStart = p.powLimitLegacy = U256::FromHex("7fffffffffffffffffffffffffffffffffffffffffffffffffffffffffffffff");
        p.fPowAllowMinDifficultyBlocks = true; p.fPowNoRetargeting = true;
        p.ehN = 48; p.ehK = 5;
        return p;
    }
};

class HeaderChain {
public:
    struct Entry { // the CBlockIndex fields the header rules read
        int32_t nVersion;
        uint32_t nTime, nBits;
        U256 nChainWork;
        uint8_t hash[32];
    };
    explicit HeaderChain(const PowParams &p) : params(p) {}

    //! Appends a header that is already known to be valid (block index loaded from disk, checkpoints, tests).
    void AddIndexEntry(int32_t nVersion, uint32_t nTime, uint32_t nBits, const uint8_t hash[32]) {
        Entry e;
        e.nVersion = nVersion; e.nTime = nTime; e.nBits = nBits;
        e.nChainWork = (chain.empty() ? U256() : chain.back().nChainWork) + BlockProof(nBits);
        memcpy(e.hash, hash, 32);
        chain.push_back(e);
    }
    int Height() const { return (int)chain.size() - 1; } // -1: empty
    const Entry &At(int h) const { return chain[(size_t)h]; }

    //! CBlockIndex::GetMedianTimePast of the block at height h (src/chain.h:312-325)
    int64_t MedianTimePast(int h) const {
        int64_t t[11];
        int n = 0;
        for (int i = h; i >= 0 && n < 11; i--) t[n++] = chain[(size_t)i].nTime;
        std::sort(t, t + n);
        return t[n / 2];
    }

    //! GetNextWorkRequired(pindexPrev = tip, pblock->nTime = new_time) (src/pow.cpp:72-105)
    uint32_t NextWorkRequired(int64_t new_time) const {
        if (chain.empty()) return params.PowLimit(false).GetCompact();
        const int prev = Height();
        if (params.fPowNoRetargeting) return chain[(size_t)prev].nBits;
        const int nHeight = prev + 1;
        const bool postfork = nHeight >= params.BCPHeight;
        if (!postfork) return LegacyNextWork(prev, new_time);
        if (nHeight < params.BCPHeight + params.BCPPremineWindow) return params.PowLimit(true).GetCompact();
        if (nHeight < params.BCPHeight + params.BCPPremineWindow + params.nPowAveragingWindow) return params.powLimitStart.GetCompact();
        return CashPlusNextWork(prev, new_time);
    }

    //! CalculateNextWorkRequired (src/pow.cpp:110-134)
    uint32_t CalculateNextWork(int prev, int64_t nFirstBlockTime) const {
        if (params.fPowNoRetargeting) return chain[(size_t)prev].nBits;
        int64_t span = (int64_t)chain[(size_t)prev].nTime - nFirstBlockTime;
        span = std::max(span, params.nPowTargetTimespanLegacy / 4);
        span = std::min(span, params.nPowTargetTimespanLegacy * 4);
        U256 n = U256::FromCompact(chain[(size_t)prev].nBits).mul32((uint32_t)span).div(U256((uint64_t)params.nPowTargetTimespanLegacy));
        if (n > params.PowLimit(false)) n = params.PowLimit(false);
        return n.GetCompact();
    }

    //! GetNextCashPlusWorkRequired (src/pow.cpp:252-295); needs prev >= 147 like the reference's asserts
    uint32_t CashPlusNextWork(int prev, int64_t new_time) const {
        const bool postfork = prev >= params.BCPHeight;
        if (params.fPowAllowMinDifficultyBlocks && new_time > (int64_t)chain[(size_t)prev].nTime + 2 * params.nPowTargetSpacing)
            return params.PowLimit(postfork).GetCompact();
        const int last = SuitableBlock(prev), first = SuitableBlock(prev - 144);
        // ComputeTarget (src/pow.cpp:168-200)
        U256 work = (chain[(size_t)last].nChainWork - chain[(size_t)first].nChainWork).mul32((uint32_t)params.nPowTargetSpacing);
        int64_t span = (int64_t)chain[(size_t)last].nTime - (int64_t)chain[(size_t)first].nTime;
        span = std::min(span, 288 * params.nPowTargetSpacing);
        span = std::max(span, 72 * params.nPowTargetSpacing);
        work = work.div(U256((uint64_t)span));
        const U256 next = work.neg().div(work);
        if (next > params.PowLimit(postfork)) return params.PowLimit(postfork).GetCompact();
        return next.GetCompact();
    }

    //! Headers-first acceptance of `count` serialized post-fork headers (140 bytes || compact size || solution) laid out
    //! at `stride`: one GPU pass for the context-free checks, then the sequential rules.  Returns how many were
    //! appended; on the first failure `reason` holds the reference's reject string and `failed` its index in the batch.
    long AcceptHeaders(int device, const uint8_t *hdrs, size_t stride, size_t hdr_len, size_t count, int64_t nAdjustedTime,
                       std::string &reason, long &failed) {
        failed = -1;
        if (count == 0) return 0;
        std::vector<uint8_t> okSol(count), okPow(count), linked(count), hash(count * 32);
        uint8_t lim[32], limLegacy[32];
        params.powLimit.ToBytesLE(lim);
        params.powLimitLegacy.ToBytesLE(limLegacy);
        if (bcp_gpu_header_check_batch(device, params.ehN, params.ehK, hdrs, stride, hdr_len, count, lim, limLegacy,
                                       (uint32_t)std::max(params.BCPHeight, 0), okSol.data(), okPow.data(), linked.data(),
                                       hash.data()) != BCP_OK) {
            reason = std::string("device-error: ") + bcp_gpu_last_error();
            failed = 0;
            return 0;
        }
        for (size_t i = 0; i < count; i++) {
            const uint8_t *h = hdrs + i * stride;
            auto fail = [&](const std::string &r) { reason = r; failed = (long)i; return (long)i; };
            auto le32 = [](const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; };
            const int32_t nVersion = (int32_t)le32(h);
            const uint32_t hdrHeight = le32(h + 68), nTime = le32(h + 100), nBits = le32(h + 104);
            // CheckBlockHeader (src/validation.cpp:3202-3220); its postfork is taken from the header's own height field
            if ((int64_t)hdrHeight >= params.BCPHeight && !okSol[i]) return fail("invalid-solution");
            if (!okPow[i]) return fail("high-hash");
            // prev lookup (:3548-3553)
            const bool prevOk = i == 0 ? (!chain.empty() && memcmp(h + 4, chain.back().hash, 32) == 0) : linked[i] != 0;
            if (!prevOk) return fail("bad-prevblk");
            // ContextualCheckBlockHeader (:3362-3411)
            const int nHeight = Height() + 1;
            if (nBits != NextWorkRequired((int64_t)nTime)) return fail("bad-diffbits");
            if (nHeight >= params.BCPHeight && hdrHeight != (uint32_t)nHeight) return fail("bad-height");
            if ((int64_t)nTime <= MedianTimePast(Height())) return fail("time-too-old");
            if ((int64_t)nTime > nAdjustedTime + 2 * 60 * 60) return fail("time-too-new");
            if ((nVersion < 2 && nHeight >= params.BIP34Height) || (nVersion < 3 && nHeight >= params.BIP66Height) ||
                (nVersion < 4 && nHeight >= params.BIP65Height)) {
                char buf[32];
                snprintf(buf, sizeof buf, "bad-version(0x%08x)", (unsigned)nVersion);
                return fail(buf);
            }
            AddIndexEntry(nVersion, nTime, nBits, &hash[i * 32]);
        }
        return (long)count;
    }

private:
    // BitcoinGetNextWorkRequired (src/pow.cpp:27-67)
    uint32_t LegacyNextWork(int prev, int64_t new_time) const {
        const uint32_t limit = params.PowLimit(false).GetCompact();
        const int64_t interval = params.DifficultyAdjustmentInterval();
        if ((prev + 1) % interval != 0) {
            if (params.fPowAllowMinDifficultyBlocks) {
                if (new_time > (int64_t)chain[(size_t)prev].nTime + params.nPowTargetSpacing * 2) return limit;
                int i = prev;
                while (i > 0 && i % interval != 0 && chain[(size_t)i].nBits == limit) i--;
                return chain[(size_t)i].nBits;
            }
            return chain[(size_t)prev].nBits;
        }
        const int first = prev - (int)(interval - 1);
        return CalculateNextWork(prev, (int64_t)chain[(size_t)first].nTime);
    }
    // GetSuitableBlock (src/pow.cpp:206-237): the median by time of heights h-2, h-1, h, same tie behaviour as the
    // reference's three conditional swaps
    int SuitableBlock(int h) const {
        int b[3] = {h - 2, h - 1, h};
        auto t = [&](int i) { return chain[(size_t)i].nTime; };
        if (t(b[0]) > t(b[2])) std::swap(b[0], b[2]);
        if (t(b[0]) > t(b[1])) std::swap(b[0], b[1]);
        if (t(b[1]) > t(b[2])) std::swap(b[1], b[2]);
        return b[1];
    }

    PowParams params;
    std::vector<Entry> chain;
};

} // namespace bcp
#endif
