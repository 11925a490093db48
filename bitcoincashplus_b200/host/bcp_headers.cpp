// bcp_headers.cpp — C face over bcp::HeaderChain (include/compat/bcp_headers.h) for ctypes callers (tests, bench.py).
// Part of bitcoincashplus_b200/libbcp_chain.so.
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/compat/bcp_headers.h"

#define CHAIN_API extern "C" __attribute__((visibility("default")))

// Plain-C mirror of bcp::PowParams; 256-bit values as uint256 memory bytes (little-endian).
struct bcp_pow_params {
    uint8_t pow_limit[32], pow_limit_legacy[32], pow_limit_start[32];
    int32_t allow_min_difficulty, no_retargeting;
    int64_t target_spacing, target_timespan_legacy;
    int32_t bcp_height, premine_window;
    int64_t averaging_window;
    int32_t bip34_height, bip65_height, bip66_height;
    uint32_t eh_n, eh_k;
};

CHAIN_API void *bcp_headers_new(const bcp_pow_params *p) {
    bcp::PowParams q;
    q.powLimit = bcp::U256::FromBytesLE(p->pow_limit);
    q.powLimitLegacy = bcp::U256::FromBytesLE(p->pow_limit_legacy);
    q.powLimitStart = bcp::U256::FromBytesLE(p->pow_limit_start);
    q.fPowAllowMinDifficultyBlocks = p->allow_min_difficulty != 0;
    q.fPowNoRetargeting = p->no_retargeting != 0;
    q.nPowTargetSpacing = p->target_spacing;
    q.nPowTargetTimespanLegacy = p->target_timespan_legacy;
    q.BCPHeight = p->bcp_height;
    q.BCPPremineWindow = p->premine_window;
    q.nPowAveragingWindow = p->averaging_window;
    q.BIP34Height = p->bip34_height;
    q.BIP65Height = p->bip65_height;
    q.BIP66Height = p->bip66_height;
    q.ehN = p->eh_n;
    q.ehK = p->eh_k;
    return new bcp::HeaderChain(q);
}
CHAIN_API void bcp_headers_free(void *h) { delete static_cast<bcp::HeaderChain *>(h); }
CHAIN_API void bcp_headers_add_index(void *h, int32_t version, uint32_t time, uint32_t bits, const uint8_t *hash32) {
    static_cast<bcp::HeaderChain *>(h)->AddIndexEntry(version, time, bits, hash32);
}
CHAIN_API int bcp_headers_height(void *h) { return static_cast<bcp::HeaderChain *>(h)->Height(); }
CHAIN_API uint32_t bcp_headers_next_work(void *h, int64_t new_time) {
    return static_cast<bcp::HeaderChain *>(h)->NextWorkRequired(new_time);
}
CHAIN_API uint32_t bcp_headers_calculate_next_work(void *h, int64_t first_block_time) {
    auto *c = static_cast<bcp::HeaderChain *>(h);
    return c->CalculateNextWork(c->Height(), first_block_time);
}
CHAIN_API uint32_t bcp_headers_cashplus_next_work(void *h, int64_t new_time) {
    auto *c = static_cast<bcp::HeaderChain *>(h);
    return c->CashPlusNextWork(c->Height(), new_time);
}
CHAIN_API int64_t bcp_headers_median_time_past(void *h) {
    auto *c = static_cast<bcp::HeaderChain *>(h);
    return c->MedianTimePast(c->Height());
}
CHAIN_API void bcp_headers_chain_work(void *h, uint8_t out32[32]) {
    auto *c = static_cast<bcp::HeaderChain *>(h);
    c->At(c->Height()).nChainWork.ToBytesLE(out32);
}
CHAIN_API long bcp_headers_accept(void *h, int device, const uint8_t *hdrs, size_t stride, size_t hdr_len, size_t count,
                                  int64_t adjusted_time, char *reason, size_t reasonlen, long *failed) {
    std::string r;
    long f = -1;
    long n = static_cast<bcp::HeaderChain *>(h)->AcceptHeaders(device, hdrs, stride, hdr_len, count, adjusted_time, r, f);
    if (reason && reasonlen) snprintf(reason, reasonlen, "%s", r.c_str());
    if (failed) *failed = f;
    return n;
}
CHAIN_API int bcp_compact_to_target(uint32_t compact, uint8_t out32[32], int *negative, int *overflow) {
    bool n, o;
    bcp::U256::FromCompact(compact, &n, &o).ToBytesLE(out32);
    if (negative) *negative = n;
    if (overflow) *overflow = o;
    return 0;
}
CHAIN_API uint32_t bcp_compact_from_target(const uint8_t t32[32]) { return bcp::U256::FromBytesLE(t32).GetCompact(); }
CHAIN_API void bcp_block_proof(uint32_t nbits, uint8_t out32[32]) { bcp::BlockProof(nbits).ToBytesLE(out32); }
