// pow.cuh — proof-of-work target arithmetic on 256-bit little-endian 32-bit limbs (device side of row N4).
// Follows arith_uint256::SetCompact (src/arith_uint256.cpp:181-199) and CheckProofOfWork (src/pow.cpp:141-161).
#pragma once
#include "bcp_common.cuh"

// target <- compact; bits shifted past 2^256 are dropped exactly like base_uint::operator<<= (arith_uint256.cpp:21-33)
BCP_DEV void pow_set_compact(uint32_t compact, uint32_t t[8], bool &negative, bool &overflow) {
    const int size = (int)(compact >> 24);
    uint32_t word = compact & 0x007fffffu;
    int k = 0, b = 0;
    if (size <= 3) {
        word >>= 8 * (3 - size);
    } else {
        const int sh = 8 * (size - 3);
        k = sh >> 5;
        b = sh & 31;
    }
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        uint32_t v = 0;
        if (i == k) v |= word << b;
        if (b != 0 && i == k + 1) v |= word >> (32 - b);
        t[i] = v;
    }
    negative = word != 0 && (compact & 0x00800000u) != 0;
    overflow = word != 0 && (size > 34 || (word > 0xffu && size > 33) || (word > 0xffffu && size > 32));
}

// a > b on 8 little-endian limbs (base_uint::CompareTo, arith_uint256.cpp:106-113)
BCP_DEV bool pow_gt(const uint32_t a[8], const uint32_t b[8]) {
    BCP_UNROLL
    for (int i = 7; i >= 0; i--) {
        if (a[i] > b[i]) return true;
        if (a[i] < b[i]) return false;
    }
    return false;
}

// CheckProofOfWork(hash, nBits, postfork, params): range checks on the target, then hash <= target.
// `hash` as 8 little-endian limbs of the uint256 (UintToArith256, arith_uint256.cpp:256-261).
BCP_DEV bool pow_check(const uint32_t hash[8], uint32_t nbits, const uint32_t limit[8]) {
    uint32_t t[8];
    bool neg, ovf;
    pow_set_compact(nbits, t, neg, ovf);
    uint32_t any = 0;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) any |= t[i];
    if (neg || any == 0 || ovf || pow_gt(t, limit)) return false;
    return !pow_gt(hash, t);
}
