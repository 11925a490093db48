// script.cuh — device side of the P2PKH fast path: what CScriptCheck::operator() (src/validation.cpp:1350-1359)
// does for an input whose scriptSig is `<sig> <pubkey>` and whose spent output is
// `OP_DUP OP_HASH160 <20> OP_EQUALVERIFY OP_CHECKSIG`, under the post-fork flag set
// (GetBlockScriptFlags, src/validation.cpp:1803-1852: P2SH|DERSIG|CLTV|CSV|FORKID|STRICTENC|LOW_S|NULLFAIL).
//
// Decision chain mirrored, in the interpreter's order (src/script/interpreter.cpp):
//   OP_HASH160 / OP_EQUALVERIFY (:994-1005)                  -> SCRIPT_ERR_EQUALVERIFY
//   CheckSignatureEncoding (:221-254): IsValidSignatureEncoding (:107-174) -> SIG_DER, IsLowDERSignature (:176-186)
//     -> SIG_HIGH_S, IsDefinedHashtypeSignature (:208-219) -> SIG_HASHTYPE, fork id rule (:244-251) -> MUST_USE_FORKID
//   CheckPubKeyEncoding (:256-269, :62-82)                    -> PUBKEYTYPE
//   empty signature: CheckSig false, FALSE left on the stack  -> EVAL_FALSE
//   SignatureHash, FORKID branch (:1357-1404) with GetPrevoutHash/GetSequenceHash/GetOutputsHash (:1321-1343)
//   CPubKey::Verify (src/pubkey.cpp:170-193) is NOT done here: the (sighash, pubkey, r||s) triple is written out for
//   the batched ECDSA kernels and the input is reported as PENDING; a failed verification becomes SIG_NULLFAIL.
// Anything that is not exactly this shape (other scriptPubKey, scriptSig with other opcodes or push forms) is
// reported as BCP_SCRIPT_NOT_TEMPLATE and must go through the host interpreter.
#pragma once
#include "ripemd160.cuh"
#include "sha256.cuh"

#define BCP_SCRIPT_OK 0u
#define BCP_SCRIPT_ERR_EVAL_FALSE 2u
#define BCP_SCRIPT_ERR_EQUALVERIFY 11u
#define BCP_SCRIPT_ERR_SIG_HASHTYPE 22u
#define BCP_SCRIPT_ERR_SIG_DER 23u
#define BCP_SCRIPT_ERR_SIG_HIGH_S 26u
#define BCP_SCRIPT_ERR_PUBKEYTYPE 28u
#define BCP_SCRIPT_ERR_SIG_NULLFAIL 31u
#define BCP_SCRIPT_ERR_MUST_USE_FORKID 36u
#define BCP_SCRIPT_PENDING 254u      /* every host-visible rule passed; the verdict is the ECDSA result */
#define BCP_SCRIPT_NOT_TEMPLATE 255u

#define BCP_TXMID_BYTES 112 /* hashPrevouts[32] hashSequence[32] hashOutputs[32] outs_off(u32) nout(u32) pad */

// compact size as the block indexer already validated it (canonical, in range)
BCP_DEV uint64_t script_read_compact(const uint8_t *p, uint32_t &pos) {
    uint8_t c = p[pos++];
    if (c < 253) return c;
    int nb = c == 253 ? 2 : c == 254 ? 4 : 8;
    uint64_t v = 0;
    for (int i = 0; i < nb; i++) v |= (uint64_t)p[pos + i] << (8 * i);
    pos += nb;
    return v;
}

// streaming SHA-256 fed a few bytes at a time (outpoints and sequence numbers are scattered through the tx)
struct sha_stream {
    uint32_t s[8];
    uint32_t w[16];
    uint32_t len;
};
BCP_DEV void sha_stream_init(sha_stream &c) {
    sha256_init(c.s);
    for (int i = 0; i < 16; i++) c.w[i] = 0;
    c.len = 0;
}
BCP_DEV void sha_stream_feed(sha_stream &c, const uint8_t *p, uint32_t n) {
    for (uint32_t i = 0; i < n; i++) {
        uint32_t pos = c.len & 63u;
        c.w[pos >> 2] |= (uint32_t)p[i] << (24 - 8 * (pos & 3u));
        c.len++;
        if ((c.len & 63u) == 0) {
            uint32_t t[16];
            for (int k = 0; k < 16; k++) { t[k] = c.w[k]; c.w[k] = 0; }
            sha256_compress(c.s, t);
        }
    }
}
// finishes SHA256d; digest bytes to out32
BCP_DEV void sha_stream_final_d(sha_stream &c, uint8_t *out32) {
    uint32_t pos = c.len & 63u;
    c.w[pos >> 2] |= 0x80u << (24 - 8 * (pos & 3u));
    uint32_t t[16];
    if (pos >= 56) {
        for (int k = 0; k < 16; k++) { t[k] = c.w[k]; c.w[k] = 0; }
        sha256_compress(c.s, t);
    }
    for (int k = 0; k < 16; k++) t[k] = c.w[k];
    t[14] = 0;
    t[15] = c.len * 8u;
    sha256_compress(c.s, t);
    sha256_second_pass(c.s);
    sha256_store_be(out32, c.s);
}

// PrecomputedTransactionData (src/script/interpreter.cpp:1347-1352) for one serialized transaction
BCP_DEV void tx_midstate(uint8_t *out /*BCP_TXMID_BYTES*/, const uint8_t *tx, uint32_t txlen) {
    uint32_t pos = 4;
    const uint64_t nin = script_read_compact(tx, pos);
    sha_stream hp, hs;
    sha_stream_init(hp);
    sha_stream_init(hs);
    for (uint64_t i = 0; i < nin; i++) {
        sha_stream_feed(hp, tx + pos, 36);
        pos += 36;
        const uint64_t sl = script_read_compact(tx, pos);
        pos += (uint32_t)sl;
        sha_stream_feed(hs, tx + pos, 4);
        pos += 4;
    }
    sha_stream_final_d(hp, out);
    sha_stream_final_d(hs, out + 32);
    const uint64_t nout = script_read_compact(tx, pos);
    const uint32_t outs_off = pos;
    dev_sha256d_bytes(out + 64, tx + outs_off, txlen - 4 - outs_off); /* all outputs are contiguous */
    for (int b = 0; b < 4; b++) {
        out[96 + b] = (uint8_t)(outs_off >> (8 * b));
        out[100 + b] = (uint8_t)((uint32_t)nout >> (8 * b));
    }
    for (int b = 104; b < BCP_TXMID_BYTES; b++) out[b] = 0;
}

// IsValidSignatureEncoding (src/script/interpreter.cpp:107-174) on sig[0..len) (DER plus the hashtype byte)
BCP_DEV bool script_is_valid_sig_encoding(const uint8_t *sig, uint32_t len) {
    if (len < 9 || len > 73) return false;
    if (sig[0] != 0x30) return false;
    if (sig[1] != len - 3) return false;
    const uint32_t lenR = sig[3];
    if (5 + lenR >= len) return false;
    const uint32_t lenS = sig[5 + lenR];
    if (lenR + lenS + 7 != len) return false;
    if (sig[2] != 0x02) return false;
    if (lenR == 0) return false;
    if (sig[4] & 0x80) return false;
    if (lenR > 1 && sig[4] == 0x00 && !(sig[5] & 0x80)) return false;
    if (sig[lenR + 4] != 0x02) return false;
    if (lenS == 0) return false;
    if (sig[lenR + 6] & 0x80) return false;
    if (lenS > 1 && sig[lenR + 6] == 0x00 && !(sig[lenR + 7] & 0x80)) return false;
    return true;
}
// big-endian compare of two 32-byte strings: a > b ?
BCP_DEV bool be32_gt(const uint8_t *a, const uint8_t *b) {
    for (int i = 0; i < 32; i++)
        if (a[i] != b[i]) return a[i] > b[i];
    return false;
}
// The (r, s) the lax parser + parse_compact produce for a strictly valid DER signature
// (src/pubkey.cpp:133-166, src/secp256k1/src/secp256k1.c:228-247): leading zero bytes skipped, a component longer
// than 32 bytes or >= n zeroes the whole signature.
BCP_DEV void script_der_to_rs(uint8_t *rs /*64*/, const uint8_t *sig) {
    const uint8_t N_BE[32] = {0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFE,
                              0xBA, 0xAE, 0xDC, 0xE6, 0xAF, 0x48, 0xA0, 0x3B, 0xBF, 0xD2, 0x5E, 0x8C, 0xD0, 0x36, 0x41, 0x41};
    for (int i = 0; i < 64; i++) rs[i] = 0;
    bool overflow = false;
    uint32_t p = 4, l = sig[3];
    for (int k = 0; k < 2; k++) {
        while (l > 0 && sig[p] == 0) { p++; l--; }
        if (l > 32) overflow = true;
        else for (uint32_t i = 0; i < l; i++) rs[32 * k + 32 - l + i] = sig[p + i];
        if (k == 0) { p = 6 + sig[3]; l = sig[5 + sig[3]]; }
    }
    for (int k = 0; k < 2 && !overflow; k++)
        if (!be32_gt(N_BE, rs + 32 * k)) overflow = true; /* component >= n */
    if (overflow) for (int i = 0; i < 64; i++) rs[i] = 0;
}

// One input.  Writes the ECDSA triple (zeros unless the result is PENDING) and returns the status.
BCP_DEV uint32_t p2pkh_extract_one(const uint8_t *tx, uint32_t txlen, uint32_t in_pos /*offset of the outpoint in tx*/,
                                   uint32_t nin_idx, const uint8_t *spk, uint32_t spk_len, int64_t amount,
                                   const uint8_t *mid /*BCP_TXMID_BYTES of this tx*/, uint8_t *msg32, uint8_t *pub65,
                                   uint8_t *sig64) {
    const uint8_t NHALF_BE[32] = {0x7F, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF,
                                  0x5D, 0x57, 0x6E, 0x73, 0x57, 0xA4, 0x50, 0x1D, 0xDF, 0xE9, 0x2F, 0x46, 0x68, 0x1B, 0x20, 0xA0};
    for (int i = 0; i < 32; i++) msg32[i] = 0;
    for (int i = 0; i < 65; i++) pub65[i] = 0;
    for (int i = 0; i < 64; i++) sig64[i] = 0;
    if (spk_len != 25 || spk[0] != 0x76 || spk[1] != 0xa9 || spk[2] != 0x14 || spk[23] != 0x88 || spk[24] != 0xac)
        return BCP_SCRIPT_NOT_TEMPLATE;
    uint32_t pos = in_pos + 36;
    const uint32_t sl = (uint32_t)script_read_compact(tx, pos);
    const uint8_t *ss = tx + pos;
    if (sl < 2) return BCP_SCRIPT_NOT_TEMPLATE;
    const uint32_t siglen = ss[0];
    if (siglen > 75 || 1 + siglen >= sl) return BCP_SCRIPT_NOT_TEMPLATE;
    const uint32_t publen = ss[1 + siglen];
    if (publen == 0 || publen > 75 || 1 + siglen + 1 + publen != sl) return BCP_SCRIPT_NOT_TEMPLATE;
    uint8_t sg[80], pk[80];
    for (uint32_t i = 0; i < siglen; i++) sg[i] = ss[1 + i];
    for (uint32_t i = 0; i < publen; i++) pk[i] = ss[2 + siglen + i];
    // OP_DUP OP_HASH160 <20> OP_EQUALVERIFY
    {
        uint32_t s8[8], r5[5];
        dev_sha256_words(s8, pk, publen);
        ripemd160_of_sha256(r5, s8);
        bool eq = true;
        for (int i = 0; i < 20; i++) eq = eq && (spk[3 + i] == (uint8_t)(r5[i >> 2] >> (8 * (i & 3))));
        if (!eq) return BCP_SCRIPT_ERR_EQUALVERIFY;
    }
    uint8_t rs[64];
    uint32_t hashtype = 0;
    if (siglen) {
        if (!script_is_valid_sig_encoding(sg, siglen)) return BCP_SCRIPT_ERR_SIG_DER;
        script_der_to_rs(rs, sg);
        if (be32_gt(rs + 32, NHALF_BE)) return BCP_SCRIPT_ERR_SIG_HIGH_S;
        hashtype = sg[siglen - 1];
        const uint32_t base = hashtype & ~(0x80u | 0x40u);
        if (base < 1 || base > 3) return BCP_SCRIPT_ERR_SIG_HASHTYPE;
        if (!(hashtype & 0x40u)) return BCP_SCRIPT_ERR_MUST_USE_FORKID;
    }
    if (publen < 33) return BCP_SCRIPT_ERR_PUBKEYTYPE;
    if (pk[0] == 0x04) { if (publen != 65) return BCP_SCRIPT_ERR_PUBKEYTYPE; }
    else if (pk[0] == 0x02 || pk[0] == 0x03) { if (publen != 33) return BCP_SCRIPT_ERR_PUBKEYTYPE; }
    else return BCP_SCRIPT_ERR_PUBKEYTYPE;
    if (!siglen) return BCP_SCRIPT_ERR_EVAL_FALSE;
    // SignatureHash, FORKID branch: 182-byte preimage for a 25-byte scriptCode
    uint8_t pre[184];
    const uint32_t base = hashtype & 0x1fu;
    const bool acp = (hashtype & 0x80u) != 0;
    for (int i = 0; i < 4; i++) pre[i] = tx[i];
    for (int i = 0; i < 32; i++) pre[4 + i] = acp ? 0 : mid[i];
    for (int i = 0; i < 32; i++) pre[36 + i] = (!acp && base != 2 && base != 3) ? mid[32 + i] : 0;
    for (int i = 0; i < 36; i++) pre[68 + i] = tx[in_pos + i];
    pre[104] = 25;
    for (int i = 0; i < 25; i++) pre[105 + i] = spk[i];
    for (int i = 0; i < 8; i++) pre[130 + i] = (uint8_t)((uint64_t)amount >> (8 * i));
    for (int i = 0; i < 4; i++) pre[138 + i] = tx[pos + sl + i];
    if (base != 2 && base != 3) {
        for (int i = 0; i < 32; i++) pre[142 + i] = mid[64 + i];
    } else {
        for (int i = 0; i < 32; i++) pre[142 + i] = 0;
        uint32_t outs_off = 0, nout = 0;
        for (int b = 0; b < 4; b++) { outs_off |= (uint32_t)mid[96 + b] << (8 * b); nout |= (uint32_t)mid[100 + b] << (8 * b); }
        if (base == 3 && nin_idx < nout) { /* SIGHASH_SINGLE: the output with the same index */
            uint32_t p = outs_off;
            for (uint32_t k = 0; k < nin_idx; k++) {
                p += 8;
                const uint32_t l = (uint32_t)script_read_compact(tx, p);
                p += l;
            }
            uint32_t q = p + 8;
            const uint32_t l = (uint32_t)script_read_compact(tx, q);
            dev_sha256d_bytes(pre + 142, tx + p, q + l - p);
        }
    }
    for (int i = 0; i < 4; i++) pre[174 + i] = tx[txlen - 4 + i];
    pre[178] = (uint8_t)hashtype; pre[179] = 0; pre[180] = 0; pre[181] = 0;
    dev_sha256d_bytes(msg32, pre, 182);
    for (uint32_t i = 0; i < publen; i++) pub65[i] = pk[i];
    for (int i = 0; i < 64; i++) sig64[i] = rs[i];
    return BCP_SCRIPT_PENDING;
}
