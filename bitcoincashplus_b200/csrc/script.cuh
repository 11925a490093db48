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

// 4 bytes at p <- word w, most significant byte first.  On the device p is 4-byte aligned wherever this is used (digest and
// signature arrays are carved at multiples of 32 / 64 bytes from 256-byte aligned allocations).
BCP_DEV uint32_t bswap32(uint32_t x) {
#if BCP_PTX
    return __byte_perm(x, 0, 0x0123);
#else
    return __builtin_bswap32(x);
#endif
}
BCP_DEV void put_be32(uint8_t *p, uint32_t w) {
#if BCP_PTX
    *reinterpret_cast<uint32_t *>(p) = __byte_perm(w, 0, 0x0123);
#else
    p[0] = (uint8_t)(w >> 24); p[1] = (uint8_t)(w >> 16); p[2] = (uint8_t)(w >> 8); p[3] = (uint8_t)w;
#endif
}

// streaming SHA-256 fed a few bytes at a time (outpoints and sequence numbers are scattered through the tx).  The block
// being filled is addressed with a run-time word index, so it lives where such indexing is cheap: the caller hands in 16
// words at a stride (shared memory on the device, one column per thread), never registers spilled to local memory.
struct sha_stream {
    uint32_t s[8];
    uint32_t *w;
    uint32_t stride;
    uint32_t len;
};
BCP_DEV void sha_stream_init(sha_stream &c, uint32_t *wbuf, uint32_t stride) {
    sha256_init(c.s);
    c.w = wbuf;
    c.stride = stride;
    for (int i = 0; i < 16; i++) c.w[i * stride] = 0;
    c.len = 0;
}
BCP_DEV void sha_stream_flush(sha_stream &c) {
    uint32_t t[16];
    BCP_UNROLL
    for (int k = 0; k < 16; k++) { t[k] = c.w[k * c.stride]; c.w[k * c.stride] = 0; }
    sha256_compress(c.s, t);
}
BCP_DEV void sha_stream_feed(sha_stream &c, const uint8_t *p, uint32_t n) {
    for (uint32_t i = 0; i < n; i++) {
        uint32_t pos = c.len & 63u;
        c.w[(pos >> 2) * c.stride] |= (uint32_t)p[i] << (24 - 8 * (pos & 3u));
        c.len++;
        if ((c.len & 63u) == 0) sha_stream_flush(c);
    }
}
// finishes SHA256d; digest bytes to out32
BCP_DEV void sha_stream_final_d(sha_stream &c, uint8_t *out32) {
    uint32_t pos = c.len & 63u;
    c.w[(pos >> 2) * c.stride] |= 0x80u << (24 - 8 * (pos & 3u));
    if (pos >= 56) sha_stream_flush(c);
    uint32_t t[16];
    BCP_UNROLL
    for (int k = 0; k < 16; k++) t[k] = c.w[k * c.stride];
    t[14] = 0;
    t[15] = c.len * 8u;
    sha256_compress(c.s, t);
    sha256_second_pass(c.s);
    BCP_UNROLL
    for (int k = 0; k < 8; k++) put_be32(out32 + 4 * k, c.s[k]);
}

// PrecomputedTransactionData (src/script/interpreter.cpp:1347-1352) for one serialized transaction.  wbuf: 32 words of
// scratch for the two streaming hashes, word k of stream j at wbuf[(16 j + k) * stride]; nullptr = a local array.
BCP_DEV void tx_midstate(uint8_t *out /*BCP_TXMID_BYTES*/, const uint8_t *tx, uint32_t txlen, uint32_t *wbuf = nullptr,
                         uint32_t stride = 1) {
#if !BCP_PTX
    uint32_t local_w[32]; /* host build of this header only; the kernels always pass shared memory */
    if (!wbuf) { wbuf = local_w; stride = 1; }
#endif
    uint32_t pos = 4;
    const uint64_t nin = script_read_compact(tx, pos);
    sha_stream hp, hs;
    sha_stream_init(hp, wbuf, stride);
    sha_stream_init(hs, wbuf + 16 * stride, stride);
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
    uint32_t ho[8];
    dev_sha256_words(ho, tx + outs_off, txlen - 4 - outs_off); /* all outputs are contiguous */
    sha256_second_pass(ho);
    BCP_UNROLL
    for (int k = 0; k < 8; k++) put_be32(out + 64 + 4 * k, ho[k]);
    put_be32(out + 96, bswap32(outs_off));  /* little-endian u32 */
    put_be32(out + 100, bswap32((uint32_t)nout));
    put_be32(out + 104, 0);
    put_be32(out + 108, 0);
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
// a > b for 256-bit numbers held as 8 words, most significant first
BCP_DEV bool be256_gt(const uint32_t *a, const uint32_t *b) {
    bool gt = false, decided = false;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        if (!decided && a[i] != b[i]) { gt = a[i] > b[i]; decided = true; }
    }
    return gt;
}
// One DER integer (l bytes at p, l <= 33 for a valid encoding) as 8 big-endian words, leading zero bytes skipped like the
// lax parser does (src/pubkey.cpp:133-166); false when more than 32 significant bytes remain.
BCP_DEV bool der_int_words(uint32_t *w8, const uint8_t *p, uint32_t l) {
    while (l > 0 && p[0] == 0) { p++; l--; }
    BCP_UNROLL
    for (int i = 0; i < 8; i++) w8[i] = 0;
    if (l > 32) return false;
    const uint32_t pad = 32 - l;
    BCP_UNROLL
    for (uint32_t k = 0; k < 32; k++) { /* k = position in the 32-byte big-endian string: static word index */
        const uint32_t byte = k >= pad ? p[k - pad] : 0u;
        w8[k >> 2] |= byte << (24 - 8 * (k & 3));
    }
    return true;
}
// The (r, s) the lax parser + parse_compact produce for a strictly valid DER signature
// (src/pubkey.cpp:133-166, src/secp256k1/src/secp256k1.c:228-247): a component longer than 32 bytes or >= n zeroes
// the whole signature.  r, s: 8 big-endian words each.
BCP_DEV void script_der_to_rs_words(uint32_t *r, uint32_t *sw, const uint8_t *sig) {
    const uint32_t N_W[8] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFEu, 0xBAAEDCE6u, 0xAF48A03Bu, 0xBFD25E8Cu, 0xD0364141u};
    const uint32_t lenR = sig[3];
    bool ok = der_int_words(r, sig + 4, lenR);
    ok = der_int_words(sw, sig + 6 + lenR, sig[5 + lenR]) && ok;
    if (ok) ok = be256_gt(N_W, r) && be256_gt(N_W, sw); /* both below n */
    if (!ok) {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) { r[i] = 0; sw[i] = 0; }
    }
}

// One input.  Writes the ECDSA triple (zeros unless the result is PENDING) and returns the status.  msg32 and sig64 must
// be 4-byte aligned on the device.  Everything is assembled in registers as words; the only byte-granular traffic left
// is reading the transaction (shared memory in the kernels) and the 65 pubkey bytes going out.
BCP_DEV uint32_t p2pkh_extract_inner(const uint8_t *tx, uint32_t txlen, uint32_t in_pos, uint32_t nin_idx, const uint8_t *spk,
                                     uint32_t spk_len, int64_t amount, const uint8_t *mid, uint8_t *msg32, uint8_t *pub65,
                                     uint8_t *sig64) {
    const uint32_t NHALF_W[8] = {0x7FFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0x5D576E73u, 0x57A4501Du, 0xDFE92F46u, 0x681B20A0u};
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
    const uint8_t *sg = ss + 1, *pk = ss + 2 + siglen;
    // OP_DUP OP_HASH160 <20> OP_EQUALVERIFY
    {
        uint32_t s8[8], r5[5];
        dev_sha256_words(s8, pk, publen);
        ripemd160_of_sha256(r5, s8);
        bool eq = true;
        BCP_UNROLL
        for (int i = 0; i < 20; i++) eq = eq && (spk[3 + i] == (uint8_t)(r5[i >> 2] >> (8 * (i & 3))));
        if (!eq) return BCP_SCRIPT_ERR_EQUALVERIFY;
    }
    uint32_t rw[8], sw[8];
    uint32_t hashtype = 0;
    if (siglen) {
        if (!script_is_valid_sig_encoding(sg, siglen)) return BCP_SCRIPT_ERR_SIG_DER;
        script_der_to_rs_words(rw, sw, sg);
        if (be256_gt(sw, NHALF_W)) return BCP_SCRIPT_ERR_SIG_HIGH_S;
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
    // SignatureHash, FORKID branch: the 182-byte preimage for a 25-byte scriptCode, built as the 48 big-endian words of
    // its three SHA-256 blocks (every byte position below is a compile-time constant after unrolling)
    const uint32_t base = hashtype & 0x1fu;
    const bool acp = (hashtype & 0x80u) != 0;
    uint32_t mw[24]; /* the three midstate hashes as stored (byte order of memory) */
#if BCP_PTX
    {
        const uint4 *m4 = reinterpret_cast<const uint4 *>(mid);
        BCP_UNROLL
        for (int q = 0; q < 6; q++) {
            const uint4 v = __ldg(m4 + q);
            mw[4 * q] = v.x; mw[4 * q + 1] = v.y; mw[4 * q + 2] = v.z; mw[4 * q + 3] = v.w;
        }
    }
#else
    for (int q = 0; q < 24; q++) mw[q] = (uint32_t)mid[4 * q] | (uint32_t)mid[4 * q + 1] << 8 | (uint32_t)mid[4 * q + 2] << 16 | (uint32_t)mid[4 * q + 3] << 24;
#endif
    uint32_t w[48];
    BCP_UNROLL
    for (int i = 0; i < 48; i++) w[i] = 0;
#define BCP_PUTB(posn, val) w[(posn) >> 2] |= (uint32_t)(val) << (24 - 8 * ((posn) & 3))
    BCP_UNROLL
    for (int i = 0; i < 4; i++) BCP_PUTB(i, tx[i]);
    BCP_UNROLL
    for (int j = 0; j < 8; j++) { /* hashPrevouts at 4, hashSequence at 36: word aligned */
        w[1 + j] = acp ? 0u : bswap32(mw[j]);
        w[9 + j] = (!acp && base != 2 && base != 3) ? bswap32(mw[8 + j]) : 0u;
    }
    BCP_UNROLL
    for (int i = 0; i < 36; i++) BCP_PUTB(68 + i, tx[in_pos + i]);
    BCP_PUTB(104, 25u);
    BCP_UNROLL
    for (int i = 0; i < 25; i++) BCP_PUTB(105 + i, spk[i]);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) BCP_PUTB(130 + i, (uint8_t)((uint64_t)amount >> (8 * i)));
    BCP_UNROLL
    for (int i = 0; i < 4; i++) BCP_PUTB(138 + i, tx[pos + sl + i]);
    uint32_t ho[8]; /* hashOutputs, big-endian words of the digest bytes */
    if (base != 2 && base != 3) {
        BCP_UNROLL
        for (int j = 0; j < 8; j++) ho[j] = bswap32(mw[16 + j]);
    } else {
        BCP_UNROLL
        for (int j = 0; j < 8; j++) ho[j] = 0;
        const uint32_t outs_off = mid[96] | (uint32_t)mid[97] << 8 | (uint32_t)mid[98] << 16 | (uint32_t)mid[99] << 24;
        const uint32_t nout = mid[100] | (uint32_t)mid[101] << 8 | (uint32_t)mid[102] << 16 | (uint32_t)mid[103] << 24;
        if (base == 3 && nin_idx < nout) { /* SIGHASH_SINGLE: the output with the same index */
            uint32_t p = outs_off;
            for (uint32_t k = 0; k < nin_idx; k++) {
                p += 8;
                const uint32_t l = (uint32_t)script_read_compact(tx, p);
                p += l;
            }
            uint32_t q = p + 8;
            const uint32_t l = (uint32_t)script_read_compact(tx, q);
            dev_sha256_words(ho, tx + p, q + l - p);
            sha256_second_pass(ho);
        }
    }
    BCP_UNROLL
    for (int i = 0; i < 32; i++) BCP_PUTB(142 + i, (ho[i >> 2] >> (24 - 8 * (i & 3))) & 0xffu);
    BCP_UNROLL
    for (int i = 0; i < 4; i++) BCP_PUTB(174 + i, tx[txlen - 4 + i]);
    BCP_PUTB(178, hashtype & 0xffu);
    BCP_PUTB(182, 0x80u);
    w[47] = 182u * 8u;
#undef BCP_PUTB
    uint32_t st[8];
    sha256_init(st);
    sha256_compress(st, w);
    sha256_compress(st, w + 16);
    sha256_compress(st, w + 32);
    sha256_second_pass(st);
    BCP_UNROLL
    for (int k = 0; k < 8; k++) put_be32(msg32 + 4 * k, st[k]);
    for (uint32_t i = 0; i < 65; i++) pub65[i] = i < publen ? pk[i] : (uint8_t)0;
    BCP_UNROLL
    for (int k = 0; k < 8; k++) { put_be32(sig64 + 4 * k, rw[k]); put_be32(sig64 + 32 + 4 * k, sw[k]); }
    return BCP_SCRIPT_PENDING;
}
BCP_DEV uint32_t p2pkh_extract_one(const uint8_t *tx, uint32_t txlen, uint32_t in_pos /*offset of the outpoint in tx*/,
                                   uint32_t nin_idx, const uint8_t *spk, uint32_t spk_len, int64_t amount,
                                   const uint8_t *mid /*BCP_TXMID_BYTES of this tx*/, uint8_t *msg32, uint8_t *pub65,
                                   uint8_t *sig64) {
    const uint32_t st = p2pkh_extract_inner(tx, txlen, in_pos, nin_idx, spk, spk_len, amount, mid, msg32, pub65, sig64);
    if (st != BCP_SCRIPT_PENDING) { /* rare: the triple of an input that is already decided is all zeros */
        BCP_UNROLL
        for (int k = 0; k < 8; k++) put_be32(msg32 + 4 * k, 0);
        for (int i = 0; i < 65; i++) pub65[i] = 0;
        BCP_UNROLL
        for (int k = 0; k < 16; k++) put_be32(sig64 + 4 * k, 0);
    }
    return st;
}
