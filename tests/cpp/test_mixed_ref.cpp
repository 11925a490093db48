// tests/cpp/test_mixed_ref.cpp — TEST INFRASTRUCTURE ONLY (never part of the product).
//
// A block that mixes the script shapes of a real chain — P2PKH (the device fast path), pay-to-pubkey, bare m-of-n
// multisig, P2SH-wrapped multisig, P2SH-wrapped P2PKH — connected by BOTH block connectors of this project
// (bcp::ChainValidator, include/compat/bcp_chain.h, and bcp::DeviceChain, include/compat/bcp_chain_gpu.h) with the inputs the
// fast path hands back (script_err 255) evaluated the way a node would: the REFERENCE's own interpreter
// (src/script/interpreter.cpp, compiled in place into oracle/_ref/libbcp_ref.so) inside this project's CCheckQueue drop-in
// with the recording signature checker, signatures verified on the GPU in Wait().  The two results are AND-ed by the
// connector like `fAllOk &= fOk` (src/checkqueue.h:77).  The expected verdict of every block is the reference's plain
// VerifyScript over all of its inputs on the CPU.
//
// Scenarios: all valid; one P2SH-multisig signature corrupted; multisig signatures in the wrong order; a bad P2PKH
// signature next to valid multisig inputs; no callback installed (the block is refused as "script-not-in-fast-path").
// After every refusal the coin table must be unchanged: the valid version of the same block connects afterwards.
//
// Built by oracle/Makefile (target ref) into oracle/_ref/test_mixed_ref; run on the GPU box by tests/test_gpu_replay.py.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "consensus/merkle.h"
#include "crypto/sha256.h"
#include "hash.h"
#include "primitives/transaction.h"
#include "pubkey.h"
#include "script/interpreter.h"
#include "script/script.h"
#include "streams.h"
#include "uint256.h"
#include "version.h"

#include "compat/bcp_chain.h"
#include "compat/bcp_chain_gpu.h"
#include "compat/checkqueue.h"

extern "C" {
int ref_init(void);
int ref_pubkey_create(const uint8_t *seckey32, int compressed, uint8_t *out, size_t *outlen);
int ref_sign(const uint8_t *seckey32, const uint8_t *msg32, uint8_t *sig64);
int ref_sig_to_der(const uint8_t *sig64, uint8_t *der, size_t *derlen);
}

#define CHECK(x)                                                           \
    do {                                                                   \
        if (!(x)) {                                                        \
            fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #x); \
            exit(1);                                                       \
        }                                                                  \
    } while (0)

static const uint32_t FLAGS = SCRIPT_VERIFY_P2SH | SCRIPT_VERIFY_STRICTENC | SCRIPT_VERIFY_DERSIG | SCRIPT_VERIFY_LOW_S |
                              SCRIPT_VERIFY_NULLFAIL | SCRIPT_VERIFY_CHECKLOCKTIMEVERIFY | SCRIPT_VERIFY_CHECKSEQUENCEVERIFY |
                              SCRIPT_ENABLE_SIGHASH_FORKID; // GetBlockScriptFlags post fork (src/validation.cpp:1803-1852)
static const uint8_t HT = SIGHASH_ALL | SIGHASH_FORKID;

class RecordingChecker : public TransactionSignatureChecker {
public:
    using TransactionSignatureChecker::TransactionSignatureChecker;

protected:
    bool VerifySignature(const std::vector<uint8_t> &vchSig, const CPubKey &pubkey, const uint256 &sighash) const override {
        bcp::Recorder *r = bcp::Recorder::current();
        if (!r) abort();
        return r->record(sighash.begin(), pubkey.begin(), pubkey.size(), vchSig.data(), vchSig.size(), nullptr);
    }
};
struct ScriptCheck { // the shape of the reference's CScriptCheck (src/validation.h:499-537)
    CScript scriptPubKey;
    Amount amount;
    const CTransaction *ptxTo = nullptr;
    unsigned nIn = 0;
    uint32_t nFlags = 0;
    ScriptError error = SCRIPT_ERR_UNKNOWN_ERROR;
    ScriptCheck() : amount(0) {}
    ScriptCheck(const CScript &spk, Amount a, const CTransaction *tx, unsigned n, uint32_t f) : scriptPubKey(spk), amount(a), ptxTo(tx), nIn(n), nFlags(f) {}
    bool operator()() { return VerifyScript(ptxTo->vin[nIn].scriptSig, scriptPubKey, nFlags, RecordingChecker(ptxTo, nIn, amount), &error); }
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
static Key make_key(uint32_t i, bool compressed) {
    Key k;
    uint8_t in[8] = {'m', 'i', 'x', 0};
    memcpy(in + 4, &i, 4);
    CSHA256().Write(in, 8).Finalize(k.sk);
    uint8_t out[65];
    size_t ol = 65;
    if (!ref_pubkey_create(k.sk, compressed ? 1 : 0, out, &ol)) abort();
    k.pub.assign(out, out + ol);
    return k;
}
static std::vector<uint8_t> sign(const Key &k, const uint256 &h) {
    uint8_t s64[64], der[80];
    size_t dl = sizeof der;
    if (!ref_sign(k.sk, h.begin(), s64) || !ref_sig_to_der(s64, der, &dl)) abort();
    std::vector<uint8_t> v(der, der + dl);
    v.push_back(HT);
    return v;
}
static CScript p2pkh(const Key &k) {
    const uint160 id = Hash160(k.pub.begin(), k.pub.end());
    return CScript() << OP_DUP << OP_HASH160 << std::vector<uint8_t>(id.begin(), id.end()) << OP_EQUALVERIFY << OP_CHECKSIG;
}
static CScript p2sh(const CScript &redeem) {
    const uint160 id = Hash160(redeem.begin(), redeem.end());
    return CScript() << OP_HASH160 << std::vector<uint8_t>(id.begin(), id.end()) << OP_EQUAL;
}
static CScript multisig(int m, const std::vector<Key> &keys) {
    CScript s;
    s << CScript::EncodeOP_N(m);
    for (const Key &k : keys) s << k.pub;
    s << CScript::EncodeOP_N((int)keys.size()) << OP_CHECKMULTISIG;
    return s;
}

struct Coin {
    uint256 txid;
    uint32_t n;
    CScript spk;
    Amount amount{0};
    int kind; // 0 p2pkh, 1 p2pk, 2 bare 2-of-3, 3 p2sh 2-of-3, 4 p2sh-p2pkh
    std::vector<Key> keys;
    CScript redeem;
};
enum Defect { NONE = 0, BAD_P2SH_MULTISIG_SIG, MULTISIG_WRONG_ORDER, BAD_P2PKH_SIG };

// one transaction per coin (a few with two inputs), every input signed; `defect` damages exactly one input
static std::vector<CTransactionRef> spend_all(const std::vector<Coin> &coins, Defect defect, size_t &bad_input) {
    std::vector<CTransactionRef> txs;
    bad_input = (size_t)-1;
    size_t input_no = 0;
    bool done = false;
    for (size_t c = 0; c < coins.size();) {
        const size_t take = (c % 5 == 3 && c + 1 < coins.size()) ? 2 : 1;
        CMutableTransaction m;
        m.nVersion = 2;
        m.vin.resize(take);
        m.vout.resize(1);
        Amount in_total(0);
        for (size_t k = 0; k < take; k++) {
            m.vin[k].prevout.hash = coins[c + k].txid;
            m.vin[k].prevout.n = coins[c + k].n;
            m.vin[k].nSequence = 0xffffffff;
            in_total += coins[c + k].amount;
        }
        m.vout[0].nValue = in_total - Amount(1000);
        m.vout[0].scriptPubKey = p2pkh(coins[c].keys[0]);
        const CTransaction pre(m);
        for (size_t k = 0; k < take; k++, input_no++) {
            const Coin &coin = coins[c + k];
            const CScript &code = (coin.kind == 3 || coin.kind == 4) ? coin.redeem : coin.spk;
            const uint256 h = SignatureHash(code, pre, (unsigned)k, HT, coin.amount, nullptr, FLAGS);
            CScript s;
            if (coin.kind == 0 || coin.kind == 4) {
                std::vector<uint8_t> sg = sign(coin.keys[0], h);
                if (defect == BAD_P2PKH_SIG && coin.kind == 0 && !done && input_no > 6) { sg[9] ^= 1; done = true; bad_input = input_no; }
                s << sg << coin.keys[0].pub;
            } else if (coin.kind == 1) {
                s << sign(coin.keys[0], h);
            } else {
                std::vector<uint8_t> a = sign(coin.keys[0], h), b = sign(coin.keys[2], h);
                if (defect == BAD_P2SH_MULTISIG_SIG && coin.kind == 3 && !done && input_no > 4) { b[10] ^= 2; done = true; bad_input = input_no; }
                if (defect == MULTISIG_WRONG_ORDER && coin.kind == 2 && !done && input_no > 2) { std::swap(a, b); done = true; bad_input = input_no; }
                s << OP_0 << a << b;
            }
            if (coin.kind == 3 || coin.kind == 4) s << std::vector<uint8_t>(coin.redeem.begin(), coin.redeem.end());
            m.vin[k].scriptSig = s;
        }
        txs.push_back(MakeTransactionRef(CTransaction(m)));
        c += take;
    }
    CHECK(defect == NONE || done);
    return txs;
}

static std::vector<uint8_t> make_block(const std::vector<CTransactionRef> &txs, uint32_t height) {
    CMutableTransaction cb;
    cb.nVersion = 2;
    cb.vin.resize(1);
    cb.vin[0].prevout.SetNull();
    cb.vin[0].scriptSig = CScript() << (int64_t)height << OP_0;
    cb.vin[0].nSequence = 0xffffffff;
    cb.vout.resize(1);
    cb.vout[0].nValue = Amount(1250000000);
    cb.vout[0].scriptPubKey = CScript() << OP_1;
    std::vector<CTransactionRef> all{MakeTransactionRef(CTransaction(cb))};
    all.insert(all.end(), txs.begin(), txs.end());
    std::vector<uint256> leaves;
    for (const auto &t : all) leaves.push_back(t->GetId());
    const uint256 root = ComputeMerkleRoot(leaves, nullptr);
    // post-fork header layout (src/primitives/block.h:49-73) with an EMPTY Equihash solution: the connectors are run with
    // the header check off (it has its own tests); 140 bytes + compact size 0
    std::vector<uint8_t> blk(141, 0);
    blk[0] = 4;
    memcpy(&blk[36], root.begin(), 32);
    memcpy(&blk[68], &height, 4);
    CDataStream ss(SER_NETWORK, PROTOCOL_VERSION);
    WriteCompactSize(ss, all.size());
    for (const auto &t : all) ss << *t;
    blk.insert(blk.end(), (const uint8_t *)ss.data(), (const uint8_t *)ss.data() + ss.size());
    return blk;
}

// the reference's verdict for the whole block: VerifyScript with the plain CPU checker over every input
static bool reference_verdict(const std::vector<CTransactionRef> &txs, const std::vector<Coin> &coins) {
    bool all = true;
    size_t c = 0;
    for (const auto &t : txs)
        for (unsigned k = 0; k < t->vin.size(); k++, c++) {
            ScriptError e;
            all = VerifyScript(t->vin[k].scriptSig, coins[c].spk, FLAGS, TransactionSignatureChecker(t.get(), k, coins[c].amount), &e) && all;
        }
    return all;
}

static size_t g_other_calls = 0, g_other_inputs = 0;
static std::mutex g_mu;
// what a node installs: CScriptCheck objects through the check queue (host interpreter, signatures on the GPU)
static bool other_scripts(CCheckQueue<ScriptCheck> *queue, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                          const bcp::OtherInput *in, size_t n, uint32_t flags, std::string &reason) {
    std::vector<CTransactionRef> keep;
    std::vector<ScriptCheck> checks;
    uint32_t last_tx = 0xFFFFFFFFu;
    for (size_t i = 0; i < n; i++) {
        if (in[i].tx != last_tx) {
            CDataStream ss((const char *)blk + tx_off[in[i].tx], (const char *)blk + tx_off[in[i].tx + 1], SER_NETWORK, PROTOCOL_VERSION);
            keep.push_back(MakeTransactionRef(CTransaction(deserialize, ss)));
            last_tx = in[i].tx;
        }
        checks.emplace_back(CScript(in[i].spk, in[i].spk + in[i].spk_len), Amount(in[i].amount), keep.back().get(), in[i].n, flags);
    }
    {
        std::lock_guard<std::mutex> l(g_mu);
        g_other_calls++;
        g_other_inputs += n;
    }
    CCheckQueueControl<ScriptCheck> control(queue);
    control.Add(checks);
    if (control.Wait()) return true;
    reason = "blk-bad-inputs (an input outside the fast path failed its script check)";
    return false;
}

template <class Chain> static void run(const char *name, Chain &chain, CCheckQueue<ScriptCheck> &queue, const std::vector<Coin> &coins) {
    std::string reason;
    size_t bad;
    // 1. no callback: refused, names the first input that is not P2PKH
    {
        auto txs = spend_all(coins, NONE, bad);
        auto blk = make_block(txs, 1000);
        CHECK(!chain.ConnectBlock(blk.data(), blk.size(), reason));
        CHECK(reason.find("script-not-in-fast-path (input 1)") == 0);
    }
    chain.SetOtherScripts([&queue](const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx, const bcp::OtherInput *in, size_t n,
                                   uint32_t flags, std::string &why) { return other_scripts(&queue, blk, len, tx_off, ntx, in, n, flags, why); });
    // 2. every defect: refused, exactly like the reference's CPU verdict; the table is untouched
    const Defect defects[] = {BAD_P2SH_MULTISIG_SIG, MULTISIG_WRONG_ORDER, BAD_P2PKH_SIG};
    for (Defect d : defects) {
        auto txs = spend_all(coins, d, bad);
        CHECK(!reference_verdict(txs, coins));
        auto blk = make_block(txs, 1000);
        CHECK(!chain.ConnectBlock(blk.data(), blk.size(), reason));
        CHECK(reason.find("blk-bad-inputs") == 0);
        if (d == BAD_P2PKH_SIG) CHECK(reason == "blk-bad-inputs (input " + std::to_string(bad) + ", script error 31)");
        printf("  %s: defect %d refused: %s\n", name, (int)d, reason.c_str());
    }
    // 3. the valid block connects (the refusals above left every coin in place), and its outputs are the new coins
    auto txs = spend_all(coins, NONE, bad);
    CHECK(reference_verdict(txs, coins));
    auto blk = make_block(txs, 1000);
    const size_t before = (size_t)chain.UtxoSize();
    CHECK(chain.ConnectBlock(blk.data(), blk.size(), reason));
    CHECK((size_t)chain.UtxoSize() == before - coins.size() + txs.size() + 1);
    // 4. connecting it again: its coins are spent
    CHECK(!chain.ConnectBlock(blk.data(), blk.size(), reason));
    CHECK(reason == "bad-txns-BIP30" || reason == "bad-txns-inputs-missingorspent");
    printf("  %s ok\n", name);
}

int main() {
    ref_init();
    if (bcp_gpu_init(0) != BCP_OK) {
        fprintf(stderr, "no CUDA device: %s\n", bcp_gpu_last_error());
        return 2;
    }
    // the coins: 60 outputs of the five kinds
    std::vector<Coin> coins;
    for (uint32_t i = 0; i < 60; i++) {
        Coin c;
        uint8_t seed[8] = {'c', 'o', 'i', 'n'};
        memcpy(seed + 4, &i, 4);
        CSHA256().Write(seed, 8).Finalize(c.txid.begin());
        c.n = i % 3;
        c.amount = Amount(int64_t(500000 + i));
        c.kind = i == 0 ? 0 : i % 5;
        for (uint32_t k = 0; k < 3; k++) c.keys.push_back(make_key(10 * i + k, (i + k) % 4 != 0));
        if (c.kind == 0) c.spk = p2pkh(c.keys[0]);
        else if (c.kind == 1) c.spk = CScript() << c.keys[0].pub << OP_CHECKSIG;
        else if (c.kind == 2) c.spk = multisig(2, c.keys);
        else if (c.kind == 3) { c.redeem = multisig(2, c.keys); c.spk = p2sh(c.redeem); }
        else { c.redeem = p2pkh(c.keys[0]); c.spk = p2sh(c.redeem); }
        coins.push_back(c);
    }
    CCheckQueue<ScriptCheck> queue(16);
    std::vector<std::thread> workers;
    for (int i = 0; i < 3; i++) workers.emplace_back([&] { queue.Thread(); });
    std::string err;
    {
        bcp::ChainValidator chain(std::vector<int>{0}, FLAGS, 0, 0, 4);
        CHECK(chain.Init(err));
        for (const Coin &c : coins) chain.AddCoin(c.txid.begin(), c.n, c.spk.data(), c.spk.size(), c.amount.GetSatoshis());
        run("bcp::ChainValidator", chain, queue, coins);
    }
    {
        bcp::DeviceChain chain(std::vector<int>{0}, FLAGS, 0, 0, 12, 2, 2);
        CHECK(chain.Init(err));
        for (const Coin &c : coins) chain.AddCoin(c.txid.begin(), c.n, c.spk.data(), c.spk.size(), c.amount.GetSatoshis());
        run("bcp::DeviceChain", chain, queue, coins);
    }
    queue.Quit();
    for (auto &t : workers) t.join();
    printf("%zu calls handed %zu inputs to the host interpreter\n", g_other_calls, g_other_inputs);
    CHECK(g_other_calls >= 6 && g_other_inputs >= 6 * 48);
    printf("mixed ok\n");
    return 0;
}
