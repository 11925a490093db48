// include/compat/bcp_other_scripts.h — where the block connectors (bcp_chain.h, bcp_chain_gpu.h) hand over the inputs the
// device fast path does not evaluate.  The fast path answers per input either a ScriptError or 255 = "not the P2PKH shape"
// (bare and P2SH multisig, P2PK, anything else).  The reference runs EVERY input through the same queue
// (CheckInputs -> CScriptCheck -> CCheckQueue, src/validation.cpp:1474-1477, 2080-2100); so does a node built on this
// library: the callback below is where it makes CScriptCheck objects for those inputs and pushes them through CCheckQueue
// (include/compat/checkqueue.h: host interpreter, signatures recorded and verified on the GPU in Wait()).  The connector
// ANDs the two results like `fAllOk &= fOk` (src/checkqueue.h:77).  tests/cpp/test_mixed_ref.cpp does exactly that with the
// reference's own interpreter.
#ifndef BCP_COMPAT_OTHER_SCRIPTS_H
#define BCP_COMPAT_OTHER_SCRIPTS_H
#include <cstddef>
#include <cstdint>
#include <functional>
#include <string>

namespace bcp {

struct OtherInput {
    uint32_t input_no; // ordinal among the block's non-coinbase inputs
    uint32_t tx;       // transaction number in the block
    uint32_t n;        // input index inside that transaction
    const uint8_t *spk; // the spent output's scriptPubKey ...
    uint32_t spk_len;
    int64_t amount;     // ... and value
};

// Called at most once per block, with every such input in block order.  true: all of them verify.  false: at least one does
// not; `reason` is what the connector reports (ConnectBlock's "blk-bad-inputs ..." wording is the callee's to fill in).
typedef std::function<bool(const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx, const OtherInput *inputs, size_t n,
                           uint32_t script_flags, std::string &reason)>
    OtherScripts;

} // namespace bcp
#endif
