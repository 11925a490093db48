"""-m gpu: the host-side block connector (include/compat/bcp_chain.h) over a synthetic chain with real UTXO
dependencies and valid regtest Equihash (48,5) headers — BASELINE.json configs[4] in miniature."""
import ctypes as C
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAGIC = bytes.fromhex("e48f1d1c")


def chain_lib():
    L = C.CDLL(os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_chain.so"))
    L.bcp_chain_new.restype = C.c_void_p
    L.bcp_chain_new.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_char_p, C.c_size_t]
    L.bcp_chain_free.argtypes = [C.c_void_p]
    L.bcp_chain_add_coin.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_size_t, C.c_int64]
    L.bcp_chain_utxo_size.restype = C.c_size_t
    L.bcp_chain_utxo_size.argtypes = [C.c_void_p]
    L.bcp_chain_inputs_checked.restype = C.c_size_t
    L.bcp_chain_inputs_checked.argtypes = [C.c_void_p]
    L.bcp_chain_connect_block.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.bcp_chain_connect_file.restype = C.c_long
    L.bcp_chain_connect_file.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_long)]
    L.bcp_chain_last_hash.argtypes = [C.c_void_p, C.c_char_p]
    return L


def new_chain(L, seeds, devices=(0,), eh=(48, 5)):
    devs = (C.c_int * len(devices))(*devices)
    err = C.create_string_buffer(256)
    h = L.bcp_chain_new(devs, len(devices), synth.POSTFORK_FLAGS, eh[0], eh[1], err, 256)
    assert h, err.value
    for txid, vout, spk, amount in seeds:
        L.bcp_chain_add_coin(h, txid, vout, spk, len(spk), amount)
    return h


def test_chain_connects_and_tracks_the_utxo_set(oracle):
    L = chain_lib()
    blocks, seeds, prevs = synth.synthetic_chain(8, 30)
    # the generator's own claim, checked with the oracle: every block's inputs verify against the coins they spend
    for b, (spk, amt) in zip(blocks, prevs):
        assert not oracle.block_script_errors(b, False, spk, amt, synth.POSTFORK_FLAGS, os.cpu_count() or 1).any()
    h = new_chain(L, seeds)
    reason = C.create_string_buffer(256)
    for i, b in enumerate(blocks):
        assert L.bcp_chain_connect_block(h, b, len(b), reason, 256) == 1, (i, reason.value)
        out = C.create_string_buffer(32)
        L.bcp_chain_last_hash(h, out)
        assert out.raw == oracle.header_hash(b[:177])
        # 30 seed coins spent and replaced each block, plus one coinbase output per block
        assert L.bcp_chain_utxo_size(h) == 30 + (i + 1)
    assert L.bcp_chain_inputs_checked(h) == 8 * 30
    # replaying the last block: its inputs are gone
    assert L.bcp_chain_connect_block(h, blocks[-1], len(blocks[-1]), reason, 256) == 0
    assert reason.value == b"bad-txns-inputs-missingorspent"
    L.bcp_chain_free(h)


@pytest.mark.parametrize("kind,expect", [("bad_sig", b"blk-bad-inputs (input 3, script error 31)"),
                                         ("double_spend", b"bad-txns-inputs-missingorspent"),
                                         ("missing_input", b"bad-txns-inputs-missingorspent"),
                                         ("bad_merkle", b"bad-txnmrklroot"), ("bad_equihash", b"invalid-solution")])
def test_chain_rejects_the_damaged_block_with_the_reference_reason(kind, expect):
    L = chain_lib()
    blocks, seeds, _ = synth.synthetic_chain(5, 12, corrupt=(3, kind))
    h = new_chain(L, seeds)
    data = b"".join(MAGIC + len(b).to_bytes(4, "little") + b for b in blocks)
    data = b"\x00\x01junk" + data  # the reader resynchronises on the magic like LoadExternalBlockFile
    reason = C.create_string_buffer(256)
    failed = C.c_long(-2)
    n = L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed))
    assert n == 3 and failed.value == 3 and reason.value == expect, (n, failed.value, reason.value)
    size_before = L.bcp_chain_utxo_size(h)
    assert size_before == 12 + 3  # a rejected block leaves the UTXO set untouched
    L.bcp_chain_free(h)


def test_chain_shards_inputs_over_two_devices_when_present():
    import bitcoincashplus_b200 as bcp

    if bcp.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    L = chain_lib()
    blocks, seeds, _ = synth.synthetic_chain(4, 50, corrupt=(2, "bad_sig"))
    h = new_chain(L, seeds, devices=(0, 1))
    reason = C.create_string_buffer(256)
    res = [L.bcp_chain_connect_block(h, b, len(b), reason, 256) for b in blocks[:3]]
    assert res == [1, 1, 0] and reason.value.startswith(b"blk-bad-inputs (input 3")
    L.bcp_chain_free(h)
