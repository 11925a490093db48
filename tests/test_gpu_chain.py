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


def chain_lib(path=None, impl="host"):
    """ctypes face of libbcp_chain.so.  impl = "host": bcp::ChainValidator (coin map on host threads); "device":
    bcp::DeviceChain (coin table in HBM, whole blocks dealt out to the GPUs) behind the same function names."""
    L = C.CDLL(path or os.path.join(ROOT, "bitcoincashplus_b200", "libbcp_chain.so"))
    L.bcp_chain_new.restype = C.c_void_p
    L.bcp_chain_new.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_char_p, C.c_size_t]
    L.bcp_chain_new_threads.restype = C.c_void_p
    L.bcp_chain_new_threads.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_int, C.c_char_p, C.c_size_t]
    L.bcp_dchain_new.restype = C.c_void_p
    L.bcp_dchain_new.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_char_p, C.c_size_t]
    for pre in ("bcp_chain_", "bcp_dchain_"):
        getattr(L, pre + "free").argtypes = [C.c_void_p]
        getattr(L, pre + "add_coin").argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_size_t, C.c_int64]
        getattr(L, pre + "utxo_size").restype = C.c_size_t
        getattr(L, pre + "utxo_size").argtypes = [C.c_void_p]
        getattr(L, pre + "inputs_checked").restype = C.c_size_t
        getattr(L, pre + "inputs_checked").argtypes = [C.c_void_p]
        getattr(L, pre + "connect_block").argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        getattr(L, pre + "connect_file").restype = C.c_long
        getattr(L, pre + "connect_file").argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_long)]
        getattr(L, pre + "last_hash").argtypes = [C.c_void_p, C.c_char_p]
    if impl == "host":
        L.impl = "host"
        return L

    class DeviceFace:  # the DeviceChain entry points under the names the tests use
        impl = "device"
        bcp_dchain_new = L.bcp_dchain_new

    for name in ("free", "add_coin", "utxo_size", "inputs_checked", "connect_block", "connect_file", "last_hash"):
        setattr(DeviceFace, "bcp_chain_" + name, staticmethod(getattr(L, "bcp_dchain_" + name)))
    return DeviceFace


def new_chain(L, seeds, devices=(0,), eh=(48, 5), walk_threads=None):
    devs = (C.c_int * len(devices))(*devices)
    err = C.create_string_buffer(256)
    if L.impl == "device":  # walk_threads has no meaning there: vary the check threads per GPU instead
        lanes = 2 if walk_threads is None else 1 + (walk_threads > 1) + (walk_threads > 4)
        h = L.bcp_dchain_new(devs, len(devices), synth.POSTFORK_FLAGS, eh[0], eh[1], 12, lanes, 2, err, 256)
    elif walk_threads is None:
        h = L.bcp_chain_new(devs, len(devices), synth.POSTFORK_FLAGS, eh[0], eh[1], err, 256)
    else:
        h = L.bcp_chain_new_threads(devs, len(devices), synth.POSTFORK_FLAGS, eh[0], eh[1], walk_threads, err, 256)
    assert h, err.value
    for txid, vout, spk, amount in seeds:
        L.bcp_chain_add_coin(h, txid, vout, spk, len(spk), amount)
    return h


@pytest.fixture(params=["host", "device"])
def impl(request):
    return request.param


def test_chain_connects_and_tracks_the_utxo_set(oracle, impl):
    L = chain_lib(impl=impl)
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
    # replaying the last block: its outputs are still unspent — the BIP30 rule (src/validation.cpp:1975-1990), which
    # ConnectBlock applies before it looks at the inputs (which are gone as well)
    assert L.bcp_chain_connect_block(h, blocks[-1], len(blocks[-1]), reason, 256) == 0
    assert reason.value == b"bad-txns-BIP30"
    assert L.bcp_chain_utxo_size(h) == 30 + 8
    # an older block: its outputs were spent by its successor, so BIP30 is silent and the missing inputs are reported
    assert L.bcp_chain_connect_block(h, blocks[2], len(blocks[2]), reason, 256) == 0
    assert reason.value == b"bad-txns-BIP30" or reason.value == b"bad-txns-inputs-missingorspent"
    assert L.bcp_chain_utxo_size(h) == 30 + 8
    L.bcp_chain_free(h)


@pytest.mark.parametrize("kind,expect", [("bad_sig", b"blk-bad-inputs (input 3, script error 31)"),
                                         ("double_spend", b"bad-txns-inputs-missingorspent"),
                                         ("missing_input", b"bad-txns-inputs-missingorspent"),
                                         ("bad_merkle", b"bad-txnmrklroot"), ("bad_equihash", b"invalid-solution")])
def test_chain_rejects_the_damaged_block_with_the_reference_reason(kind, expect, impl):
    L = chain_lib(impl=impl)
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


@pytest.mark.parametrize("kind", sorted(synth.STRUCTURE_KINDS))
def test_chain_applies_checkblock_structure_rules(kind, impl):
    """CheckBlock's context-free transaction rules (src/validation.cpp:3255-3330 -> CheckCoinbase / CheckRegularTransaction
    :476-570): a second coinbase-shaped transaction, a missing coinbase, null prevouts, empty input / output lists and
    out-of-range amounts are rejected with the reference's reason and leave the UTXO set untouched."""
    L = chain_lib(impl=impl)
    blocks, seeds, _ = synth.synthetic_chain(3, 10, corrupt=(1, kind))
    good, _, _ = synth.synthetic_chain(3, 10)
    reason = C.create_string_buffer(256)
    for threads in (1, 4):
        h = new_chain(L, seeds, walk_threads=threads)
        assert L.bcp_chain_connect_block(h, blocks[0], len(blocks[0]), reason, 256) == 1, reason.value
        assert L.bcp_chain_connect_block(h, blocks[1], len(blocks[1]), reason, 256) == 0
        assert reason.value == synth.STRUCTURE_KINDS[kind], (kind, reason.value)
        assert L.bcp_chain_utxo_size(h) == 10 + 1
        assert L.bcp_chain_connect_block(h, good[1], len(good[1]), reason, 256) == 1, (kind, reason.value)
        L.bcp_chain_free(h)


def test_chain_in_block_index_survives_txid_prefix_collisions(oracle):
    """The in-block transaction index is keyed by the first 8 bytes of the txid; two transactions of a block sharing them
    (2^32 hashes to craft) must not hide one another.  A test build of the connector narrows the key to 4 bits
    (BCP_CHAIN_TEST_TXKEY_BITS), so nearly every pair of transactions collides, and the chains of in-block spends must
    connect exactly as with the full key — and the two failure modes must still be caught."""
    path = os.path.join(ROOT, "tests", "cpp", "libbcp_chain_txkey4.so")
    if not os.path.exists(path):
        pytest.skip("tests/cpp/libbcp_chain_txkey4.so not built (__graft_entry__.build())")
    L = chain_lib(path)
    inb = (2, 7, 8, 9, 15, 16, 17, 30, 31)
    blocks, seeds, _ = synth.synthetic_chain(3, 40, inblock=inb)
    reason = C.create_string_buffer(256)
    for threads in (1, 8):
        h = new_chain(L, seeds, walk_threads=threads)
        for i, b in enumerate(blocks):
            assert L.bcp_chain_connect_block(h, b, len(b), reason, 256) == 1, (threads, i, reason.value)
            assert L.bcp_chain_utxo_size(h) == 40 + (i + 1)
        L.bcp_chain_free(h)
    for kind in ("inblock_reversed", "inblock_double"):
        bad, seeds2, _ = synth.synthetic_chain(2, 20, corrupt=(1, kind))
        h = new_chain(L, seeds2)
        assert L.bcp_chain_connect_block(h, bad[0], len(bad[0]), reason, 256) == 1
        assert L.bcp_chain_connect_block(h, bad[1], len(bad[1]), reason, 256) == 0
        assert reason.value == b"bad-txns-inputs-missingorspent", (kind, reason.value)
        L.bcp_chain_free(h)


def test_chain_shards_inputs_over_two_devices_when_present(impl):
    import bitcoincashplus_b200 as bcp

    if bcp.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    L = chain_lib(impl=impl)
    blocks, seeds, _ = synth.synthetic_chain(4, 50, corrupt=(2, "bad_sig"))
    h = new_chain(L, seeds, devices=(0, 1))
    reason = C.create_string_buffer(256)
    res = [L.bcp_chain_connect_block(h, b, len(b), reason, 256) for b in blocks[:3]]
    assert res == [1, 1, 0] and reason.value.startswith(b"blk-bad-inputs (input 3")
    L.bcp_chain_free(h)


def test_device_chain_leaves_out_a_gpu_without_peer_access(monkeypatch):
    """bcp::DeviceChain deals blocks only to GPUs that reach the coin table's GPU by a direct device-to-device copy (a
    staged one per block costs more than the GPU brings: measured 340 instead of 1,200 blocks/s on a 2-GPU box without
    peer access).  BCP_CHAIN_NO_PEER=1 treats every other GPU as such a one: the chain must then run on the table's GPU
    alone, with the same verdicts."""
    import bitcoincashplus_b200 as bcp

    if bcp.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import ctypes

    raw = ctypes.CDLL(os.path.join(os.path.dirname(bcp.__file__), "libbcp_chain.so"))
    L = chain_lib(impl="device")
    blocks, seeds, _ = synth.synthetic_chain(4, 50, corrupt=(2, "bad_sig"))
    for no_peer, want in (("1", 1), ("0", None)):
        monkeypatch.setenv("BCP_CHAIN_NO_PEER", no_peer)
        h = new_chain(L, seeds, devices=(0, 1))
        in_use = raw.bcp_dchain_gpus_in_use(C.c_void_p(h))
        assert in_use == (want if want else (2 if bcp._lib.lib().bcp_gpu_peer_access(0, 1) == 1 else 1))
        reason = C.create_string_buffer(256)
        res = [L.bcp_chain_connect_block(h, b, len(b), reason, 256) for b in blocks[:3]]
        assert res == [1, 1, 0] and reason.value.startswith(b"blk-bad-inputs (input 3")
        L.bcp_chain_free(h)


def test_chain_shards_inputs_over_two_lane_pairs_of_one_gpu():
    """The multi-device path on a one-GPU box: lanes 0-1 and lanes 2-3 of GPU 0 stand in for two devices, each block's
    inputs are split in two index ranges, and a bad signature in the SECOND range (owned by the second "device") must be
    found there and reported with its block-wide input number; valid blocks connect, and the ConnectFile pipeline (two
    blocks in flight) reports the same."""
    L = chain_lib()
    devs = (0, 2 << 8)
    for ntx, bad_at in ((50, 40), (7, 6)):
        blocks, seeds, _ = synth.synthetic_chain(4, ntx, corrupt=(2, "bad_sig@%d" % bad_at))
        h = new_chain(L, seeds, devices=devs)
        reason = C.create_string_buffer(256)
        res = [L.bcp_chain_connect_block(h, b, len(b), reason, 256) for b in blocks[:3]]
        assert res == [1, 1, 0] and reason.value == b"blk-bad-inputs (input %d, script error 31)" % bad_at, reason.value
        assert L.bcp_chain_utxo_size(h) == ntx + 2
        L.bcp_chain_free(h)
        h = new_chain(L, seeds, devices=devs)
        data = b"".join(MAGIC + len(b).to_bytes(4, "little") + b for b in blocks)
        failed = C.c_long(-2)
        n = L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed))
        assert n == 2 and failed.value == 2 and reason.value.startswith(b"blk-bad-inputs (input %d," % bad_at)
        L.bcp_chain_free(h)


def test_device_chain_deals_blocks_over_several_check_threads_and_unwinds_a_failure_in_the_middle():
    """bcp::DeviceChain with GPU 0 listed twice (two "devices" x two lanes = four check threads, six blocks in flight): a
    long chain connects in order; with a bad signature in block 5 of 12 the blocks applied on top of it — already in the
    table when its verdict arrives — are rolled back newest first, the table is exactly as after block 4 (the undamaged
    rest of the chain connects from there), and the first failure is the one reported even though later blocks finish
    their checks earlier or later than it."""
    L = chain_lib(impl="device")
    good, seeds, _ = synth.synthetic_chain(12, 30, inblock=(3, 4, 11))
    bad, _, _ = synth.synthetic_chain(12, 30, inblock=(3, 4, 11), corrupt=(5, "bad_sig@17"))
    reason = C.create_string_buffer(256)
    failed = C.c_long(-2)
    frame = lambda blocks: b"".join(MAGIC + len(b).to_bytes(4, "little") + b for b in blocks)
    h = new_chain(L, seeds, devices=(0, 0))
    data = frame(good)
    assert L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed)) == 12, reason.value
    assert failed.value == -1 and L.bcp_chain_utxo_size(h) == 30 + 12 and L.bcp_chain_inputs_checked(h) == 12 * 30
    L.bcp_chain_free(h)
    h = new_chain(L, seeds, devices=(0, 0))
    data = frame(bad)
    assert L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed)) == 5
    assert failed.value == 5 and reason.value == b"blk-bad-inputs (input 17, script error 31)", reason.value
    assert L.bcp_chain_utxo_size(h) == 30 + 5
    data = frame(good[5:])
    assert L.bcp_chain_connect_file(h, data, len(data), MAGIC, reason, 256, C.byref(failed)) == 7, reason.value
    assert L.bcp_chain_utxo_size(h) == 30 + 12
    out = C.create_string_buffer(32)
    L.bcp_chain_last_hash(h, out)
    from oracle import pyoracle as po
    assert out.raw == po.oracle().header_hash(good[-1][:177])
    L.bcp_chain_free(h)


def test_chain_in_block_spends_and_their_failure_modes(oracle, impl):
    """Transactions that spend outputs created earlier in the SAME block (chains of them), over several blocks; then the
    two ways this can go wrong: spending an output of a LATER transaction, and spending an in-block output twice.  After a
    rejection the UTXO set is exactly what it was, so the undamaged version of the same block still connects."""
    L = chain_lib(impl=impl)
    inb = (2, 7, 8, 9, 15)
    blocks, seeds, prevs = synth.synthetic_chain(4, 20, inblock=inb)
    for b, (spk, amt) in zip(blocks, prevs):
        assert not oracle.block_script_errors(b, False, spk, amt, synth.POSTFORK_FLAGS, os.cpu_count() or 1).any()
    reason = C.create_string_buffer(256)
    for threads in (1, 8):
        h = new_chain(L, seeds, walk_threads=threads)
        for i, b in enumerate(blocks):
            assert L.bcp_chain_connect_block(h, b, len(b), reason, 256) == 1, (threads, i, reason.value)
            assert L.bcp_chain_utxo_size(h) == 20 + (i + 1)
        L.bcp_chain_free(h)
    for kind in ("inblock_reversed", "inblock_double"):
        bad, seeds2, _ = synth.synthetic_chain(3, 20, corrupt=(1, kind))
        good, _, _ = synth.synthetic_chain(3, 20)
        assert bad[0] == good[0] and bad[1] != good[1]
        for threads in (1, 4):
            h = new_chain(L, seeds2, walk_threads=threads)
            assert L.bcp_chain_connect_block(h, bad[0], len(bad[0]), reason, 256) == 1
            assert L.bcp_chain_connect_block(h, bad[1], len(bad[1]), reason, 256) == 0
            assert reason.value == b"bad-txns-inputs-missingorspent", (kind, reason.value)
            assert L.bcp_chain_utxo_size(h) == 20 + 1
            assert L.bcp_chain_connect_block(h, good[1], len(good[1]), reason, 256) == 1, (kind, reason.value)
            assert L.bcp_chain_connect_block(h, good[2], len(good[2]), reason, 256) == 1, (kind, reason.value)
            assert L.bcp_chain_utxo_size(h) == 20 + 3
            L.bcp_chain_free(h)


def test_chain_rollback_restores_coins_with_long_scripts(impl):
    """Coins whose scriptPubKey does not fit the staging slot (> 64 bytes) take the one-thread path; a block spending one
    is refused by the P2PKH fast path (script-not-in-fast-path), and the rollback must put every coin back, long ones
    included — shown by the same block failing the same way again and by the count.  (In the device table such scripts
    live in the arena behind the 64-byte slots.)"""
    L = chain_lib(impl=impl)
    blocks, seeds, _ = synth.synthetic_chain(2, 16)
    long_spk = bytes([0x51, 0x4c, 100]) + bytes(100)  # 103 bytes, not the P2PKH shape
    seeds = list(seeds)
    seeds[4] = (seeds[4][0], seeds[4][1], long_spk, seeds[4][3])
    seeds[9] = (seeds[9][0], seeds[9][1], long_spk + b"\x00", seeds[9][3])
    reason = C.create_string_buffer(256)
    for threads in (1, 8):
        h = new_chain(L, seeds, walk_threads=threads)
        for _ in range(2):
            assert L.bcp_chain_connect_block(h, blocks[0], len(blocks[0]), reason, 256) == 0
            assert reason.value.startswith(b"script-not-in-fast-path (input 4)"), reason.value
            assert L.bcp_chain_utxo_size(h) == 16
        L.bcp_chain_free(h)
