"""-m gpu, collected last on purpose: the allocation-free Merkle entry point.

bcp_gpu_merkle_root_dev_static serves callers that cannot allocate device memory — the validation
thread, which must not hit cudaMalloc's device-wide synchronisation while holding the node's main lock,
or an embedder with a fixed VRAM budget.  Its inner-node scratch is a static device array inside the
library, shared by every call.  The test drives it through the C ABI and requires bit-identical roots and
`mutated` flags to the existing bcp_gpu_merkle_root_dev (caller-provided scratch), on two calls in a row
— a large tree first, then a smaller one that finds the first call's stale nodes still in the array —
and checks that an over-size request is refused with a message instead of being computed."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


_STREAM = {}


def _roots(L, dev, torch, leaves, static, dev_id=0):
    # one explicit (non-default) stream orders the buffer fills, both entry points and the read-back
    st = _STREAM.setdefault("s", torch.cuda.Stream(dev))
    n = leaves.shape[0]
    with torch.cuda.stream(st):
        d_leaves = torch.from_numpy(leaves).to(dev)
        d_root = torch.full((32,), 0xAA, dtype=torch.uint8, device=dev)
        d_mut = torch.full((1,), 7, dtype=torch.int32, device=dev)
        d_scr = torch.empty(L.bcp_gpu_merkle_scratch_bytes(n), dtype=torch.uint8, device=dev)
        if static:
            rc = L.bcp_gpu_merkle_root_dev_static(dev_id, d_leaves.data_ptr(), n, d_root.data_ptr(), d_mut.data_ptr(),
                                                  st.cuda_stream)
        else:
            rc = L.bcp_gpu_merkle_root_dev(dev_id, d_leaves.data_ptr(), n, d_root.data_ptr(), d_mut.data_ptr(),
                                           d_scr.data_ptr(), st.cuda_stream)
        assert rc == 0, L.bcp_gpu_last_error()
        st.synchronize()
        return d_root.cpu().numpy().tobytes(), int(d_mut.cpu()[0])


def test_static_scratch_merkle_matches_existing_entry_point_twice_in_a_row():
    import torch

    import bitcoincashplus_b200 as bcp

    bcp.Engine(0)
    L = bcp.lib()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(20261106)
    big = rng.integers(0, 256, size=(70001, 32), dtype=np.uint8)
    small = rng.integers(0, 256, size=(4016, 32), dtype=np.uint8)
    small[1001] = small[1000]  # a real identical pair: the mutated flag must come out 1
    for leaves in (big, small):  # two calls in a row on the shared array, second one smaller
        got = _roots(L, dev, torch, leaves, static=True)
        want = _roots(L, dev, torch, leaves, static=False)
        assert got == want
    assert _roots(L, dev, torch, small, static=True)[1] == 1
    assert _roots(L, dev, torch, big, static=True)[1] == 0
    # edge sizes through the same shared array
    for n in (0, 1, 2, 3, 257):
        leaves = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
        assert _roots(L, dev, torch, leaves, static=True) == _roots(L, dev, torch, leaves, static=False)


def test_static_scratch_merkle_from_other_lanes_and_against_the_oracle():
    """Every `device` argument of the library may carry a lane (BCP_DEVICE_LANE(ordinal, lane) = ordinal | lane << 8).
    All lanes of a GPU share the one built-in array; a call through lane 1, 2 or 3 must find it (it is looked up by
    the GPU ordinal, not by the raw id) and give the root the CPU oracle gives."""
    import torch

    import bitcoincashplus_b200 as bcp
    from oracle import pyoracle as po

    bcp.Engine(0)
    L = bcp.lib()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(7)
    for lane, n in ((1, 41651), (2, 777), (3, 1 << 16), (0, 12345)):
        bcp.Engine(0, lane=lane)
        leaves = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
        if lane == 2:
            leaves[501] = leaves[500]
        want_root, want_mut = po.oracle().merkle_root(leaves)
        got = _roots(L, dev, torch, leaves, static=True, dev_id=lane << 8)
        assert got == (want_root, int(want_mut)), "lane %d" % lane
        assert got == _roots(L, dev, torch, leaves, static=False, dev_id=0)


def test_static_scratch_merkle_rejects_oversize():
    import torch

    import bitcoincashplus_b200 as bcp

    bcp.Engine(0)
    L = bcp.lib()
    dev = torch.device("cuda", 0)
    limit = 1 << 20  # BCP_MERKLE_STATIC_MAX_LEAVES in include/bcp_gpu.h
    d_root = torch.zeros(32, dtype=torch.uint8, device=dev)
    d_mut = torch.zeros(1, dtype=torch.int32, device=dev)
    d_leaves = torch.zeros(64, dtype=torch.uint8, device=dev)  # never read: the size check comes first
    rc = L.bcp_gpu_merkle_root_dev_static(0, d_leaves.data_ptr(), limit + 1, d_root.data_ptr(), d_mut.data_ptr(), 0)
    assert rc == -4  # BCP_ERR_INVALID
    assert b"exceed" in L.bcp_gpu_last_error()
