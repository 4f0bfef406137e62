"""GPU: ggnn_compact_adjacency_static -- the compaction entry point that allocates nothing (its scratch is a fixed
device array inside the library; no counterpart in the reference).

Called through the C ABI and compared with ggnn_compact_adjacency: integer work, so bit-exact.  Two calls in a row
share the scratch array (the second sees the first one's leftovers).  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import Propagator, _lib, synth

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _compact(prop, fn_name, adj, dtype_code):
    """One call of `fn_name` through the C ABI; outputs are caller-allocated BEFORE the call."""
    b, c, v, _ = adj.shape
    rows = b * c * v
    cap = int(adj.numel())                                   # generous: no second sizing call needed
    rowptr = torch.full((rows + 1,), -7, dtype=torch.int32, device=adj.device)
    col = torch.full((max(cap, 1),), -7, dtype=torch.int32, device=adj.device)
    nnz = torch.full((1,), -7, dtype=torch.int64, device=adj.device)
    torch.cuda.synchronize()
    fn = getattr(prop._l, fn_name)
    rc = fn(prop._ctx, _ptr(adj), dtype_code, b, c, v, _ptr(rowptr), _ptr(col), cap, _ptr(nnz),
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, rowptr, col, nnz


def test_static_scratch_compaction_equals_the_allocating_one():
    dev = torch.device("cuda:0")
    prop = Propagator(ggnn_b200.from_dict(synth.make_weights(10, 16, 0)), dev)
    g = torch.Generator().manual_seed(5)
    cases = [
        # first a large call (close to the limit: 14.4 M elements, 112 k rows), then smaller ones that see its leftovers
        ((synth.dense_adjacency(synth.random_trees(9, 127, 49, seed=1), 128, 49)).to(dev), _lib.ADJ_F32),
        ((torch.rand(5, 10, 33, 33, generator=g) < 0.2).float().to(dev), _lib.ADJ_F32),          # unaligned rows, dense-ish
        ((torch.rand(3, 10, 20, 20, generator=g) < 0.5).to(torch.uint8).to(dev), _lib.ADJ_U8),
        (torch.zeros(2, 10, 8, 8, device=dev), _lib.ADJ_F32),                                    # no edges at all
    ]
    for adj, code in cases:
        assert adj.shape[1] in (10, 98)
        rc0, rp0, col0, n0 = _compact(prop, "ggnn_compact_adjacency", adj, code)
        assert rc0 == _lib.OK
        for k in range(2):                                   # two calls in a row on the shared scratch
            rc1, rp1, col1, n1 = _compact(prop, "ggnn_compact_adjacency_static", adj, code)
            torch.cuda.synchronize()
            assert rc1 == _lib.OK, prop._l.ggnn_last_error(prop._ctx)
            if k == 1:      # the driver's free-memory figure did not move across the call (outputs were allocated before it)
                free_before = torch.cuda.mem_get_info(dev)[0]
                b_, c_, v_, _ = adj.shape
                fn = prop._l.ggnn_compact_adjacency_static
                assert fn(prop._ctx, _ptr(adj), code, b_, c_, v_, _ptr(rp1), _ptr(col1), col1.numel(), _ptr(n1),
                          C.c_void_p(torch.cuda.current_stream().cuda_stream)) == _lib.OK
                torch.cuda.synchronize()
                assert torch.cuda.mem_get_info(dev)[0] == free_before
            assert int(n1) == int(n0)
            assert torch.equal(rp1, rp0)
            assert torch.equal(col1[:int(n0)], col0[:int(n0)])
        # and against NumPy directly (np.argwhere order)
        nz = np.argwhere(adj.cpu().numpy() != 0)
        assert int(n0) == len(nz)
        assert np.array_equal(col1[:len(nz)].cpu().numpy(), nz[:, 3].astype(np.int32))


def test_static_scratch_limits_are_enforced():
    dev = torch.device("cuda:0")
    prop = Propagator(ggnn_b200.from_dict(synth.make_weights(10, 16, 0)), dev)
    # 17 x 10 x 320 x 320 = 17.4 M elements > GGNN_STATIC_MAX_ELEMS (16 777 216)
    adj = torch.zeros(17, 10, 320, 320, dtype=torch.uint8, device=dev)
    rc, *_ = _compact(prop, "ggnn_compact_adjacency_static", adj, _lib.ADJ_U8)
    assert rc == _lib.E_UNSUPPORTED
    msg = prop._l.ggnn_last_error(prop._ctx).decode()
    assert "static scratch" in msg and "16777216" in msg
    # too many rows: 2 M rows of 2 elements
    adj = torch.zeros(2 ** 16, 16, 2, 2, dtype=torch.uint8, device=dev)
    prop2 = Propagator(ggnn_b200.from_dict(synth.make_weights(16, 16, 0)), dev)
    rc, *_ = _compact(prop2, "ggnn_compact_adjacency_static", adj, _lib.ADJ_U8)
    assert rc == _lib.E_UNSUPPORTED
    # the allocating entry point serves both
    rc, _, _, n = _compact(prop2, "ggnn_compact_adjacency", adj, _lib.ADJ_U8)
    assert rc == _lib.OK and int(n) == 0
