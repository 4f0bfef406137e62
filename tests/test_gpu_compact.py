"""GPU parity: adjacency compaction (A1) is integer-exact against the reference's edge set
(golden vectors from graph_to_adj_mat_bd) and against np.argwhere (the oracle)."""
import os

import numpy as np
import pytest
import torch

from ggnn_b200 import Propagator, synth, weights as W
from oracle import ggnn_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prop():
    return Propagator(W.from_dict(O.make_weights(4, 8, 0)), "cuda:0")


def _check(prop, adj_np):
    b, c, v, _ = adj_np.shape
    adj = torch.from_numpy(adj_np).cuda()
    rowptr, col, nnz = prop.compact_adjacency(adj)
    rp, cl = O.compact_adjacency(adj_np)
    assert nnz == len(cl)
    assert np.array_equal(rowptr.cpu().numpy(), rp)
    assert np.array_equal(col.cpu().numpy(), cl)
    e4 = O.csr_to_edges(rowptr.cpu().numpy(), col.cpu().numpy(), b, c, v)
    assert np.array_equal(e4, np.argwhere(adj_np != 0).astype(np.int32))
    return e4


def test_c1_golden(prop, golden_dir):
    d = np.load(os.path.join(golden_dir, "c1_small_dev.npz"))
    e4 = _check(prop, d["adj"])
    assert np.array_equal(e4, d["edges_bets"])
    assert [int((e4[:, 0] == g).sum()) for g in range(3)] == [16, 20, 20]


def test_reference_built_random_trees(prop, golden_dir):
    d = np.load(os.path.join(golden_dir, "random_trees_ref_adj.npz"))
    ci = 0
    while f"case{ci}_shape" in d:
        b, c, v = d[f"case{ci}_shape"].tolist()
        adj = np.zeros((b, c, v, v), np.float32)
        e = d[f"case{ci}_edges_bets"]
        adj[e[:, 0], e[:, 1], e[:, 2], e[:, 3]] = 1
        assert np.array_equal(_check(prop, adj), e)
        ci += 1


@pytest.mark.parametrize("dtype", [np.float32, np.uint8])
@pytest.mark.parametrize("shape", [(1, 2, 4), (3, 10, 6), (2, 4, 10), (5, 6, 33), (2, 8, 128),
                                   (7, 92, 128), (1, 2, 198), (3, 3, 5)])
def test_random_dense_patterns(prop, dtype, shape):
    """Arbitrary 0/1 (and non-0/1) patterns, all bucket alignments: v*rows not a multiple of 32,
    rows straddling bitmask words, v = 128 fast path."""
    b, c, v = shape
    rng = np.random.default_rng(hash(shape) % 1000)
    adj = (rng.random((b, c, v, v)) < 0.03).astype(dtype)
    if dtype == np.float32:
        adj[rng.random(adj.shape) < 0.002] = -0.0          # negative zero is still "no edge"
        adj.flat[:: 97] *= 2.5                             # any non-zero value is an edge
    _check(prop, adj)


def test_all_zero_all_one_and_empty(prop):
    _check(prop, np.zeros((2, 4, 6, 6), np.float32))
    _check(prop, np.ones((2, 4, 6, 6), np.float32))
    _check(prop, np.ones((1, 2, 128, 128), np.uint8))
    rowptr, col, nnz = prop.compact_adjacency(torch.zeros((0, 4, 6, 6), device="cuda"))
    assert nnz == 0 and rowptr.cpu().tolist() == [0]


def test_unaligned_base_pointer(prop):
    rng = np.random.default_rng(3)
    raw = torch.from_numpy((rng.random(2 * 6 * 10 * 10 + 1) < 0.1).astype(np.float32)).cuda()
    adj = raw[1:].view(2, 6, 10, 10)                      # 4-byte aligned only
    assert adj.data_ptr() % 16 != 0
    rowptr, col, nnz = prop.compact_adjacency(adj)
    rp, cl = O.compact_adjacency(adj.cpu().numpy())
    assert np.array_equal(rowptr.cpu().numpy(), rp) and np.array_equal(col.cpu().numpy(), cl)


def test_capacity_is_reported_not_overrun(prop):
    adj = torch.ones((1, 2, 8, 8), device="cuda")
    with pytest.raises(RuntimeError, match="CAPACITY"):
        prop.compact_adjacency(adj, col_capacity=10)


def test_full_size_c3_graphs_properties(prop):
    """BASELINE config 3 shape (v=128, 2E=92) at a size the box handles in seconds: nnz = 4n per
    graph, sortedness, and agreement with the edge list built straight from the triples."""
    E, v, b = 46, 128, 64
    tri = synth.random_trees(b, v - 1, E, seed=11).cuda()
    adj = synth.dense_adjacency(tri, v, E)
    rowptr, col, nnz = prop.compact_adjacency(adj)
    assert nnz == 4 * (v - 1) * b
    rp = rowptr.cpu().numpy().astype(np.int64)
    assert np.all(np.diff(rp) >= 0) and rp[-1] == nnz
    e4 = O.csr_to_edges(rp, col.cpu().numpy(), b, 2 * E, v)
    want = synth.edges_from_triples(tri.cpu(), v, E).numpy()
    want = np.unique(want, axis=0)                         # lexicographic (g, ch, t, s)
    assert np.array_equal(e4, want)
    # uint8 twin gives the same CSR
    rowptr8, col8, nnz8 = prop.compact_adjacency(adj.to(torch.uint8))
    assert nnz8 == nnz and torch.equal(rowptr8, rowptr) and torch.equal(col8, col)
