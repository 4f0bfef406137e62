"""GPU: host-resident dense adjacency packed to its bitmask on the host cores (ggnn_host_pack_adjacency) and propagated
through ggnn_propagate_dense_bits -- the bitmask equals the GPU's own first pass bit for bit, and the results equal the
device-resident route bit for bit (the reference feeds the dense tensor from host memory, chem_tensorflow.py:262-268)."""
import ctypes as C

import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import Propagator, _lib, api, synth

pytestmark = pytest.mark.gpu


def _pack(adj_host: torch.Tensor, threads=0):
    l = _lib.lib()
    n = adj_host.numel()
    bits = torch.full(((n + 31) // 32,), -1, dtype=torch.int32)
    code = _lib.ADJ_F32 if adj_host.dtype == torch.float32 else _lib.ADJ_U8
    assert l.ggnn_host_pack_adjacency(C.c_void_p(adj_host.data_ptr()), code, n, C.c_void_p(bits.data_ptr()), threads) == _lib.OK
    return bits


def test_host_bitmask_equals_numpy_and_the_gpu_pass():
    g = torch.Generator().manual_seed(3)
    for shape, dt in (((7, 10, 33, 33), torch.float32), ((5, 6, 128, 128), torch.float32), ((3, 10, 20, 20), torch.uint8)):
        adj = (torch.rand(shape, generator=g) < 0.05).to(dt)
        if dt == torch.float32:
            adj.view(-1)[3] = -0.0                       # not an edge
            adj.view(-1)[5] = float("nan")               # NaN != 0: an edge, as on the GPU
            adj.view(-1)[9] = 1e-38                      # a denormal-range value is an edge
        for threads in (1, 3, 0):
            bits = _pack(adj, threads)
            flat = (adj.view(-1).numpy() != 0)
            ref = np.packbits(np.pad(flat, (0, (-flat.size) % 32)).astype(np.uint8), bitorder="little").view(np.uint32)
            assert np.array_equal(bits.numpy().view(np.uint32), ref)
    # the edge set the GPU compaction finds in the same tensor
    dev = torch.device("cuda:0")
    prop = Propagator(ggnn_b200.from_dict(synth.make_weights(10, 16, 0)), dev)
    adj = (torch.rand(4, 10, 32, 32, generator=g) < 0.1).float()
    bits = _pack(adj).numpy().view(np.uint32)
    rowptr, col, nnz = prop.compact_adjacency(adj.to(dev))
    nz = np.flatnonzero(np.unpackbits(bits.view(np.uint8), bitorder="little")[:adj.numel()])
    assert np.array_equal(nz % 32, col.cpu().numpy())
    assert int(nnz) == len(nz) == int(rowptr[-1])


@pytest.mark.parametrize("precision", ["fp32", "f16"])
def test_packed_host_route_equals_the_device_route(precision, monkeypatch):
    dev = torch.device("cuda:0")
    b, v, E, h, T = 37, 128, 5, 128, 4
    lengths = torch.tensor(([127, 127, 90, 127, 1] * 8)[:b])
    tri = synth.random_trees(b, v - 1, E, seed=5, lengths=lengths)
    adj = synth.dense_adjacency(tri, v, E)
    w = synth.make_weights(2 * E, h, 2)
    prop = Propagator(ggnn_b200.from_dict(w), dev)
    g = torch.Generator().manual_seed(1)
    h0 = (torch.rand(b, v, h, generator=g) - 0.5) * 0.2
    ref = prop.propagate_dense(h0.to(dev), adj.to(dev), num_timesteps=T, precision=precision).cpu()
    # through the bitmask entry point directly
    bits = _pack(adj).to(dev)
    out = prop.propagate_dense_bits(h0.to(dev), bits, num_timesteps=T, precision=precision).cpu()
    assert torch.equal(out, ref)
    # through the public host call: several chunks (7 graphs each), pageable NumPy and pinned torch inputs, uint8 adjacency
    monkeypatch.setattr(api, "_HOST_PACK_CHUNK_BYTES", 7 * adj[0].numel() * 4)
    for a, s in ((adj.numpy(), h0.numpy()), (adj.pin_memory(), h0.pin_memory()), (adj.to(torch.uint8).numpy(), h0.numpy())):
        got = ggnn_b200.compute_final_node_representations(s, a, weights=prop, num_timesteps=T, precision=precision)
        got = torch.from_numpy(got) if isinstance(got, np.ndarray) else got
        assert torch.equal(got, ref)
    # the hybrid route (pinned inputs: packed and dense units alternate), small units so that both kinds occur several times
    monkeypatch.setattr(api, "_HYBRID_PACK_GRAPH_BYTES", 5 * adj[0].numel() * 4)
    monkeypatch.setattr(api, "_HYBRID_DENSE_GRAPH_BYTES", 3 * adj[0].numel() * 4)
    monkeypatch.setattr(api, "_HOST_HYBRID", True)
    got = ggnn_b200.compute_final_node_representations(h0.pin_memory(), adj.pin_memory(), weights=prop, num_timesteps=T, precision=precision)
    assert torch.equal(got, ref)
    monkeypatch.setattr(api, "_HOST_HYBRID", False)
    got = ggnn_b200.compute_final_node_representations(h0.pin_memory(), adj.pin_memory(), weights=prop, num_timesteps=T, precision=precision)
    assert torch.equal(got, ref)
    # the dense-over-the-bus pipeline (GGNN_HOST_PACK=0) gives the same bits
    monkeypatch.setattr(api, "_HOST_PACK", False)
    got = ggnn_b200.compute_final_node_representations(h0.numpy(), adj.numpy(), weights=prop, num_timesteps=T, precision=precision)
    assert np.array_equal(got, ref.numpy())


def test_shapes_the_bitmask_route_does_not_serve_fall_back():
    dev = torch.device("cuda:0")
    assert not api._bits_route_ok(10, 300)
    prop = Propagator(ggnn_b200.from_dict(synth.make_weights(4, 16, 0)), dev)
    with pytest.raises(RuntimeError, match="bitmask route"):
        prop.propagate_dense_bits(torch.zeros(1, 300, 16, device=dev), torch.zeros(4 * 300 * 300 // 32 + 1, dtype=torch.int32, device=dev),
                                  num_timesteps=1)
    # v = 20: 2E*v*v is a multiple of 32 here (4*400 = 1600), the packed route serves it; odd sizes take the dense pipeline
    h0 = np.random.default_rng(0).random((3, 20, 16), dtype=np.float32)
    adj = (np.random.default_rng(1).random((3, 4, 20, 20)) < 0.2).astype(np.float32)
    a = ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=2)
    bdev = prop.propagate_dense(torch.from_numpy(h0).to(dev), torch.from_numpy(adj).to(dev), num_timesteps=2).cpu().numpy()
    assert np.array_equal(a, bdev)
    adj3 = (np.random.default_rng(2).random((3, 4, 21, 21)) < 0.2).astype(np.float32)      # 4*441 = 1764: not a multiple of 32
    h3 = np.random.default_rng(3).random((3, 21, 16), dtype=np.float32)
    a3 = ggnn_b200.compute_final_node_representations(h3, adj3, weights=prop, num_timesteps=2)
    b3 = prop.propagate_dense(torch.from_numpy(h3).to(dev), torch.from_numpy(adj3).to(dev), num_timesteps=2).cpu().numpy()
    assert np.array_equal(a3, b3)
