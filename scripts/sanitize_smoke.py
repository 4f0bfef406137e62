"""Small run of every round-2 kernel for compute-sanitizer (memcheck): F16 dense route with the word-order fold and an
announced next batch, the strict tensor-core mode, the TF32 mode, the edge-list route and the static-scratch compaction."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth, _lib
b, E, v, h, T = 6, 5, 128, 128, 2
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
lengths = torch.tensor([127, 127, 60, 127, 5, 127])
tri = synth.random_trees(b, v - 1, E, seed=1, lengths=lengths)
adj = synth.dense_adjacency(tri, v, E).cuda()
h0 = ((torch.rand(b, v, h) - 0.5) * 0.2).cuda()
outs = {}
for prec in ("f16", "fp32", "tf32"):
    prop.prefetch_dense(adj)
    outs[prec] = prop.propagate_dense(h0, adj, num_timesteps=T, precision=prec)
edges = synth.global_edge_list(tri, v, E).cuda()
sp = prop.propagate_edges(h0.view(-1, h), edges, num_timesteps=T, precision="f16")
rows = b * 2 * E * v
rowptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
col = torch.empty(adj.numel(), dtype=torch.int32, device="cuda")
nnz = torch.zeros(1, dtype=torch.int64, device="cuda")
rc = prop._l.ggnn_compact_adjacency_static(prop._ctx, C.c_void_p(adj.data_ptr()), _lib.ADJ_F32, b, 2 * E, v,
                                           C.c_void_p(rowptr.data_ptr()), C.c_void_p(col.data_ptr()), col.numel(),
                                           C.c_void_p(nnz.data_ptr()), C.c_void_p(0))
torch.cuda.synchronize()
assert rc == 0
err = float((outs["f16"] - outs["fp32"]).abs().max() / outs["fp32"].abs().max())
print("ok: f16 vs fp32", err, "edges vs dense", float((sp.view(b, v, h) - outs["f16"]).abs().max()), "nnz", int(nnz))
assert err < 2e-3
