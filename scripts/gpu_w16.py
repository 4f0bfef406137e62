"""GGNN_PREC_F16 at hidden sizes above 128 (wide16.cu): error vs the FP64 oracle on small cases (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth, Propagator, weights as W
from oracle import ggnn_oracle as O

def trees(b, v, E, h, T, seed=0):
    tri = synth.random_trees(b, v - 1, E, seed=seed)
    adj = synth.dense_adjacency(tri, v, E).numpy()
    w = O.make_weights(2 * E, h, seed)
    h0 = np.random.default_rng(seed + 1).uniform(-0.1, 0.1, (b, v, h)).astype(np.float32)
    ref = O.propagate_dense(h0, adj, w, T)
    for p in ("tf32", "f16"):
        out = ggnn_b200.compute_final_node_representations(h0, adj, weights=w, num_timesteps=T, precision=p)
        print(f"trees b={b} v={v} 2E={2*E} h={h} T={T} {p}: {O.normwise_err(out, ref):.3e}", flush=True)

def multigraph(n, nodes, base, ch, h, T, seed=0):
    edges = synth.random_multigraph(n, nodes, base, ch // 2, seed=seed)
    w = O.make_weights(ch, h, seed)
    h0 = np.random.default_rng(seed + 1).uniform(-0.1, 0.1, (n * nodes, h)).astype(np.float32)
    ref = O.propagate_edges(h0, edges.numpy(), w, T)
    prop = Propagator(W.from_dict(w), "cuda:0")
    for p in ("tf32", "f16"):
        out = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision=p).cpu().numpy()
        print(f"multigraph n={n} nodes={nodes} nnz/graph={2*base} ch={ch} h={h} T={T} {p}: {O.normwise_err(out, ref):.3e}", flush=True)

if __name__ == "__main__":
    trees(4, 32, 10, 256, 4)
    trees(3, 20, 6, 400, 4)
    trees(5, 16, 6, 132, 2)
    trees(2, 130, 12, 192, 3)
    trees(6, 40, 46, 320, 4)
    trees(3, 24, 5, 512, 2)
    multigraph(2, 300, 2400, 8, 256, 4)      # dense-row plan (nnz >= N*2E): rows by TMA
    multigraph(2, 300, 600, 8, 256, 4)       # gathered rows
    multigraph(1, 1000, 8000, 8, 384, 2)
