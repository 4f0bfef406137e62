"""Norm-wise error of each arithmetic mode vs the FP64 oracle on C3-shaped graphs (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth
from oracle import ggnn_oracle as O
b, v, E, h, T = 6, 128, 46, 128, 8
for seed in (0, 1):
    tri = synth.random_trees(b, v - 1, E, seed=seed)
    adj = synth.dense_adjacency(tri, v, E).numpy()
    w = O.make_weights(2 * E, h, seed)
    h0 = np.random.default_rng(seed + 1).uniform(-0.1, 0.1, (b, v, h)).astype(np.float32)
    ref = O.propagate_dense(h0, adj, w, T)
    for p in ("fp32", "tf32", "f16"):
        out = ggnn_b200.compute_final_node_representations(h0, adj, weights=w, num_timesteps=T, precision=p)
        print(f"seed {seed} {p}: {O.normwise_err(out, ref):.3e}")
