"""e2e sweep of the host pipeline chunk size (run under gpurun)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import ggnn_b200
from ggnn_b200 import api, synth
from oracle import ggnn_oracle as O

V, E, H, T, ne = 128, 46, 128, 8, 256
w = O.make_weights(2 * E, H, 0)
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(w), "cuda:0")
tri = synth.random_trees(ne, V - 1, E, seed=1)
adj = synth.dense_adjacency(tri, V, E).pin_memory()
h0 = ((torch.rand(ne, V, H) - 0.5) * 0.2).pin_memory()
for mb in (16, 24, 32, 48, 64, 96, 128):
    api._HOST_CHUNK_BYTES = mb << 20
    for _ in range(2):
        ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 4
    print(mb, "MB chunks:", round(ne / dt), "graphs/s", round(dt * 1e3, 2), "ms", flush=True)
