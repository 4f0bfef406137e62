"""Small C3-shaped run for ncu captures: 2 048 graphs, v=128, 2E=92, h=128, T=4, F16 mode, with the next call's
adjacency announced (ggnn_dense_prefetch) so the work-queue bitmask kernel runs too."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
b, E, v, h, T = 2048, 46, 128, 128, 4
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
adj = synth.dense_adjacency(synth.random_trees(b, v - 1, E, seed=1).cuda(), v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
for i in range(3):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
    if i < 2:
        prop.prefetch_dense(adj)
torch.cuda.synchronize()
print("ok", float(out.abs().max()))
