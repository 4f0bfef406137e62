"""Development aid: error of the reduced-precision modes over a larger sample of C3-shaped graphs."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import Propagator, synth, weights as W
from oracle import ggnn_oracle as O
b = int(sys.argv[1]) if len(sys.argv) > 1 else 128
v, E, h = 128, 46, 128
tri = synth.random_trees(b, v - 1, E, seed=77)
adj = synth.dense_adjacency(tri, v, E)
edges = synth.global_edge_list(tri, v, E)
w = O.make_weights(2 * E, h, 0)
h0 = np.random.default_rng(9).uniform(-0.1, 0.1, (b, v, h)).astype(np.float32)
prop = Propagator(W.from_dict(w), "cuda:0")
d_h0, d_adj = torch.from_numpy(h0).cuda(), adj.cuda()
for T in (4, 8, 16):
    t0 = time.time()
    ref = O.propagate_edges(h0.reshape(b * v, h), edges.numpy(), w, T).reshape(b, v, h)
    line = f"T={T} (oracle {time.time() - t0:.1f}s, max|ref| {np.abs(ref).max():.3f}):"
    for prec in ("f16", "tf32", "fp32"):
        out = prop.propagate_dense(d_h0, d_adj, num_timesteps=T, precision=prec).cpu().numpy()
        pg = np.array([O.normwise_err(out[g], ref[g]) for g in range(b)])
        line += f"  {prec}: batch {O.normwise_err(out, ref):.2e} per-graph mean {pg.mean():.2e} max {pg.max():.2e}"
    sp = prop.propagate_edges(d_h0.view(-1, h), edges.cuda(), num_timesteps=T, precision="f16").view(b, v, h).cpu().numpy()
    pg = np.array([O.normwise_err(sp[g], ref[g]) for g in range(b)])
    line += f"  f16-unfolded: batch {O.normwise_err(sp, ref):.2e} per-graph mean {pg.mean():.2e} max {pg.max():.2e}"
    print(line, flush=True)
