"""Development aid: per-graph error of the F16 mode with and without the word-order fold."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import Propagator, synth, weights as W
from oracle import ggnn_oracle as O
b, v, E, h, T = 12, 128, 7, 128, 8
lengths = torch.tensor([127, 127, 60, 127, 1, 127, 126, 127, 127, 100, 127, 127])
tri = synth.random_trees(b, v - 1, E, seed=21, lengths=lengths)
adj = synth.dense_adjacency(tri, v, E)
w = O.make_weights(2 * E, h, 4)
h0 = np.random.default_rng(9).uniform(-0.1, 0.1, (b, v, h)).astype(np.float32)
ref = O.propagate_dense(h0, adj.numpy(), w, T)
prop = Propagator(W.from_dict(w), "cuda:0")
d_h0, d_adj = torch.from_numpy(h0).cuda(), adj.cuda()
for prec in ("f16", "tf32"):
    out = prop.propagate_dense(d_h0, d_adj, num_timesteps=T, precision=prec).cpu().numpy()
    print(prec, "dense:", O.normwise_err(out, ref), [round(O.normwise_err(out[g], ref[g]) * 1e3, 3) for g in range(b)])
edges = torch.from_numpy(O.dense_to_edges(adj.numpy())).cuda()
sp = prop.propagate_edges(d_h0.view(-1, h), edges, num_timesteps=T, precision="f16").view(b, v, h).cpu().numpy()
print("f16 edges:", O.normwise_err(sp, ref), [round(O.normwise_err(sp[g], ref[g]) * 1e3, 3) for g in range(b)])
print("max|ref| per graph", [round(float(np.abs(ref[g]).max()), 3) for g in range(b)])
out = prop.propagate_dense(d_h0, d_adj, num_timesteps=T, precision="f16")
for g in (0, 10, 11):
    one = prop.propagate_dense(d_h0[g:g + 1].contiguous(), d_adj[g:g + 1].contiguous(), num_timesteps=T, precision="f16")
    print("graph", g, "alone == in batch:", bool(torch.equal(one[0], out[g])), "err alone", O.normwise_err(one[0].cpu().numpy(), ref[g]))
rev = torch.arange(b - 1, -1, -1).cuda()
o2 = prop.propagate_dense(d_h0[rev].contiguous(), d_adj[rev].contiguous(), num_timesteps=T, precision="f16")
print("reversed batch equal:", bool(torch.equal(o2, out[rev])))
d = (out[11].cpu().numpy() - ref[11])
i = np.unravel_index(np.abs(d).argmax(), d.shape)
print("worst element of graph 11:", i, d[i], ref[11][i], "row errors:", [round(float(np.abs(d[r]).max()) * 1e3, 2) for r in (0, 1, 2, 63, 125, 126, 127)])
for T2 in (1, 2, 4):
    r2 = O.propagate_dense(h0, adj.numpy(), w, T2)
    o = prop.propagate_dense(d_h0, d_adj, num_timesteps=T2, precision="f16").cpu().numpy()
    print("T", T2, [round(O.normwise_err(o[g], r2[g]) * 1e3, 3) for g in range(b)])
