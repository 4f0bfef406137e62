"""Development aid: F16-mode error of the word-order test case, shipped GRU kernel vs the CTA-pair kernel, over seeds."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth, Propagator, weights as W
from oracle import ggnn_oracle as O
b, v, E, h, T = 12, 128, 7, 128, 8
for seed in range(4, 10):
    tri = synth.random_trees(b, v - 1, E, seed=21 + seed)
    adj = synth.dense_adjacency(tri, v, E)
    w = O.make_weights(2 * E, h, seed)
    rng = np.random.default_rng(seed)
    h0 = ((rng.random((b, v, h), dtype=np.float32) - 0.5) * 0.2).astype(np.float32)
    ref = O.propagate_dense(h0, adj.numpy(), w, T)
    prop = Propagator(W.from_dict(w), "cuda:0")
    res = []
    for mode in ("0", "1"):
        os.environ["GGNN_GRU_PAIR"] = mode
        out = prop.propagate_dense(torch.from_numpy(h0).cuda(), adj.cuda(), num_timesteps=T, precision="f16")
        o = out.cpu().numpy()
        res.append(O.normwise_err(o, ref))
        d = np.abs(o.astype(np.float64) - ref) / np.max(np.abs(ref))
        k = np.unravel_index(np.argmax(d), d.shape)
        big = np.argwhere(d > 1.6e-3)
        print(f"   mode {mode}: argmax (graph,row,col)={k} ref={ref[k]:.5f} out={o[k]:.5f}; {len(big)} elements above 1.6e-3; rows of those: {sorted(set(int(x[1]) for x in big))[:20]}")
    print(f"seed {seed}: shipped {res[0]:.3e}  pair {res[1]:.3e}")
