"""Quick GPU check of the TF32 training path against the FP64 autograd oracle (run under gpurun)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import ggnn_b200
from ggnn_b200 import Propagator, synth
from oracle import ggnn_oracle as O, grad_oracle as G

def case(b, v, E, h, T, seed=0):
    tri = synth.random_trees(b, v - 1, E, seed=seed)
    edges = synth.global_edge_list(tri, v, E)
    w = O.make_weights(2 * E, h, seed)
    rng = np.random.default_rng(seed + 1)
    h0 = rng.uniform(-0.3, 0.3, (b * v, h)).astype(np.float32)
    d = rng.standard_normal((b * v, h)).astype(np.float32)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    for prec in ("fp32", "tf32"):
        hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision=prec, save_tape=True)
        dh0, flat = prop.backward(torch.from_numpy(d).cuda())
        torch.cuda.synchronize()
        rT, rh0, rg = G.grads(h0, edges.numpy(), w, T, d)
        g = prop.split_grads(flat)
        errs = {"hT": O.normwise_err(hT.cpu().numpy(), rT), "dh0": O.normwise_err(dh0.cpu().numpy(), rh0)}
        for k, r in rg.items():
            errs[k] = O.normwise_err(g[k].cpu().numpy().reshape(r.shape), r)
        print((b, v, E, h, T), prec, {k: f"{e:.2e}" for k, e in errs.items()}, flush=True)

case(4, 16, 5, 64, 3)
case(3, 32, 10, 100, 4)
case(2, 128, 46, 128, 2)
case(6, 24, 8, 400, 2)
case(64, 40, 46, 400, 4)
