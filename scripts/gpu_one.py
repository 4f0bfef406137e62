"""One small TF32 propagate (for ncu captures of a single kernel) -- development aid."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
from oracle import ggnn_oracle as O
b = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2
E, v, h = 46, 128, 128
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(O.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
prec = sys.argv[3] if len(sys.argv) > 3 else "tf32"
out = prop.propagate_dense(h0, adj, num_timesteps=T, precision=prec)
torch.cuda.synchronize()
print("ok", float(out.abs().max()))
