"""Quick timing of the stages on one GPU (development aid, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth
from oracle import ggnn_oracle as O

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
b = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
E, v, h, T = 46, 128, 128, 8
w = ggnn_b200.from_dict(O.make_weights(2 * E, h, 0))
prop = ggnn_b200.Propagator(w, "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
for it in range(2):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision=prec, out=out)
torch.cuda.synchronize()
prop.profile(True)
t0 = time.time()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
iters = 3
for it in range(iters):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision=prec, out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
st = prop.profile_read()
print(f"{prec} b={b}: {ms:.3f} ms/call -> {b / ms * 1e3:.0f} graphs/s")
for k, (m, l) in st.items():
    print(f"  {k:14s} {m / iters:9.3f} ms  {l // iters:5d} launches")
adj_bytes = adj.numel() * 4
print(f"  compact_mask: {adj_bytes / (st['compact_mask'][0] / iters * 1e-3) / 1e9:.0f} GB/s on the dense read")
