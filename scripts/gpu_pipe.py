"""Experiment: two ctxs on two streams, alternating chunks (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
from oracle import ggnn_oracle as O
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nchunks = int(sys.argv[2]) if len(sys.argv) > 2 else 8
prio = len(sys.argv) > 3
E, v, h, T = 46, 128, 128, 8
w = ggnn_b200.from_dict(O.make_weights(2 * E, h, 0))
props = [ggnn_b200.Propagator(w, "cuda:0") for _ in range(2)]
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
outs = [torch.empty_like(h0) for _ in range(2)]
streams = [torch.cuda.Stream(priority=(-1 if (prio and i == 0) else 0)) for i in range(2)]

def run(two):
    for c in range(nchunks):
        k = c & 1 if two else 0
        with torch.cuda.stream(streams[k]):
            props[k].propagate_dense(h0, adj, num_timesteps=T, precision="tf32", out=outs[k])
for two in (False, True, False, True):
    run(two); torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(two); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"two_streams={two}: {dt / nchunks * 1e3:.3f} ms per {b}-graph chunk -> {b * nchunks / dt:.0f} graphs/s")
print("same result:", torch.equal(outs[0], outs[1]))
