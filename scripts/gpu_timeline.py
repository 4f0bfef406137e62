"""Development aid: kernel timeline (CUPTI through torch.profiler) of the pipelined dense call: where the time between
the per-stage kernel durations goes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import ggnn_b200
from ggnn_b200 import synth
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
E, v, h, T = 46, 128, 128, 8
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
def piped():
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
    prop.prefetch_dense(adj)
prop.prefetch_dense(adj)
for _ in range(3):
    piped()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        piped()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
print(f"{len(ev)} device activities")
prev_end = {}
for e in ev:
    st, en = e.time_range.start - t0, e.time_range.end - t0
    name = e.name[:60]
    print(f"{st:10.1f} {en - st:9.1f} us  {name}")
