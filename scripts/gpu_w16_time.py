"""Forward propagation time at the WSJ hidden size (h=400, 2E=92, T=4) on tree batches: TF32 generic kernels vs the
wide FP16 kernels (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth, Propagator, weights as W
from oracle import ggnn_oracle as O

b, v, E, h, T = int(os.environ.get("B", 2048)), 64, 46, int(os.environ.get("H", 400)), 4
w = O.make_weights(2 * E, h, 0)
prop = Propagator(W.from_dict(w), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1)
edges = synth.global_edge_list(tri, v, E).cuda()
h0 = (torch.rand(b * v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
res = {}
for p in ("tf32", "f16"):
    for _ in range(3):
        prop.propagate_edges(h0, edges, num_timesteps=T, precision=p, out=out)
    torch.cuda.synchronize()
    prop.profile(True)
    prop.profile_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        prop.propagate_edges(h0, edges, num_timesteps=T, precision=p, out=out)
    e1.record()
    torch.cuda.synchronize()
    st = prop.profile_read()
    prop.profile(False)
    ms = e0.elapsed_time(e1) / 5
    res[p] = out.clone()
    print(f"{p}: {ms:.3f} ms per {b}-graph call ({b / ms * 1e3:.0f} graphs/s); stages",
          {k: round(x[0] / 5, 3) for k, x in st.items() if x[0] > 0}, flush=True)
print("f16 vs tf32 norm-wise:", float((res["f16"] - res["tf32"]).abs().max() / res["tf32"].abs().max()))
