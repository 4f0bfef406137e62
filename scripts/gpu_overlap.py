"""Development aid: how much of the adjacency bitmask pass hides behind the propagation kernels (prefetch)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
PREC = sys.argv[2] if len(sys.argv) > 2 else "f16"
E, v, h, T = 46, 128, 128, 8
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
gb = adj.numel() * 4 / 1e9
sync = torch.cuda.synchronize

def timed(f, n=5):
    f(); sync()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    sync()
    return (time.perf_counter() - t0) / n * 1e3

def plain(): prop.propagate_dense(h0, adj, num_timesteps=T, precision=PREC, out=out)
def bg_only(): prop.prefetch_dense(adj); sync()
def piped():
    prop.propagate_dense(h0, adj, num_timesteps=T, precision=PREC, out=out)
    prop.prefetch_dense(adj)
t_plain = timed(plain)
t_bg = timed(bg_only)
prop.prefetch_dense(adj)
t_pipe = timed(piped)
sync()
print(f"b={b} adjacency {gb:.1f} GB; plain call {t_plain:.3f} ms; background pass alone {t_bg:.3f} ms = {gb / t_bg:.2f} TB/s; "
      f"pipelined call {t_pipe:.3f} ms -> {b / t_pipe:.1f} k graphs/s (plain {b / t_plain:.1f} k)")
for name, f in (("plain", plain), ("piped", piped)):
    if name == "piped":
        prop.prefetch_dense(adj)
    f(); sync()
    prop.profile(True)
    for _ in range(5):
        f()
    st = prop.profile_read()
    prop.profile(False)
    sync()
    print(name, {k: round(v[0] / 5, 3) for k, v in st.items() if v[0] > 0})
