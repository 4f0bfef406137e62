"""Needs a library built with GGNN_NVCC_EXTRA=-DGGNN_PAIR_WIDE_TRACE python ggnn_b200/build.py --force (the hooks are compiled out otherwise).
Whole-grid timing of gru_pair_kernel (GP_WIDE): per CTA entry / set-up / weights-in / pair starts / exit -- development aid."""
import sys, os, ctypes
os.environ["GGNN_PAIR_TRACE_WIDE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth, _lib
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
E, v, h, T = 46, 128, 128, 2
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
for _ in range(2):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16")
nsm = torch.cuda.get_device_properties(0).multi_processor_count
buf = torch.zeros(2 * 12 * 32 + nsm * 48, dtype=torch.int64, device="cuda")
lib = _lib.lib()
lib.ggnn_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
lib.ggnn_debug_trace(prop._ctx, ctypes.c_void_p(buf.data_ptr()))
prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16")
torch.cuda.synchronize()
lib.ggnn_debug_trace(prop._ctx, None)
w = buf.cpu().numpy()[2 * 12 * 32:].reshape(nsm, 48)
w = w[w[:, 0] > 0]
t0 = w[:, 0].min()
print(f"CTAs traced {len(w)}; kernel span (first entry -> last exit) {(w[:, 3].max() - t0) / 1e3:.1f} us")
def stat(name, x):
    print(f"{name:34s} min {x.min() / 1e3:8.1f}  median {np.median(x) / 1e3:8.1f}  max {x.max() / 1e3:8.1f} us")
stat("entry (vs first entry)", w[:, 0] - t0)
stat("set-up done - entry", w[:, 1] - w[:, 0])
lead = w[w[:, 2] > 0]
stat("weights in - entry", lead[:, 2] - lead[:, 0])
stat("first pair starts - entry", lead[:, 8] - lead[:, 0])
n_it = (lead[:, 8:] > 0).sum(axis=1)
print("pairs per cluster: min", n_it.min(), "max", n_it.max())
last = np.array([lead[i, 8 + n_it[i] - 1] for i in range(len(lead))])
stat("exit - last pair start", lead[:, 3] - last)
stat("exit (vs first entry)", w[:, 3] - t0)
per = np.array([(lead[i, 8 + n_it[i] - 1] - lead[i, 8]) / max(1, n_it[i] - 1) for i in range(len(lead))])
stat("period per pair (first..last)", per)
for c in (0, len(lead) // 2, len(lead) - 1):
    d = np.diff(lead[c, 8:8 + n_it[c]]) / 1e3
    print(f"cluster {c}: periods us", " ".join(f"{x:.1f}" for x in d))
print("cluster: smid(rank0) smid(rank1) pairs period_us exit_us")
for i in range(0, len(w), 2):
    a = w[i]
    n = int((a[8:] > 0).sum())
    per_i = (a[8 + n - 1] - a[8 + 1]) / max(1, n - 2) / 1e3
    print(f"{i // 2:3d}: {int(a[4]):3d} {int(w[i + 1][4]):3d} {n:2d} {per_i:6.2f} {(a[3] - t0) / 1e3:7.1f}")
