"""Needs a library built with GGNN_NVCC_EXTRA=-DGGNN_GRU_FUSED_TRACE python ggnn_b200/build.py --force.
Phase trace of the fused GRU kernel (CTA 0, first iterations) -- development aid."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth, _lib
from oracle import ggnn_oracle as O
b = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
E, v, h, T = 46, 128, 128, 1
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(O.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
for _ in range(2):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="tf32")
buf = torch.zeros(16 * 64, dtype=torch.int64, device="cuda")
lib = _lib.lib()
lib.ggnn_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
lib.ggnn_debug_trace(prop._ctx, ctypes.c_void_p(buf.data_ptr()))
prop.propagate_dense(h0, adj, num_timesteps=T, precision="tf32")
torch.cuda.synchronize()
t = buf.cpu().numpy().reshape(16, 64)
t0 = t[t > 0].min()
names = {16: "mma:tmem_free", 17: "mma:x_full", 18: "mma:x_issued", 19: "mma:h_full", 20: "mma:gates_issued",
         21: "mma:rh_full", 22: "mma:cand_issued", 32: "epi:gates_done", 33: "epi:rh_written",
         34: "epi:cand_done", 35: "epi:tile_done"}
for it in range(6):
    row = t[it]
    agg_s = [row[2 * w] for w in range(8) if row[2 * w]]
    agg_e = [row[2 * w + 1] for w in range(8) if row[2 * w + 1]]
    if not agg_s:
        break
    print(f"iter {it}: agg start {min(agg_s) - t0:7d}..{max(agg_s) - t0:7d}  done {min(agg_e) - t0:7d}..{max(agg_e) - t0:7d} ns")
    print("   w0: start=%d done=%d " % (row[0]-t0, row[1]-t0) + " ".join(f"e{k}={row[k]-t0}" for k in range(40,48) if row[k]))
    print("   " + "  ".join(f"{names[k]}={row[k] - t0}" for k in sorted(names) if row[k]))
