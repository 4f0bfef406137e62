"""Phase trace of gru_pair_kernel (cluster 0, first iterations) -- development aid.
Needs a library built with GGNN_NVCC_EXTRA=-DGGNN_PAIR_TRACE python ggnn_b200/build.py --force (hooks compiled out otherwise)."""
import sys, os, ctypes
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
buf = torch.zeros(2 * 12 * 32, dtype=torch.int64, device="cuda")
lib = _lib.lib()
lib.ggnn_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
lib.ggnn_debug_trace(prop._ctx, ctypes.c_void_p(buf.data_ptr()))
prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16")
torch.cuda.synchronize()
lib.ggnn_debug_trace(prop._ctx, None)
t = buf.cpu().numpy().reshape(2, 12, 32)
t0 = t[t > 0].min()
names = {0: "mma:tmem_free", 1: "mma:x_full", 2: "mma:ru+cx_issued", 3: "mma:rh_full", 4: "mma:cand_issued", 5: "mma:z_next",
         8: "epi:ru_done", 9: "epi:rh_arrived", 10: "epi:u_done", 11: "epi:cand_done", 12: "epi:tile_done",
         16: "agg:rows_requested", 17: "agg:x_free", 18: "agg:sums_written", 19: "epi:z_next_done", 20: "agg:x_full_arrived", 21: "cmp:z_ready", 22: "cmp:iter1", 23: "cmp:ld_issued", 25: "cmp:ld_done", 26: "cmp:math_done",
         27: "agg:b0_start", 28: "agg:b1_start", 29: "agg:b2_start", 30: "agg:b0_summed", 31: "agg:b1_summed", 24: "tma:h_store", 25: "tma:store_read"}
for cta in range(2):
    print(f"--- CTA {cta}")
    for it in range(int(os.environ.get('TRACE_IT0', 0)), int(os.environ.get('TRACE_IT1', 8))):
        row = t[cta, it]
        ev = sorted((int(row[k] - t0), names[k]) for k in names if row[k])
        print(f"it {it}: " + "  ".join(f"{n}={ts}" for ts, n in ev))
