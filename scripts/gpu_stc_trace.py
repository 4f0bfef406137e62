"""Needs a library built with GGNN_NVCC_EXTRA=-DGGNN_STC_TRACE python ggnn_b200/build.py --force.
Phase trace of stc::split_gemm_kernel<GATES> (CTA 0) -- development aid."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth, _lib
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
E, v, h, T = 46, 128, 128, 1
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
for _ in range(2):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="fp32")
buf = torch.zeros(6 * 64, dtype=torch.int64, device="cuda")
lib = _lib.lib()
lib.ggnn_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
lib.ggnn_debug_trace(prop._ctx, ctypes.c_void_p(buf.data_ptr()))
prop.propagate_dense(h0, adj, num_timesteps=T, precision="fp32")
torch.cuda.synchronize()
lib.ggnn_debug_trace(prop._ctx, None)
t = buf.cpu().numpy().reshape(6, 64)
t0 = t[t > 0].min()
names = ["prod:chunk_top", "prod:raw_landed", "prod:stage_free", "mma:stage_full", "epi:wait_acc", "epi:acc_full"]
for k in range(6):
    print(names[k], " ".join(str(int(x - t0)) if x else "-" for x in t[k][:40]))
