"""Per-tile event timeline of CTA 0 of the fused F16 GRU kernel (development aid; run under gpurun).
Needs the instrumented kernel: GGNN_NVCC_EXTRA=-DGGNN_GRU_TRACE python ggnn_b200/build.py --force; rebuild without it afterwards."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
b, T = 4096, 2
E, v, h = 46, 128, 128
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
torch.cuda.synchronize()
tr = torch.zeros(64 * 16, dtype=torch.int64, device="cuda")
os.environ["GGNN_GRU_TRACE_PTR"] = str(tr.data_ptr())
prop.propagate_dense(h0, adj, num_timesteps=1, precision="f16", out=out)
torch.cuda.synchronize()
t = tr.cpu().view(64, 16)
names = ["agg_start", "agg_batch1", "agg_xempty", "agg_done", "mma_tmem_free", "mma_x_full", "mma_h_full", "mma_rh_issued",
         "epi_r_done", "epi_1a_done", "epi_uc_done", "epi_1b_done", "epi_cand_done", "epi_2_done", "store_issue", "tma_h_issue"]
t0 = int(t[0][t[0] > 0].min())
for it in range(2, 12):
    row = t[it]
    print(it, " ".join(f"{n}={(int(row[i]) - t0) / 1000:.2f}" for i, n in enumerate(names) if int(row[i]) > 0))
