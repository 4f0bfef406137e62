"""Development aid: GRU kernel time with parts switched off (GGNN_GRU_DBG)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
b, E, v, h, T = 4096, 46, 128, 128, 8
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
adj = synth.dense_adjacency(synth.random_trees(b, v - 1, E, seed=1).cuda(), v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
for _ in range(2):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
prop.profile(True)
for _ in range(5):
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
st = prop.profile_read()
print("GGNN_GRU_DBG=%s" % os.environ.get("GGNN_GRU_DBG", "0"), {k: round(v[0] / 5, 3) for k, v in st.items() if v[0] > 0})
