"""Development aid: planning time of an edge list with one hub node (plan_edges.cu target_sort_hub_kernel)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ggnn_b200
from ggnn_b200 import synth
N, two_e, h = 200_000, 4, 32
hub = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
rng = np.random.default_rng(0)
tri = np.stack([np.arange(1, hub + 1), np.arange(1, hub + 1) % two_e, np.zeros(hub, dtype=np.int64)], 1)
rest = np.stack([rng.integers(0, N, 400_000), rng.integers(0, two_e, 400_000), rng.integers(1, N, 400_000)], 1)
edges = torch.from_numpy(np.concatenate([tri, rest]).astype(np.int32)).cuda()
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(two_e, h, 0)), "cuda:0")
h0 = torch.zeros(N, h, device="cuda")
for T in (0, 1):
    prop.propagate_edges(h0, edges, num_timesteps=T)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        prop.propagate_edges(h0, edges, num_timesteps=T)
    torch.cuda.synchronize()
    print(f"hub in-degree {hub}, {edges.shape[0]} edges, T={T}: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms per call")
