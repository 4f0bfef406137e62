"""Development aid: shipped F16 GRU kernel (GGNN_GRU_PAIR=0) vs the CTA-pair kernel (=1), same process, same box, interleaved."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
E, v, h, T = 46, 128, 128, 8
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(b, v - 1, E, seed=1).cuda()
adj = synth.dense_adjacency(tri, v, E)
h0 = (torch.rand(b, v, h, device="cuda") - 0.5) * 0.2
out = torch.empty_like(h0)
sync = torch.cuda.synchronize

def timed(f, n=8):
    f(); sync()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    sync()
    return (time.perf_counter() - t0) / n * 1e3

def plain(): prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
def piped():
    prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16", out=out)
    prop.prefetch_dense(adj)

for rnd in range(3):
    for mode in ("0", "1"):
        os.environ["GGNN_GRU_PAIR"] = mode
        t_plain = timed(plain)
        prop.prefetch_dense(adj)
        t_pipe = timed(piped)
        sync()
        plain(); sync()
        prop.profile(True)
        for _ in range(5):
            plain()
        st = prop.profile_read()
        prop.profile(False)
        sync()
        print(f"round {rnd} pair={mode}: plain call {t_plain:.3f} ms, pipelined call {t_pipe:.3f} ms ({b / t_pipe:.1f} k graphs/s), "
              f"gru alone {st['gru_gates'][0] / 5:.3f} ms, msg {st['msg'][0] / 5:.3f} ms")
