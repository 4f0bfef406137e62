"""GPU: ggnn_dense_prefetch -- the adjacency bitmask pass of the NEXT batch as a background work queue.

A prefetched call must return exactly what the plain call returns (the bitmask is integer work; the plan built
from it is the same), whatever happens between the announcement and the call."""
import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import Propagator, synth
from oracle import ggnn_oracle as O

pytestmark = pytest.mark.gpu


def _case(b, v, E, h, seed, dtype=torch.float32):
    tri = synth.random_trees(b, v - 1, E, seed=seed)
    adj = synth.dense_adjacency(tri, v, E).cuda()
    if dtype == torch.uint8:
        adj = (adj != 0).to(torch.uint8)
    h0 = (torch.rand(b, v, h, generator=torch.Generator().manual_seed(seed)) - 0.5).mul(0.2).cuda()
    return adj, h0


@pytest.mark.parametrize("precision", ["fp32", "tf32", "f16"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.uint8])
def test_prefetched_call_equals_plain_call(precision, dtype):
    b, v, E, h, T = 96, 64, 10, 128, 3
    w = synth.make_weights(2 * E, h, 0)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    adj, h0 = _case(b, v, E, h, 3, dtype)
    plain = prop.propagate_dense(h0, adj, num_timesteps=T, precision=precision)
    for _ in range(3):                                   # both prefetch buffers get used
        prop.prefetch_dense(adj)
        out = prop.propagate_dense(h0, adj, num_timesteps=T, precision=precision)
        assert torch.equal(out, plain)
    ref = O.propagate_dense(h0.cpu().numpy(), adj.float().cpu().numpy(), w, T)
    assert O.normwise_err(plain.cpu().numpy(), ref) <= (1e-5 if precision == "fp32" else 2e-3)


def test_prefetch_pipeline_over_batches():
    """The bench's pattern: while batch k propagates, batch k+1's adjacency is announced."""
    v, E, h, T = 128, 6, 128, 2
    w = synth.make_weights(2 * E, h, 1)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    batches = [_case(256, v, E, h, 10 + i) for i in range(4)]
    plain = [prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16") for adj, h0 in batches]
    prop.prefetch_dense(batches[0][0])
    for k, (adj, h0) in enumerate(batches):
        out = prop.propagate_dense(h0, adj, num_timesteps=T, precision="f16")
        if k + 1 < len(batches):
            prop.prefetch_dense(batches[k + 1][0])
        assert torch.equal(out, plain[k]), k


def test_background_counts_follow_the_last_calls_setting():
    """The side stream also runs the counting half of the plan, assuming the next call folds the word-order channels
    like the last one did (api.cu): v = 128 F16 calls fold them, strict calls do not.  Alternating precisions makes the
    consuming call reuse the counts (same setting) or re-count from the bitmask (different setting); either way the
    result is the plain call's."""
    v, E, h, T = 128, 6, 128, 2
    prop = Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 4)), "cuda:0")
    adj, h0 = _case(192, v, E, h, 30)
    adj2, h2 = _case(192, v, E, h, 31)
    adj2[5, E - 1, 7, 9] = 1.0                            # one graph whose word-order channel is not regular
    want = {(p, i): prop.propagate_dense(hh, aa, num_timesteps=T, precision=p)
            for p in ("f16", "fp32") for i, (aa, hh) in enumerate(((adj, h0), (adj2, h2)))}
    order = ["f16", "f16", "fp32", "fp32", "f16", "fp32", "f16", "f16"]
    for k, p in enumerate(order):
        aa, hh = ((adj, h0), (adj2, h2))[k & 1]
        prop.prefetch_dense(aa)
        assert torch.equal(prop.propagate_dense(hh, aa, num_timesteps=T, precision=p), want[(p, k & 1)]), (k, p)


def test_prefetch_dropped_or_replaced_is_harmless():
    b, v, E, h, T = 64, 32, 5, 128, 2
    w = synth.make_weights(2 * E, h, 2)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    adj_a, h_a = _case(b, v, E, h, 20)
    adj_b, h_b = _case(b, v, E, h, 21)
    want_a = prop.propagate_dense(h_a, adj_a, num_timesteps=T)
    want_b = prop.propagate_dense(h_b, adj_b, num_timesteps=T)
    prop.prefetch_dense(adj_a)                          # announced, then a different batch arrives
    assert torch.equal(prop.propagate_dense(h_b, adj_b, num_timesteps=T), want_b)
    prop.prefetch_dense(adj_a)                          # replaced before use
    prop.prefetch_dense(adj_b)
    assert torch.equal(prop.propagate_dense(h_b, adj_b, num_timesteps=T), want_b)
    prop.prefetch_dense(adj_a)                          # other calls in between do not disturb it
    e = synth.global_edge_list(synth.random_trees(4, v - 1, E, seed=5), v, E).cuda()
    prop.propagate_edges(h_a[:4].reshape(4 * v, h), e, num_timesteps=1)
    assert torch.equal(prop.propagate_dense(h_a, adj_a, num_timesteps=T), want_a)
    # shapes the fast route does not serve are accepted and ignored: element-aligned view, v > 255
    buf = torch.zeros(adj_a.numel() + 1, device="cuda")
    buf[1:] = adj_a.reshape(-1)
    odd = buf[1:].view_as(adj_a)
    prop.prefetch_dense(odd)
    assert torch.equal(prop.propagate_dense(h_a, odd, num_timesteps=T), want_a)
    ev = torch.cuda.Event()
    ev.record()
    prop.prefetch_dense(adj_b, ready_event=ev)
    assert torch.equal(prop.propagate_dense(h_b, adj_b, num_timesteps=T), want_b)
    prop.prefetch_dense(adj_a)                          # left unconsumed: destroy must cope
    prop.close()
