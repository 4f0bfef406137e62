"""GPU parity of the backward pass (BASELINE config 5) against the torch-CPU FP64 autograd oracle,
through the C ABI (ggnn_propagate_* with GGNN_FLAG_SAVE_TAPE + ggnn_propagate_backward).

Tolerance: strict FP32 arithmetic -> norm-wise max|g - ref| / max|ref| <= 2e-5 per gradient tensor
(measured ~1e-6; float atomics make the summation order vary run to run)."""
import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import Propagator, synth
from oracle import ggnn_oracle as O
from oracle import grad_oracle as G

pytestmark = pytest.mark.gpu
TOL = 2e-5


def _case(b, v, E, h, seed, lengths=None, bias=True):
    tri = synth.random_trees(b, v - 1, E, seed=seed, lengths=lengths)
    edges = synth.global_edge_list(tri, v, E)
    w = O.make_weights(2 * E, h, seed)
    if not bias:
        w.pop("edge_biases")
    rng = np.random.default_rng(seed + 1)
    h0 = rng.uniform(-0.3, 0.3, (b * v, h)).astype(np.float32)
    d = rng.standard_normal((b * v, h)).astype(np.float32)
    return tri, edges, w, h0, d


def _check(prop, w, hT, dh0, flat, ref):
    rT, rh0, rg = ref
    assert O.normwise_err(hT.cpu().numpy(), rT) <= 1e-5
    assert O.normwise_err(dh0.cpu().numpy(), rh0) <= TOL
    g = prop.split_grads(flat)
    for k, r in rg.items():
        got = g[k].cpu().numpy().reshape(r.shape)
        if np.abs(r).max() == 0:
            assert np.abs(got).max() == 0, k
            continue
        assert O.normwise_err(got, r) <= TOL, (k, O.normwise_err(got, r))
    if "edge_biases" not in w:
        pass  # dB segment is still produced (gradient of a zero bias); nothing to compare it with


@pytest.mark.parametrize("b,v,E,h,T", [(4, 16, 5, 64, 3), (3, 32, 10, 100, 4), (2, 128, 46, 128, 2), (6, 24, 8, 400, 2)])
def test_backward_edges(b, v, E, h, T):
    tri, edges, w, h0, d = _case(b, v, E, h, seed=b + T)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check(prop, w, hT, dh0, flat, G.grads(h0, edges.numpy(), w, T, d))


def test_backward_dense_adjacency_ragged_no_bias_and_accumulation():
    b, v, E, h, T = 5, 20, 6, 96, 3
    lengths = torch.tensor([19, 3, 11, 1, 7])
    tri, edges, w, h0, d = _case(b, v, E, h, seed=9, lengths=lengths, bias=False)
    adj = synth.dense_adjacency(tri, v, E).cuda()
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    hT = prop.propagate_dense(torch.from_numpy(h0).cuda().view(b, v, h), adj, num_timesteps=T, save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda().view(b, v, h))
    ref = G.grads(h0, O.dense_to_edges(adj.cpu().numpy()), w, T, d)
    _check(prop, w, hT.view(b * v, h), dh0.view(b * v, h), flat, ref)
    # a second backward over the same tape accumulates into the same buffer: exactly doubles
    _, flat2 = prop.backward(torch.from_numpy(d).cuda().view(b, v, h), flat.clone())
    assert O.normwise_err(flat2.cpu().numpy(), 2 * flat.cpu().numpy()) <= 1e-5


def test_backward_edge_cases():
    w = O.make_weights(6, 32, 0)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    h0 = torch.rand(7, 32, device="cuda") - 0.5
    d = torch.randn(7, 32, device="cuda")
    with pytest.raises(RuntimeError):                      # no tape yet
        prop.backward(d)
    # no edges at all: GRU-only gradients
    e0 = torch.zeros(0, 3, dtype=torch.int32, device="cuda")
    hT = prop.propagate_edges(h0, e0, num_timesteps=2, save_tape=True)
    dh0, flat = prop.backward(d)
    ref = G.grads(h0.cpu().numpy(), np.zeros((0, 3), np.int64), w, 2, d.cpu().numpy())
    _check(prop, w, hT, dh0, flat, ref)
    # T = 0: identity
    prop.propagate_edges(h0, e0, num_timesteps=0, save_tape=True)
    dh0, flat = prop.backward(d)
    assert torch.equal(dh0, d) and float(flat.abs().max()) == 0.0
    # the tape is dropped by a non-tape call; the FP16 mode cannot tape
    prop.propagate_edges(h0, e0, num_timesteps=1)
    with pytest.raises(RuntimeError):
        prop.backward(d)
    with pytest.raises(RuntimeError):
        prop.propagate_edges(h0, e0, num_timesteps=1, precision="f16", save_tape=True)


def test_train_step_helper_single_process():
    from ggnn_b200 import dist as gdist
    tri, edges, w, h0, d = _case(3, 12, 4, 64, seed=2)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    dt = torch.from_numpy(d).cuda()
    hT, dh0, flat = gdist.train_step(prop, torch.from_numpy(h0).cuda(), edges.cuda(), lambda hT: dt, 3)
    _check(prop, w, hT, dh0, flat, G.grads(h0, edges.numpy(), w, 3, d))


@pytest.mark.parametrize("use_w,use_s", [(True, True), (True, False), (False, True)])
def test_training_dropout_with_supplied_masks(use_w, use_s):
    """A8: tf.nn.dropout on the edge weights (fresh mask per unrolled step) and DropoutWrapper on the new
    state, keep = 0.9 as the reference feeds in training (chem_tensorflow_dense.py:800, 860-861), with
    the masks supplied by the caller; forward and gradients against the autograd oracle."""
    b, v, E, h, T, keep = 4, 16, 5, 64, 3, 0.9
    tri, edges, w, h0, d = _case(b, v, E, h, seed=21)
    rng = np.random.default_rng(3)
    wm = (rng.random((T, 2 * E, h, h)) < keep).astype(np.float32) if use_w else None
    sm = (rng.random((T, b * v, h)) < keep).astype(np.float32) if use_s else None
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    dev = lambda a: None if a is None else torch.from_numpy(a).cuda()
    prop.set_dropout(dev(wm), dev(sm), keep, T)
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    ref = G.grads(h0, edges.numpy(), w, T, d, w_mask=wm, state_mask=sm, keep=keep)
    _check(prop, w, hT, dh0, flat, ref)
    nodrop = G.grads(h0, edges.numpy(), w, T, d)
    assert O.normwise_err(hT.cpu().numpy(), nodrop[0]) > 1e-3         # the masks really did something
    # the masks apply to ONE forward only: the next tape forward is dropout-free again
    hT2 = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, save_tape=True)
    assert O.normwise_err(hT2.cpu().numpy(), nodrop[0]) <= 1e-5
    with pytest.raises(ValueError):
        prop.set_dropout(dev(wm), dev(sm), keep, T + 1)


def test_optimizer_step_clip_and_adam_against_oracle():
    """N4: three steps of per-variable clip_by_norm(1.0) + Adam(0.003) on the ctx-owned weights against the
    NumPy FP64 restatement of make_train_step (chem_tensorflow.py:483-506); gradients of very different
    norms so some variables are clipped and some are not.  Tolerance: FP32 update arithmetic, 1e-6 absolute
    on weights of magnitude ~0.1 (lr_t * m / (sqrt(v) + eps) is at most ~lr per step)."""
    E2, h = 6, 64
    w = O.make_weights(E2, h, 3)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    keys = G.KEYS
    P = {k: np.asarray(w[k], dtype=np.float64) for k in keys}
    M = {k: np.zeros_like(P[k]) for k in keys}
    V = {k: np.zeros_like(P[k]) for k in keys}
    rng = np.random.default_rng(0)
    scales = {"edge_weights": 0.05, "edge_biases": 1e-3, "gru_gate_kernel": 0.5, "gru_gate_bias": 3.0,
              "gru_cand_kernel": 1e-4, "gru_cand_bias": 0.0}
    for t in range(1, 4):
        g = {k: (rng.standard_normal(P[k].shape) * scales[k]).astype(np.float32) for k in keys}
        flat = torch.from_numpy(np.concatenate([g[k].ravel() for k in keys])).cuda()
        assert flat.numel() == prop.grad_floats
        prop.optimizer_step(flat, lr=0.003, clip_norm=1.0)
        G.adam_step(P, g, M, V, t, lr=0.003, clip=1.0)
        got = prop.get_weights()
        for k in keys:
            assert np.abs(got[k].cpu().numpy().reshape(P[k].shape) - P[k]).max() <= 1e-6, (t, k)
    # the packed tensor-core copies follow the master weights: all three modes agree with the oracle
    # evaluated on the UPDATED weights
    w2 = {k: P[k].astype(np.float32) for k in keys}
    tri = synth.random_trees(3, 15, 3, seed=5)
    edges = synth.global_edge_list(tri, 16, 3)
    h0 = rng.uniform(-0.3, 0.3, (48, h)).astype(np.float32)
    ref = O.propagate_edges(h0, edges.numpy(), w2, 2)
    for prec, tol in (("fp32", 1e-5), ("tf32", 2e-3)):
        out = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=2, precision=prec)
        assert O.normwise_err(out.cpu().numpy(), ref) <= tol, prec


def test_training_loop_reduces_loss():
    """forward(tape) -> backward -> optimizer_step, 30 steps on one fixed batch with L = 0.5 |h_T - y|^2:
    the loss must fall (end-to-end sign/consistency check of the N4 row with the backward of config 5)."""
    b, v, E, h, T = 8, 16, 4, 64, 3
    tri, edges, w, h0, _ = _case(b, v, E, h, seed=21)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    y = torch.from_numpy(np.random.default_rng(2).uniform(-0.5, 0.5, (b * v, h)).astype(np.float32)).cuda()
    h0d, ed = torch.from_numpy(h0).cuda(), edges.cuda()
    losses = []
    for _ in range(30):
        hT = prop.propagate_edges(h0d, ed, num_timesteps=T, save_tape=True)
        diff = hT - y
        losses.append(float(0.5 * (diff * diff).sum()))
        _, flat = prop.backward(diff.contiguous())
        prop.optimizer_step(flat)
    assert losses[-1] < 0.8 * losses[0], losses[::5]
    assert all(np.isfinite(losses))


# ---------------------------------------------------------------------------------------------------
# tensor-core training path: forward with the tape in GGNN_PREC_TF32, backward on tcgen05 kind::tf32
# (NT contractions on the forward kernel, TN contractions with MN-major operands).  Tolerance: the
# north_star bound for TF32-input / FP32-accumulate arithmetic, 2e-3 norm-wise per tensor against the
# FP64 autograd oracle (measured 3e-4 .. 9e-4).
TOL_TF32 = 2e-3


def _check_tf32(prop, hT, dh0, flat, ref):
    rT, rh0, rg = ref
    assert O.normwise_err(hT.cpu().numpy(), rT) <= TOL_TF32
    assert O.normwise_err(dh0.cpu().numpy(), rh0) <= TOL_TF32
    g = prop.split_grads(flat)
    for k, r in rg.items():
        got = g[k].cpu().numpy().reshape(r.shape)
        if np.abs(r).max() == 0:
            assert np.abs(got).max() == 0, k
            continue
        e = O.normwise_err(got, r)
        assert e <= TOL_TF32, (k, e)
        assert e > 1e-6, (k, "suspiciously exact: did the FP32 path run?")


@pytest.mark.parametrize("b,v,E,h,T", [(4, 16, 5, 64, 3), (3, 32, 10, 100, 4), (2, 128, 46, 128, 2),
                                       (6, 24, 8, 400, 2), (40, 40, 46, 400, 4), (5, 20, 3, 36, 1)])
def test_backward_tf32_edges(b, v, E, h, T):
    tri, edges, w, h0, d = _case(b, v, E, h, seed=b + T)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="tf32",
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check_tf32(prop, hT, dh0, flat, G.grads(h0, edges.numpy(), w, T, d))


def test_backward_tf32_dense_ragged_no_bias_and_no_edges():
    b, v, E, h, T = 5, 20, 6, 96, 3
    lengths = torch.tensor([19, 3, 11, 1, 7])
    tri, edges, w, h0, d = _case(b, v, E, h, seed=9, lengths=lengths, bias=False)
    adj = synth.dense_adjacency(tri, v, E).cuda()
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    hT = prop.propagate_dense(torch.from_numpy(h0).cuda().view(b, v, h), adj, num_timesteps=T, precision="tf32",
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda().view(b, v, h))
    ref = G.grads(h0, O.dense_to_edges(adj.cpu().numpy()), w, T, d)
    _check_tf32(prop, hT.view(b * v, h), dh0.view(b * v, h), flat, ref)
    # no edges at all: GRU-only gradients, edge gradients exactly zero
    e0 = torch.zeros(0, 3, dtype=torch.int32, device="cuda")
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), e0, num_timesteps=2, precision="tf32", save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check_tf32(prop, hT, dh0, flat, G.grads(h0, np.zeros((0, 3), np.int64), w, 2, d))


def test_training_dropout_tf32():
    b, v, E, h, T, keep = 4, 16, 5, 64, 3, 0.9
    tri, edges, w, h0, d = _case(b, v, E, h, seed=21)
    rng = np.random.default_rng(3)
    wm = (rng.random((T, 2 * E, h, h)) < keep).astype(np.float32)
    sm = (rng.random((T, b * v, h)) < keep).astype(np.float32)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    prop.set_dropout(torch.from_numpy(wm).cuda(), torch.from_numpy(sm).cuda(), keep, T)
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="tf32",
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check_tf32(prop, hT, dh0, flat, G.grads(h0, edges.numpy(), w, T, d, w_mask=wm, state_mask=sm, keep=keep))
    # the packed weights are restored after the dropped forward: a plain TF32 call matches the oracle again
    out = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="tf32")
    assert O.normwise_err(out.cpu().numpy(), O.propagate_edges(h0, edges.numpy(), w, T)) <= TOL_TF32


@pytest.mark.parametrize("b,v,E,h,T", [(6, 24, 8, 400, 2), (40, 40, 46, 400, 4), (4, 16, 5, 256, 3), (3, 20, 4, 192, 2)])
def test_backward_f16_forward_edges(b, v, E, h, T):
    """Training forward on the FP16 kernels of wide16.cu (h > 128) with the FP32 tape, backward on the TF32 tensor-core
    path: same tolerance as the TF32 mode against the FP64 autograd oracle."""
    tri, edges, w, h0, d = _case(b, v, E, h, seed=b + T)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="f16",
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check_tf32(prop, hT, dh0, flat, G.grads(h0, edges.numpy(), w, T, d))


def test_training_dropout_f16_forward_and_unsupported_sizes():
    b, v, E, h, T, keep = 4, 16, 5, 256, 3, 0.9
    tri, edges, w, h0, d = _case(b, v, E, h, seed=21)
    rng = np.random.default_rng(3)
    wm = (rng.random((T, 2 * E, h, h)) < keep).astype(np.float32)
    sm = (rng.random((T, b * v, h)) < keep).astype(np.float32)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    prop.set_dropout(torch.from_numpy(wm).cuda(), torch.from_numpy(sm).cuda(), keep, T)
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="f16",
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    _check_tf32(prop, hT, dh0, flat, G.grads(h0, edges.numpy(), w, T, d, w_mask=wm, state_mask=sm, keep=keep))
    # the packed FP16 weights are restored after the dropped forward
    out = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision="f16")
    assert O.normwise_err(out.cpu().numpy(), O.propagate_edges(h0, edges.numpy(), w, T)) <= TOL_TF32
    # h <= 128 has no FP16 training forward
    tri, edges, w, h0, d = _case(2, 8, 3, 128, seed=1)
    p128 = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    with pytest.raises(RuntimeError, match="SAVE_TAPE"):
        p128.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=2, precision="f16", save_tape=True)


def test_training_loop_tf32_tracks_fp32():
    """20 optimiser steps on one batch in both modes from the same start: the TF32 loss curve follows the
    FP32 one (relative difference of the final loss below 2 %), and both fall."""
    b, v, E, h, T = 8, 16, 4, 64, 3
    tri, edges, w, h0, _ = _case(b, v, E, h, seed=21)
    y = torch.from_numpy(np.random.default_rng(2).uniform(-0.5, 0.5, (b * v, h)).astype(np.float32)).cuda()
    h0d, ed = torch.from_numpy(h0).cuda(), edges.cuda()
    final = {}
    for prec in ("fp32", "tf32"):
        prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
        for _ in range(20):
            hT = prop.propagate_edges(h0d, ed, num_timesteps=T, precision=prec, save_tape=True)
            diff = hT - y
            loss = float(0.5 * (diff * diff).sum())
            _, flat = prop.backward(diff.contiguous())
            prop.optimizer_step(flat)
        final[prec] = loss
    assert abs(final["tf32"] - final["fp32"]) <= 0.02 * final["fp32"], final


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_device_generated_dropout_matches_oracle_with_same_masks(precision):
    """A8 with the masks drawn on the device (Philox4x32-10): the device masks equal the NumPy Philox restatement
    bit for bit (itself pinned by Random123's known answers), and forward + gradients equal the autograd oracle
    fed with exactly those masks.  A second call without re-arming is dropout-free."""
    b, v, E, h, T, keep, seed = 4, 16, 5, 64, 3, 0.9, 12345678901234567
    tri, edges, w, h0, d = _case(b, v, E, h, seed=21)
    prop = Propagator(ggnn_b200.from_dict(w), "cuda:0")
    wn, nh = 2 * E * h * h, b * v * h
    wm = np.stack([G.philox_mask(wn, seed, 2 * t, keep) for t in range(T)]).reshape(T, 2 * E, h, h)
    sm = np.stack([G.philox_mask(nh, seed, 2 * t + 1, keep) for t in range(T)]).reshape(T, b * v, h)
    for t in range(T):
        assert np.array_equal(prop.dropout_mask(wn, seed, 2 * t, keep).cpu().numpy(), wm[t].ravel())
        assert np.array_equal(prop.dropout_mask(nh, seed, 2 * t + 1, keep).cpu().numpy(), sm[t].ravel())
    assert abs(float(wm.mean()) - keep) < 5e-3 and abs(float(sm.mean()) - keep) < 2e-2
    prop.set_dropout_seed(keep, seed)
    hT = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision=precision,
                              save_tape=True)
    dh0, flat = prop.backward(torch.from_numpy(d).cuda())
    ref = G.grads(h0, edges.numpy(), w, T, d, w_mask=wm, state_mask=sm, keep=keep)
    if precision == "fp32":
        _check(prop, w, hT, dh0, flat, ref)
    else:
        _check_tf32(prop, hT, dh0, flat, ref)
    hT2 = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision=precision,
                               save_tape=True)
    assert O.normwise_err(hT2.cpu().numpy(), G.grads(h0, edges.numpy(), w, T, d)[0]) <= 2e-3
    # only one of the two dropouts
    prop.set_dropout_seed(keep, seed, drop_weights=False)
    hT3 = prop.propagate_edges(torch.from_numpy(h0).cuda(), edges.cuda(), num_timesteps=T, precision=precision,
                               save_tape=True)
    ref3 = G.grads(h0, edges.numpy(), w, T, d, state_mask=sm, keep=keep)
    assert O.normwise_err(hT3.cpu().numpy(), ref3[0]) <= (1e-5 if precision == "fp32" else 2e-3)
