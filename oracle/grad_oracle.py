"""Gradient oracle for the training configuration (BASELINE config 5).

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE (same rule as ggnn_oracle.py).

The reference obtains gradients from TensorFlow's autodiff over the unrolled path
(``make_train_step``, chem_tensorflow.py:483-506: ``tf.gradients`` via
``AdamOptimizer.compute_gradients``); there is no hand-written backward to restate.  The
oracle therefore is torch-CPU FP64 autograd over the same restatement of the forward that
ggnn_oracle.py holds (edge-list form, chem_tensorflow_dense.py:391-437 + TF GRUCell), with both
dropouts at keep-probability 1.0 (or with externally supplied edge-weight masks).

PINNING: the FP64 forward here must equal ggnn_oracle.propagate_edges to 1e-12, which in turn is held to
the outputs of the reference's own propagation functions (tests/golden/ref_propagation.npz, see
ggnn_oracle.py); the gradients are autograd of that forward, cross-checked by central finite differences
(tests/test_oracle.py).  The dropout semantics at keep < 1 (tf.nn.dropout, DropoutWrapper) are restated from
TensorFlow's documentation and are not pinnable without TensorFlow.
"""
from __future__ import annotations

import numpy as np
import torch

KEYS = ("edge_weights", "edge_biases", "gru_gate_kernel", "gru_gate_bias", "gru_cand_kernel", "gru_cand_bias")


def forward_torch(h0, edges, wt, T, use_bias=True, w_mask=None, state_mask=None, keep=1.0):
    """h0 [N,H] tensor, edges int64 tensor [nnz,3] (src, ch, dst), wt dict of tensors.
    Training-time dropout with supplied masks: tf.nn.dropout(W, rate=1-keep) draws a fresh mask per
    unrolled step -> W * w_mask[t] / keep (chem_tensorflow_dense.py:397-403); DropoutWrapper(
    state_keep_prob=keep) drops the NEW state -> h' * state_mask[t] / keep (:239-240, :333 takes [1])."""
    H = h0.shape[1]
    src, ch, dst = edges[:, 0], edges[:, 1], edges[:, 2]
    h = h0
    Wg, bg, Wc, bc = wt["gru_gate_kernel"], wt["gru_gate_bias"], wt["gru_cand_kernel"], wt["gru_cand_bias"]
    for t in range(T):
        x = torch.zeros_like(h)
        W = wt["edge_weights"] if w_mask is None else wt["edge_weights"] * w_mask[t] / keep
        if len(edges):
            # m_j = h[src_j] @ W[ch_j] + b[ch_j]   (bias per edge)
            m = torch.bmm(h[src].unsqueeze(1), W[ch]).squeeze(1)
            if use_bias:
                m = m + wt["edge_biases"].reshape(-1, H)[ch]
            x = x.index_add(0, dst, m)
        g = torch.sigmoid(torch.cat([x, h], 1) @ Wg + bg)
        r, u = g[:, :H], g[:, H:]
        c = torch.tanh(torch.cat([x, r * h], 1) @ Wc + bc)
        h = u * h + (1 - u) * c
        if state_mask is not None:
            h = h * state_mask[t] / keep
    return h


def grads(h0, edges, w, T, d_hT, use_bias=True, dtype=torch.float64, w_mask=None, state_mask=None, keep=1.0):
    """Returns (h_T, d_h0, {key: gradient}) as numpy, for L = sum(h_T * d_hT)."""
    wt = {k: torch.tensor(np.asarray(w[k]), dtype=dtype, requires_grad=True) for k in KEYS if k in w}
    if "edge_biases" not in wt:
        use_bias = False
    h0t = torch.tensor(np.asarray(h0), dtype=dtype, requires_grad=True)
    e = torch.tensor(np.asarray(edges), dtype=torch.int64).reshape(-1, 3)
    tm = lambda a: None if a is None else torch.tensor(np.asarray(a), dtype=dtype)
    hT = forward_torch(h0t, e, wt, T, use_bias, tm(w_mask), tm(state_mask), keep)
    loss = (hT * torch.tensor(np.asarray(d_hT), dtype=dtype)).sum()
    loss.backward()
    out = {}
    for k, t in wt.items():
        g = t.grad if t.grad is not None else torch.zeros_like(t)
        out[k] = g.detach().numpy()
    return hT.detach().numpy(), h0t.grad.detach().numpy(), out


def adam_step(params, grads, m, v, t, lr=0.003, beta1=0.9, beta2=0.999, eps=1e-8, clip=1.0):
    """make_train_step (chem_tensorflow.py:483-506) in NumPy FP64: per-variable tf.clip_by_norm
    (g * clip / max(||g||_2, clip)) then TF1 AdamOptimizer's update (epsilon outside the square root,
    bias correction folded into the step size).  Dicts keyed by variable; updates in place; t is 1-based."""
    lr_t = lr * np.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    for k in params:
        g = np.asarray(grads[k], dtype=np.float64)
        if clip and clip > 0:
            g = g * clip / max(np.sqrt((g * g).sum()), clip)
        m[k] = beta1 * m[k] + (1 - beta1) * g
        v[k] = beta2 * v[k] + (1 - beta2) * g * g
        params[k] = params[k] - lr_t * m[k] / (np.sqrt(v[k]) + eps)
    return params


def philox4x32_10(counter, key):
    """Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11).
    counter uint32 [n, 4], key uint32 [2] -> uint32 [n, 4].  Restated from the paper / Random123's
    published round function; pinned by Random123's known-answer vectors in tests/test_oracle.py."""
    c = np.asarray(counter, dtype=np.uint64).copy()
    k0, k1 = int(key[0]), int(key[1])
    M0, M1, MASK = 0xD2511F53, 0xCD9E8D57, 0xFFFFFFFF
    for _ in range(10):
        p0 = c[:, 0] * np.uint64(M0)
        p1 = c[:, 2] * np.uint64(M1)
        hi0, lo0 = p0 >> np.uint64(32), p0 & np.uint64(MASK)
        hi1, lo1 = p1 >> np.uint64(32), p1 & np.uint64(MASK)
        n0 = hi1 ^ c[:, 1] ^ np.uint64(k0)
        n2 = hi0 ^ c[:, 3] ^ np.uint64(k1)
        c = np.stack([n0, lo1, n2, lo0], axis=1)
        k0 = (k0 + 0x9E3779B9) & MASK
        k1 = (k1 + 0xBB67AE85) & MASK
    return c.astype(np.uint32)


def philox_mask(n, seed, stream, keep):
    """0/1 float32 mask of ggnn_dropout_mask: element i = word i % 4 of the block with counter
    (i // 4 as 64 bit, stream as 64 bit) and key = seed as 64 bit; 1 where float32(u32) * 2^-32 < keep."""
    nb = (n + 3) // 4
    b = np.arange(nb, dtype=np.uint64)
    ctr = np.stack([b & np.uint64(0xFFFFFFFF), b >> np.uint64(32),
                    np.full(nb, stream & 0xFFFFFFFF, np.uint64), np.full(nb, (stream >> 32) & 0xFFFFFFFF, np.uint64)], axis=1)
    r = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).reshape(-1)[:n]
    u = r.astype(np.float32) * np.float32(2.3283064365386963e-10)
    return (u < np.float32(keep)).astype(np.float32)
