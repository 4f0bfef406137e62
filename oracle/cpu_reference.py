"""Timed CPU baseline: the reference's propagation step restated op-for-op in torch-CPU FP32.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/ggnn_oracle.py header).  TensorFlow cannot be
installed here or on the GPU box, so the reference's own TF graph cannot be timed; this is
"kind": "port" -- the same dense operation order as compute_timestep_fast
(/root/reference/chem_tensorflow_dense.py:391-437: matmul over ALL 2E edge types, two
transposes, batched matmul with the dense [b, v, 2E*v] adjacency) followed by TF GRUCell
(:237-241, :333), run with torch's multithreaded CPU GEMMs on all host cores -- the closest
stand-in for TF-CPU's Eigen/MKL kernels.
"""
from __future__ import annotations

import os
import time

import numpy as np
import torch


def propagate_dense_torch_cpu(h0, adj, w, T, use_bias=True):
    """h0 [b, v, h] float32, adj [b, 2E, v, v] float32 (torch CPU tensors) -> [b, v, h]."""
    b, v, H = h0.shape
    c = adj.shape[1]
    W = torch.as_tensor(w["edge_weights"])
    B = torch.as_tensor(w["edge_biases"]) if use_bias and w.get("edge_biases") is not None else None
    Wg, bg = torch.as_tensor(w["gru_gate_kernel"]), torch.as_tensor(w["gru_gate_bias"])
    Wc, bc = torch.as_tensor(w["gru_cand_kernel"]), torch.as_tensor(w["gru_cand_bias"])
    h = h0.reshape(b * v, H)
    adj_t = adj.permute(1, 0, 2, 3)                                   # :195  [2E, b, v, v]
    for _ in range(T):
        m = torch.matmul(h.unsqueeze(0), W)                           # :397  [2E, b*v, h]
        if B is not None:
            m = m + B                                                 # :407-412
        m = m.reshape(c, b, v, H).permute(1, 0, 2, 3).reshape(b, c * v, H)      # :414-419
        adj_m = adj_t.permute(1, 2, 0, 3).reshape(b, v, c * v)                  # :422-423
        acts = torch.matmul(adj_m, m).reshape(b * v, H)               # :428-431
        g = torch.sigmoid(torch.cat([acts, h], 1) @ Wg + bg)          # GRUCell
        r, u = g[:, :H], g[:, H:]
        cand = torch.tanh(torch.cat([acts, r * h], 1) @ Wc + bc)
        h = u * h + (1 - u) * cand
    return h.reshape(b, v, H)


def propagate_edges_torch_cpu(h0, edges, w, T, use_bias=True):
    """Edge-list restatement for shapes whose dense adjacency is impractical (C4):
    index_select -> matmul per channel -> index_add_."""
    N, H = h0.shape
    W = torch.as_tensor(w["edge_weights"])
    B = torch.as_tensor(w["edge_biases"]) if use_bias and w.get("edge_biases") is not None else None
    Wg, bg = torch.as_tensor(w["gru_gate_kernel"]), torch.as_tensor(w["gru_gate_bias"])
    Wc, bc = torch.as_tensor(w["gru_cand_kernel"]), torch.as_tensor(w["gru_cand_bias"])
    e = edges.long()
    order = torch.argsort(e[:, 1], stable=True)
    e = e[order]
    chans, counts = torch.unique_consecutive(e[:, 1], return_counts=True)
    h = h0
    for _ in range(T):
        acts = torch.zeros_like(h)
        pos = 0
        for ch, n in zip(chans.tolist(), counts.tolist()):
            sel = e[pos:pos + n]
            m = h.index_select(0, sel[:, 0]) @ W[ch]
            if B is not None:
                m = m + B[ch, 0]
            acts.index_add_(0, sel[:, 2], m)
            pos += n
        g = torch.sigmoid(torch.cat([acts, h], 1) @ Wg + bg)
        r, u = g[:, :H], g[:, H:]
        cand = torch.tanh(torch.cat([acts, r * h], 1) @ Wc + bc)
        h = u * h + (1 - u) * cand
    return h


def time_dense(h0, adj, w, T, chunk=16, threads=None):
    """Seconds to propagate all graphs of (h0, adj), `chunk` graphs per call (bounds memory:
    the dense form materialises [2E, chunk*v, h] messages)."""
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    b = h0.shape[0]
    t0 = time.perf_counter()
    with torch.no_grad():
        for c in range(0, b, chunk):
            propagate_dense_torch_cpu(h0[c:c + chunk], adj[c:c + chunk], w, T)
    return time.perf_counter() - t0, threads
