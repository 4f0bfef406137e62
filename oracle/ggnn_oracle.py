"""CPU oracle for the GGNN propagation path of crismolav/ggnn.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  ``ggnn_b200`` never does: the product path is
CUDA only and fails loudly when the extension is missing.

What it restates (all citations relative to /root/reference):

* edge set           -- ``graph_to_adj_mat_bd``         chem_tensorflow_dense.py:65-83
* driver             -- ``compute_final_node_representations``        :312-340
* fast (default) form-- ``compute_timestep_fast``                     :391-437
* loop form (--old)  -- ``compute_timestep_normal``                   :350-389
* parameters         -- ``prepare_specific_graph_model``       :202-212, 237-241
* GRU                -- ``tf.compat.v1.nn.rnn_cell.GRUCell.call`` (TensorFlow
  ``rnn_cell_impl.py``; third-party, NOT under /root/reference; the reference
  pins ``tensorflow==1.3.0`` in requirements.txt:2 but needs a TF 2.x API).
  Published algorithm restated here:
      g = [x, h] @ Wg + bg ; r, u = split(sigmoid(g), 2)
      c = tanh([x, r*h] @ Wc + bc) ; h' = u*h + (1-u)*c
* edge-list ("sparse") form -- the reference has none
  (chem_tensorflow.py:522-523 raises); defined as the same sum over the
  non-zeros of the adjacency.

PINNING STATUS
  integer part (edge set, compaction order): PINNED against golden vectors made
  by the reference's own ``graph_to_adj_mat_bd`` / ``process_sentence`` run in
  the authoring container (scripts/make_golden.py -> tests/golden/*.npz) and
  against the conventions fixed by tests_chem.py:29-48.
  floating-point part (message transform, aggregation, GRU): PINNED AT THE
  REFERENCE'S OWN CALL SITES -- scripts/make_golden_fp.py executes the reference's
  compute_final_node_representations / compute_timestep_fast / _normal
  (chem_tensorflow_dense.py:312-437) over NumPy-backed stand-ins for the TF
  primitives they call (GRUCell restated from TF's published GRUCell.call) and
  commits tests/golden/ref_propagation.npz; tests/test_oracle.py holds every form
  below to it at 1e-13 in FP64.  Not pinnable without TensorFlow itself: its
  kernels' summation order (inside the norm-wise tolerance) and GRUCell's source.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# A1: the edge set
# --------------------------------------------------------------------------

def graph_to_adj_mat_bd(graph, max_n_vertices, num_edge_types, dtype=np.float64):
    """Dense [2E, v, v] adjacency, convention A[e, target, source].

    Follows chem_tensorflow_dense.py:65-83: for a triple (src, e, dest), e 1-based:
    head->child on channel e-1, child->head on e-1+E, previous word -> word on E-1,
    word -> previous word on 2E-1.  Assignment (not accumulation): values are 0/1.
    """
    E = int(num_edge_types)
    amat = np.zeros((2 * E, max_n_vertices, max_n_vertices), dtype=dtype)
    for (src, e, dest) in graph:
        amat[e - 1, dest, src] = 1
        amat[e - 1 + E, src, dest] = 1
        amat[E - 1, dest, dest - 1] = 1
        amat[2 * E - 1, dest - 1, dest] = 1
    return amat


def edges_from_triples(graph, num_edge_types):
    """The same edge set as ``graph_to_adj_mat_bd`` but as a sorted, de-duplicated
    int32 [nnz, 3] array of (channel, target, source) rows, lexicographic order --
    i.e. exactly ``np.argwhere(A != 0)`` of the dense matrix, without building it."""
    E = int(num_edge_types)
    s = set()
    for (src, e, dest) in graph:
        s.add((e - 1, dest, src))
        s.add((e - 1 + E, src, dest))
        s.add((E - 1, dest, dest - 1))
        s.add((2 * E - 1, dest - 1, dest))
    out = np.array(sorted(s), dtype=np.int32).reshape(-1, 3)
    return out


def compact_adjacency(adj):
    """Dense [b, 2E, v, v] -> canonical CSR.

    rows are (graph, channel, target) in row-major order, columns are sources:
    ``rowptr`` int32 [b*2E*v + 1], ``col`` int32 [nnz].  Identical to
    ``np.argwhere(adj != 0)`` read row-major."""
    adj = np.asarray(adj)
    b, c, v, v2 = adj.shape
    assert v == v2
    nz = adj != 0
    counts = nz.reshape(b * c * v, v).sum(axis=1)
    rowptr = np.zeros(b * c * v + 1, dtype=np.int64)
    np.cumsum(counts, out=rowptr[1:])
    col = np.nonzero(nz.reshape(b * c * v, v))[1].astype(np.int32)
    return rowptr.astype(np.int32), col


def csr_to_edges(rowptr, col, b, two_e, v):
    """Canonical CSR -> int32 [nnz, 4] rows (graph, channel, target, source)."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    counts = np.diff(rowptr)
    rows = np.repeat(np.arange(b * two_e * v, dtype=np.int64), counts)
    g = rows // (two_e * v)
    e = (rows // v) % two_e
    t = rows % v
    return np.stack([g, e, t, np.asarray(col, dtype=np.int64)], axis=1).astype(np.int32)


# --------------------------------------------------------------------------
# parameters
# --------------------------------------------------------------------------

def glorot_init(rng, shape):
    """utils.py:11-14 (np.random.uniform there; a Generator here so seeds are local)."""
    r = np.sqrt(6.0 / (shape[-2] + shape[-1]))
    return rng.uniform(low=-r, high=r, size=shape).astype(np.float32)


def make_weights(two_e, h, seed=0, bias_std=0.05):
    """Synthetic parameters in the reference's shapes (chem_tensorflow_dense.py:202-212,
    237-241; TF GRUCell init: gate bias 1.0, candidate bias 0).  Edge biases are
    N(0, bias_std) instead of the reference's zeros so the per-edge bias rule (F9) is
    exercised."""
    rng = np.random.default_rng(seed)
    return {
        "edge_weights": glorot_init(rng, (two_e, h, h)),
        "edge_biases": (rng.standard_normal((two_e, 1, h)) * bias_std).astype(np.float32),
        "gru_gate_kernel": glorot_init(rng, (2 * h, 2 * h)),
        "gru_gate_bias": np.ones((2 * h,), np.float32),
        "gru_cand_kernel": glorot_init(rng, (2 * h, h)),
        "gru_cand_bias": np.zeros((h,), np.float32),
    }


# --------------------------------------------------------------------------
# A6: GRU (TensorFlow GRUCell semantics)
# --------------------------------------------------------------------------

def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def gru_cell(x, h, w, dtype):
    """One TF-GRUCell step.  x = acts [n, H], h = state [n, H].
    Gate kernel rows 0:H multiply x, rows H:2H multiply h; columns 0:H are the
    reset gate r, H:2H the update gate u (tf.split(value, 2, axis=1))."""
    H = h.shape[1]
    Wg = w["gru_gate_kernel"].astype(dtype)
    bg = w["gru_gate_bias"].astype(dtype)
    Wc = w["gru_cand_kernel"].astype(dtype)
    bc = w["gru_cand_bias"].astype(dtype)
    g = np.concatenate([x, h], axis=1) @ Wg + bg
    g = _sigmoid(g)
    r, u = g[:, :H], g[:, H:]
    c = np.tanh(np.concatenate([x, r * h], axis=1) @ Wc + bc)
    return (u * h + (1.0 - u) * c).astype(dtype)


# --------------------------------------------------------------------------
# A4 / A5 / A9: one timestep of message passing, three formulations
# --------------------------------------------------------------------------

def timestep_fast(h, adj, w, use_bias, dtype):
    """compute_timestep_fast, chem_tensorflow_dense.py:391-437.
    h [b*v, H]; adj [b, 2E, v, v] (A[b, e, target, source])."""
    b, c, v, _ = adj.shape
    H = h.shape[1]
    W = w["edge_weights"].astype(dtype)
    m = np.matmul(h[None, :, :], W)                      # [2E, b*v, H]   :397-399
    if use_bias:
        m = m + w["edge_biases"].astype(dtype)           # [2E, 1, H]     :407-412
    m = m.reshape(c, b, v, H).transpose(1, 0, 2, 3).reshape(b, c * v, H)      # :414-419
    adj_m = adj.astype(dtype).transpose(0, 2, 1, 3).reshape(b, v, c * v)     # :422-423
    acts = np.matmul(adj_m, m)                           # [b, v, H]      :428
    return acts.reshape(b * v, H)


def timestep_loop(h, adj, w, use_bias, dtype):
    """compute_timestep_normal, chem_tensorflow_dense.py:350-389 (the --old form)."""
    b, c, v, _ = adj.shape
    H = h.shape[1]
    acts = None
    for e in range(c):
        m = h @ w["edge_weights"][e].astype(dtype)
        m = m.reshape(b, v, H)
        if use_bias:
            m = m + w["edge_biases"][e].astype(dtype)
        a = np.matmul(adj[:, e].astype(dtype), m)
        acts = a if acts is None else acts + a
    return acts.reshape(b * v, H)


def timestep_edges(h, edges, w, use_bias, dtype):
    """Edge-list form (A9).  h [N, H]; edges int [nnz, 3] rows (src, channel, dst),
    channel 0-based in [0, 2E).  acts[dst] = sum over edges (h[src] @ W_ch + b_ch)."""
    N, H = h.shape
    acts = np.zeros((N, H), dtype=dtype)
    if len(edges) == 0:
        return acts
    edges = np.asarray(edges)
    W = w["edge_weights"].astype(dtype)
    B = w["edge_biases"].astype(dtype) if use_bias else None
    for ch in np.unique(edges[:, 1]):
        sel = edges[edges[:, 1] == ch]
        m = h[sel[:, 0]] @ W[ch]
        if use_bias:
            m = m + B[ch, 0]
        np.add.at(acts, sel[:, 2], m)
    return acts


# --------------------------------------------------------------------------
# A2: the driver
# --------------------------------------------------------------------------

def _select_weights(w, fixed_ts):
    """fixed_ts selects the *_fixed twins (chem_tensorflow_dense.py:400-403, 411-412)."""
    if fixed_ts is None:
        return w
    w2 = dict(w)
    w2["edge_weights"] = w["edge_weights_fixed"]
    if "edge_biases_fixed" in w:
        w2["edge_biases"] = w["edge_biases_fixed"]
    return w2


def propagate_dense(h0, adj, w, num_timesteps, fixed_ts=None, use_bias=True,
                    dtype=np.float64, form="fast"):
    """compute_final_node_representations, chem_tensorflow_dense.py:312-340, at
    evaluation (both dropouts are identities, :800-801).  h0 [b, v, H] -> [b, v, H].
    Padded rows are updated like any other."""
    b, v, H = h0.shape
    w = _select_weights(w, fixed_ts)
    T = num_timesteps if fixed_ts is None else fixed_ts
    step = timestep_fast if form == "fast" else timestep_loop
    h = h0.reshape(b * v, H).astype(dtype)
    adj = np.asarray(adj)
    for _ in range(T):
        acts = step(h, adj, w, use_bias, dtype)
        h = gru_cell(acts, h, w, dtype)
    return h.reshape(b, v, H)


def propagate_edges(h0, edges, w, num_timesteps, fixed_ts=None, use_bias=True,
                    dtype=np.float64):
    """Edge-list twin of ``propagate_dense``: h0 [N, H], edges [nnz, 3] (src, ch, dst)."""
    w = _select_weights(w, fixed_ts)
    T = num_timesteps if fixed_ts is None else fixed_ts
    h = h0.astype(dtype)
    for _ in range(T):
        acts = timestep_edges(h, edges, w, use_bias, dtype)
        h = gru_cell(acts, h, w, dtype)
    return h


def dense_to_edges(adj):
    """[b, 2E, v, v] dense -> int32 [nnz, 3] (src, ch, dst) with global node ids
    g*v + node, in canonical (g, ch, dst, src) order."""
    adj = np.asarray(adj)
    b, c, v, _ = adj.shape
    idx = np.argwhere(adj != 0)
    g, ch, t, s = idx[:, 0], idx[:, 1], idx[:, 2], idx[:, 3]
    return np.stack([g * v + s, ch, g * v + t], axis=1).astype(np.int32)


def normwise_err(out, ref):
    """max|out - ref| / max|ref| -- the parity norm."""
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(np.asarray(out, dtype=np.float64) - ref)) / np.max(np.abs(ref)))


# --------------------------------------------------------------------------
# reduced-precision emulation (used to set / justify the TF32 tolerance)
# --------------------------------------------------------------------------

def round_mantissa(x, bits=10):
    """Round FP32 values to `bits` explicit mantissa bits, nearest-even (TF32 / FP16: 10,
    BF16: 7)."""
    x = np.asarray(x, np.float32).copy()
    u = x.view(np.uint32)
    drop = 23 - bits
    u += (np.uint32(1) << np.uint32(drop - 1)) - np.uint32(1) + ((u >> np.uint32(drop)) & np.uint32(1))
    u &= ~np.uint32((1 << drop) - 1)
    return u.view(np.float32)


def propagate_edges_emulated(h0, edges, w, num_timesteps, bits=10, use_bias=True):
    """Edge-list propagation with every matmul operand rounded to `bits` mantissa bits and
    FP32 accumulation -- a CPU model of a tensor-core TF32 implementation.  Shows what error
    TF32 rounding alone produces on a given problem, independent of any GPU kernel."""
    q = lambda a: round_mantissa(a, bits)
    W = q(w["edge_weights"])
    Wg, Wc = q(w["gru_gate_kernel"]), q(w["gru_cand_kernel"])
    h = np.asarray(h0, np.float32).copy()
    H = h.shape[1]
    edges = np.asarray(edges)
    for _ in range(num_timesteps):
        hq = q(h)
        acts = np.zeros_like(h)
        for ch in np.unique(edges[:, 1]):
            sel = edges[edges[:, 1] == ch]
            m = hq[sel[:, 0]] @ W[ch]
            if use_bias:
                m = m + w["edge_biases"][ch, 0]
            np.add.at(acts, sel[:, 2], m.astype(np.float32))
        xq = q(acts)
        g = _sigmoid(np.concatenate([xq, hq], 1) @ Wg + w["gru_gate_bias"])
        r, u = g[:, :H], g[:, H:]
        c = np.tanh(np.concatenate([xq, q(r * h)], 1) @ Wc + w["gru_cand_bias"])
        h = (u * h + (1 - u) * c).astype(np.float32)
    return h
