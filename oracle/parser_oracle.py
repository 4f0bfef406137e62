"""CPU oracle for the rows either side of the propagation path -- the
front end, the output heads, the masks, the decode and the LAS/UAS scorer of crismolav/ggnn.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE (same rule as ggnn_oracle.py).

Restated from (citations relative to /root/reference):
* get_initial_node_representation     chem_tensorflow_dense.py:264-310  (evaluation: dropouts identity)
* gated_regression, btb branch        chem_tensorflow_dense.py:439-505 ; utils.MLP utils.py:40-84
* get_mask                            chem_tensorflow_dense.py:633-661
* target_to_adj_mat                   chem_tensorflow_dense.py:92-104
* get_target_values_(edges_)formatted chem_tensorflow_dense.py:884-909
* get_results_reshaped (btb)          chem_tensorflow_dense.py:1054-1080
* adj_mat_to_target                   chem_tensorflow_dense.py:106-131
* merge_head_and_edge_graph           chem_tensorflow_dense.py:1279-1280
* humanize_batch_results_btb          chem_tensorflow_dense.py:1160-1215
* get_las_uas                         chem_tensorflow_dense.py:1304-1319

PINNING: the NumPy part (masks, target formatting, decode, scorer) is PINNED against
tests/golden/parser_decode.npz, produced by scripts/make_golden_parser.py running the reference's
own functions.  The TF part (embedding lookup, concat / pad, the one-layer MLPs, softmax) is PINNED at the
reference's own call sites: scripts/make_golden_heads.py executes get_initial_node_representation and
gated_regression (with the reference's utils.MLP) over NumPy-backed TF primitives and commits
tests/golden/ref_heads.npz; tests/test_oracle.py holds initial_node_representation() and heads() to it.
"""
from __future__ import annotations

import numpy as np


# ---- N1 ------------------------------------------------------------------------------------
def initial_node_representation(word_inputs, loc_tab, pos_tab, word_tab, h, dtype=np.float64):
    """word_inputs int [b, v, >=4] columns (loc, pos, word, head_loc, ...) -> [b, v, h]."""
    wi = np.asarray(word_inputs).astype(np.int64)
    parts = [loc_tab[wi[..., 0]], pos_tab[wi[..., 1]], word_tab[wi[..., 2]], loc_tab[wi[..., 3]]]
    x = np.concatenate(parts, axis=-1).astype(dtype)
    pad = h - x.shape[-1]
    if pad < 0:
        raise ValueError("embeddings wider than hidden size (tf.pad with a negative amount fails)")
    return np.pad(x, [(0, 0), (0, 0), (0, pad)])


# ---- N2: heads -----------------------------------------------------------------------------
def _softmax(x):
    e = np.exp(x - x.max(-1, keepdims=True))
    return e / e.sum(-1, keepdims=True)


def heads(h_T, h_0, W_head, b_head, W_lab, b_lab, dtype=np.float64):
    """[b, v, h] x2 -> (softmax over the O head positions [b, v*O], softmax over labels [b, v*E])."""
    b, v, _ = h_T.shape
    gi = np.concatenate([h_T, h_0], axis=-1).astype(dtype).reshape(b * v, -1)
    ph = _softmax(gi @ W_head.astype(dtype) + b_head.astype(dtype)).reshape(b, -1)
    pe = _softmax(gi @ W_lab.astype(dtype) + b_lab.astype(dtype)).reshape(b, -1)
    return ph, pe


# ---- masks / targets -----------------------------------------------------------------------
def get_mask(n_active_nodes, v, e_o, o, is_edge=False):
    mask = np.ones((e_o, n_active_nodes))
    if v - n_active_nodes > 0:
        mask = np.concatenate([mask, np.zeros((e_o, v - n_active_nodes))], axis=-1)
    if is_edge:
        return np.reshape(np.transpose(mask), [-1])
    fm = np.tile(np.reshape(mask, [-1, 1]), [1, n_active_nodes])
    fm = np.concatenate([fm, np.zeros((fm.shape[0], o - n_active_nodes))], axis=-1)
    fm = np.transpose(np.reshape(fm, [e_o, v, o]), [1, 0, 2])
    return np.reshape(np.mean(fm, axis=1), [-1])


def target_to_adj_mat(target, v, e_o, output_size):
    amat = np.zeros((e_o, v, output_size))
    for i, (src, e) in enumerate(target):
        amat[e - 1, i + 1, src] = 1
    return amat


def target_values(labels, o):
    lab = np.array(labels)
    b, e, v, _ = lab.shape
    nl = np.transpose(lab, [0, 2, 1, 3])
    nl = np.pad(nl, [[0, 0], [0, 0], [0, 0], [0, o - v]])
    return np.reshape(np.sum(nl, axis=2), [b, v * o])


def target_values_edges(labels):
    lab = np.array(labels)
    b, e, v, _ = lab.shape
    return np.reshape(np.sum(np.transpose(lab, [0, 2, 1, 3]), axis=3), [b, v * e])


# ---- decode / score ------------------------------------------------------------------------
def adj_mat_to_target(adj_mat, is_probability=False):
    num_e, num_v, num_o = adj_mat.shape
    graph = []
    for node in range(1, num_v):
        a = adj_mat[:, node, :]
        if np.amax(a) == 0:
            continue
        e_, src_ = np.where(a == np.amax(a)) if is_probability else np.where(a == 1)
        if len(e_) == 0:
            continue
        graph.append([int(src_[0]), int(e_[0]) + 1])
    return graph


def get_las_uas(target_graph, result_graph, is_edge=False):
    las = uas = 0
    k = 1 if is_edge else 0
    for i, r in enumerate(result_graph):
        t = target_graph[i]
        if t == r:
            las += 1
            uas += 1
        elif t[k] == r[k]:
            uas += 1
    return las / len(result_graph), uas / len(result_graph)


def score_batch(tv, ph, mask, tve, pe, mask_e, v, e_o, o):
    """humanize_batch_results_btb: returns (per-sentence [las, uas, uas_e] array, result graphs)."""
    res = np.reshape(ph * mask, [-1, 1, v, o])
    tgt = np.reshape(tv, [-1, 1, v, o])
    res_e = np.transpose(np.reshape(pe * mask_e, [-1, v, e_o, 1]), [0, 2, 1, 3])
    tgt_e = np.transpose(np.reshape(tve, [-1, v, e_o, 1]), [0, 2, 1, 3])
    scores, graphs = [], []
    for i in range(res.shape[0]):
        tg_h, tg_e = adj_mat_to_target(tgt[i]), adj_mat_to_target(tgt_e[i])
        tg = [[x[0], y[1]] for x, y in zip(tg_h, tg_e)]
        rg_h = adj_mat_to_target(res[i], is_probability=True)
        rg_e = adj_mat_to_target(res_e[i], is_probability=True)
        rg = [[x[0], y[1]] for x, y in zip(rg_h, rg_e)]
        las, uas = get_las_uas(tg, rg)
        _, uas_e = get_las_uas(tg_e, rg_e, is_edge=True)
        scores.append([las, uas, uas_e])
        graphs.append(rg)
    return np.array(scores), graphs
