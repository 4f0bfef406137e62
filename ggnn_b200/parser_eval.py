"""Evaluation pipeline around the propagation path (BASELINE config 2):
embedding front end -> GGNN propagation -> output heads -> decode -> LAS/UAS, mirroring the
evaluation branch of the reference's ``run_epoch`` (chem_tensorflow.py:528-667) for ``--pr btb``.

Host side (this file): bucketing and batch assembly as ``process_raw_graphs`` /
``make_minibatch_iterator`` do it (chem_tensorflow_dense.py:519-582, 792-875), and the per-sentence
scorer (``get_las_uas`` :1304-1319).  Device side (libggnn_b200.so): ``ggnn_embed_nodes``,
``ggnn_propagate_*``, ``ggnn_heads_decode``.  No TensorFlow, no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import re
from dataclasses import dataclass
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .engine import Propagator, _ptr, _stream_ptr
from .weights import GGNNWeights

OUTPUT_SIZE = 150            # head positions, params['output_size'] (chem_tensorflow.py:81)
BUCKET_SIZES = np.arange(4, 200, 2)   # chem_tensorflow_dense.py:584-585

# TF variable names of the tensors either side of the path
HEAD_TF_NAMES = {
    "loc_embeddings": r"graph_model/loc_embeddings(:0)?$",
    "pos_embeddings": r"graph_model/pos_embedding(:0)?$",
    "word_embeddings": r"graph_model/word_embedding(:0)?$",
    "head_W": r"out_layer_task0/regression_gate/MLP_W_layer0(:0)?$",
    "head_b": r"out_layer_task0/regression_gate/MLP_b_layer0(:0)?$",
    "label_W": r"out_layer_task0/regression_gate/MLP_W_layer0_1(:0)?$",
    "label_b": r"out_layer_task0/regression_gate/MLP_b_layer0_1(:0)?$",
}
HEAD_SAVE_NAMES = {
    "loc_embeddings": "graph_model/loc_embeddings:0", "pos_embeddings": "graph_model/pos_embedding:0",
    "word_embeddings": "graph_model/word_embedding:0",
    "head_W": "out_layer_task0/regression_gate/MLP_W_layer0:0",
    "head_b": "out_layer_task0/regression_gate/MLP_b_layer0:0",
    "label_W": "out_layer_task0/regression_gate/MLP_W_layer0_1:0",
    "label_b": "out_layer_task0/regression_gate/MLP_b_layer0_1:0",
}


@dataclass
class ParserHeads:
    """Parameters of the front end (dense :217-235) and of the two output layers
    (chem_tensorflow.py:329-341; utils.MLP with no hidden layer = one linear map)."""
    loc_embeddings: np.ndarray     # [max_nodes, 80]
    pos_embeddings: np.ndarray     # [pos_size, 50]
    word_embeddings: np.ndarray    # [vocab, 100]
    head_W: np.ndarray             # [2h, 150]
    head_b: np.ndarray             # [150]
    label_W: np.ndarray            # [2h, E_out]
    label_b: np.ndarray            # [E_out]

    def validate(self, h: int) -> "ParserHeads":
        if self.head_W.shape != (2 * h, self.head_b.shape[0]) or self.label_W.shape != (2 * h, self.label_b.shape[0]):
            raise ValueError(f"output layers must be [2h, n] with h={h}: got {self.head_W.shape}, {self.label_W.shape}")
        width = 2 * self.loc_embeddings.shape[1] + self.pos_embeddings.shape[1] + self.word_embeddings.shape[1]
        if width > h:
            raise ValueError(f"embeddings are {width} wide, hidden size is {h} (SURVEY.md F7)")
        return self


def heads_from_checkpoint(weights: Dict[str, Any]) -> ParserHeads:
    """Pick the front-end / head tensors out of a reference ``weights`` dict (TF variable names)."""
    got = {}
    for key, pat in HEAD_TF_NAMES.items():
        hits = [k for k in weights if re.search(pat, k)]
        if len(hits) != 1:
            raise ValueError(f"checkpoint: expected exactly one variable matching {pat!r} for {key}, found {hits}")
        got[key] = np.ascontiguousarray(np.asarray(weights[hits[0]], dtype=np.float32))
    return ParserHeads(**got)


def bucket_size(graph: Sequence[Sequence[int]]) -> int:
    """First bucket size strictly greater than the largest node id (dense :529-532)."""
    mx = max(max(e[0], e[2]) for e in graph)
    return int(BUCKET_SIZES[np.argmax(BUCKET_SIZES > mx)])


def make_batches(records: Iterable[Dict[str, Any]], batch_size: int = 20):
    """Bucket by size, then consecutive slices of ``batch_size`` per bucket (dense :519-582, 803-807).
    Yields (v, list of records); records with an empty graph are skipped as the reference does."""
    buckets: Dict[int, List[Dict[str, Any]]] = {}
    for d in records:
        if len(d["graph"]) == 0:
            continue
        buckets.setdefault(bucket_size(d["graph"]), []).append(d)
    for v, items in buckets.items():
        for i in range(0, len(items), batch_size):
            yield v, items[i:i + batch_size]


def batch_arrays(v: int, items: Sequence[Dict[str, Any]], num_edge_types: int):
    """Host arrays of one batch: idx int32 [b, v, 4] (loc, pos, word, head_loc -- columns 0-3 of the
    reference's word_inputs, dense :825-842), edges int32 [nnz, 3] (src, channel, dst) with global
    node ids g*v + node built by the four rules of graph_to_adj_mat_bd (:65-83), n_active [b],
    targets int32 [b, v, 2] rows (head, label) for nodes 1..n (row 0 and padding = -1)."""
    b, E = len(items), num_edge_types
    idx = np.zeros((b, v, 4), np.int32)
    n_active = np.zeros(b, np.int32)
    targets = np.full((b, v, 2), -1, np.int32)
    edges = []
    for g, d in enumerate(items):
        na = len(d["node_features"])
        n_active[g] = na
        idx[g, :na, 0] = np.arange(na)
        idx[g, :na, 1] = d["node_features"]
        idx[g, :len(d["words_index"]), 2] = d["words_index"]
        heads = [0] + [e[0] for e in d["graph"]]
        idx[g, :len(heads), 3] = heads
        tr = np.asarray(d["graph"], np.int64)
        src, lab, dst = tr[:, 0], tr[:, 1], tr[:, 2]
        ch = np.concatenate([lab - 1, lab - 1 + E, np.full_like(lab, E - 1), np.full_like(lab, 2 * E - 1)])
        tg = np.concatenate([dst, src, dst, dst - 1])
        sc = np.concatenate([src, dst, dst - 1, dst])
        key = np.unique((ch * v + tg) * v + sc)          # assignment semantics: duplicates collapse
        edges.append(np.stack([g * v + key % v, key // (v * v), g * v + (key // v) % v], 1))
        t = np.asarray(d["targets"], np.int32)
        targets[g, 1:1 + len(t)] = t
    return idx, np.concatenate(edges).astype(np.int32), n_active, targets


def dense_adjacency_from_edges(edges: np.ndarray, b: int, v: int, two_e: int) -> np.ndarray:
    """The reference's feed format [b, 2E, v, v] float32 from the edge list of ``batch_arrays``."""
    adj = np.zeros((b, two_e, v, v), np.float32)
    g = edges[:, 0] // v
    adj[g, edges[:, 1], edges[:, 2] % v, edges[:, 0] % v] = 1.0
    return adj


def las_uas(pred: np.ndarray, targets: np.ndarray, n_active: np.ndarray) -> np.ndarray:
    """Per-sentence [LAS, UAS, UAS_e] (get_las_uas, dense :1304-1319): pred / targets int [b, v, 2]
    rows (head, label); nodes 1..n_active-1 of each sentence are scored."""
    out = np.zeros((len(n_active), 3))
    for g, na in enumerate(n_active):
        p, t = pred[g, 1:na], targets[g, 1:na]
        hm, lm = p[:, 0] == t[:, 0], p[:, 1] == t[:, 1]
        out[g] = [(hm & lm).mean(), hm.mean(), lm.mean()]
    return out


class ParserEvaluator:
    """Front end + propagation + heads + decode on one device."""

    def __init__(self, weights: GGNNWeights, heads: ParserHeads, device="cuda:0", num_timesteps: Optional[int] = None):
        self.prop = Propagator(weights, device)
        self.device = self.prop.device
        h = weights.hidden_size
        self.h, self.heads = h, heads.validate(h)
        self.T = int(num_timesteps if num_timesteps is not None else weights.params.get("num_timesteps", 4))
        self.n_head, self.n_lab = heads.head_b.shape[0], heads.label_b.shape[0]
        self.lab_off = (self.n_head + 3) // 4 * 4
        self.ncat = (self.lab_off + self.n_lab + 3) // 4 * 4
        Wcat = np.zeros((2 * h, self.ncat), np.float32)
        bcat = np.zeros(self.ncat, np.float32)
        Wcat[:, :self.n_head], bcat[:self.n_head] = heads.head_W, heads.head_b
        Wcat[:, self.lab_off:self.lab_off + self.n_lab] = heads.label_W
        bcat[self.lab_off:self.lab_off + self.n_lab] = heads.label_b
        to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(self.device)
        self.Wcat, self.bcat = to(Wcat), to(bcat)
        self.loc, self.pos, self.word = to(heads.loc_embeddings), to(heads.pos_embeddings), to(heads.word_embeddings)

    def embed(self, idx: torch.Tensor) -> torch.Tensor:
        """idx int32 [b, v, 4] on the device -> initial node representations [b, v, h]."""
        b, v, _ = idx.shape
        h0 = torch.empty(b, v, self.h, dtype=torch.float32, device=self.device)
        l = self.prop._l
        with torch.cuda.device(self.device):
            rc = l.ggnn_embed_nodes(self.prop._ctx, _ptr(idx.contiguous()), b * v,
                                    _ptr(self.loc), self.loc.shape[0], self.loc.shape[1],
                                    _ptr(self.pos), self.pos.shape[0], self.pos.shape[1],
                                    _ptr(self.word), self.word.shape[0], self.word.shape[1],
                                    self.h, _ptr(h0), _stream_ptr(self.device))
        self.prop._check(rc, "ggnn_embed_nodes")
        return h0

    def decode(self, hT: torch.Tensor, h0: torch.Tensor, n_active: torch.Tensor, probs: bool = False):
        """-> pred int32 [b, v, 2] (head, label), -1 on padded rows (and, with probs=True, the masked
        head / label probabilities [b, v, 150] / [b, v, E_out])."""
        b, v, _ = hT.shape
        rows = torch.arange(v, device=self.device, dtype=torch.int32).unsqueeze(0)
        na = n_active.to(self.device, torch.int32).unsqueeze(1)
        na_row = torch.where(rows < na, na, torch.zeros_like(na)).expand(b, v).contiguous()
        pred = torch.empty(b, v, 2, dtype=torch.int32, device=self.device)
        ph = torch.empty(b, v, self.n_head, device=self.device) if probs else None
        pl = torch.empty(b, v, self.n_lab, device=self.device) if probs else None
        with torch.cuda.device(self.device):
            rc = self.prop._l.ggnn_heads_decode(self.prop._ctx, _ptr(hT.contiguous()), _ptr(h0.contiguous()), b * v,
                                                self.h, _ptr(self.Wcat), _ptr(self.bcat), self.ncat, self.n_head,
                                                self.n_lab, self.lab_off, _ptr(na_row), _ptr(pred), _ptr(ph),
                                                _ptr(pl), _stream_ptr(self.device))
        self.prop._check(rc, "ggnn_heads_decode")
        return (pred, ph, pl) if probs else pred

    def predict_batch(self, v: int, items, precision: str = "fp32", feed: str = "edges"):
        """One batch of records -> (pred [b, v, 2] numpy, targets, n_active)."""
        two_e = self.prop.weights.two_e
        idx, edges, n_active, targets = batch_arrays(v, items, two_e // 2)
        b = len(items)
        d_idx = torch.from_numpy(idx).to(self.device)
        h0 = self.embed(d_idx)
        if feed == "dense":      # the reference's feed format (dense :854)
            adj = torch.from_numpy(dense_adjacency_from_edges(edges, b, v, two_e)).to(self.device)
            hT = self.prop.propagate_dense(h0, adj, num_timesteps=self.T, precision=precision)
        else:                    # N3: straight from the triples, no dense adjacency
            hT = self.prop.propagate_edges(h0.view(b * v, self.h), torch.from_numpy(edges).to(self.device),
                                           num_timesteps=self.T, precision=precision).view(b, v, self.h)
        pred = self.decode(hT, h0, torch.from_numpy(n_active)).cpu().numpy()
        pred[..., 1] = np.where(pred[..., 1] >= 0, pred[..., 1] + 1, -1)   # labels are 1-based (dense :127)
        return pred, targets, n_active

    def evaluate(self, records, batch_size: int = 20, precision: str = "fp32", feed: str = "edges"):
        """-> {"LAS", "UAS", "UAS_e"} macro-averaged over sentences (chem_tensorflow.py:631-633, 661-663)
        and the per-sentence scores in input-bucket order."""
        per = []
        for v, items in make_batches(records, batch_size):
            pred, targets, n_active = self.predict_batch(v, items, precision, feed)
            per.append(las_uas(pred, targets, n_active))
        per = np.concatenate(per) if per else np.zeros((0, 3))
        m = per.mean(0) if len(per) else np.zeros(3)
        return {"LAS": float(m[0]), "UAS": float(m[1]), "UAS_e": float(m[2]), "sentences": len(per)}, per
