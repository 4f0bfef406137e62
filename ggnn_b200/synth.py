"""Synthetic inputs in the reference's formats: dependency trees as
``(head, label, word)`` triples (the ``graph`` field of the reference's JSON,
parser/to_graph.py:36-41), their dense ``[b, 2E, v, v]`` adjacency built by the four rules of
``graph_to_adj_mat_bd`` (chem_tensorflow_dense.py:65-83), and random multi-relational graphs
for the edge-list path.  Pure torch so the same code runs on the host (tests) and on the
device (bench: the 6 MB/graph dense tensors are generated in HBM, never shipped over PCIe).
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch


def glorot_init(rng, shape):
    """The reference's initialiser (utils.py:11-14), with a local Generator."""
    r = np.sqrt(6.0 / (shape[-2] + shape[-1]))
    return rng.uniform(low=-r, high=r, size=shape).astype(np.float32)


def make_weights(two_e: int, h: int, seed: int = 0, bias_std: float = 0.05) -> dict:
    """Synthetic parameters in the reference's shapes (chem_tensorflow_dense.py:202-212, 237-241; TF
    GRUCell init: gate bias 1.0, candidate bias 0).  Edge biases are N(0, bias_std) rather than the
    reference's zeros so that the per-edge bias rule is exercised.  Same stream of
    draws as the generator the tests use, so a seed names the same weights on both sides."""
    rng = np.random.default_rng(seed)
    return {
        "edge_weights": glorot_init(rng, (two_e, h, h)),
        "edge_biases": (rng.standard_normal((two_e, 1, h)) * bias_std).astype(np.float32),
        "gru_gate_kernel": glorot_init(rng, (2 * h, 2 * h)),
        "gru_gate_bias": np.ones((2 * h,), np.float32),
        "gru_cand_kernel": glorot_init(rng, (2 * h, h)),
        "gru_cand_bias": np.zeros((h,), np.float32),
    }


def zipf_probs(n: int, s: float = 1.1) -> torch.Tensor:
    p = torch.arange(1, n + 1, dtype=torch.float64).pow(-s)
    return (p / p.sum()).float()


def random_trees(b: int, n_words: int, num_edge_types: int, seed: int = 0,
                 device="cpu", lengths: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Random recursive trees.  Returns int32 [b, n_words, 3] rows (head, label, word) with
    word = 1..n (node 0 is the artificial ROOT), label in 1..E-1 drawn Zipf(1.1); rows past a
    graph's own length (``lengths`` [b], optional) are (-1, -1, -1)."""
    E = num_edge_types
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = n_words
    # tree over ranks 1..n: rank 1 is the root (head 0), rank r>1 attaches to a uniform earlier rank
    u = torch.rand(b, n, generator=g)
    ranks = torch.arange(1, n + 1).unsqueeze(0).expand(b, n)
    head_rank = (u * (ranks - 1).clamp(min=1)).floor().long() + 1          # in [1, r-1]
    head_rank[:, 0] = 0
    if lengths is None:
        lengths = torch.full((b,), n, dtype=torch.long)
    lengths = lengths.long().clamp(1, n)
    # random word position for every rank: a permutation of 1..len per graph (ranks > len unused)
    keys = torch.rand(b, n, generator=g)
    keys = torch.where(ranks <= lengths.unsqueeze(1), keys, torch.full_like(keys, 2.0))
    perm = keys.argsort(dim=1) + 1               # perm[:, j] = rank placed at word position j+1
    pos_of_rank = torch.empty(b, n + 1, dtype=torch.long)
    pos_of_rank[:, 0] = 0
    pos_of_rank.scatter_(1, perm, torch.arange(1, n + 1).unsqueeze(0).expand(b, n).contiguous())
    word = pos_of_rank.gather(1, ranks)                      # word position of rank r
    head = pos_of_rank.gather(1, head_rank)                  # word position of its head (0 = ROOT)
    label = torch.multinomial(zipf_probs(E - 1), b * n, replacement=True, generator=g).view(b, n) + 1
    tri = torch.stack([head, label, word], dim=2)
    valid = (ranks <= lengths.unsqueeze(1)).unsqueeze(2)
    tri = torch.where(valid, tri, torch.full_like(tri, -1))
    # order rows by word position like the CoNLL reader does (padding rows last)
    order = torch.where(valid.squeeze(2), word, torch.full_like(word, n + 1)).argsort(dim=1)
    tri = tri.gather(1, order.unsqueeze(2).expand(b, n, 3))
    return tri.to(torch.int32).to(device)


def edges_from_triples(triples: torch.Tensor, v: int, num_edge_types: int) -> torch.Tensor:
    """The four rules of graph_to_adj_mat_bd as an edge list: int32 [4*m, 4] rows
    (graph, channel, target, source) for the m valid triples (not de-duplicated, unsorted)."""
    E = num_edge_types
    b, n, _ = triples.shape
    t = triples.long()
    gidx = torch.arange(b, device=t.device).view(b, 1).expand(b, n)
    valid = t[..., 2] >= 0
    src, lab, dst = t[..., 0][valid], t[..., 1][valid], t[..., 2][valid]
    g = gidx[valid]
    ch = torch.cat([lab - 1, lab - 1 + E, torch.full_like(lab, E - 1), torch.full_like(lab, 2 * E - 1)])
    tg = torch.cat([dst, src, dst, dst - 1])
    sc = torch.cat([src, dst, dst - 1, dst])
    return torch.stack([g.repeat(4), ch, tg, sc], dim=1).to(torch.int32)


def dense_adjacency(triples: torch.Tensor, v: int, num_edge_types: int,
                    dtype=torch.float32, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[b, 2E, v, v] adjacency, A[b, e, target, source] = 1 (assignment, so duplicates are 1)."""
    b = triples.shape[0]
    c = 2 * num_edge_types
    if out is None:
        out = torch.zeros(b, c, v, v, dtype=dtype, device=triples.device)
    else:
        out.zero_()
    e = edges_from_triples(triples, v, num_edge_types).long()
    flat = ((e[:, 0] * c + e[:, 1]) * v + e[:, 2]) * v + e[:, 3]
    out.view(-1)[flat] = 1
    return out


def global_edge_list(triples: torch.Tensor, v: int, num_edge_types: int) -> torch.Tensor:
    """int32 [nnz, 3] rows (src, channel, dst) with global node ids g*v + node, de-duplicated
    -- the input format of the edge-list path for the same graphs."""
    e = edges_from_triples(triples, v, num_edge_types).long()
    c = 2 * num_edge_types
    key = ((e[:, 0] * c + e[:, 1]) * v + e[:, 2]) * v + e[:, 3]
    key = torch.unique(key)
    s = key % v
    t = (key // v) % v
    ch = (key // (v * v)) % c
    g = key // (v * v * c)
    return torch.stack([g * v + s, ch, g * v + t], dim=1).to(torch.int32)


def packed_edge_list(triples: torch.Tensor, rows_per_graph: torch.Tensor, num_edge_types: int):
    """Edge list for ragged graphs packed back to back: graph g owns node rows
    [offsets[g], offsets[g+1]) with rows_per_graph[g] rows (its bucket size, padded rows included
    as in the reference).  Returns (int32 [nnz, 3] (src, channel, dst), int64 offsets [b+1])."""
    e = edges_from_triples(triples, 0, num_edge_types).long()
    off = torch.zeros(triples.shape[0] + 1, dtype=torch.long, device=triples.device)
    off[1:] = torch.cumsum(rows_per_graph.long().to(triples.device), 0)
    base = off[e[:, 0]]
    return torch.stack([base + e[:, 3], e[:, 1], base + e[:, 2]], dim=1).to(torch.int32), off


def random_multigraph(n_graphs: int, nodes: int, base_edges: int, num_types: int, seed: int = 0,
                      device="cpu") -> torch.Tensor:
    """C4-style graphs: per graph `base_edges` uniform random (src, dst, type) edges plus their
    reverses on channel type + num_types.  Returns int32 [n_graphs*2*base_edges, 3]
    (src, channel, dst) with global node ids."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    src = torch.randint(0, nodes, (n_graphs, base_edges), generator=g)
    dst = torch.randint(0, nodes, (n_graphs, base_edges), generator=g)
    typ = torch.randint(0, num_types, (n_graphs, base_edges), generator=g)
    off = (torch.arange(n_graphs) * nodes).view(-1, 1)
    fwd = torch.stack([src + off, typ, dst + off], dim=2)
    bwd = torch.stack([dst + off, typ + num_types, src + off], dim=2)
    return torch.cat([fwd, bwd], dim=1).view(-1, 3).to(torch.int32).to(device)


def wsj_like_lengths(n: int, seed: int = 0, max_len: int = 140) -> torch.Tensor:
    """Sentence lengths ~ round(lognormal(3.1, 0.45)) clipped to [2, max_len]."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    z = torch.randn(n, generator=g)
    return torch.exp(3.1 + 0.45 * z).round().clamp(2, max_len).long()


def bucket_size(max_node_id: int) -> int:
    """First even bucket size in 4, 6, ..., 198 strictly greater than the largest node id
    (chem_tensorflow_dense.py:529-532, 584-585)."""
    for s in range(4, 200, 2):
        if s > max_node_id:
            return s
    raise ValueError(f"node id {max_node_id} exceeds the reference's largest bucket (198)")


def wsj_like_corpus(n_sentences: int, num_edge_types: int, num_out_labels: int, pos_size: int = 45,
                    vocab: int = 5000, seed: int = 0, max_len: int = 140, agree: float = 0.8):
    """Synthetic sentences in the reference's JSON record format (parser/to_graph.py:68-85): 'graph'
    = input-bank tree as (head, label, word) triples, 'targets' = output-bank (head, label) per word
    (agreeing with the input head with probability `agree`), 'node_features' = POS id per node (ROOT
    first), 'words_index' = vocabulary id per node.  Lengths follow wsj_like_lengths."""
    lengths = wsj_like_lengths(n_sentences, seed=seed, max_len=max_len)
    tri = random_trees(n_sentences, int(lengths.max()), num_edge_types, seed=seed, lengths=lengths)
    alt = random_trees(n_sentences, int(lengths.max()), num_out_labels + 1, seed=seed + 1, lengths=lengths)
    g = torch.Generator().manual_seed(seed + 2)
    recs = []
    for i in range(n_sentences):
        n = int(lengths[i])
        graph = tri[i, :n].tolist()
        keep = torch.rand(n, generator=g) < agree
        targets = [[graph[j][0] if keep[j] else alt[i, j, 0].item(), int(alt[i, j, 1])] for j in range(n)]
        recs.append({"graph": graph, "targets": targets,
                     "node_features": torch.randint(0, pos_size, (n + 1,), generator=g).tolist(),
                     "words_index": torch.randint(0, vocab, (n + 1,), generator=g).tolist(),
                     "id": i})
    return recs
