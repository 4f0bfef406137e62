#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE's own NumPy helpers.

Runs only in the authoring container (needs /root/reference).  TensorFlow and
docopt are not installed, so both are stubbed in sys.modules: the helpers used
here (graph_to_adj_mat_bd, graph_to_adj_mat_dir, adj_mat_to_target, get_las_uas,
parser.to_graph.process_sentence) never touch tf.* at run time.
The committed .npz files are what travels to the GPU box.

    python scripts/make_golden.py
"""
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def import_reference():
    tf = types.ModuleType("tensorflow")
    tf.Tensor = object
    tf.float32 = "float32"
    sys.modules["tensorflow"] = tf
    docopt = types.ModuleType("docopt")
    docopt.docopt = lambda *a, **k: {}
    sys.modules["docopt"] = docopt
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "parser"))
    import chem_tensorflow_dense as dense   # noqa: E402
    import to_graph                          # noqa: E402
    return dense, to_graph


def read_conll(path):
    sents, cur = [], []
    with open(path) as f:
        for line in f:
            if line.strip() == "":
                if cur:
                    sents.append(cur)
                    cur = []
            else:
                cur.append(line.rstrip("\n"))
    if cur:
        sents.append(cur)
    return sents


def main():
    dense, to_graph = import_reference()
    os.makedirs(OUT, exist_ok=True)

    # ---- C1: the three small_dev.conll sentences through process_sentence('btb')
    deps = ["ROOT", "aux", "det", "nsubj"]
    E = len(deps) + 1                                # chem_tensorflow.py:198
    sents = read_conll(os.path.join(REF, "small_dev.conll"))
    word_dict = {"NA": 0}
    graphs = []
    for i, s in enumerate(sents):
        d = to_graph.process_sentence(
            s, deps, problem="btb", sentence_list_out=s, dep_list_out=deps,
            sentence_id=i, pos_list=to_graph.POS_list, word_dict=word_dict)
        graphs.append(d["graph"])
    bucket_sizes = np.array(list(range(4, 200, 2)))  # chem_tensorflow_dense.py:584-585
    vmax = max(max(max(e[0], e[2]) for e in g) for g in graphs)
    v = int(bucket_sizes[np.argmax(bucket_sizes > vmax)])     # :529-532
    adj = np.stack([dense.graph_to_adj_mat_bd(g, v, E) for g in graphs])  # [b, 2E, v, v]
    flat = {"adj": adj.astype(np.float32), "E": E, "v": v,
            "edges_bets": np.argwhere(adj != 0).astype(np.int32)}
    for i, g in enumerate(graphs):
        flat[f"triples_{i}"] = np.array(g, dtype=np.int32)
    np.savez_compressed(os.path.join(OUT, "c1_small_dev.npz"), **flat)
    print("C1: v=%d 2E=%d nnz per graph %s" % (v, 2 * E, [int((a != 0).sum()) for a in adj]))

    # ---- random trees through the reference builder, several shapes
    rng = np.random.default_rng(1234)
    cases = {}
    for ci, (n_graphs, vv, EE) in enumerate([(4, 8, 3), (5, 16, 6), (3, 32, 10), (2, 64, 46), (2, 128, 46)]):
        adjs, trip = [], []
        for gi in range(n_graphs):
            n = int(rng.integers(1, vv))             # words; nodes = n+1 <= vv
            g = []
            root = int(rng.integers(1, n + 1))
            for w in range(1, n + 1):
                if w == root:
                    head = 0
                else:
                    head = int(rng.integers(1, n + 1))
                    while head == w:
                        head = int(rng.integers(1, n + 1)) if n > 1 else 0
                g.append([head, int(rng.integers(1, EE)), w])     # label in 1..E-1
            adjs.append(dense.graph_to_adj_mat_bd(g, vv, EE))
            t = np.full((vv, 3), -1, np.int32)
            t[: len(g)] = np.array(g, np.int32)
            trip.append(t)
        a = np.stack(adjs)
        cases[f"case{ci}_shape"] = np.array([n_graphs, 2 * EE, vv], np.int32)
        cases[f"case{ci}_edges_bets"] = np.argwhere(a != 0).astype(np.int32)
        cases[f"case{ci}_triples"] = np.stack(trip)
    np.savez_compressed(os.path.join(OUT, "random_trees_ref_adj.npz"), **cases)

    # ---- conventions pinned by tests_chem.py:9-48 (run through the reference functions)
    g = [[3, 1, 1], [3, 1, 2], [0, 2, 3]]
    a_dir = dense.graph_to_adj_mat_dir(graph=g, max_n_vertices=4, num_edge_types=2)
    tgt = dense.adj_mat_to_target(adj_mat=np.pad(a_dir, ((0, 0), (0, 0), (0, 1))))
    np.savez_compressed(os.path.join(OUT, "conventions.npz"),
                        graph=np.array(g, np.int32), adj_dir=a_dir,
                        decoded=np.array(tgt, np.int32))
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
