#!/usr/bin/env python
"""Golden vectors for the floating-point rows either side of the path, made by
EXECUTING the reference's own functions over the NumPy-backed TF stand-in of make_golden_fp.py:

  N1  DenseGGNNChemModel.get_initial_node_representation   chem_tensorflow_dense.py:264-310 (btb branch)
  N2  DenseGGNNChemModel.gated_regression, btb branch       chem_tensorflow_dense.py:439-505
      with the reference's own utils.MLP (utils.py:40-84) as regression_gate, called twice as
      chem_tensorflow.py:343-355 does (head positions, then is_edge_regr=True for the labels)

Extra TF calls these bodies make, added to the shim here: nn.embedding_lookup, concat, pad, nn.softmax,
nn.tanh, Variable, zeros.  Dropouts run at keep-prob 1 (evaluation).  Output: tests/golden/ref_heads.npz.

    python scripts/make_golden_heads.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden_fp as G  # noqa: E402


def extend_shim(tf):
    tf.nn.embedding_lookup = lambda table, ids: np.asarray(table)[np.asarray(ids)]
    tf.nn.tanh = np.tanh
    tf.concat = lambda parts, axis: np.concatenate(parts, axis=axis)
    tf.zeros = lambda shape: np.zeros(shape, np.float32)
    tf.Variable = lambda init, **_: np.asarray(init)

    def pad(t, paddings, mode="constant"):
        assert mode == "constant"
        return np.pad(t, [(int(a), int(b)) for a, b in paddings])
    tf.pad = pad

    def softmax(x, axis=-1):
        e = np.exp(x - x.max(axis=axis, keepdims=True))
        return e / e.sum(axis=axis, keepdims=True)
    tf.nn.softmax = softmax


def main():
    dense = G.import_reference()
    tf = sys.modules["tensorflow"]
    extend_shim(tf)
    import utils as ref_utils                      # the reference's utils.py (MLP)
    rng = np.random.default_rng(77)
    np.random.seed(77)                             # utils.MLP.init_weights draws from np.random
    out = {}
    #        name  b   v   h    O   E_out  loc  pos  vocab
    cases = [("a", 3,  8, 320,  40,  7,    40,   9,   50),
             ("b", 2, 14, 312,  64, 46,    60,  48,  120)]
    for name, b, v, h, O, E_out, n_loc, n_pos, vocab in cases:
        for dtype in (np.float64, np.float32):
            m = object.__new__(dense.DenseGGNNChemModel)
            m.args = {"--pr": "btb"}
            m.params = {"hidden_size": h, "output_size": O}
            m.output_size_edges = E_out
            m.num_edge_types = E_out + 1
            m.pos_embedding_size, m.loc_embedding_size, m.word_embedding_size = 50, 80, 100   # chem_tensorflow.py:184-186
            if dtype is np.float64:                # draw once per case, evaluate in both precisions
                tabs = {"loc_embeddings": G.glorot(rng, n_loc, 80), "pos_embeddings": G.glorot(rng, n_pos, 50),
                        "word_embeddings": G.glorot(rng, vocab, 100), "edge_embeddings": G.glorot(rng, E_out + 1, 20)}
                wi = np.stack([rng.integers(0, n_loc, (b, v)), rng.integers(0, n_pos, (b, v)),
                               rng.integers(0, vocab, (b, v)), rng.integers(0, n_loc, (b, v)),
                               rng.integers(0, n_pos, (b, v)), rng.integers(0, E_out + 1, (b, v))], axis=-1).astype(np.int32)
                gate_heads = ref_utils.MLP(2 * h, O, [], 1.0)
                gate_edges = ref_utils.MLP(2 * h, E_out, [], 1.0)
                for mlp in (gate_heads, gate_edges):       # FP32 variables; non-zero biases so the bias term is exercised
                    mlp.params["weights"] = [np.asarray(w, np.float32) for w in mlp.params["weights"]]
                    mlp.params["biases"] = [(0.1 * rng.standard_normal(bb.shape)).astype(np.float32)
                                            for bb in mlp.params["biases"]]
                h_T = rng.standard_normal((b, v, h)).astype(np.float32)
            m.weights = {k: a.astype(dtype) for k, a in tabs.items()}
            m.placeholders = {"emb_dropout_keep_prob": 1.0, "word_inputs": wi, "num_graphs": b, "num_vertices": v,
                              "node_mask": None, "softmax_mask": None}
            h0 = dense.DenseGGNNChemModel.get_initial_node_representation(m)
            assert h0.shape == (b, v, h)

            def typed(mlp):
                t = object.__new__(ref_utils.MLP)
                t.dropout_keep_prob = 1.0
                t.params = {"weights": [w.astype(dtype) for w in mlp.params["weights"]],
                            "biases": [bb.astype(dtype) for bb in mlp.params["biases"]]}
                return t
            dense.DenseGGNNChemModel.gated_regression(m, h_T.astype(dtype), h0, typed(gate_heads), None, None)
            ph = m.output
            dense.DenseGGNNChemModel.gated_regression(m, h_T.astype(dtype), h0, typed(gate_edges), None, None,
                                                      is_edge_regr=True)
            pe = m.output
            assert ph.shape == (b, v * O) and pe.shape == (b, v * E_out)
            tag = np.dtype(dtype).name
            out[f"{name}_h0_{tag}"], out[f"{name}_p_head_{tag}"], out[f"{name}_p_label_{tag}"] = h0, ph, pe
        out[f"{name}_meta"] = np.array([b, v, h, O, E_out], np.int32)
        out[f"{name}_word_inputs"] = wi
        out[f"{name}_h_T"] = h_T
        for k in ("loc_embeddings", "pos_embeddings", "word_embeddings"):
            out[f"{name}_{k}"] = tabs[k]
        out[f"{name}_head_W"], out[f"{name}_head_b"] = gate_heads.params["weights"][0], gate_heads.params["biases"][0]
        out[f"{name}_label_W"], out[f"{name}_label_b"] = gate_edges.params["weights"][0], gate_edges.params["biases"][0]
        print(f"case {name}: b={b} v={v} h={h} O={O} E_out={E_out}  sum p_head row0 = {ph[0, :O].sum():.6f}")
    path = os.path.join(G.OUT, "ref_heads.npz")
    np.savez_compressed(path, **out)
    print("wrote ref_heads.npz,", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
