#!/usr/bin/env python
"""Generate tests/golden/ref_propagation.npz by EXECUTING the reference's own
`DenseGGNNChemModel.compute_final_node_representations` (chem_tensorflow_dense.py:312-340) and the
two timestep bodies it calls (`compute_timestep_fast` :391-437, `compute_timestep_normal` :350-389).

TensorFlow is not installable here, so the handful of TF primitives those bodies call are backed by
NumPy in the shim below (reshape / transpose / matmul / identity / dropout at keep-prob 1 / the
variable-scope context managers) and `GRUCell` is restated from TensorFlow's published
`rnn_cell_impl.GRUCell.call`:

    gate_inputs = matmul(concat([inputs, state], 1), gate_kernel) + gate_bias
    r, u        = split(sigmoid(gate_inputs), 2, axis=1)
    candidate   = matmul(concat([inputs, r * state], 1), candidate_kernel) + candidate_bias
    new_h       = u * state + (1 - u) * tanh(candidate)          # returned as (new_h, new_h)

Everything else -- the order of reshapes, transposes, the [2E, b, v', v] adjacency view (:195), the
bias broadcast, which operand is the target axis, the `[1]` taken from the cell's result, the
fixed_ts weight twins -- is the reference's own code running on these arrays.  The adjacency matrices
come from the reference's own `graph_to_adj_mat_bd`.

Runs only in the authoring container (needs /root/reference); the committed .npz travels.

    python scripts/make_golden_fp.py
"""
import contextlib
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


# ------------------------------------------------------------------------------------------
# NumPy-backed stand-in for the TF calls on the path
# ------------------------------------------------------------------------------------------
def make_tf_shim():
    tf = types.ModuleType("tensorflow")
    tf.Tensor = np.ndarray
    tf.float32 = np.float32
    tf.reshape = lambda t, shape: np.reshape(t, [int(s) for s in shape])
    tf.identity = lambda t: np.array(t, copy=True)
    tf.matmul = lambda a, b: np.matmul(a, b)

    def transpose(a, perm=None):
        return np.transpose(a, perm)
    tf.transpose = transpose

    nn = types.ModuleType("tensorflow.nn")

    def dropout(x, rate=0.0, **_):
        assert float(rate) == 0.0, "golden vectors are generated at keep-prob 1 (evaluation, :800-801)"
        return x
    nn.dropout = dropout
    tf.nn = nn

    class _Scope:
        def reuse_variables(self):
            pass

    @contextlib.contextmanager
    def variable_scope(name, *a, **k):
        yield _Scope()

    v1 = types.ModuleType("tensorflow.compat.v1")
    v1.variable_scope = variable_scope
    v1.get_variable_scope = lambda: _Scope()
    compat = types.ModuleType("tensorflow.compat")
    compat.v1 = v1
    tf.compat = compat
    return tf


class NumpyGRUCell:
    """tf.compat.v1.nn.rnn_cell.GRUCell (chem_tensorflow_dense.py:238) restated from
    tensorflow/python/ops/rnn_cell_impl.py GRUCell.call; default activation tanh."""

    def __init__(self, gate_kernel, gate_bias, cand_kernel, cand_bias):
        self.gk, self.gb, self.ck, self.cb = gate_kernel, gate_bias, cand_kernel, cand_bias

    def __call__(self, inputs, state):
        gate_inputs = np.matmul(np.concatenate([inputs, state], 1), self.gk) + self.gb
        value = 1.0 / (1.0 + np.exp(-gate_inputs))
        r, u = np.split(value, 2, axis=1)
        candidate = np.matmul(np.concatenate([inputs, r * state], 1), self.ck) + self.cb
        new_h = u * state + (1 - u) * np.tanh(candidate)
        return new_h, new_h


def import_reference():
    sys.modules["tensorflow"] = make_tf_shim()
    docopt = types.ModuleType("docopt")
    docopt.docopt = lambda *a, **k: {}
    sys.modules["docopt"] = docopt
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "parser"))
    import chem_tensorflow_dense as dense   # noqa: E402
    return dense


def bare_model(dense, tf, adj_placeholder, weights, params, fast, dtype):
    """A DenseGGNNChemModel without __init__ (which would build the whole TF graph): only the
    attributes the three functions on the path read."""
    m = object.__new__(dense.DenseGGNNChemModel)
    b, two_e, v, _ = adj_placeholder.shape
    m.placeholders = {"num_vertices": v, "num_graphs": b, "edge_weight_dropout_keep_prob": 1.0}
    m.num_edge_types = two_e // 2
    m.params = params
    m.pos_embedding_size = m.loc_embedding_size = m.word_embedding_size = 0
    m.ops = {}
    m.args = {"--pr": "btb", "--old": not fast}
    cast = lambda a: np.asarray(a, dtype=dtype)
    m.weights = {k: cast(a) for k, a in weights.items() if k.startswith("edge_")}
    m.weights["node_gru"] = NumpyGRUCell(cast(weights["gru_gate_kernel"]), cast(weights["gru_gate_bias"]),
                                         cast(weights["gru_cand_kernel"]), cast(weights["gru_cand_bias"]))
    # chem_tensorflow_dense.py:195 -- [b, 2E, v', v] placeholder -> [2E, b, v', v]
    m._DenseGGNNChemModel__adjacency_matrix = tf.transpose(a=cast(adj_placeholder), perm=[1, 0, 2, 3])
    return m


def random_trees(dense, rng, n_graphs, v, E):
    adjs = []
    for _ in range(n_graphs):
        n = int(rng.integers(1, v))                      # words; nodes = n + 1 <= v
        root = int(rng.integers(1, n + 1))
        g = []
        for w in range(1, n + 1):
            head = 0
            if w != root and n > 1:
                head = int(rng.integers(1, n + 1))
                while head == w:
                    head = int(rng.integers(1, n + 1))
            g.append([head, int(rng.integers(1, E)), w])
        adjs.append(dense.graph_to_adj_mat_bd(g, v, E))
    return np.stack(adjs).astype(np.float32)             # [b, 2E, v, v]


def glorot(rng, *shape):
    r = np.sqrt(6.0 / (shape[-2] + shape[-1]))
    return rng.uniform(-r, r, shape).astype(np.float32)


def main():
    dense = import_reference()
    tf = sys.modules["tensorflow"]
    rng = np.random.default_rng(20240611)
    out = {}
    #        name      b   v   E   h   T  bias  fixed_ts
    cases = [("a",     3,  8,  3, 12,  4, True,  None),
             ("b",     2, 16,  6, 20,  8, False, None),
             ("c",     4,  6,  4, 16,  4, True,  2),       # fixed_ts: *_fixed twins, 2 steps
             ("d",     2, 32, 10, 36,  8, True,  None),
             ("e",     2, 16,  3, 100, 8, True,  None)]    # h in (96, 128]: also served by the F16 kernels
    for name, b, v, E, h, T, bias, fixed_ts in cases:
        adj = random_trees(dense, rng, b, v, E)
        w = {"edge_weights": glorot(rng, 2 * E, h, h),
             "edge_biases": (0.05 * rng.standard_normal((2 * E, 1, h))).astype(np.float32),
             "edge_weights_fixed": glorot(rng, 2 * E, h, h),
             "edge_biases_fixed": (0.05 * rng.standard_normal((2 * E, 1, h))).astype(np.float32),
             "gru_gate_kernel": glorot(rng, 2 * h, 2 * h),
             "gru_gate_bias": (1.0 + 0.1 * rng.standard_normal(2 * h)).astype(np.float32),
             "gru_cand_kernel": glorot(rng, 2 * h, h),
             "gru_cand_bias": (0.1 * rng.standard_normal(h)).astype(np.float32)}
        h0 = rng.standard_normal((b, v, h)).astype(np.float32)
        if name == "e":
            h0 *= 0.5            # keeps h * max|W| * max|h0| under the F16 mode's range guard (include/ggnn_b200.h)
        params = {"hidden_size": h, "num_timesteps": T, "use_edge_bias": bias}
        out[f"{name}_meta"] = np.array([b, v, E, h, T, int(bias), -1 if fixed_ts is None else fixed_ts], np.int32)
        out[f"{name}_adj_nz"] = np.argwhere(adj != 0).astype(np.int32)       # (b, channel, target, source)
        out[f"{name}_h0"] = h0
        for k, a in w.items():
            if fixed_ts is not None or not k.endswith("_fixed"):
                out[f"{name}_{k}"] = a
        for fast in (True, False):
            if fixed_ts is not None and not fast:
                continue                                   # compute_timestep_normal has no fixed_ts form
            for dtype in (np.float64, np.float32):
                m = bare_model(dense, tf, adj, w, params, fast, dtype)
                res = dense.DenseGGNNChemModel.compute_final_node_representations(
                    m, np.asarray(h0, dtype=dtype), fixed_ts=fixed_ts)
                assert res.shape == (b, v, h) and res.dtype == dtype
                out[f"{name}_out_{'fast' if fast else 'normal'}_{np.dtype(dtype).name}"] = res
                if fast and dtype is np.float64:
                    out[f"{name}_acts_last_float64"] = m.ops["acts"]        # activations of the last step
        d = np.abs(out[f"{name}_out_fast_float64"] - out.get(f"{name}_out_normal_float64",
                                                              out[f"{name}_out_fast_float64"])).max()
        print(f"case {name}: b={b} v={v} 2E={2 * E} h={h} T={T} bias={bias} fixed_ts={fixed_ts} "
              f"nnz={len(out[f'{name}_adj_nz'])} |fast-normal|_max={d:.2e}")
    np.savez_compressed(os.path.join(OUT, "ref_propagation.npz"), **out)
    print("wrote ref_propagation.npz,", os.path.getsize(os.path.join(OUT, "ref_propagation.npz")), "bytes")


if __name__ == "__main__":
    main()
