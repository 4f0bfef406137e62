"""CPU tests: the oracle against the golden vectors made by the reference's own NumPy helpers
(scripts/make_golden.py), and its internal consistency (dense == loop == edge-list)."""
import os

import numpy as np
import pytest

from oracle import ggnn_oracle as O


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_c1_adjacency_matches_reference(golden_dir):
    d = _load(golden_dir, "c1_small_dev.npz")
    E, v = int(d["E"]), int(d["v"])
    assert (E, v) == (5, 6)
    for i in range(3):
        tri = d[f"triples_{i}"].tolist()
        ours = O.graph_to_adj_mat_bd(tri, v, E)
        assert np.array_equal(ours, d["adj"][i].astype(np.float64))
        # edge list straight from the triples == argwhere of the reference's dense matrix
        ref_edges = d["edges_bets"][d["edges_bets"][:, 0] == i][:, 1:]
        assert np.array_equal(O.edges_from_triples(tri, E), ref_edges)
    nnz = [(d["adj"][i] != 0).sum() for i in range(3)]
    assert nnz == [16, 20, 20]                      # 4n non-zeros
    # graph 0, spelled out in SURVEY.md 8(a) A1
    g0 = d["edges_bets"][d["edges_bets"][:, 0] == 0][:, 1:].tolist()
    assert g0[:4] == [[0, 3, 0], [2, 1, 2], [2, 4, 3], [3, 2, 3]]


def test_random_trees_match_reference_adjacency(golden_dir):
    d = _load(golden_dir, "random_trees_ref_adj.npz")
    ci = 0
    while f"case{ci}_shape" in d:
        b, c, v = d[f"case{ci}_shape"].tolist()
        E = c // 2
        adj = np.zeros((b, c, v, v), np.float32)
        for g in range(b):
            tri = [t for t in d[f"case{ci}_triples"][g].tolist() if t[2] >= 0]
            adj[g] = O.graph_to_adj_mat_bd(tri, v, E)
        assert np.array_equal(np.argwhere(adj != 0).astype(np.int32), d[f"case{ci}_edges_bets"])
        rowptr, col = O.compact_adjacency(adj)
        e4 = O.csr_to_edges(rowptr, col, b, c, v)
        assert np.array_equal(e4, d[f"case{ci}_edges_bets"])
        ci += 1
    assert ci == 5


def test_conventions_pinned_by_reference_tests(golden_dir):
    d = _load(golden_dir, "conventions.npz")
    # tests_chem.py:29-48: amat[e-1, dest, src] = 1
    a = np.zeros((2, 4, 4))
    for (src, e, dest) in d["graph"].tolist():
        a[e - 1, dest, src] = 1
    assert np.array_equal(a, d["adj_dir"])
    assert d["decoded"].tolist() == [[3, 1], [3, 1], [0, 2]]      # tests_chem.py:9-27


def _random_batch(rng, b, v, E):
    adjs = []
    for _ in range(b):
        n = v - 1
        tri = []
        for w in range(1, n + 1):
            head = 0 if w == 1 else int(rng.integers(1, w))
            tri.append([head, int(rng.integers(1, E)), w])
        adjs.append(O.graph_to_adj_mat_bd(tri, v, E))
    return np.stack(adjs).astype(np.float32)


def test_forms_agree_and_probe_numbers():
    """FP64 dense == FP64 edge-list to ~1e-15; FP32 within ~5e-7 of FP64."""
    rng = np.random.default_rng(0)
    b, v, E, h, T = 8, 32, 10, 128, 8
    adj = _random_batch(rng, b, v, E)
    w = O.make_weights(2 * E, h, 0)
    h0 = rng.uniform(-0.1, 0.1, (b, v, h)).astype(np.float32)
    r64 = O.propagate_dense(h0, adj, w, T, dtype=np.float64)
    edges = O.dense_to_edges(adj)
    e64 = O.propagate_edges(h0.reshape(-1, h), edges, w, T, dtype=np.float64).reshape(b, v, h)
    l64 = O.propagate_dense(h0, adj, w, T, dtype=np.float64, form="loop")
    assert O.normwise_err(e64, r64) < 1e-13
    assert O.normwise_err(l64, r64) < 1e-13
    r32 = O.propagate_dense(h0, adj, w, T, dtype=np.float32)
    e32 = O.propagate_edges(h0.reshape(-1, h), edges, w, T, dtype=np.float32).reshape(b, v, h)
    assert O.normwise_err(r32, r64) < 2e-6
    assert O.normwise_err(e32, r64) < 2e-6
    assert 0.3 < np.abs(r64).max() < 1.0


def test_bias_counted_per_edge_and_isolated_nodes_get_zero():
    """F9: acts[t] = sum over edges (h[s] W_e + b_e); a node without in-edges gets exactly 0."""
    E, v, h = 3, 6, 8
    tri = [[0, 1, 1], [1, 2, 2], [1, 2, 3]]        # nodes 4, 5 are padding
    adj = O.graph_to_adj_mat_bd(tri, v, E)[None].astype(np.float32)
    w = O.make_weights(2 * E, h, 1)
    hh = np.random.default_rng(2).standard_normal((v, h))
    acts = O.timestep_fast(hh, adj, w, True, np.float64)
    assert np.all(acts[4] == 0) and np.all(acts[5] == 0)
    # node 1: in-edges: head 0 (ch 0), children 2, 3 (ch 1+E each), prev 0 (ch E-1), next 2 (ch 2E-1)
    W, B = w["edge_weights"].astype(np.float64), w["edge_biases"].astype(np.float64)
    want = (hh[0] @ W[0] + B[0, 0]) + (hh[2] @ W[1 + E] + B[1 + E, 0]) + (hh[3] @ W[1 + E] + B[1 + E, 0]) \
        + (hh[0] @ W[E - 1] + B[E - 1, 0]) + (hh[2] @ W[2 * E - 1] + B[2 * E - 1, 0])
    assert np.allclose(acts[1], want, rtol=1e-12, atol=1e-12)


def test_fixed_ts_uses_fixed_weights():
    rng = np.random.default_rng(5)
    E, v, h = 4, 8, 16
    adj = _random_batch(rng, 2, v, E)
    w = O.make_weights(2 * E, h, 0)
    w2 = O.make_weights(2 * E, h, 9)
    w["edge_weights_fixed"], w["edge_biases_fixed"] = w2["edge_weights"], w2["edge_biases"]
    h0 = rng.uniform(-0.1, 0.1, (2, v, h)).astype(np.float32)
    a = O.propagate_dense(h0, adj, w, 4, fixed_ts=1)
    wf = dict(w, edge_weights=w2["edge_weights"], edge_biases=w2["edge_biases"])
    bb = O.propagate_dense(h0, adj, wf, 1)
    assert np.array_equal(a, bb)
    assert not np.allclose(a, O.propagate_dense(h0, adj, w, 1))


def test_padded_rows_are_updated():
    """F8: node_mask is never read inside the path; padded rows go through GRU(0, h)."""
    E, v, h = 3, 6, 8
    tri = [[0, 1, 1], [1, 2, 2]]
    adj = O.graph_to_adj_mat_bd(tri, v, E)[None].astype(np.float32)
    w = O.make_weights(2 * E, h, 3)
    h0 = np.random.default_rng(0).uniform(-0.5, 0.5, (1, v, h)).astype(np.float32)
    out = O.propagate_dense(h0, adj, w, 2)
    assert not np.allclose(out[0, 5], h0[0, 5])
    want = O.gru_cell(np.zeros((1, h)), O.gru_cell(np.zeros((1, h)), h0[0, 5:6].astype(np.float64), w, np.float64), w, np.float64)
    assert np.allclose(out[0, 5], want[0], atol=1e-12)


def test_tf32_emulation_sets_the_tolerance():
    """Operand rounding to TF32 (10 mantissa bits, FP32 accumulate) emulated on the CPU: on the
    tree workload it lands at ~7e-4 after T=8 (inside the 2e-3 bound), BF16 (7 bits) does not."""
    rng = np.random.default_rng(0)
    b, v, E, h, T = 4, 32, 10, 128, 8
    adj = _random_batch(rng, b, v, E)
    w = O.make_weights(2 * E, h, 0)
    h0 = rng.uniform(-0.1, 0.1, (b * v, h)).astype(np.float32)
    edges = O.dense_to_edges(adj)
    ref = O.propagate_edges(h0, edges, w, T)
    e_tf32 = O.normwise_err(O.propagate_edges_emulated(h0, edges, w, T, bits=10), ref)
    e_bf16 = O.normwise_err(O.propagate_edges_emulated(h0, edges, w, T, bits=7), ref)
    assert e_tf32 < 2e-3 < e_bf16
    assert np.array_equal(O.round_mantissa(np.float32([1.0, 1.0 + 2 ** -11, 1.0 + 3 * 2 ** -11])),
                          np.float32([1.0, 1.0, 1.0 + 2 ** -9]))


def test_grad_oracle_forward_and_finite_differences():
    """The autograd twin (oracle/grad_oracle.py) reproduces the forward oracle and its gradients
    agree with central finite differences of the forward oracle."""
    from oracle import grad_oracle as G
    from ggnn_b200 import synth
    b, v, E, h, T = 2, 6, 3, 8, 3
    tri = synth.random_trees(b, v - 1, E, seed=3)
    edges = synth.global_edge_list(tri, v, E).numpy()
    w = O.make_weights(2 * E, h, 5)
    rng = np.random.default_rng(0)
    h0 = rng.uniform(-0.5, 0.5, (b * v, h))
    d = rng.standard_normal((b * v, h))
    hT, dh0, g = G.grads(h0, edges, w, T, d)
    ref = O.propagate_edges(h0, edges, w, T)
    assert np.abs(hT - ref).max() < 1e-12
    loss = lambda ww, hh: float((O.propagate_edges(hh, edges, ww, T) * d).sum())
    eps = 1e-6
    for key, idx in (("edge_weights", (1, 2, 3)), ("edge_biases", (4, 0, 5)), ("gru_gate_kernel", (3, 9)),
                     ("gru_gate_bias", (11,)), ("gru_cand_kernel", (12, 2)), ("gru_cand_bias", (7,))):
        wp = {k: np.array(a, dtype=np.float64) for k, a in w.items()}
        wm = {k: np.array(a, dtype=np.float64) for k, a in w.items()}
        wp[key][idx] += eps
        wm[key][idx] -= eps
        fd = (loss(wp, h0) - loss(wm, h0)) / (2 * eps)
        assert abs(fd - g[key][idx]) < 1e-6 * max(1.0, abs(fd)), (key, fd, g[key][idx])
    hp, hm = h0.copy(), h0.copy()
    hp[5, 2] += eps
    hm[5, 2] -= eps
    fd = (loss(w, hp) - loss(w, hm)) / (2 * eps)
    assert abs(fd - dh0[5, 2]) < 1e-6 * max(1.0, abs(fd))


def test_parser_oracle_pinned_by_reference_decode(golden_dir):
    """masks, target formatting, decode and LAS/UAS against the reference's own NumPy code
    (tests/golden/parser_decode.npz, scripts/make_golden_parser.py), incl. first-maximum ties."""
    from oracle import parser_oracle as P
    d = np.load(os.path.join(golden_dir, "parser_decode.npz"))
    E_out, Osz = int(d["E_out"]), int(d["O"])
    for ci in range(3):
        v = int(d[f"c{ci}_v"])
        na, tg = d[f"c{ci}_n_active"], d[f"c{ci}_targets"]
        masks = np.stack([P.get_mask(int(n), v, E_out, Osz) for n in na])
        masks_e = np.stack([P.get_mask(int(n), v, E_out, Osz, is_edge=True) for n in na])
        assert np.array_equal(masks.astype(np.float32), d[f"c{ci}_mask"])
        assert np.array_equal(masks_e.astype(np.float32), d[f"c{ci}_mask_edges"])
        labels = [P.target_to_adj_mat([list(t) for t in tg[g][: na[g] - 1]], v, E_out, v) for g in range(len(na))]
        tv, tve = P.target_values(labels, Osz), P.target_values_edges(labels)
        assert np.array_equal(tv.astype(np.float32), d[f"c{ci}_target_values"])
        assert np.array_equal(tve.astype(np.float32), d[f"c{ci}_target_values_edges"])
        b = len(na)
        scores, _ = P.score_batch(tv, d[f"c{ci}_ph"].reshape(b, -1), masks, tve, d[f"c{ci}_pe"].reshape(b, -1),
                                  masks_e, v, E_out, Osz)
        assert np.allclose(scores.mean(0), d[f"c{ci}_scores"], rtol=0, atol=1e-15)
    wi = np.zeros((2, 6)); wi[0, :3] = [3, 1, 2]; wi[1, :1] = [5]
    assert np.array_equal(wi, d["word_inputs_padded"])


def test_host_scorer_and_batch_builder_match_the_oracle(golden_dir):
    """ggnn_b200.parser_eval's vectorised scorer == the reference-shaped decode+scorer, and its
    batch builder == graph_to_adj_mat_bd / the reference's word_inputs columns."""
    from oracle import parser_oracle as P
    from ggnn_b200 import parser_eval as PE, synth
    d = np.load(os.path.join(golden_dir, "parser_decode.npz"))
    E_out, Osz = int(d["E_out"]), int(d["O"])
    for ci in range(3):
        v, na, tg = int(d[f"c{ci}_v"]), d[f"c{ci}_n_active"], d[f"c{ci}_targets"]
        b = len(na)
        ph, pe = d[f"c{ci}_ph"], d[f"c{ci}_pe"]
        pred = np.full((b, v, 2), -1, np.int32)
        targets = np.full((b, v, 2), -1, np.int32)
        for g in range(b):
            n = int(na[g])
            pred[g, :n, 0] = ph[g, :n, :n].argmax(-1)      # first maximum among s < n_active
            pred[g, :n, 1] = pe[g, :n].argmax(-1) + 1
            targets[g, 1:n] = tg[g, : n - 1]
        per = PE.las_uas(pred, targets, na)
        assert np.allclose(per.mean(0), d[f"c{ci}_scores"], rtol=0, atol=1e-15)
    E = 6
    recs = synth.wsj_like_corpus(12, E, 5, pos_size=9, vocab=50, seed=4, max_len=20)
    for v, items in PE.make_batches(recs, batch_size=5):
        idx, edges, n_active, targets = PE.batch_arrays(v, items, E)
        adj = PE.dense_adjacency_from_edges(edges, len(items), v, 2 * E)
        for g, r in enumerate(items):
            assert v == synth.bucket_size(max(max(e[0], e[2]) for e in r["graph"]))
            assert np.array_equal(adj[g], O.graph_to_adj_mat_bd(r["graph"], v, E).astype(np.float32))
            n = len(r["node_features"])
            assert n_active[g] == n and list(idx[g, :n, 0]) == list(range(n))
            assert list(idx[g, :n, 1]) == r["node_features"] and list(idx[g, :n, 2]) == r["words_index"]
            assert list(idx[g, :n, 3]) == [0] + [e[0] for e in r["graph"]]
            assert np.array_equal(targets[g, 1:n], np.array(r["targets"]))


def test_adam_oracle_matches_torch_adam_with_per_tensor_clip():
    """The NumPy restatement of make_train_step (chem_tensorflow.py:483-506) against torch.optim.Adam (FP64).
    TF1 puts epsilon outside the UNcorrected root ("epsilon hat"): lr_t m / (sqrt(v) + eps) equals torch's
    update with eps_torch = eps / sqrt(1 - beta2^t), set per step here, so the two agree to rounding.
    tf.clip_by_norm is g * clip / max(||g||, clip), applied per variable."""
    import torch
    from oracle import grad_oracle as G
    rng = np.random.default_rng(0)
    shapes = {"a": (7, 5), "b": (11,)}
    P = {k: rng.standard_normal(s) for k, s in shapes.items()}
    M = {k: np.zeros(s) for k, s in shapes.items()}
    V = {k: np.zeros(s) for k, s in shapes.items()}
    tp = {k: torch.tensor(P[k].copy(), requires_grad=True) for k in shapes}
    opt = torch.optim.Adam(list(tp.values()), lr=0.003, betas=(0.9, 0.999), eps=1e-8)
    for t in range(1, 6):
        g = {"a": rng.standard_normal(shapes["a"]) * 2.0, "b": rng.standard_normal(shapes["b"]) * 0.01}
        assert np.linalg.norm(g["a"]) > 1.0 > np.linalg.norm(g["b"])      # one clipped, one not
        for k in shapes:
            tp[k].grad = torch.tensor(g[k] * 1.0 / max(np.linalg.norm(g[k]), 1.0))
        opt.param_groups[0]["eps"] = 1e-8 / np.sqrt(1.0 - 0.999 ** t)
        opt.step()
        G.adam_step(P, g, M, V, t)
        for k in shapes:
            np.testing.assert_allclose(P[k], tp[k].detach().numpy(), rtol=0, atol=1e-12)


def test_philox_known_answers():
    """Random123's published known-answer vectors for philox4x32-10 (kat_vectors: counter, key -> output)."""
    from oracle import grad_oracle as G
    z = G.philox4x32_10(np.zeros((1, 4), np.uint32), (0, 0))[0]
    assert [hex(int(x)) for x in z] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    f = G.philox4x32_10(np.full((1, 4), 0xFFFFFFFF, np.uint32), (0xFFFFFFFF, 0xFFFFFFFF))[0]
    assert [hex(int(x)) for x in f] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    p = G.philox4x32_10(np.array([[0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344]], np.uint32), (0xa4093822, 0x299f31d0))[0]
    assert [hex(int(x)) for x in p] == ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]
    m = G.philox_mask(100003, seed=7, stream=3, keep=0.9)
    assert m.shape == (100003,) and abs(float(m.mean()) - 0.9) < 5e-3
    assert not np.array_equal(m, G.philox_mask(100003, seed=7, stream=4, keep=0.9))


def _ref_case(d, name):
    b, v, E, h, T, bias, fixed_ts = (int(x) for x in d[f"{name}_meta"])
    adj = np.zeros((b, 2 * E, v, v), np.float32)
    nz = d[f"{name}_adj_nz"]
    adj[nz[:, 0], nz[:, 1], nz[:, 2], nz[:, 3]] = 1.0
    w = {k: d[f"{name}_{k}"] for k in ("edge_weights", "edge_biases", "edge_weights_fixed", "edge_biases_fixed",
                                        "gru_gate_kernel", "gru_gate_bias", "gru_cand_kernel", "gru_cand_bias")
         if f"{name}_{k}" in d.files}
    return b, v, E, h, T, bool(bias), (None if fixed_ts < 0 else fixed_ts), adj, w, d[f"{name}_h0"]


@pytest.mark.parametrize("name", ["a", "b", "c", "d", "e"])
def test_oracle_pinned_by_the_reference_functions_run_on_a_numpy_tf_shim(golden_dir, name):
    """tests/golden/ref_propagation.npz holds outputs of the reference's OWN
    compute_final_node_representations / compute_timestep_fast / compute_timestep_normal
    (chem_tensorflow_dense.py:312-437) executed over NumPy-backed TF primitives
    (scripts/make_golden_fp.py; GRUCell restated from TF's rnn_cell_impl).  Every form of the oracle
    has to reproduce them."""
    d = np.load(os.path.join(golden_dir, "ref_propagation.npz"))
    b, v, E, h, T, bias, fixed_ts, adj, w, h0 = _ref_case(d, name)
    ref = d[f"{name}_out_fast_float64"]
    got = O.propagate_dense(h0, adj, w, T, fixed_ts=fixed_ts, use_bias=bias, dtype=np.float64, form="fast")
    assert np.abs(got - ref).max() <= 1e-13
    if fixed_ts is None:
        assert np.abs(d[f"{name}_out_normal_float64"] - ref).max() <= 1e-13     # the reference's two forms agree
        got = O.propagate_dense(h0, adj, w, T, use_bias=bias, dtype=np.float64, form="loop")
        assert np.abs(got - d[f"{name}_out_normal_float64"]).max() <= 1e-13
    edges = O.dense_to_edges(adj)
    got = O.propagate_edges(h0.reshape(b * v, h), edges, w, T, fixed_ts=fixed_ts, use_bias=bias, dtype=np.float64)
    assert np.abs(got.reshape(b, v, h) - ref).max() <= 1e-13
    # FP32 evaluation of the reference's code and of the oracle: both inside the strict tolerance
    assert O.normwise_err(d[f"{name}_out_fast_float32"], ref) <= 1e-5
    got32 = O.propagate_dense(h0, adj, w, T, fixed_ts=fixed_ts, use_bias=bias, dtype=np.float32)
    assert O.normwise_err(got32, ref) <= 1e-5
    # torch-CPU restatement used as the timed CPU baseline
    from oracle import cpu_reference as C
    import torch
    wt = O._select_weights(w, fixed_ts)
    Tt = T if fixed_ts is None else fixed_ts
    got_t = C.propagate_dense_torch_cpu(torch.from_numpy(h0), torch.from_numpy(adj), wt, Tt, use_bias=bias)
    assert O.normwise_err(got_t.numpy(), ref) <= 1e-5
    got_e = C.propagate_edges_torch_cpu(torch.from_numpy(h0.reshape(b * v, h)), torch.from_numpy(edges), wt, Tt,
                                        use_bias=bias)
    assert O.normwise_err(got_e.numpy().reshape(b, v, h), ref) <= 1e-5


@pytest.mark.parametrize("name", ["a", "b"])
def test_front_end_and_heads_pinned_by_the_reference_functions(golden_dir, name):
    """N1 / N2 floating-point part: tests/golden/ref_heads.npz holds the outputs of the reference's own
    get_initial_node_representation (chem_tensorflow_dense.py:264-310) and gated_regression (:439-505, with
    the reference's utils.MLP) executed over NumPy-backed TF primitives (scripts/make_golden_heads.py)."""
    from oracle import parser_oracle as P
    d = np.load(os.path.join(golden_dir, "ref_heads.npz"))
    b, v, h, Osz, E_out = (int(x) for x in d[f"{name}_meta"])
    h0 = P.initial_node_representation(d[f"{name}_word_inputs"], d[f"{name}_loc_embeddings"],
                                       d[f"{name}_pos_embeddings"], d[f"{name}_word_embeddings"], h)
    assert np.array_equal(h0, d[f"{name}_h0_float64"])                      # a gather + zero padding: exact
    ph, pe = P.heads(d[f"{name}_h_T"], h0, d[f"{name}_head_W"], d[f"{name}_head_b"],
                     d[f"{name}_label_W"], d[f"{name}_label_b"])
    assert ph.shape == (b, v * Osz) and pe.shape == (b, v * E_out)
    assert np.abs(ph - d[f"{name}_p_head_float64"]).max() <= 1e-14
    assert np.abs(pe - d[f"{name}_p_label_float64"]).max() <= 1e-14
    assert np.abs(d[f"{name}_p_head_float32"] - ph).max() <= 1e-6
