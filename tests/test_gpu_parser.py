"""BASELINE config 2 on synthetic WSJ-shaped data (the real WSJ files and the .pickle are not
shipped): front end -> propagation -> heads -> decode -> LAS/UAS through the CUDA
pipeline must give IDENTICAL per-sentence scores to the oracle pipeline on the same inputs and
weights (read from a synthetic checkpoint in the reference's pickle format)."""
import os
import pickle

import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import parser_eval as PE, synth
from oracle import ggnn_oracle as O
from oracle import parser_oracle as P

pytestmark = pytest.mark.gpu


def _checkpoint(tmp_path, E, E_out, h, pos_size, vocab, max_nodes, seed=0):
    w = O.make_weights(2 * E, h, seed)
    rng = np.random.default_rng(seed + 7)
    g = lambda *s: O.glorot_init(rng, s)
    heads = PE.ParserHeads(loc_embeddings=g(max_nodes, 80), pos_embeddings=g(pos_size, 50),
                           word_embeddings=g(vocab, 100), head_W=g(2 * h, PE.OUTPUT_SIZE) * 4,
                           head_b=(rng.standard_normal(PE.OUTPUT_SIZE) * 0.1).astype(np.float32),
                           label_W=g(2 * h, E_out) * 4, label_b=(rng.standard_normal(E_out) * 0.1).astype(np.float32))
    path = os.path.join(tmp_path, "synthetic_model_best.pickle")
    extra = {PE.HEAD_SAVE_NAMES[k]: getattr(heads, k) for k in PE.HEAD_SAVE_NAMES}
    ggnn_b200.save_checkpoint(path, ggnn_b200.from_dict(w), params={"num_timesteps": 4, "hidden_size": h}, extra=extra)
    return path, w, heads


def _oracle_scores(recs, w, heads, E, E_out, h, T, batch_size):
    per, graphs = [], []
    for v, items in PE.make_batches(recs, batch_size):
        idx, edges, n_active, _ = PE.batch_arrays(v, items, E)
        b = len(items)
        h0 = P.initial_node_representation(idx, heads.loc_embeddings, heads.pos_embeddings, heads.word_embeddings, h)
        adj = np.stack([O.graph_to_adj_mat_bd(r["graph"], v, E) for r in items])
        hT = O.propagate_dense(h0, adj, w, T)
        ph, pe = P.heads(hT, h0, heads.head_W, heads.head_b, heads.label_W, heads.label_b)
        masks = np.stack([P.get_mask(int(n), v, E_out, PE.OUTPUT_SIZE) for n in n_active])
        masks_e = np.stack([P.get_mask(int(n), v, E_out, PE.OUTPUT_SIZE, is_edge=True) for n in n_active])
        labels = [P.target_to_adj_mat(r["targets"], v, E_out, v) for r in items]
        sc, gr = P.score_batch(P.target_values(labels, PE.OUTPUT_SIZE), ph, masks, P.target_values_edges(labels), pe,
                               masks_e, v, E_out, PE.OUTPUT_SIZE)
        per.append(sc)
        graphs += gr
    return np.concatenate(per), graphs


@pytest.mark.parametrize("feed", ["edges", "dense"])
def test_las_uas_identical_to_oracle(tmp_path, feed):
    E, E_out, h, T = 12, 11, 400, 4
    path, w, heads = _checkpoint(str(tmp_path), E, E_out, h, pos_size=20, vocab=300, max_nodes=150)
    data = pickle.load(open(path, "rb"))
    ev = PE.ParserEvaluator(ggnn_b200.load_checkpoint(path), PE.heads_from_checkpoint(data["weights"]), "cuda:0")
    assert ev.T == T
    recs = synth.wsj_like_corpus(60, E, E_out, pos_size=20, vocab=300, seed=3, max_len=60)
    ref_per, ref_graphs = _oracle_scores(recs, w, heads, E, E_out, h, T, batch_size=20)
    res, per = ev.evaluate(recs, batch_size=20, precision="fp32", feed=feed)
    assert res["sentences"] == 60
    assert np.array_equal(per, ref_per), np.abs(per - ref_per).max()
    assert res["LAS"] == pytest.approx(ref_per[:, 0].mean(), abs=1e-15)
    # decoded graphs, not only the scores
    k = 0
    for v, items in PE.make_batches(recs, 20):
        pred, _, n_active = ev.predict_batch(v, items, "fp32", feed)
        for g in range(len(items)):
            assert pred[g, 1:n_active[g]].tolist() == ref_graphs[k]
            k += 1
    # TF32 propagation: scores may differ only through near-ties; report, and bound the drift
    res32, _ = ev.evaluate(recs, batch_size=20, precision="tf32", feed=feed)
    assert abs(res32["UAS"] - res["UAS"]) <= 0.02 and abs(res32["LAS"] - res["LAS"]) <= 0.02
    # F16 propagation (h = 400: the wide FP16 kernels): the same bound, and what is actually true is recorded --
    # the decoded heads / labels of the 11-bit-operand modes may flip only where two soft-max scores nearly tie
    res16, per16 = ev.evaluate(recs, batch_size=20, precision="f16", feed=feed)
    assert abs(res16["UAS"] - res["UAS"]) <= 0.02 and abs(res16["LAS"] - res["LAS"]) <= 0.02
    flipped = int((np.abs(per16 - ref_per).max(axis=1) > 0).sum())
    print(f"f16 ({feed}): {flipped} of {len(ref_per)} sentences score differently from the oracle pipeline; "
          f"LAS {res16['LAS']:.4f} vs {res['LAS']:.4f}, UAS {res16['UAS']:.4f} vs {res['UAS']:.4f}")
    assert flipped <= len(ref_per) // 10


def test_front_end_and_heads_numerics():
    E, E_out, h = 6, 5, 400
    w = O.make_weights(2 * E, h, 1)
    rng = np.random.default_rng(0)
    g = lambda *s: O.glorot_init(rng, s)
    heads = PE.ParserHeads(g(40, 80), g(9, 50), g(50, 100), g(2 * h, 150), g(1, 150)[0], g(2 * h, E_out), g(1, E_out)[0])
    ev = PE.ParserEvaluator(ggnn_b200.from_dict(w), heads, "cuda:0", num_timesteps=2)
    recs = synth.wsj_like_corpus(7, E, E_out, pos_size=9, vocab=50, seed=1, max_len=30)
    v, items = next(PE.make_batches(recs, 20))
    idx, edges, n_active, _ = PE.batch_arrays(v, items, E)
    h0 = ev.embed(torch.from_numpy(idx).cuda())
    ref0 = P.initial_node_representation(idx, heads.loc_embeddings, heads.pos_embeddings, heads.word_embeddings, h)
    assert np.array_equal(h0.cpu().numpy(), ref0.astype(np.float32))         # a gather: bit-exact
    hT = torch.from_numpy(rng.standard_normal(ref0.shape).astype(np.float32)).cuda()
    pred, ph, pl = ev.decode(hT, h0, torch.from_numpy(n_active), probs=True)
    rph, rpl = P.heads(hT.cpu().numpy(), ref0, heads.head_W, heads.head_b, heads.label_W, heads.label_b)
    b = len(items)
    m = np.stack([P.get_mask(int(n), v, E_out, 150) for n in n_active]).reshape(b, v, 150)
    me = np.stack([P.get_mask(int(n), v, E_out, 150, is_edge=True) for n in n_active]).reshape(b, v, E_out)
    assert np.abs(ph.cpu().numpy() - rph.reshape(b, v, 150) * m).max() <= 1e-6
    assert np.abs(pl.cpu().numpy() - rpl.reshape(b, v, E_out) * me).max() <= 1e-6
    with pytest.raises(ValueError):                                         # index outside its table
        bad = idx.copy(); bad[0, 0, 1] = 99
        ev.embed(torch.from_numpy(bad).cuda())


@pytest.mark.parametrize("name", ["a", "b"])
def test_front_end_and_heads_against_the_reference_functions_own_output(golden_dir, name):
    """ggnn_embed_nodes / ggnn_heads_decode against tests/golden/ref_heads.npz: outputs of the reference's own
    get_initial_node_representation and gated_regression (btb branch, utils.MLP) run over NumPy-backed TF
    primitives by scripts/make_golden_heads.py."""
    d = np.load(os.path.join(golden_dir, "ref_heads.npz"))
    b, v, h, Osz, E_out = (int(x) for x in d[f"{name}_meta"])
    heads = PE.ParserHeads(d[f"{name}_loc_embeddings"], d[f"{name}_pos_embeddings"], d[f"{name}_word_embeddings"],
                           d[f"{name}_head_W"], d[f"{name}_head_b"], d[f"{name}_label_W"], d[f"{name}_label_b"])
    ev = PE.ParserEvaluator(ggnn_b200.from_dict(O.make_weights(4, h, 0)), heads, "cuda:0", num_timesteps=1)
    idx = np.ascontiguousarray(d[f"{name}_word_inputs"][..., :4])
    h0 = ev.embed(torch.from_numpy(idx).cuda())
    assert np.array_equal(h0.cpu().numpy(), d[f"{name}_h0_float32"])
    hT = torch.from_numpy(d[f"{name}_h_T"]).cuda()
    n_active = torch.full((b,), v, dtype=torch.int32)                      # every row active
    pred, ph, pl = ev.decode(hT, h0, n_active, probs=True)
    rph = d[f"{name}_p_head_float64"].reshape(b, v, Osz)
    rpl = d[f"{name}_p_label_float64"].reshape(b, v, E_out)
    # the product masks head positions >= n_active (get_mask, chem_tensorflow_dense.py:633-661)
    assert np.abs(ph.cpu().numpy()[..., :v] - rph[..., :v]).max() <= 1e-6
    assert np.abs(pl.cpu().numpy() - rpl).max() <= 1e-6
    p = pred.cpu().numpy()
    assert np.array_equal(p[:, 1:, 0], rph[:, 1:, :v].argmax(-1))
    assert np.array_equal(p[:, 1:, 1], rpl[:, 1:].argmax(-1))
