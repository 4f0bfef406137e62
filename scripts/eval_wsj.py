#!/usr/bin/env python
"""BASELINE config 2: LAS/UAS of a restored model on the WSJ Stanford->Matsumoto btb test set,
forward propagation on one B200 -- the evaluation branch of the reference
(`chem_tensorflow_dense.py --pr btb --evaluate --restore F.pickle`, dense :1330-1340, 978-986).

    python scripts/eval_wsj.py --data en-wsj-std-test-stanford-3.3.0-tagged_btb.json --restore best.pickle
    python scripts/eval_wsj.py            # synthetic WSJ-shaped twin

The WSJ files are LDC-licensed and the reference ships no checkpoint: without
--data/--restore this prints "WSJ data / pickle not present" and runs the same pipeline on 2 416
synthetic sentences with a synthetic checkpoint written in the reference's pickle format.  With
--check the oracle pipeline (CPU, slow) is run beside it and per-sentence scores are compared.
"""
import argparse
import json
import os
import pickle
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ggnn_b200
from ggnn_b200 import parser_eval as PE, synth


def synthetic(tmp, n, h, E, E_out, seed=0):
    rng = np.random.default_rng(seed)
    r = lambda *s: (np.sqrt(6.0 / (s[-2] + s[-1])) * rng.uniform(-1, 1, s)).astype(np.float32)
    w = ggnn_b200.GGNNWeights(edge_weights=r(2 * E, h, h), edge_biases=np.zeros((2 * E, 1, h), np.float32),
                              gru_gate_kernel=r(2 * h, 2 * h), gru_gate_bias=np.ones(2 * h, np.float32),
                              gru_cand_kernel=r(2 * h, h), gru_cand_bias=np.zeros(h, np.float32))
    heads = {"loc_embeddings": r(150, 80), "pos_embeddings": r(48, 50), "word_embeddings": r(20000, 100),
             "head_W": r(2 * h, 150), "head_b": np.zeros(150, np.float32),
             "label_W": r(2 * h, E_out), "label_b": np.zeros(E_out, np.float32)}
    path = os.path.join(tmp, "synthetic_model_best.pickle")
    ggnn_b200.save_checkpoint(path, w, params={"num_timesteps": 4, "hidden_size": h},
                              extra={PE.HEAD_SAVE_NAMES[k]: a for k, a in heads.items()})
    recs = synth.wsj_like_corpus(n, E, E_out, pos_size=48, vocab=20000, seed=seed)
    return path, recs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--data"); ap.add_argument("--restore")
    ap.add_argument("--sentences", type=int, default=2416)
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32", "f16"])
    ap.add_argument("--feed", default="edges", choices=["edges", "dense"])
    ap.add_argument("--check", type=int, default=0, help="also run the CPU oracle on this many sentences")
    ap.add_argument("--batch", type=int, default=20, help="sentences per batch within a bucket (reference: 20)")
    a = ap.parse_args()
    tmp = tempfile.mkdtemp()
    if a.data and a.restore and os.path.exists(a.data) and os.path.exists(a.restore):
        recs, path, kind = json.load(open(a.data)), a.restore, "WSJ"
    else:
        print("WSJ data / pickle not present -- running the synthetic WSJ-shaped twin (SURVEY.md F6)")
        path, recs = synthetic(tmp, a.sentences, 400, 46, 45)
        kind = "synthetic"
    data = pickle.load(open(path, "rb"))
    ev = PE.ParserEvaluator(ggnn_b200.load_checkpoint(path), PE.heads_from_checkpoint(data["weights"]), "cuda:0")
    ev.evaluate(recs[:64], precision=a.precision, feed=a.feed)          # warm-up
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res, per = ev.evaluate(recs, batch_size=a.batch, precision=a.precision, feed=a.feed)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    out = {"data": kind, "precision": a.precision, "feed": a.feed, "batch": a.batch, **res, "seconds": dt,
           "instances_per_s": res["sentences"] / dt}
    if a.batch != 20:      # batching must not change any sentence's scores
        _, per20 = ev.evaluate(recs, batch_size=20, precision=a.precision, feed=a.feed)
        out["per_sentence_scores_equal_to_batch_20"] = bool(np.array_equal(np.sort(per20, 0), np.sort(per, 0)))
    if a.check:
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
        from test_gpu_parser import _oracle_scores
        from ggnn_b200.weights import PATH_KEYS
        w = ggnn_b200.load_checkpoint(path)
        wd = {k: getattr(w, k) for k in PATH_KEYS if getattr(w, k) is not None}
        ref, _ = _oracle_scores(recs[:a.check], wd, ev.heads, w.two_e // 2, ev.n_lab, w.hidden_size, ev.T, 20)
        _, got = ev.evaluate(recs[:a.check], precision=a.precision, feed=a.feed)
        out["oracle_sentences"] = int(len(ref))
        out["per_sentence_scores_identical"] = bool(np.array_equal(ref, got))
        out["oracle_LAS_UAS"] = [float(ref[:, 0].mean()), float(ref[:, 1].mean())]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
