#!/usr/bin/env python
"""Golden vectors for the rows either side of the path (SURVEY.md 8(f) N1/N2 host logic), made by
running the REFERENCE's own NumPy code: masks (get_mask), target formatting
(target_to_adj_mat, get_target_values_formatted, get_target_values_edges_formatted), word-input
padding, and the decode + LAS/UAS scorer (humanize_batch_results_btb -> adj_mat_to_target,
merge_head_and_edge_graph, get_las_uas) on random head / label probability tensors.

Runs only in the authoring container (needs /root/reference); TensorFlow and docopt are stubbed
exactly as in make_golden.py.  Output: tests/golden/parser_decode.npz.

    python scripts/make_golden_parser.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import import_reference, OUT  # noqa: E402


def softmax(x):
    e = np.exp(x - x.max(-1, keepdims=True))
    return e / e.sum(-1, keepdims=True)


def main():
    dense, _ = import_reference()
    m = dense.DenseGGNNChemModel({"dummy": True, "--pr": "btb"})
    E_out, O = 7, 150
    m.params = {"output_size": O, "tie_fwd_bkwd": True}
    m.output_size_edges = E_out
    m.num_edge_types = E_out + 1
    rng = np.random.default_rng(42)
    out = {"E_out": E_out, "O": O}
    for ci, (b, v) in enumerate([(5, 8), (4, 14), (3, 6)]):
        labels, masks, masks_e, n_active, targets_all = [], [], [], [], []
        for g in range(b):
            n = int(rng.integers(1, v))                       # words; nodes n+1 <= v
            tgt = [[int(rng.integers(0, n + 1)), int(rng.integers(1, E_out + 1))] for _ in range(n)]
            labels.append(dense.target_to_adj_mat(target=tgt, max_n_vertices=v, num_edge_types=E_out,
                                                  output_size=v))
            masks.append(m.get_mask(n_active_nodes=n + 1, chosen_bucket_size=v))
            masks_e.append(m.get_mask(n_active_nodes=n + 1, chosen_bucket_size=v, is_edge=True))
            n_active.append(n + 1)
            t = np.full((v, 2), -1, np.int32)
            t[:n] = np.array(tgt, np.int32)
            targets_all.append(t)
        tv = m.get_target_values_formatted(labels=labels)                 # [b, v*O]
        tve = m.get_target_values_edges_formatted(labels=labels)          # [b, v*E_out]
        # probabilities as the btb heads produce them: softmax over O head positions / E_out labels
        ph = softmax(rng.standard_normal((b, v, O)) * 2).astype(np.float32)
        pe = softmax(rng.standard_normal((b, v, E_out)) * 2).astype(np.float32)
        # exact ties: the reference takes the FIRST maximum in (e, src) order
        ph[0, 1, 0] = ph[0, 1, 1] = ph[0, 1].max() + 0.1
        pe[0, 1, 2] = pe[0, 1, 4] = pe[0, 1].max() + 0.1
        # make roughly half of the predictions correct so the scores are not trivially 0
        for g in range(b):
            for i in range(n_active[g] - 1):
                if rng.random() < 0.5:
                    ph[g, i + 1, targets_all[g][i, 0]] = 2.0
                if rng.random() < 0.5:
                    pe[g, i + 1, targets_all[g][i, 1] - 1] = 2.0
        adms = [np.zeros((2 * (E_out + 1), v, v)) for _ in range(b)]
        las, uas, uas_e = m.humanize_batch_results_btb(
            labels=tv, computed_values=ph.reshape(b, v * O), num_vertices=v, mask=np.array(masks),
            ids=list(range(b)), adms=adms, labels_e=tve, computed_values_e=pe.reshape(b, v * E_out),
            mask_edges=np.array(masks_e), out_file=None)
        out.update({f"c{ci}_v": v, f"c{ci}_n_active": np.array(n_active, np.int32),
                    f"c{ci}_targets": np.stack(targets_all), f"c{ci}_mask": np.array(masks, np.float32),
                    f"c{ci}_mask_edges": np.array(masks_e, np.float32),
                    f"c{ci}_target_values": tv.astype(np.float32), f"c{ci}_target_values_edges": tve.astype(np.float32),
                    f"c{ci}_ph": ph, f"c{ci}_pe": pe, f"c{ci}_scores": np.array([las, uas, uas_e], np.float64)})
        print(f"case {ci}: b={b} v={v} LAS={las:.4f} UAS={uas:.4f} UAS_e={uas_e:.4f}")
    # word-input padding (get_word_inputs_padded, chem_tensorflow_dense.py:877-882)
    wi = m.get_word_inputs_padded(words_pos=[[3, 1, 2], [5]], b=2, v=6)
    out["word_inputs_padded"] = wi
    np.savez_compressed(os.path.join(OUT, "parser_decode.npz"), **out)
    print("wrote parser_decode.npz")


if __name__ == "__main__":
    main()
