"""CPU tests of the host side: checkpoint reader, C-ABI exports, loud failure without CUDA."""
import ctypes
import os
import pickle
import re

import numpy as np
import pytest
import torch

import ggnn_b200
from ggnn_b200 import _lib, weights as W, synth
from oracle import ggnn_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mk(two_e=6, h=8, seed=0, fixed=True):
    w = O.make_weights(two_e, h, seed)
    if fixed:
        w2 = O.make_weights(two_e, h, seed + 100)
        w["edge_weights_fixed"], w["edge_biases_fixed"] = w2["edge_weights"], w2["edge_biases"]
    return W.from_dict(w, {"num_timesteps": 4, "hidden_size": h})


def test_library_exports_every_declared_symbol():
    """include/ggnn_b200.h is the contract: every function it declares must be exported."""
    hdr = open(os.path.join(ROOT, "include", "ggnn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ggnn_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = _lib.lib()                                  # builds must exist in-tree
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), name
    assert lib.ggnn_version() >= 100


def test_sass_is_sm100a_only():
    import subprocess
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", _lib.LIB_PATH],
                         capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_cuda():
    w = _mk()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ggnn_b200.compute_final_node_representations(
            np.zeros((1, 4, 8), np.float32), np.zeros((1, 6, 4, 4), np.float32), weights=w, num_timesteps=1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ggnn_b200.Propagator(w)
    p = ctypes.c_void_p()
    assert _lib.lib().ggnn_create(ctypes.byref(p), 0) == _lib.E_CUDA
    assert not p.value


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "ggnn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("FP64 oracle", "").replace("the oracle", ""), f


def test_checkpoint_roundtrip_reference_format(tmp_path):
    w = _mk()
    path = str(tmp_path / "model_best.pickle")
    extra = {"graph_model/Variable_4:0": np.zeros((16, 16), np.float32),          # att_weights (unused)
             "graph_model/loc_embeddings:0": np.zeros((150, 80), np.float32),
             "beta1_power:0": np.float32(0.9)}
    W.save_checkpoint(path, w, {"num_timesteps": 4}, extra)
    raw = pickle.load(open(path, "rb"))
    assert set(raw) == {"params", "weights", "train_step", "valid_step"}          # chem_tensorflow.py:802-807
    assert "graph_model/gru_scope/gru_cell/gates/kernel:0" in raw["weights"]
    back = W.load_checkpoint(path)
    for k in W.PATH_KEYS:
        assert np.array_equal(getattr(back, k), getattr(w, k)), k
    assert back.params["num_timesteps"] == 4
    assert back.two_e == 6 and back.hidden_size == 8


def test_checkpoint_without_bias_or_fixed():
    w = O.make_weights(6, 8, 0)
    names = {W.TF_NAMES[k]: w[k] for k in ("edge_weights", "gru_gate_kernel", "gru_gate_bias",
                                           "gru_cand_kernel", "gru_cand_bias")}
    g = W.from_dict(names)
    assert g.edge_biases is None and g.edge_weights_fixed is None


def test_checkpoint_ambiguity_is_an_error():
    w = O.make_weights(6, 8, 0)
    names = {W.TF_NAMES[k]: w[k] for k in w}
    bad = dict(names)
    bad["other/gru_cell/gates/kernel:0"] = w["gru_gate_kernel"]
    with pytest.raises(ValueError, match="exactly one"):
        W.from_dict(bad)
    bad = dict(names)
    del bad[W.TF_NAMES["gru_cand_bias"]]
    with pytest.raises(ValueError, match="exactly one"):
        W.from_dict(bad)
    bad = dict(names)
    bad["graph_model/Variable_5:0"] = w["edge_weights"]
    bad["graph_model/Variable_6:0"] = w["edge_weights"]
    with pytest.raises(ValueError, match="1 or 2 anonymous"):
        W.from_dict(bad)
    bad = dict(names)
    bad[W.TF_NAMES["gru_gate_kernel"]] = np.zeros((8, 8), np.float32)
    with pytest.raises(ValueError, match="inconsistent"):
        W.from_dict(bad)


def test_weights_shape_validation():
    w = O.make_weights(6, 8, 0)
    w["edge_biases"] = np.zeros((5, 1, 8), np.float32)
    with pytest.raises(ValueError, match="edge_biases"):
        W.from_dict(w)


def test_synth_matches_reference_rules():
    E, v = 6, 16
    tri = synth.random_trees(5, 15, E, seed=3, lengths=torch.tensor([15, 3, 1, 9, 12]))
    adj = synth.dense_adjacency(tri, v, E).numpy()
    for g in range(5):
        t = [r for r in tri[g].tolist() if r[2] >= 0]
        assert sorted(r[2] for r in t) == list(range(1, len(t) + 1))
        assert np.array_equal(O.graph_to_adj_mat_bd(t, v, E), adj[g])
        assert int((adj[g] != 0).sum()) == 4 * len(t)
    el = synth.global_edge_list(tri, v, E).numpy()
    de = O.dense_to_edges(adj)
    assert sorted(map(tuple, el)) == sorted(map(tuple, de))
    assert synth.bucket_size(5) == 6 and synth.bucket_size(4) == 6 and synth.bucket_size(3) == 4


def test_host_adjacency_packer_is_bit_exact():
    """ggnn_host_pack_adjacency is a host function (no GPU): word w, bit i <=> element 32 w + i != 0, for FP32 (with
    -0.0, NaN and a denormal-range value) and uint8 inputs, any thread count, with a partial last word."""
    import ctypes as C
    from ggnn_b200 import _lib
    l = _lib.lib()
    rng = np.random.default_rng(0)
    n = 3 * 10 * 33 * 33 + 5
    a = (rng.random(n) < 0.03).astype(np.float32)
    a[3], a[5], a[9] = -0.0, np.nan, 1e-38
    ref = np.packbits(np.pad(a != 0, (0, (-n) % 32)).astype(np.uint8), bitorder="little").view(np.uint32)
    for arr, code in ((a, _lib.ADJ_F32), ((a != 0).astype(np.uint8), _lib.ADJ_U8)):
        for threads in (1, 2, 0):
            bits = np.full((n + 31) // 32, 0xFFFFFFFF, dtype=np.uint32)
            assert l.ggnn_host_pack_adjacency(C.c_void_p(arr.ctypes.data), code, n, C.c_void_p(bits.ctypes.data), threads) == _lib.OK
            assert np.array_equal(bits, ref)
    assert l.ggnn_host_pack_adjacency(None, _lib.ADJ_F32, 64, None, 0) == _lib.E_ARG
    assert l.ggnn_host_pack_adjacency(C.c_void_p(a.ctypes.data), 7, 64, C.c_void_p(ref.ctypes.data), 0) == _lib.E_ARG
