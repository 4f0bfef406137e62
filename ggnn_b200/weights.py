"""Weights of the propagation path, read from the reference's checkpoint format.

The reference pickles ``{"params", "weights", "train_step", "valid_step"}`` where
``weights`` maps TF variable names to ndarrays (chem_tensorflow.py:796-809).  The six
tensors of the path are created at chem_tensorflow_dense.py:202-212 (edge weights /
biases and their ``*_fixed`` twins, unnamed ``tf.Variable``s inside
``variable_scope("graph_model")``, chem_tensorflow.py:314) and :237-241 (GRUCell inside
``"gru_scope"``).  TensorFlow is not available to confirm the auto-generated names, so
entries are matched by name pattern AND shape, and ambiguity is an error.
"""
from __future__ import annotations

import pickle
import re
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

import numpy as np

PATH_KEYS = ("edge_weights", "edge_biases", "edge_weights_fixed", "edge_biases_fixed",
             "gru_gate_kernel", "gru_gate_bias", "gru_cand_kernel", "gru_cand_bias")

# names the reference's graph would give these variables
TF_NAMES = {
    "edge_weights": "graph_model/Variable:0",
    "edge_biases": "graph_model/Variable_1:0",
    "edge_weights_fixed": "graph_model/Variable_2:0",
    "edge_biases_fixed": "graph_model/Variable_3:0",
    "gru_gate_kernel": "graph_model/gru_scope/gru_cell/gates/kernel:0",
    "gru_gate_bias": "graph_model/gru_scope/gru_cell/gates/bias:0",
    "gru_cand_kernel": "graph_model/gru_scope/gru_cell/candidate/kernel:0",
    "gru_cand_bias": "graph_model/gru_scope/gru_cell/candidate/bias:0",
}


@dataclass
class GGNNWeights:
    """Host copy of the path's parameters (float32 ndarrays, reference layouts)."""
    edge_weights: np.ndarray                 # [2E, h, h]
    gru_gate_kernel: np.ndarray              # [2h, 2h]
    gru_gate_bias: np.ndarray                # [2h]
    gru_cand_kernel: np.ndarray              # [2h, h]
    gru_cand_bias: np.ndarray                # [h]
    edge_biases: Optional[np.ndarray] = None          # [2E, 1, h] (use_edge_bias)
    edge_weights_fixed: Optional[np.ndarray] = None
    edge_biases_fixed: Optional[np.ndarray] = None
    params: Dict[str, Any] = field(default_factory=dict)

    @property
    def two_e(self) -> int:
        return int(self.edge_weights.shape[0])

    @property
    def hidden_size(self) -> int:
        return int(self.edge_weights.shape[1])

    def validate(self) -> "GGNNWeights":
        c, h, h2 = self.edge_weights.shape
        if h != h2:
            raise ValueError(f"edge_weights must be [2E, h, h], got {self.edge_weights.shape}")
        def chk(name, arr, shape):
            if arr is not None and tuple(arr.shape) != shape:
                raise ValueError(f"{name} has shape {tuple(arr.shape)}, expected {shape}")
        chk("edge_biases", self.edge_biases, (c, 1, h))
        chk("edge_weights_fixed", self.edge_weights_fixed, (c, h, h))
        chk("edge_biases_fixed", self.edge_biases_fixed, (c, 1, h))
        chk("gru_gate_kernel", self.gru_gate_kernel, (2 * h, 2 * h))
        chk("gru_gate_bias", self.gru_gate_bias, (2 * h,))
        chk("gru_cand_kernel", self.gru_cand_kernel, (2 * h, h))
        chk("gru_cand_bias", self.gru_cand_bias, (h,))
        return self


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def from_dict(d: Dict[str, Any], params: Optional[Dict[str, Any]] = None) -> GGNNWeights:
    """Accepts either our own key names (PATH_KEYS) or a reference ``weights`` dict keyed
    by TF variable names."""
    if "edge_weights" in d:
        if d["edge_weights"] is None:
            raise ValueError("edge_weights is required")
        bias = d.get("edge_biases")
        if bias is not None and np.asarray(bias).ndim == 2:
            bias = np.asarray(bias)[:, None, :]
        bias_f = d.get("edge_biases_fixed")
        if bias_f is not None and np.asarray(bias_f).ndim == 2:
            bias_f = np.asarray(bias_f)[:, None, :]
        return GGNNWeights(
            edge_weights=_f32(d["edge_weights"]), edge_biases=_f32(bias),
            edge_weights_fixed=_f32(d.get("edge_weights_fixed")), edge_biases_fixed=_f32(bias_f),
            gru_gate_kernel=_f32(d["gru_gate_kernel"]), gru_gate_bias=_f32(d["gru_gate_bias"]),
            gru_cand_kernel=_f32(d["gru_cand_kernel"]), gru_cand_bias=_f32(d["gru_cand_bias"]),
            params=dict(params or {})).validate()
    return _from_tf_names(d, params)


def _from_tf_names(w: Dict[str, Any], params) -> GGNNWeights:
    def find_one(pattern, what):
        hits = [k for k in w if re.search(pattern, k)]
        if len(hits) != 1:
            raise ValueError(f"checkpoint: expected exactly one variable matching {pattern!r} "
                             f"for {what}, found {hits}")
        return np.asarray(w[hits[0]])

    gk = find_one(r"gru_cell/gates/kernel(:0)?$", "GRU gate kernel")
    gb = find_one(r"gru_cell/gates/bias(:0)?$", "GRU gate bias")
    ck = find_one(r"gru_cell/candidate/kernel(:0)?$", "GRU candidate kernel")
    cb = find_one(r"gru_cell/candidate/bias(:0)?$", "GRU candidate bias")
    h = ck.shape[1]
    if gk.shape != (2 * h, 2 * h):
        raise ValueError(f"GRU gate kernel shape {gk.shape} inconsistent with hidden size {h}")

    # unnamed tf.Variables in creation order: Variable, Variable_1, ...
    anon = []
    for k in w:
        m = re.fullmatch(r"graph_model/Variable(?:_(\d+))?(?::0)?", k)
        if m:
            anon.append((int(m.group(1) or 0), k))
    anon.sort()
    w3 = [np.asarray(w[k]) for _, k in anon
          if np.asarray(w[k]).ndim == 3 and np.asarray(w[k]).shape[1:] == (h, h)]
    b3 = [np.asarray(w[k]) for _, k in anon
          if np.asarray(w[k]).ndim == 3 and np.asarray(w[k]).shape[1:] == (1, h)]
    if len(w3) not in (1, 2):
        raise ValueError(f"checkpoint: expected 1 or 2 anonymous [2E,{h},{h}] variables "
                         f"(edge_weights[, edge_weights_fixed]), found {len(w3)}")
    if len(b3) not in (0, 1, 2) or (b3 and len(b3) != len(w3)):
        raise ValueError(f"checkpoint: {len(b3)} edge-bias tensors for {len(w3)} edge-weight tensors")
    if len(w3) == 2 and w3[0].shape != w3[1].shape:
        raise ValueError("checkpoint: edge_weights and edge_weights_fixed differ in shape")
    return GGNNWeights(
        edge_weights=_f32(w3[0]), edge_biases=_f32(b3[0]) if b3 else None,
        edge_weights_fixed=_f32(w3[1]) if len(w3) == 2 else None,
        edge_biases_fixed=_f32(b3[1]) if len(b3) == 2 else None,
        gru_gate_kernel=_f32(gk), gru_gate_bias=_f32(gb),
        gru_cand_kernel=_f32(ck), gru_cand_bias=_f32(cb),
        params=dict(params or {})).validate()


def load_checkpoint(path: str) -> GGNNWeights:
    """Read a reference ``*_model_best.pickle`` (chem_tensorflow.py:796-809)."""
    with open(path, "rb") as f:
        data = pickle.load(f)
    if not isinstance(data, dict) or "weights" not in data:
        raise ValueError(f"{path}: not a reference checkpoint (no 'weights' entry)")
    return from_dict(data["weights"], data.get("params"))


def save_checkpoint(path: str, w: GGNNWeights, params: Optional[Dict[str, Any]] = None,
                    extra: Optional[Dict[str, np.ndarray]] = None) -> None:
    """Write a pickle in the reference's format (used to make synthetic checkpoints)."""
    weights = {}
    for key, name in TF_NAMES.items():
        arr = getattr(w, key)
        if arr is not None:
            weights[name] = np.asarray(arr, dtype=np.float32)
    weights.update(extra or {})
    p = dict(w.params)
    p.update(params or {})
    p.setdefault("hidden_size", w.hidden_size)
    with open(path, "wb") as f:
        pickle.dump({"params": p, "weights": weights, "train_step": 0, "valid_step": 0},
                    f, pickle.HIGHEST_PROTOCOL)


def as_weights(obj) -> GGNNWeights:
    if isinstance(obj, GGNNWeights):
        return obj
    if isinstance(obj, str):
        return load_checkpoint(obj)
    if isinstance(obj, dict):
        if "weights" in obj and isinstance(obj["weights"], dict):
            return from_dict(obj["weights"], obj.get("params"))
        return from_dict(obj)
    raise ValueError(f"cannot interpret {type(obj).__name__} as GGNN weights")
