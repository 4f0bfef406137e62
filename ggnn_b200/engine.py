"""Host-side mirror of the reference's propagation interface, on top of the C ABI.

``Propagator`` owns one ``ggnn_ctx`` (one per device and host thread).  torch is used only
for device memory and streams; all arithmetic happens in libggnn_b200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple, Union

import numpy as np
import torch

from . import _lib
from .weights import GGNNWeights, as_weights

_PREC = {"fp32": _lib.PREC_FP32, "tf32": _lib.PREC_TF32, "f16": _lib.PREC_F16}


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _flags(save_tape: bool) -> int:
    return _lib.FLAG_SAVE_TAPE if save_tape else 0


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Propagator:
    """One ctx: packed weights + scratch on one CUDA device."""

    def __init__(self, weights: Union[GGNNWeights, dict, str], device: Union[int, str, torch.device, None] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("ggnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if dev.type != "cuda":
            raise ValueError(f"device must be a CUDA device, got {dev}")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self._l = _lib.lib()
        self._ctx = C.c_void_p()
        rc = self._l.ggnn_create(C.byref(self._ctx), dev.index)
        if rc != _lib.OK:
            self._ctx = C.c_void_p()
            raise RuntimeError(f"ggnn_create failed: {_lib.ERR_NAMES.get(rc, rc)} "
                               "(needs an sm_100 device)")
        self.weights: Optional[GGNNWeights] = None
        self.set_weights(weights)

    # -- lifetime ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._l.ggnn_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc == _lib.OK:
            return
        msg = self._l.ggnn_last_error(self._ctx)
        msg = msg.decode() if msg else ""
        name = _lib.ERR_NAMES.get(rc, str(rc))
        if rc in (_lib.E_ARG, _lib.E_SHAPE):
            raise ValueError(f"{what}: {name}: {msg}")
        raise RuntimeError(f"{what}: {name}: {msg}")

    @property
    def launches(self) -> int:
        return int(self._l.ggnn_launch_count(self._ctx))

    def profile(self, on: bool = True):
        """Turn the per-stage CUDA-event instrumentation on or off."""
        self._check(self._l.ggnn_profile_enable(self._ctx, int(on)), "ggnn_profile_enable")

    def profile_read(self):
        """{stage: (milliseconds, launches)} since the last read; synchronises the device."""
        n = len(_lib.STAGES)
        ms = (C.c_double * n)()
        ln = (C.c_int64 * n)()
        self._check(self._l.ggnn_profile_read(self._ctx, ms, ln), "ggnn_profile_read")
        return {name: (float(ms[i]), int(ln[i])) for i, name in enumerate(_lib.STAGES)}

    # -- parameters -------------------------------------------------------------------
    def set_weights(self, weights):
        w = as_weights(weights)
        dev = self.device

        def up(a):
            return None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)

        with torch.cuda.device(dev):
            st = _stream_ptr(dev)
            keep = [up(w.edge_weights), up(w.edge_biases), up(w.gru_gate_kernel), up(w.gru_gate_bias),
                    up(w.gru_cand_kernel), up(w.gru_cand_bias)]
            rc = self._l.ggnn_set_weights(self._ctx, _lib.SLOT_MAIN, w.two_e, w.hidden_size,
                                          *[_ptr(t) for t in keep], st)
            self._check(rc, "ggnn_set_weights")
            if w.edge_weights_fixed is not None:
                kf = [up(w.edge_weights_fixed), up(w.edge_biases_fixed)]
                rc = self._l.ggnn_set_weights(self._ctx, _lib.SLOT_FIXED, w.two_e, w.hidden_size,
                                              _ptr(kf[0]), _ptr(kf[1]), None, None, None, None, st)
                self._check(rc, "ggnn_set_weights(fixed)")
                keep += kf
            torch.cuda.current_stream(dev).synchronize()   # staging tensors may now be freed
        self.weights = w

    # -- A1 ---------------------------------------------------------------------------
    def compact_adjacency(self, adj: torch.Tensor, col_capacity: Optional[int] = None
                          ) -> Tuple[torch.Tensor, torch.Tensor, int]:
        """Dense [b, 2E, v, v] (float32 or uint8, CUDA) -> (rowptr int32 [b*2E*v+1],
        col int32 [nnz], nnz).  Synchronises to return nnz and trim ``col``."""
        adj, dt = self._adj(adj)
        b, c, v, _ = adj.shape
        rows = b * c * v
        with torch.cuda.device(self.device):
            rowptr = torch.empty(rows + 1, dtype=torch.int32, device=self.device)
            nnz_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
            if col_capacity is None:
                # two calls: count first (capacity 0), then fill
                rc = self._l.ggnn_compact_adjacency(self._ctx, _ptr(adj), dt, b, c, v, _ptr(rowptr),
                                                    None, 0, _ptr(nnz_dev), _stream_ptr(self.device))
                self._check(rc, "ggnn_compact_adjacency")
                col_capacity = int(nnz_dev.item())
            col = torch.empty(max(col_capacity, 1), dtype=torch.int32, device=self.device)
            rc = self._l.ggnn_compact_adjacency(self._ctx, _ptr(adj), dt, b, c, v, _ptr(rowptr),
                                                _ptr(col), col_capacity, _ptr(nnz_dev),
                                                _stream_ptr(self.device))
            self._check(rc, "ggnn_compact_adjacency")
            nnz = int(nnz_dev.item())
        if nnz > col_capacity:
            raise RuntimeError(f"compact_adjacency: GGNN_E_CAPACITY: {nnz} edges > capacity {col_capacity}")
        return rowptr, col[:nnz], nnz

    # -- A2 ---------------------------------------------------------------------------
    def _adj(self, adj: torch.Tensor):
        if adj.dim() != 4 or adj.shape[2] != adj.shape[3]:
            raise ValueError(f"adjacency_matrix must be [b, 2E, v, v], got {tuple(adj.shape)}")
        if adj.dtype == torch.float32:
            dt = _lib.ADJ_F32
        elif adj.dtype == torch.uint8:
            dt = _lib.ADJ_U8
        else:
            raise ValueError(f"adjacency_matrix must be float32 or uint8, got {adj.dtype}")
        if adj.device != self.device:
            raise ValueError(f"adjacency_matrix is on {adj.device}, ctx is on {self.device}")
        return adj.contiguous(), dt

    def _slot_T(self, num_timesteps, fixed_ts):
        if fixed_ts is not None:
            if self.weights.edge_weights_fixed is None:
                raise ValueError("fixed_ts given but the weights hold no edge_weights_fixed")
            return _lib.SLOT_FIXED, int(fixed_ts)
        if num_timesteps is None:
            num_timesteps = self.weights.params.get("num_timesteps")
            if num_timesteps is None:
                raise ValueError("num_timesteps not given and not in the checkpoint params")
        return _lib.SLOT_MAIN, int(num_timesteps)

    def _state(self, h0: torch.Tensor, nd: int):
        if h0.dim() != nd or h0.shape[-1] != self.weights.hidden_size:
            raise ValueError(f"node states must be {nd}-D with last dim {self.weights.hidden_size}, "
                             f"got {tuple(h0.shape)}")
        if h0.dtype != torch.float32:
            raise ValueError(f"node states must be float32, got {h0.dtype}")
        if h0.device != self.device:
            raise ValueError(f"node states are on {h0.device}, ctx is on {self.device}")
        return h0.contiguous()

    def propagate_dense(self, h0: torch.Tensor, adj: torch.Tensor, num_timesteps=None, fixed_ts=None,
                        precision: str = "fp32", out: Optional[torch.Tensor] = None,
                        save_tape: bool = False, validate: bool = False) -> torch.Tensor:
        """``validate=True`` checks that every adjacency value is exactly 0 or 1 (one extra pass over the tensor):
        the path treats the adjacency as an edge SET (`A != 0`), which equals the reference's
        ``tf.matmul(adj, m)`` (chem_tensorflow_dense.py:428) only for the 0/1 matrices graph_to_adj_mat_bd
        builds; a weighted or normalised adjacency would give different numbers and is refused."""
        if precision not in _PREC:
            raise ValueError(f"precision must be one of {sorted(_PREC)}, got {precision!r}")
        h0 = self._state(h0, 3)
        adj, dt = self._adj(adj)
        if validate and adj.numel() and not bool(((adj == 0) | (adj == 1)).all()):
            raise ValueError("adjacency_matrix holds values other than 0 and 1: the path implements the reference's "
                             "0/1 adjacency (graph_to_adj_mat_bd); weighted adjacencies are not supported")
        b, v, h = h0.shape
        if adj.shape[0] != b or adj.shape[2] != v or adj.shape[1] != self.weights.two_e:
            raise ValueError(f"adjacency {tuple(adj.shape)} does not match states {tuple(h0.shape)} "
                             f"and 2E={self.weights.two_e}")
        slot, T = self._slot_T(num_timesteps, fixed_ts)
        if out is None:
            out = torch.empty_like(h0)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_propagate_dense(self._ctx, slot, _ptr(h0), _ptr(adj), dt, None, b, v, T,
                                              _PREC[precision], _flags(save_tape), _ptr(out),
                                              _stream_ptr(self.device))
        self._check(rc, "ggnn_propagate_dense")
        return out

    def propagate_dense_bits(self, h0: torch.Tensor, bits: torch.Tensor, num_timesteps=None, fixed_ts=None,
                             precision: str = "fp32", out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``propagate_dense`` with the adjacency given as its bitmask (int32 device tensor of ceil(b*2E*v*v / 32)
        words as ``pack_adjacency_host`` builds it on the host cores): what a host-resident dense adjacency is
        reduced to before it crosses the bus (include/ggnn_b200.h, ggnn_propagate_dense_bits)."""
        if precision not in _PREC:
            raise ValueError(f"precision must be one of {sorted(_PREC)}, got {precision!r}")
        h0 = self._state(h0, 3)
        b, v, h = h0.shape
        words = (b * self.weights.two_e * v * v + 31) // 32
        if bits.dtype != torch.int32 or bits.device != self.device or bits.numel() < words or not bits.is_contiguous():
            raise ValueError(f"bits must be a contiguous int32 tensor of >= {words} words on {self.device}")
        slot, T = self._slot_T(num_timesteps, fixed_ts)
        if out is None:
            out = torch.empty_like(h0)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_propagate_dense_bits(self._ctx, slot, _ptr(h0), _ptr(bits), b, v, T, _PREC[precision],
                                                   _flags(False), _ptr(out), _stream_ptr(self.device))
        self._check(rc, "ggnn_propagate_dense_bits")
        return out

    def prefetch_dense(self, adj: torch.Tensor, ready_event: Optional[torch.cuda.Event] = None):
        """Announce the adjacency of the NEXT ``propagate_dense`` call: its dense -> bitmask pass starts now in the
        background of the work already queued (the reference's ThreadedIterator prepares the next batch the same
        way, utils.py:17-37).  ``adj`` must be final (or ``ready_event`` marks it) and must be the very tensor the
        next call receives."""
        adj, dt = self._adj(adj)
        b, c, v, _ = adj.shape
        if c != self.weights.two_e:
            raise ValueError(f"adjacency has {c} channels, the weights have 2E={self.weights.two_e}")
        ev = C.c_void_p(ready_event.cuda_event if ready_event is not None else 0)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_dense_prefetch(self._ctx, _ptr(adj), dt, b, v, ev)
        self._check(rc, "ggnn_dense_prefetch")

    def propagate_csr(self, h0, rowptr, col, num_timesteps=None, fixed_ts=None, precision="fp32", out=None,
                      save_tape=False):
        if precision not in _PREC:
            raise ValueError(f"precision must be one of {sorted(_PREC)}, got {precision!r}")
        h0 = self._state(h0, 3)
        b, v, h = h0.shape
        if rowptr.dtype != torch.int32 or col.dtype != torch.int32:
            raise ValueError("rowptr and col must be int32")
        if rowptr.numel() != b * self.weights.two_e * v + 1:
            raise ValueError(f"rowptr has {rowptr.numel()} entries, expected b*2E*v+1 = "
                             f"{b * self.weights.two_e * v + 1}")
        slot, T = self._slot_T(num_timesteps, fixed_ts)
        if out is None:
            out = torch.empty_like(h0)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_propagate_csr(self._ctx, slot, _ptr(h0), _ptr(rowptr.contiguous()),
                                            _ptr(col.contiguous()), b, v, T, _PREC[precision],
                                            _flags(save_tape), _ptr(out), _stream_ptr(self.device))
        self._check(rc, "ggnn_propagate_csr")
        return out

    def propagate_edges(self, h0, edges, num_timesteps=None, fixed_ts=None, precision="fp32", out=None,
                        save_tape=False):
        if precision not in _PREC:
            raise ValueError(f"precision must be one of {sorted(_PREC)}, got {precision!r}")
        h0 = self._state(h0, 2)
        if edges.dim() != 2 or edges.shape[1] != 3 or edges.dtype != torch.int32:
            raise ValueError(f"edges must be int32 [nnz, 3] (src, channel, dst), got "
                             f"{edges.dtype} {tuple(edges.shape)}")
        if edges.device != self.device:
            raise ValueError(f"edges are on {edges.device}, ctx is on {self.device}")
        slot, T = self._slot_T(num_timesteps, fixed_ts)
        if out is None:
            out = torch.empty_like(h0)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_propagate_edges(self._ctx, slot, _ptr(h0), _ptr(edges.contiguous()),
                                              edges.shape[0], h0.shape[0], T, _PREC[precision],
                                              _flags(save_tape), _ptr(out), _stream_ptr(self.device))
        self._check(rc, "ggnn_propagate_edges")
        return out

    # -- config 5: backward ---------------------------------------------------------------
    @property
    def grad_floats(self) -> int:
        return int(self._l.ggnn_grad_floats(self._ctx))

    def set_dropout(self, w_mask: Optional[torch.Tensor], state_mask: Optional[torch.Tensor], keep: float,
                    num_timesteps: int):
        """Supplied-mask dropout for the next ``save_tape=True`` propagate call and its backward
        (reference: tf.nn.dropout on the edge weights per unrolled step, DropoutWrapper on the new
        state).  w_mask float32 [T, 2E, h, h], state_mask float32 [T, N, h] (or [T, b, v, h]), 0/1."""
        for m in (w_mask, state_mask):
            if m is not None and (m.dtype != torch.float32 or m.device != self.device or not m.is_contiguous()):
                raise ValueError("dropout masks must be contiguous float32 tensors on the ctx device")
        w = self.weights
        if w_mask is not None and tuple(w_mask.shape) != (num_timesteps, w.two_e, w.hidden_size, w.hidden_size):
            raise ValueError(f"w_mask must be [T, 2E, h, h], got {tuple(w_mask.shape)}")
        if state_mask is not None and (state_mask.shape[0] != num_timesteps or state_mask.shape[-1] != w.hidden_size):
            raise ValueError(f"state_mask must be [T, N, h], got {tuple(state_mask.shape)}")
        self._drop_keep = (w_mask, state_mask)          # keep the tensors alive until the backward
        rc = self._l.ggnn_set_dropout(self._ctx, _ptr(w_mask), _ptr(state_mask), float(keep), int(num_timesteps))
        self._check(rc, "ggnn_set_dropout")

    def set_dropout_seed(self, keep: float, seed: int, drop_weights: bool = True, drop_state: bool = True):
        """Device-generated dropout (Philox4x32-10) for the next ``save_tape=True`` propagate call and its
        backward: fresh edge-weight mask per unrolled step and a mask on every new state, keep-prob ``keep``
        (reference: chem_tensorflow_dense.py:397-403, 239-240; 0.9 in training, :800)."""
        self._drop_keep = None
        rc = self._l.ggnn_set_dropout_seed(self._ctx, float(keep), int(seed) & (2 ** 64 - 1), int(drop_weights),
                                           int(drop_state))
        self._check(rc, "ggnn_set_dropout_seed")

    def dropout_mask(self, n: int, seed: int, stream_id: int, keep: float) -> torch.Tensor:
        """The 0/1 mask (float32 [n]) the kernels use for (seed, stream_id): stream 2t = edge weights of step
        t (flattened [2E, h, h]), 2t+1 = state of step t (flattened [N, h])."""
        out = torch.empty(int(n), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_dropout_mask(self._ctx, _ptr(out), int(n), int(seed) & (2 ** 64 - 1), int(stream_id),
                                           float(keep), _stream_ptr(self.device))
        self._check(rc, "ggnn_dropout_mask")
        return out

    def backward(self, d_hT: torch.Tensor, grads: Optional[torch.Tensor] = None
                 ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Backward of the last ``save_tape=True`` propagate call.  ``d_hT``: gradient w.r.t. the
        returned node representations (same shape).  Returns (d_h0, flat gradient buffer); the
        weight gradients are ACCUMULATED into ``grads`` when given (float32 [grad_floats])."""
        if d_hT.dtype != torch.float32 or d_hT.device != self.device:
            raise ValueError("d_hT must be a float32 tensor on the ctx device")
        d_hT = d_hT.contiguous()
        if grads is None:
            grads = torch.zeros(self.grad_floats, dtype=torch.float32, device=self.device)
        elif grads.dtype != torch.float32 or grads.numel() != self.grad_floats or not grads.is_contiguous():
            raise ValueError(f"grads must be contiguous float32 [{self.grad_floats}]")
        d_h0 = torch.empty_like(d_hT)
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_propagate_backward(self._ctx, _ptr(d_hT), _ptr(d_h0), _ptr(grads),
                                                 _stream_ptr(self.device))
        self._check(rc, "ggnn_propagate_backward")
        return d_h0, grads

    def optimizer_step(self, grads: torch.Tensor, lr: float = 0.003, beta1: float = 0.9, beta2: float = 0.999,
                       eps: float = 1e-8, clip_norm: float = 1.0):
        """Per-variable clip_by_norm + Adam on the ctx-owned weights (reference make_train_step,
        chem_tensorflow.py:483-506; defaults: learning_rate .003 `:91`, clamp_gradient_norm 1.0)."""
        if grads.dtype != torch.float32 or grads.numel() != self.grad_floats or not grads.is_contiguous():
            raise ValueError(f"grads must be contiguous float32 [{self.grad_floats}]")
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_optimizer_step(self._ctx, _lib.SLOT_MAIN, _ptr(grads), lr, beta1, beta2, eps,
                                             clip_norm, _stream_ptr(self.device))
        self._check(rc, "ggnn_optimizer_step")

    def get_weights(self) -> dict:
        """Current FP32 master weights of slot 0 as device tensors keyed like GGNNWeights fields."""
        w = self.weights
        e2, h = w.two_e, w.hidden_size
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.device)
        out = {"edge_weights": f(e2, h, h), "edge_biases": f(e2, 1, h) if w.edge_biases is not None else None,
               "gru_gate_kernel": f(2 * h, 2 * h), "gru_gate_bias": f(2 * h), "gru_cand_kernel": f(2 * h, h),
               "gru_cand_bias": f(h)}
        with torch.cuda.device(self.device):
            rc = self._l.ggnn_get_weights(self._ctx, _lib.SLOT_MAIN, *[_ptr(out[k]) for k in
                                          ("edge_weights", "edge_biases", "gru_gate_kernel", "gru_gate_bias",
                                           "gru_cand_kernel", "gru_cand_bias")], _stream_ptr(self.device))
        self._check(rc, "ggnn_get_weights")
        return out

    def split_grads(self, flat: torch.Tensor) -> dict:
        """Views of the flat gradient buffer, keyed like GGNNWeights fields."""
        w = self.weights
        e2, h = w.two_e, w.hidden_size
        sizes = [("edge_weights", (e2, h, h)), ("edge_biases", (e2, h)), ("gru_gate_kernel", (2 * h, 2 * h)),
                 ("gru_gate_bias", (2 * h,)), ("gru_cand_kernel", (2 * h, h)), ("gru_cand_bias", (h,))]
        out, off = {}, 0
        for name, shp in sizes:
            n = int(np.prod(shp))
            out[name] = flat[off:off + n].view(*shp)
            off += n
        assert off == flat.numel()
        return out
