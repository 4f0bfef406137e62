"""ctypes binding of libggnn_b200.so (include/ggnn_b200.h).

There is no CPU or eager fallback: if the shared library is missing or does not load,
importing this module's ``lib()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libggnn_b200.so")

OK, E_ARG, E_SHAPE, E_CAPACITY, E_CUDA, E_UNSUPPORTED, E_STATE = 0, -1, -2, -3, -4, -5, -6
ADJ_F32, ADJ_U8 = 0, 1
PREC_FP32, PREC_TF32, PREC_F16 = 0, 1, 2
SLOT_MAIN, SLOT_FIXED = 0, 1
FLAG_SAVE_TAPE = 2
STAGES = ("compact_mask", "compact_rest", "plan", "msg", "agg", "gru_gates", "gru_cand", "other")

ERR_NAMES = {E_ARG: "GGNN_E_ARG", E_SHAPE: "GGNN_E_SHAPE", E_CAPACITY: "GGNN_E_CAPACITY",
             E_CUDA: "GGNN_E_CUDA", E_UNSUPPORTED: "GGNN_E_UNSUPPORTED", E_STATE: "GGNN_E_STATE"}

# every symbol include/ggnn_b200.h declares: name -> (restype, argtypes)
_vp, _i, _i64, _u = C.c_void_p, C.c_int, C.c_int64, C.c_uint
SIGNATURES = {
    "ggnn_create": (_i, [C.POINTER(_vp), _i]),
    "ggnn_destroy": (None, [_vp]),
    "ggnn_last_error": (C.c_char_p, [_vp]),
    "ggnn_version": (_i, []),
    "ggnn_set_weights": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ggnn_compact_adjacency": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _i64, _vp, _vp]),
    "ggnn_compact_adjacency_static": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _i64, _vp, _vp]),
    "ggnn_propagate_dense": (_i, [_vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _u, _vp, _vp]),
    "ggnn_dense_prefetch": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "ggnn_host_pack_adjacency": (_i, [_vp, _i, _i64, _vp, _i]),
    "ggnn_propagate_dense_bits": (_i, [_vp, _i, _vp, _vp, _i, _i, _i, _i, _u, _vp, _vp]),
    "ggnn_propagate_csr": (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _u, _vp, _vp]),
    "ggnn_propagate_edges": (_i, [_vp, _i, _vp, _vp, _i64, _i64, _i, _i, _u, _vp, _vp]),
    "ggnn_grad_floats": (_i64, [_vp]),
    "ggnn_optimizer_step": (_i, [_vp, _i, _vp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _vp]),
    "ggnn_get_weights": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ggnn_set_dropout_seed": (_i, [_vp, C.c_float, C.c_uint64, _i, _i]),
    "ggnn_dropout_mask": (_i, [_vp, _vp, C.c_int64, C.c_uint64, C.c_uint64, C.c_float, _vp]),
    "ggnn_set_dropout": (_i, [_vp, _vp, _vp, C.c_float, _i]),
    "ggnn_propagate_backward": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "ggnn_embed_nodes": (_i, [_vp, _vp, _i64, _vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _i, _vp, _vp]),
    "ggnn_heads_decode": (_i, [_vp, _vp, _vp, _i64, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "ggnn_workspace_bytes": (C.c_size_t, [_vp, _i64, _i64, _i, _i]),
    "ggnn_launch_count": (_i64, [_vp]),
    "ggnn_profile_enable": (_i, [_vp, _i]),
    "ggnn_profile_read": (_i, [_vp, _vp, _vp]),
}

_lib = None


def lib():
    """Load the shared library (once).  Raises if it is absent -- build it with
    ``python ggnn_b200/build.py`` or ``__graft_entry__.build()``."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is not built and ggnn_b200 has no "
                "CPU fallback. Run `python ggnn_b200/build.py`.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)       # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib
