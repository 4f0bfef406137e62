"""The Python drop-in: same inputs and outputs as the reference's
``DenseGGNNChemModel.compute_final_node_representations``
(/root/reference/chem_tensorflow_dense.py:312-340), with the method's implicit inputs
(placeholders ``adjacency_matrix`` :192, ``node_mask`` :177, weights :202-241,
``params['num_timesteps']``) made explicit arguments.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional, Union

import numpy as np
import torch

from . import _lib
from .engine import Propagator
from .weights import GGNNWeights, as_weights

_cache_lock = threading.Lock()
_cache = {}          # (thread id, device index, id(weights object)) -> (weights ref, Propagator)


def get_propagator(weights, device=None) -> Propagator:
    """Propagator for these weights on this device, cached per host thread (a ctx is not
    thread-safe).  ``weights``: path to a reference .pickle, its dict, or GGNNWeights.
    The cache key is the path or the identity of the object: weights are treated as IMMUTABLE.  Mutating
    the arrays of a cached dict in place (or rewriting the pickle) is not seen; and a Propagator whose
    ``optimizer_step`` has run holds newer weights than ``Propagator.weights`` (``get_weights()`` returns
    the current ones).  Build a new Propagator for new weights."""
    if isinstance(weights, Propagator):
        return weights
    if not torch.cuda.is_available():
        raise RuntimeError("ggnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    key = (threading.get_ident(), dev.index, weights if isinstance(weights, str) else id(weights))
    with _cache_lock:
        hit = _cache.get(key)
        if hit is not None:
            return hit[1]
    prop = Propagator(as_weights(weights), dev)
    with _cache_lock:
        if len(_cache) > 16:
            _cache.clear()
        _cache[key] = (weights, prop)     # keep `weights` alive so id() stays unique
    return prop


def _to_device(x, device, dtype=None, name="tensor"):
    """numpy / torch (any device) -> contiguous torch tensor on `device`.  Host arrays go
    through pinned memory so the copy is a real async DMA."""
    if isinstance(x, np.ndarray):
        if dtype is not None and x.dtype != dtype:
            x = x.astype(dtype)
        t = torch.from_numpy(np.ascontiguousarray(x))
        return t.pin_memory().to(device, non_blocking=True), "numpy"
    if isinstance(x, torch.Tensor):
        if x.device.type == "cpu":
            return x.contiguous().pin_memory().to(device, non_blocking=True), "cpu"
        return x, "cuda"
    raise ValueError(f"{name} must be a numpy array or torch tensor, got {type(x).__name__}")


_streams = {}       # (thread id, device index) -> (copy-in, compute, copy-out) streams of the host pipeline
_HOST_CHUNK_BYTES = 96 << 20


def _as_host_tensor(x, dtype=None):
    if isinstance(x, np.ndarray):
        if dtype is not None and x.dtype != dtype:
            x = x.astype(dtype)
        x = torch.from_numpy(np.ascontiguousarray(x))
    return x.contiguous()


def _propagate_dense_from_host(prop: Propagator, h0, adj, num_timesteps, fixed_ts, precision):
    """Host-resident inputs (the reference's feed_dict arrays): the batch is cut into chunks of ~96 MB of
    adjacency and pipelined over three streams -- H2D of chunk k+1 (straight from the caller's memory when it
    is pinned, else through a pinned double buffer) overlaps compaction + propagation of chunk k and the D2H
    of chunk k-1 -- so the call runs at the speed of the bus.  Returns a pinned CPU tensor."""
    dev = prop.device
    b, v, h = h0.shape
    per_graph = adj[0].numel() * adj.element_size()
    cg = int(max(1, min(b, _HOST_CHUNK_BYTES // max(per_graph, 1))))
    key = (threading.get_ident(), dev.index)
    if key not in _streams:
        _streams[key] = tuple(torch.cuda.Stream(dev) for _ in range(3))
    s_in, s_run, s_out = _streams[key]
    cur = torch.cuda.current_stream(dev)
    d_adj = [torch.empty((cg,) + tuple(adj.shape[1:]), dtype=adj.dtype, device=dev) for _ in range(2)]
    d_h = torch.empty((b, v, h), dtype=torch.float32, device=dev)
    out_host = torch.empty((b, v, h), dtype=torch.float32, pin_memory=True)
    stage = None
    if not adj.is_pinned():
        stage = [torch.empty((cg,) + tuple(adj.shape[1:]), dtype=adj.dtype, pin_memory=True) for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    for st in (s_in, s_run, s_out):
        st.wait_stream(cur)
    nchunks = (b + cg - 1) // cg

    def enqueue_copy(k):
        slot, lo, hi = k & 1, k * cg, min(b, (k + 1) * cg)
        src = adj[lo:hi]
        if stage is not None:
            if k >= 2:
                copied[slot].synchronize()            # the DMA out of this staging buffer has finished
            stage[slot][:hi - lo].copy_(src)
            src = stage[slot][:hi - lo]
        with torch.cuda.stream(s_in):
            if k >= 2:
                s_in.wait_event(done[slot])           # chunk k-2 no longer reads this device buffer
            d_adj[slot][:hi - lo].copy_(src, non_blocking=True)
            d_h[lo:hi].copy_(h0[lo:hi], non_blocking=True)
            copied[slot].record(s_in)

    enqueue_copy(0)
    for k in range(nchunks):
        slot, lo, hi = k & 1, k * cg, min(b, (k + 1) * cg)
        if k + 1 < nchunks:
            enqueue_copy(k + 1)                        # in flight while this chunk is planned and propagated
        with torch.cuda.stream(s_run):
            s_run.wait_event(copied[slot])
            prop.propagate_dense(d_h[lo:hi], d_adj[slot][:hi - lo], num_timesteps=num_timesteps, fixed_ts=fixed_ts,
                                 precision=precision, out=d_h[lo:hi])
            done[slot].record(s_run)
        with torch.cuda.stream(s_out):
            s_out.wait_event(done[slot])
            out_host[lo:hi].copy_(d_h[lo:hi], non_blocking=True)
    s_out.synchronize()
    for t in d_adj + [d_h]:                            # the caching allocator must not hand these out too early
        t.record_stream(s_in); t.record_stream(s_run); t.record_stream(s_out)
    return out_host


_HOST_PACK = os.environ.get("GGNN_HOST_PACK", "1") != "0"      # A-B switch: 0 = ship the dense tensor over the bus
_HOST_PACK_CHUNK_BYTES = 768 << 20                          # 128 C3 graphs: few calls; the sweep is in scripts/gpu_e2e_pack.py
# packer threads: all host cores but one (the calling thread drives the GPU and spins in its planning sync)
_HOST_PACK_THREADS = int(os.environ.get("GGNN_HOST_PACK_THREADS", "0")) or max(1, (os.cpu_count() or 2) - 1)


_packer_pool = None


def _packer():
    """One background thread that calls the (multi-threaded, GIL-free) host packer ahead of the GPU calls."""
    global _packer_pool
    if _packer_pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _packer_pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="ggnn-pack")
    return _packer_pool


def _bits_route_ok(two_e: int, v: int) -> bool:
    """Shapes the bitmask -> plan route of the library serves (ggnn_plan_bits_supported, plan_bits.cu)."""
    vt = (v + 31) // 32 * 32
    return 1 <= v <= 255 and two_e <= 256 and two_e * vt * 2 <= 48 * 1024 and (vt // 32) * two_e * 4 <= 48 * 1024


def _propagate_dense_from_host_packed(prop: Propagator, h0, adj, num_timesteps, fixed_ts, precision):
    """Host-resident inputs, packed: the path uses one bit per adjacency element (A != 0), so the dense tensor is
    reduced to its bitmask ON THE HOST CORES (ggnn_host_pack_adjacency: threads + AVX, 29 k C3 graphs/s on the 16
    cores of the GPU box against 9 k graphs/s for the FP32 tensor over PCIe) and 1/32 of the bytes cross the bus.
    Chunks of ~768 MB of adjacency, packed by a background thread up to two chunks ahead: packing of chunk k+1 overlaps the GPU's work on chunk k (row counts, plan,
    messages, GRU steps -- ggnn_propagate_dense_bits) and the D2H of chunk k-1.  The bitmask is the same
    intermediate the GPU's own first pass produces for device-resident inputs (tests/test_gpu_hostpack.py).
    Works on pageable memory as it is (no pinned staging copy).  Returns a pinned CPU tensor."""
    import time as _t0m
    t_enter = _t0m.perf_counter()
    dev = prop.device
    lib = _lib.lib()
    b, v, h = h0.shape
    per_graph_elems = adj[0].numel()
    words_pg = per_graph_elems // 32                              # _bits_route_ok shapes only; checked by the caller
    cg = int(max(1, min(b, _HOST_PACK_CHUNK_BYTES // max(per_graph_elems * adj.element_size(), 1))))
    dt = _lib.ADJ_F32 if adj.dtype == torch.float32 else _lib.ADJ_U8
    key = (threading.get_ident(), dev.index)
    if key not in _streams:
        _streams[key] = tuple(torch.cuda.Stream(dev) for _ in range(3))
    s_in, s_run, s_out = _streams[key]
    cur = torch.cuda.current_stream(dev)
    h_bits = [torch.empty(cg * words_pg, dtype=torch.int32, pin_memory=True) for _ in range(2)]
    d_bits = [torch.empty(cg * words_pg, dtype=torch.int32, device=dev) for _ in range(2)]
    d_h = torch.empty((b, v, h), dtype=torch.float32, device=dev)
    out_host = torch.empty((b, v, h), dtype=torch.float32, pin_memory=True)
    h0p = h0 if h0.is_pinned() else None
    copied = [torch.cuda.Event() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    for st in (s_in, s_run, s_out):
        st.wait_stream(cur)
    nchunks = (b + cg - 1) // cg
    esz = adj.element_size()

    dbg = os.environ.get("GGNN_HOST_PACK_DEBUG")      # development aid: where a call's time goes
    import time as _t
    tl = {"pack": 0.0, "wait_pack": 0.0, "propagate": 0.0, "t0": _t.perf_counter()}

    def pack(k):                                      # runs on the packer thread (ctypes releases the GIL)
        slot, lo, hi = k & 1, k * cg, min(b, (k + 1) * cg)
        if k >= 2:
            copied[slot].synchronize()                # the DMA out of this pinned bitmask buffer has finished
        t0 = _t.perf_counter()
        try:
            return _pack_call(k, slot, lo, hi)
        finally:
            tl["pack"] += _t.perf_counter() - t0

    def _pack_call(k, slot, lo, hi):
        rc = lib.ggnn_host_pack_adjacency(C.c_void_p(adj.data_ptr() + lo * per_graph_elems * esz), dt,
                                          (hi - lo) * per_graph_elems, C.c_void_p(h_bits[slot].data_ptr()),
                                          _HOST_PACK_THREADS)
        if rc != _lib.OK:
            raise RuntimeError(f"ggnn_host_pack_adjacency failed ({rc})")

    def copy_in(k):
        slot, lo, hi = k & 1, k * cg, min(b, (k + 1) * cg)
        with torch.cuda.stream(s_in):
            if k >= 2:
                s_in.wait_event(done[slot])           # chunk k-2 no longer reads this device buffer
            d_bits[slot][:(hi - lo) * words_pg].copy_(h_bits[slot][:(hi - lo) * words_pg], non_blocking=True)
            d_h[lo:hi].copy_((h0p if h0p is not None else h0)[lo:hi], non_blocking=True)
            copied[slot].record(s_in)

    # the packer runs up to two chunks ahead of the GPU calls (whose planning step waits for the device)
    packer = _packer()
    futs = {k: packer.submit(pack, k) for k in range(min(2, nchunks))}
    for k in range(nchunks):
        slot, lo, hi = k & 1, k * cg, min(b, (k + 1) * cg)
        t0 = _t.perf_counter()
        futs.pop(k).result()
        tl["wait_pack"] += _t.perf_counter() - t0
        copy_in(k)
        if k + 2 < nchunks:
            futs[k + 2] = packer.submit(pack, k + 2)  # (waits for this chunk's DMA out of the buffer it reuses)
        t0 = _t.perf_counter()
        with torch.cuda.stream(s_run):
            s_run.wait_event(copied[slot])
            prop.propagate_dense_bits(d_h[lo:hi], d_bits[slot], num_timesteps=num_timesteps, fixed_ts=fixed_ts,
                                      precision=precision, out=d_h[lo:hi])
            done[slot].record(s_run)
        tl["propagate"] += _t.perf_counter() - t0
        with torch.cuda.stream(s_out):
            s_out.wait_event(done[slot])
            out_host[lo:hi].copy_(d_h[lo:hi], non_blocking=True)
    t0 = _t.perf_counter()
    s_out.synchronize()
    if dbg:
        import sys
        print(f"host-pack call: {nchunks} chunks of {cg} graphs; setup {(tl['t0'] - t_enter) * 1e3:.2f} ms, pack {tl['pack'] * 1e3:.2f} ms "
              f"(packer thread), waiting for the packer {tl['wait_pack'] * 1e3:.2f} ms, propagate calls {tl['propagate'] * 1e3:.2f} ms, "
              f"final sync {(_t.perf_counter() - t0) * 1e3:.2f} ms, total {(_t.perf_counter() - t_enter) * 1e3:.2f} ms", file=sys.stderr)
    for t in d_bits + [d_h]:                           # the caching allocator must not hand these out too early
        t.record_stream(s_in); t.record_stream(s_run); t.record_stream(s_out)
    return out_host


# opt-in (GGNN_HOST_HYBRID=1): measured 23.5-25.8 k graphs/s against 22.7-23.6 k for the packer alone (scripts/gpu_e2e_pack.py) --
# the copy engine and the packer read the same host memory, which is the limit either way
_HOST_HYBRID = os.environ.get("GGNN_HOST_HYBRID", "0") == "1"
_HYBRID_PACK_GRAPH_BYTES = 384 << 20       # a packed unit: ~64 C3 graphs (about 2.7 ms of the packer)
_HYBRID_DENSE_GRAPH_BYTES = 96 << 20       # a dense unit: ~16 C3 graphs (about 1.8 ms of the bus)


def _propagate_dense_from_host_hybrid(prop: Propagator, h0, adj, num_timesteps, fixed_ts, precision):
    """Host-resident PINNED inputs: the host packer (memory-bound on the host cores) and the copy engine (bound by the
    bus) are two independent ways of getting adjacency to the GPU, so both are used at once.  The batch is cut into
    units that alternate between the two routes -- packed units (bitmask built on the host, ggnn_propagate_dense_bits)
    and dense units (the FP32 tensor over the bus, ggnn_propagate_dense: the GPU builds the bitmask) -- each route
    works through its own units in order, two ahead of the GPU calls, which consume the units in batch order.
    Results are the same bits whichever route a graph takes (tests/test_gpu_hostpack.py)."""
    dev = prop.device
    lib = _lib.lib()
    b, v, h = h0.shape
    per_graph_elems = adj[0].numel()
    esz = adj.element_size()
    per_graph = per_graph_elems * esz
    words_pg = per_graph_elems // 32
    gp = int(max(1, _HYBRID_PACK_GRAPH_BYTES // per_graph))
    gd = int(max(1, _HYBRID_DENSE_GRAPH_BYTES // per_graph))
    dt = _lib.ADJ_F32 if adj.dtype == torch.float32 else _lib.ADJ_U8
    units = []                                            # (route, lo, hi); the packer's first unit is ready last: dense first
    lo, k = 0, 0
    while lo < b:
        route = "D" if k % 2 == 0 else "P"
        hi = min(b, lo + (gd if route == "D" else gp))
        units.append((route, lo, hi))
        lo, k = hi, k + 1
    key = (threading.get_ident(), dev.index)
    if key not in _streams:
        _streams[key] = tuple(torch.cuda.Stream(dev) for _ in range(3))
    s_in, s_run, s_out = _streams[key]
    cur = torch.cuda.current_stream(dev)
    DEPTH = 2
    h_bits = [torch.empty(gp * words_pg, dtype=torch.int32, pin_memory=True) for _ in range(DEPTH)]
    d_bits = [torch.empty(gp * words_pg, dtype=torch.int32, device=dev) for _ in range(DEPTH)]
    d_adj = [torch.empty((gd,) + tuple(adj.shape[1:]), dtype=adj.dtype, device=dev) for _ in range(DEPTH)]
    d_h = torch.empty((b, v, h), dtype=torch.float32, device=dev)
    out_host = torch.empty((b, v, h), dtype=torch.float32, pin_memory=True)
    for st in (s_in, s_run, s_out):
        st.wait_stream(cur)
    p_units = [i for i, u in enumerate(units) if u[0] == "P"]
    d_units = [i for i, u in enumerate(units) if u[0] == "D"]
    p_ord = {i: n for n, i in enumerate(p_units)}         # unit index -> its number on its route (ring slot = n % DEPTH)
    d_ord = {i: n for n, i in enumerate(d_units)}
    bits_copied = [torch.cuda.Event() for _ in range(DEPTH)]     # per ring slot: the bitmask DMA has left the pinned buffer
    consumed_p = [torch.cuda.Event() for _ in range(DEPTH)]      # per ring slot: the GPU call reading d_bits / d_adj is done
    consumed_d = [torch.cuda.Event() for _ in range(DEPTH)]
    ready = [torch.cuda.Event() for _ in units]                  # the unit's inputs are on the device

    def pack(i):                                          # packer thread
        n = p_ord[i]
        slot, (_, lo, hi) = n % DEPTH, units[i]
        if n >= DEPTH:
            bits_copied[slot].synchronize()
        rc = lib.ggnn_host_pack_adjacency(C.c_void_p(adj.data_ptr() + lo * per_graph), dt, (hi - lo) * per_graph_elems,
                                          C.c_void_p(h_bits[slot].data_ptr()), _HOST_PACK_THREADS)
        if rc != _lib.OK:
            raise RuntimeError(f"ggnn_host_pack_adjacency failed ({rc})")

    def copy_bits(i):                                     # after the unit's pack job: bitmask + states to the device
        n = p_ord[i]
        slot, (_, lo, hi) = n % DEPTH, units[i]
        with torch.cuda.stream(s_in):
            if n >= DEPTH:
                s_in.wait_event(consumed_p[slot])         # the call that read this device buffer has run
            d_bits[slot][:(hi - lo) * words_pg].copy_(h_bits[slot][:(hi - lo) * words_pg], non_blocking=True)
            d_h[lo:hi].copy_(h0[lo:hi], non_blocking=True)
            bits_copied[slot].record(s_in)
            ready[i].record(s_in)

    def copy_dense(i):                                    # the FP32 tensor itself + states to the device
        n = d_ord[i]
        slot, (_, lo, hi) = n % DEPTH, units[i]
        with torch.cuda.stream(s_in):
            if n >= DEPTH:
                s_in.wait_event(consumed_d[slot])
            d_adj[slot][:hi - lo].copy_(adj[lo:hi], non_blocking=True)
            d_h[lo:hi].copy_(h0[lo:hi], non_blocking=True)
            ready[i].record(s_in)

    packer = _packer()
    futs = {i: packer.submit(pack, i) for i in p_units[:DEPTH]}
    for i in d_units[:DEPTH]:
        copy_dense(i)
    for i, (route, lo, hi) in enumerate(units):
        if route == "P":
            n = p_ord[i]
            futs.pop(i).result()
            copy_bits(i)
            if n + DEPTH < len(p_units):                  # its pinned buffer is this unit's: free once the DMA above has run
                j = p_units[n + DEPTH]
                futs[j] = packer.submit(pack, j)
        with torch.cuda.stream(s_run):
            s_run.wait_event(ready[i])
            if route == "P":
                prop.propagate_dense_bits(d_h[lo:hi], d_bits[p_ord[i] % DEPTH], num_timesteps=num_timesteps, fixed_ts=fixed_ts,
                                          precision=precision, out=d_h[lo:hi])
                consumed_p[p_ord[i] % DEPTH].record(s_run)
            else:
                prop.propagate_dense(d_h[lo:hi], d_adj[d_ord[i] % DEPTH][:hi - lo], num_timesteps=num_timesteps,
                                     fixed_ts=fixed_ts, precision=precision, out=d_h[lo:hi])
                consumed_d[d_ord[i] % DEPTH].record(s_run)
            done = torch.cuda.Event()
            done.record(s_run)
        if route == "D" and d_ord[i] + DEPTH < len(d_units):
            copy_dense(d_units[d_ord[i] + DEPTH])
        with torch.cuda.stream(s_out):
            s_out.wait_event(done)
            out_host[lo:hi].copy_(d_h[lo:hi], non_blocking=True)
    s_out.synchronize()
    for t in d_bits + d_adj + [d_h]:                      # the caching allocator must not hand these out too early
        t.record_stream(s_in); t.record_stream(s_run); t.record_stream(s_out)
    return out_host


def _device_of(*xs):
    for x in xs:
        if isinstance(x, torch.Tensor) and x.device.type == "cuda":
            return x.device
    return torch.device("cuda", torch.cuda.current_device())


def _back(out: torch.Tensor, kind: str):
    """Result in the caller's kind; host results come back through a pinned buffer (a real async DMA at bus
    speed instead of a staged pageable copy)."""
    if kind == "cuda":
        return out
    host = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
    host.copy_(out, non_blocking=True)
    torch.cuda.current_stream(out.device).synchronize()
    return host.numpy() if kind == "numpy" else host


def compute_final_node_representations(initial_node_representations, adjacency_matrix,
                                       node_mask=None, weights=None, num_timesteps: Optional[int] = None,
                                       fixed_ts: Optional[int] = None, precision: str = "fp32",
                                       skip_padded: bool = False, validate: bool = False):
    """T steps of GGNN propagation (message transform per edge type + bias, aggregation over
    the adjacency, GRU update) -- reference chem_tensorflow_dense.py:312-340, 391-437.

    initial_node_representations : float32 [b, v, h]
    adjacency_matrix             : float32 or uint8 [b, 2E, v, v], A[b, e, target, source] != 0 <=> edge
    node_mask                    : [b, v] or None.  The reference never reads it inside the path
                                   (padded rows are GRU-updated like any other); it is used only
                                   when skip_padded=True, to zero the padded rows of the result.
    weights                      : path to a reference .pickle, its dict, GGNNWeights, or Propagator
    num_timesteps                : defaults to the checkpoint's params['num_timesteps']
    fixed_ts                     : k -> use edge_weights_fixed / edge_biases_fixed for k steps
    precision                    : "fp32" (strict, <= 1e-5 norm-wise), "tf32" or "f16" (tensor cores: TF32 operands /
                                   FP16 containers of the same 11-bit significand, FP32 accumulate, <= 2e-3)
    validate                     : True -> raise ValueError unless every adjacency value is exactly 0 or 1 (the
                                   path uses the adjacency as an edge set; see Propagator.propagate_dense)
    Returns float32 [b, v, h] of the same kind (numpy / torch cpu / torch cuda) as the input states.
    """
    if weights is None:
        raise ValueError("weights is required (path to a reference .pickle, dict or GGNNWeights)")
    if not torch.cuda.is_available():
        raise RuntimeError("ggnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = weights.device if isinstance(weights, Propagator) else _device_of(initial_node_representations, adjacency_matrix)
    prop = get_propagator(weights, dev)
    if isinstance(adjacency_matrix, np.ndarray) and adjacency_matrix.dtype not in (np.float32, np.uint8):
        adjacency_matrix = adjacency_matrix.astype(np.float32)     # the placeholder is tf.float32
    host = lambda x: isinstance(x, np.ndarray) or (isinstance(x, torch.Tensor) and x.device.type == "cpu")
    if validate:
        a = adjacency_matrix
        ok = bool(((a == 0) | (a == 1)).all())
        if not ok:
            raise ValueError("adjacency_matrix holds values other than 0 and 1: the path implements the reference's "
                             "0/1 adjacency (graph_to_adj_mat_bd); weighted adjacencies are not supported")
    if host(initial_node_representations) and host(adjacency_matrix) and not skip_padded:
        kind = "numpy" if isinstance(initial_node_representations, np.ndarray) else "cpu"
        h0 = _as_host_tensor(initial_node_representations, np.float32)
        adj = _as_host_tensor(adjacency_matrix)
        if adj.dtype == torch.float64:
            adj = adj.float()
        if h0.dim() != 3 or adj.dim() != 4 or h0.dtype != torch.float32 or adj.dtype not in (torch.float32, torch.uint8) \
                or adj.shape[0] != h0.shape[0] or adj.shape[2] != h0.shape[1] or adj.shape[3] != h0.shape[1] \
                or h0.shape[2] != prop.weights.hidden_size or adj.shape[1] != prop.weights.two_e:
            raise ValueError(f"states {tuple(h0.shape)} {h0.dtype} / adjacency {tuple(adj.shape)} {adj.dtype} do not "
                             f"match [b, v, {prop.weights.hidden_size}] float32 / [b, {prop.weights.two_e}, v, v]")
        if h0.shape[0] > 0:
            if _HOST_PACK and _bits_route_ok(prop.weights.two_e, h0.shape[1]) and adj[0].numel() % 32 == 0:
                if _HOST_HYBRID and adj.is_pinned() and h0.is_pinned():
                    out = _propagate_dense_from_host_hybrid(prop, h0, adj, num_timesteps, fixed_ts, precision)
                else:
                    out = _propagate_dense_from_host_packed(prop, h0, adj, num_timesteps, fixed_ts, precision)
            else:
                out = _propagate_dense_from_host(prop, h0, adj, num_timesteps, fixed_ts, precision)
            return out.numpy() if kind == "numpy" else out
    h0, kind = _to_device(initial_node_representations, dev, np.float32, "initial_node_representations")
    adj, _ = _to_device(adjacency_matrix, dev, None, "adjacency_matrix")
    if adj.dtype == torch.float64:
        adj = adj.float()
    out = prop.propagate_dense(h0, adj, num_timesteps=num_timesteps, fixed_ts=fixed_ts, precision=precision)
    if skip_padded:
        if node_mask is None:
            raise ValueError("skip_padded=True needs node_mask")
        m, _ = _to_device(node_mask, dev, np.float32, "node_mask")
        if tuple(m.shape) != tuple(out.shape[:2]):
            raise ValueError(f"node_mask must be [b, v] = {tuple(out.shape[:2])}, got {tuple(m.shape)}")
        out = out * (m != 0).to(out.dtype).unsqueeze(-1)
    return _back(out, kind)


def compute_final_node_representations_sparse(node_states, edges, graph_offsets=None, weights=None,
                                              num_timesteps: Optional[int] = None,
                                              fixed_ts: Optional[int] = None, precision: str = "fp32"):
    """Edge-list form (the reference declares it at chem_tensorflow.py:522-523 and never
    implements it).  node_states float32 [N, h]; edges int32 [nnz, 3] = (src, channel, dst),
    channel 0-based in [0, 2E), node ids global.  ``graph_offsets`` is accepted for interface
    symmetry (graphs are independent, so it is not needed for the arithmetic)."""
    del graph_offsets
    if weights is None:
        raise ValueError("weights is required")
    if not torch.cuda.is_available():
        raise RuntimeError("ggnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = weights.device if isinstance(weights, Propagator) else _device_of(node_states, edges)
    prop = get_propagator(weights, dev)
    h0, kind = _to_device(node_states, dev, np.float32, "node_states")
    e, _ = _to_device(edges, dev, np.int32, "edges")
    if e.dtype != torch.int32:
        e = e.to(torch.int32)
    out = prop.propagate_edges(h0, e, num_timesteps=num_timesteps, fixed_ts=fixed_ts, precision=precision)
    return _back(out, kind)


def compact_adjacency(adjacency_matrix, weights_or_propagator=None):
    """Dense adjacency -> (rowptr, col) canonical CSR on the device (np.argwhere order)."""
    dev = _device_of(adjacency_matrix)
    adj, kind = _to_device(adjacency_matrix, dev, None, "adjacency_matrix")
    if isinstance(weights_or_propagator, Propagator):
        prop = weights_or_propagator
    elif weights_or_propagator is not None:
        prop = get_propagator(weights_or_propagator, dev)
    else:
        raise ValueError("compact_adjacency needs a Propagator (or weights) to own the scratch")
    rowptr, col, _ = prop.compact_adjacency(adj)
    if kind == "numpy":
        return rowptr.cpu().numpy(), col.cpu().numpy()
    return rowptr, col
