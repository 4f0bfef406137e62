#!/usr/bin/env python
"""bench.py -- GGNN propagation throughput on B200 (BASELINE.json metric).

Workload (config 3 of BASELINE.json): synthetic dependency trees,
v = 128 nodes (127 words + ROOT), 2E = 92 edge channels (45 labels + word order, x2),
hidden 128, T = 8, batch 65 536 graphs sharded over N GPUs (strong scaling: 65 536 / N per
GPU, no forward collective).  One step = one pass of the hot path over the whole batch:
dense FP32 adjacency -> compaction -> T x (message transform, aggregation, GRU).

    python bench.py --gpus 1 --steps 3 --warmup 3
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference      # the reference's dense op order on the host cores

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

# ---- workload constants (C3) -----------------------------------------------------------
V, E, H, T = 128, 46, 128, 8
TWO_E = 2 * E
NNZ_PER_GRAPH = 4 * (V - 1)                                   # 508
TOTAL_GRAPHS = 65536
# algorithmic work
FLOP_MSG_PER_GRAPH_STEP = NNZ_PER_GRAPH * 2 * H * H           # 16.6 MFLOP
FLOP_GATES_PER_GRAPH_STEP = V * 2 * (2 * H) * (2 * H)         # 16.8 MFLOP
FLOP_CAND_PER_GRAPH_STEP = V * 2 * (2 * H) * H                #  8.4 MFLOP
BYTES_COMPACT_PER_GRAPH = TWO_E * V * V * 4 + NNZ_PER_GRAPH * 4 + (TWO_E * V) * 4   # 6.08 MB
# DRAM traffic of the compaction bitmask pass per graph from the ncu --set full capture
# (profiles/r1_compact_ncu.md, bitmask_f32_v8_kernel: 12.348 GB read + 0.342 GB written over 2 048 graphs)
TRAFFIC_COMPACT_PER_GRAPH = (12.348041e9 + 341.879552e6) / 2048
BYTES_AGG_PER_GRAPH_STEP = NNZ_PER_GRAPH * H * 4 + NNZ_PER_GRAPH * 4 + (V + 1) * 4 + V * H * 4


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        out = {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
               "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
               "source": "measured (MEASURED_PEAKS.json)"}
        t = os.path.join(ROOT, "profiles", "r2_tf32_peak.json")     # same protocol, TF32 (scripts/measure_tf32_peak.py)
        if os.path.exists(t):
            out["tf32_tflops_sustained"] = float(json.load(open(t))["tf32_tflops_sustained"])
        return out
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------
# CPU arm: the reference's dense operation order on the host cores (oracle port)
# ---------------------------------------------------------------------------------------
def cpu_sample(n_graphs, seed=7):
    from ggnn_b200 import synth
    tri = synth.random_trees(n_graphs, V - 1, E, seed=seed)
    adj = synth.dense_adjacency(tri, V, E)
    h0 = (torch.rand(n_graphs, V, H, generator=torch.Generator().manual_seed(seed)) - 0.5) * 0.2
    return h0, adj, synth.make_weights(TWO_E, H, 0)


def cpu_baseline(target_seconds=12.0):
    """Bounded sample of the same workload through oracle/cpu_reference.py, all host threads."""
    from oracle import cpu_reference as R
    threads = os.cpu_count() or 1
    h0, adj, w = cpu_sample(8)
    R.time_dense(h0[:2], adj[:2], w, T, chunk=2, threads=threads)          # warm-up
    t8, _ = R.time_dense(h0, adj, w, T, chunk=8, threads=threads)
    n = int(max(8, min(512, round(target_seconds / (t8 / 8) / 8) * 8)))
    if n > 8:
        h0, adj, w = cpu_sample(n)
        tn, _ = R.time_dense(h0, adj, w, T, chunk=8, threads=threads)
    else:
        tn = t8
    return {"value": n / tn, "unit": "graphs/s", "cores": threads, "kind": "port",
            "sample": f"{n} C3 graphs (v=128, 2E=92, h=128, T=8), dense op order of "
                      "compute_timestep_fast in torch-CPU FP32, 8 graphs per call",
            "node_updates_per_s": n * V * T / tn}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_reference as R
    threads = os.cpu_count() or 1
    n = args.ref_graphs
    h0, adj, w = cpu_sample(n)
    for _ in range(args.warmup):
        R.time_dense(h0[:8], adj[:8], w, T, chunk=8, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        R.time_dense(h0, adj, w, T, chunk=8, threads=threads)
    dt = (time.perf_counter() - t0) / args.steps
    val = n / dt
    line = {
        "impl": "reference", "metric": "GGNN propagation graphs/sec (T=8 steps)", "value": val,
        "unit": "graphs/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C3 synthetic dependency trees v=128 2E=92 h=128 T=8",
                   "graphs_per_step": n, "note": "bounded sample of the 65536-graph batch; "
                   "TensorFlow is not installable, so this is the oracle port of the reference's "
                   "dense op order (torch-CPU FP32, all host threads)"},
        "node_updates_per_s": val * V * T,
        "cpu_baseline": {"value": val, "unit": "graphs/s", "cores": threads, "kind": "port",
                         "sample": f"{n} C3 graphs per step, 8 graphs per call"},
        "e2e": {"value": val, "unit": "graphs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------
def init_env():
    """One process per GPU (torchrun sets RANK / LOCAL_RANK / WORLD_SIZE); NCCL only when there is more than one."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    return world, rank, local, torch.device("cuda", local)


def max_over_ranks(x, dev, world):
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, dev, world):
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def measure_c3(args, env):
    import torch.distributed as dist
    import ggnn_b200
    from ggnn_b200 import synth

    world, rank, local, dev = env
    peaks = load_peaks()

    graphs_rank = args.graphs // world
    pool = min(args.pool, graphs_rank)
    w = ggnn_b200.from_dict(synth.make_weights(TWO_E, H, 0))
    prop = ggnn_b200.Propagator(w, dev)
    precision = args.precision

    # inputs resident in HBM before the timed region.  Dense FP32 adjacency is 6.03 MB/graph
    # (395 GB for 65 536 graphs), so `pool` distinct graphs are generated on the device and the
    # batch cycles through them (every pass re-reads them from HBM: 24.7 GB >> 126 MB L2);
    # node states are distinct for all graphs.
    tri = synth.random_trees(pool, V - 1, E, seed=1000 + rank).to(dev)
    adj = synth.dense_adjacency(tri, V, E)
    gen = torch.Generator(device=dev).manual_seed(1 + rank)
    h0 = (torch.rand(graphs_rank, V, H, device=dev, generator=gen) - 0.5) * 0.2
    out = torch.empty_like(h0)

    chunks = [(c, min(pool, graphs_rank - c)) for c in range(0, graphs_rank, pool)]
    prefetch = (not args.no_prefetch) and precision in ("f16", "fp32")

    def step(last=False, pf=None):
        """One pass over this rank's graphs.  With prefetch, every call announces the next call's adjacency
        (ggnn_dense_prefetch), across the step boundary too, so its bitmask pass streams in the background of
        the propagation kernels; `last` leaves nothing announced behind."""
        pf = prefetch if pf is None else pf
        for i, (c, n) in enumerate(chunks):
            prop.propagate_dense(h0[c:c + n], adj[:n], num_timesteps=T, precision=precision, out=out[c:c + n])
            if pf and not (last and i + 1 == len(chunks)):
                prop.prefetch_dense(adj[:chunks[(i + 1) % len(chunks)][1]])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if prefetch:
        prop.prefetch_dense(adj[:chunks[0][1]])
    for i in range(args.warmup):
        step(last=i + 1 == args.warmup)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = prop.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    if prefetch:                                   # inside the timed region: K steps = K x 16 bitmask passes, all timed
        prop.prefetch_dense(adj[:chunks[0][1]])
    for i in range(args.steps):
        step(last=i + 1 == args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    launches = prop.launches - l0
    if world > 1:
        tms = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    # per-stage device times: ONE extra pass with the stages serialised on the stream (no prefetch), so every
    # kernel is timed alone by its own CUDA-event pair; the headline above is the overlapped pipeline
    prop.profile(True)
    step(pf=False)
    stages = prop.profile_read()
    prop.profile(False)
    stage_steps = 1
    ms_per_step = ms / args.steps
    value = args.graphs / (ms_per_step * 1e-3)

    # ---- end to end through the public drop-in, host buffers, copies inside the timed region
    ne = min(args.e2e_graphs, pool)
    h0_host = h0[:ne].cpu().pin_memory()
    adj_host = adj[:ne].cpu().pin_memory()
    from ggnn_b200 import api as _api

    def time_e2e(host_pack):
        """The public drop-in on HOST tensors, copies inside the timed region.  host_pack: the dense adjacency is reduced
        to its bitmask on the host cores before it crosses the bus (the default for host inputs, api.py); False: the FP32
        tensor itself travels (round 1 / 2a behaviour, GGNN_HOST_PACK=0)."""
        saved = _api._HOST_PACK
        _api._HOST_PACK = host_pack
        try:
            r = None
            for _ in range(max(3, args.warmup)):   # results are held like in the timed loop, so the pinned-host allocator
                r = ggnn_b200.compute_final_node_representations(h0_host, adj_host, weights=prop,   # caches both blocks
                                                                 num_timesteps=T, precision=precision)
            barrier()
            t0 = time.perf_counter()
            per_call = []
            for _ in range(args.steps):
                tc0 = time.perf_counter()
                r = ggnn_b200.compute_final_node_representations(h0_host, adj_host, weights=prop,
                                                                 num_timesteps=T, precision=precision)
                per_call.append(time.perf_counter() - tc0)     # the call returns host data: it has synchronised
            torch.cuda.synchronize()
            sec = (time.perf_counter() - t0) / args.steps
        finally:
            _api._HOST_PACK = saved
        if os.environ.get("GGNN_BENCH_DEBUG"):
            print("e2e per call ms:", [round(x * 1e3, 2) for x in per_call], file=sys.stderr)
        return max_over_ranks(sec, dev, world), r

    pack_threads = max(1, (os.cpu_count() or 2) // world - 1)  # the ranks of one box share its host cores; one is left to the caller
    _api._HOST_PACK_THREADS = pack_threads
    # The packer is bound by host memory bandwidth and needs cores: with fewer than 6 per rank (4 or 8 ranks on a 16-core box) the
    # copy engines, which need none, move more -- those runs send the FP32 tensor (both routes read the same host memory).
    packed = _api._HOST_PACK and _api._bits_route_ok(2 * E, V) and pack_threads >= 6
    e2e_s, res = time_e2e(packed)
    assert res.shape == h0_host.shape
    adj_bus_bytes = adj_host.numel() // 8 if packed else adj_host.numel() * 4
    e2e = {"value": ne * world / e2e_s, "unit": "graphs/s",
           "h2d_bytes_per_step": int(adj_bus_bytes + h0_host.numel() * 4),
           "d2h_bytes_per_step": int(res.numel() * 4), "graphs_per_step": ne * world,
           "host_bytes_read_per_step": int(adj_host.numel() * 4 + h0_host.numel() * 4),
           "note": ("ggnn_b200.compute_final_node_representations on pinned host tensors holding the dense FP32 adjacency "
                    "(6.03 MB/graph, the reference's feed format) and the states; the adjacency is reduced to its bitmask "
                    "(1 bit per element, the path's own first intermediate) by ggnn_host_pack_adjacency on the host cores, "
                    "and the bitmask + states go H2D, the result D2H" if packed else
                    "ggnn_b200.compute_final_node_representations on pinned host tensors: dense "
                    "FP32 adjacency (6.03 MB/graph, the reference's feed format) + states H2D, result D2H"),
           "host_pack": bool(packed), "host_pack_threads": (pack_threads if packed else 0)}
    if packed:      # the same call with the FP32 tensor itself on the bus (what rounds 1 / 2a measured)
        dense_s, _ = time_e2e(False)
        e2e["dense_over_the_bus"] = {"value": ne * world / dense_s, "unit": "graphs/s",
                                     "h2d_bytes_per_step": int(adj_host.numel() * 4 + h0_host.numel() * 4)}

    # context for the e2e figure: what a bare pinned-host -> device copy of the same adjacency reaches on this box
    d_tmp = torch.empty_like(adj_host, device=dev)
    d_tmp.copy_(adj_host, non_blocking=True)
    torch.cuda.synchronize()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(3):
        d_tmp.copy_(adj_host, non_blocking=True)
    c1.record()
    torch.cuda.synchronize()
    e2e["h2d_copy_peak_gbps"] = 3 * adj_host.numel() * 4 / (c0.elapsed_time(c1) * 1e-3) / 1e9
    e2e["achieved_bus_gbps"] = (e2e["h2d_bytes_per_step"] + e2e["d2h_bytes_per_step"]) / e2e_s / 1e9
    e2e["host_read_gbps"] = e2e["host_bytes_read_per_step"] / e2e_s / 1e9
    del d_tmp

    # same graphs handed over as (src, channel, dst) triples: no dense adjacency on the bus
    edges_host = synth.global_edge_list(tri[:ne].cpu(), V, E).pin_memory()
    h0_flat = h0_host.view(ne * V, H)
    res2 = None
    for _ in range(max(3, args.warmup)):
        res2 = ggnn_b200.compute_final_node_representations_sparse(h0_flat, edges_host, weights=prop,
                                                                   num_timesteps=T, precision=precision)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res2 = ggnn_b200.compute_final_node_representations_sparse(h0_flat, edges_host, weights=prop,
                                                                   num_timesteps=T, precision=precision)
    torch.cuda.synchronize()
    e2e_edges_s = max_over_ranks((time.perf_counter() - t0) / args.steps, dev, world)
    e2e_edges = {"value": ne * world / e2e_edges_s, "unit": "graphs/s",
                 "h2d_bytes_per_step": int(edges_host.numel() * 4 + h0_flat.numel() * 4),
                 "d2h_bytes_per_step": int(res2.numel() * 4), "graphs_per_step": ne * world,
                 "note": "compute_final_node_representations_sparse on pinned host tensors: edge list "
                         "(6 KB/graph) + states H2D, result D2H"}

    # the other arithmetic modes and the uint8 adjacency on the same workload, timed exactly like the headline
    # (same --steps / --warmup, same prefetch pipeline)
    def timed_variant(mode, adj_t, steps, warmup):
        # the background pass is sized to sit next to the F16 kernels (16 warps x 96 registers) and, since round 2b, next
        # to the strict mode's tensor-core kernels (strict_tc.cu: 9 warps capped at 128 registers, h <= 128); the TF32
        # kernels (20 warps x 96) fill the register file, so a prefetch would only serialise behind them: not used there
        prefetch = (not args.no_prefetch) and mode in ("f16", "fp32")

        def one(last):
            for i, (c, n) in enumerate(chunks):
                prop.propagate_dense(h0[c:c + n], adj_t[:n], num_timesteps=T, precision=mode, out=out[c:c + n])
                if prefetch and not (last and i + 1 == len(chunks)):
                    prop.prefetch_dense(adj_t[:chunks[(i + 1) % len(chunks)][1]])
        if prefetch:
            prop.prefetch_dense(adj_t[:chunks[0][1]])
        for i in range(warmup):
            one(i + 1 == warmup)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        if prefetch:
            prop.prefetch_dense(adj_t[:chunks[0][1]])
        for i in range(steps):
            one(i + 1 == steps)
        f1.record()
        barrier()
        fms = max_over_ranks(f0.elapsed_time(f1), dev, world) / steps
        return {"value": args.graphs / (fms * 1e-3), "unit": "graphs/s", "ms_per_step": fms, "steps": steps,
                "warmup": warmup, "node_updates_per_s": args.graphs / (fms * 1e-3) * V * T}

    other_modes = {}
    notes = {"fp32": "GGNN_PREC_FP32, strict (<= 1e-5 norm-wise vs the FP64 oracle)",
             "tf32": "GGNN_PREC_TF32, tcgen05 kind::tf32 (<= 2e-3)", "f16": "GGNN_PREC_F16, tcgen05 kind::f16 (<= 2e-3)"}
    for mode in ([] if args.no_fp32 else [m for m in ("tf32", "fp32") if m != precision]):
        other_modes[mode] = timed_variant(mode, adj, args.steps, max(1, args.warmup))
        other_modes[mode]["note"] = notes[mode] + ", same inputs"
    other_inputs = {}
    if not args.no_fp32:
        adj8 = (adj[:pool] != 0).to(torch.uint8)
        other_inputs["adjacency_uint8"] = timed_variant(precision, adj8, args.steps, max(1, args.warmup))
        other_inputs["adjacency_uint8"]["note"] = "GGNN_ADJ_U8: same graphs, states and mode; 1.51 MB of adjacency per graph"
        del adj8

    # SURVEY 8(d) C3 secondary run: the C2 length distribution (lognormal, mean ~24 words) bucketed to the reference's
    # even sizes <= 128 (chem_tensorflow_dense.py:584-585), one bucket-homogeneous batch per size like make_minibatch_iterator
    ragged = None
    if not args.no_fp32:
        ng = min(args.ragged_graphs, graphs_rank)
        lengths = synth.wsj_like_lengths(ng, seed=2000 + rank, max_len=V - 1)
        bsz = torch.tensor([synth.bucket_size(int(l)) for l in lengths])
        groups = []
        for vb in sorted(set(bsz.tolist())):
            idx = (bsz == vb).nonzero().flatten()
            tri_b = synth.random_trees(len(idx), vb - 1, E, seed=3000 + rank + vb, lengths=lengths[idx]).to(dev)
            groups.append((synth.dense_adjacency(tri_b, vb, E),
                           (torch.rand(len(idx), vb, H, device=dev, generator=gen) - 0.5) * 0.2))
        real_nodes = int((lengths + 1).sum())
        padded_nodes = int(bsz.sum())

        def ragged_pass():
            for a_b, h_b in groups:
                prop.propagate_dense(h_b, a_b, num_timesteps=T, precision=precision)
        for _ in range(max(1, args.warmup)):
            ragged_pass()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(args.steps):
            ragged_pass()
        f1.record()
        barrier()
        rms = max_over_ranks(f0.elapsed_time(f1), dev, world) / args.steps
        tot = lambda x: sum_over_ranks(float(x), dev, world)
        # the same sentences as ONE packed edge list (CSR straight from the triples, rows = bucket size):
        # one ggnn_propagate_edges call instead of one dense call per bucket
        tri_all = synth.random_trees(ng, int(lengths.max()), E, seed=4000 + rank, lengths=lengths)
        edges_p, off_p = synth.packed_edge_list(tri_all, bsz, E)
        edges_p = edges_p.to(dev)
        hp_ = (torch.rand(int(off_p[-1]), H, device=dev, generator=gen) - 0.5) * 0.2
        outp = torch.empty_like(hp_)
        for _ in range(max(1, args.warmup)):
            prop.propagate_edges(hp_, edges_p, num_timesteps=T, precision=precision, out=outp)
        barrier()
        f0.record()
        for _ in range(args.steps):
            prop.propagate_edges(hp_, edges_p, num_timesteps=T, precision=precision, out=outp)
        f1.record()
        barrier()
        pms = max_over_ranks(f0.elapsed_time(f1), dev, world) / args.steps
        ragged = {"value": tot(ng) / (rms * 1e-3), "unit": "graphs/s", "ms_per_step": rms, "graphs_per_step": int(tot(ng)),
                  "buckets": len(groups), "mean_words": float(lengths.float().mean()),
                  "node_updates_per_s_padded_rows": tot(padded_nodes) * T / (rms * 1e-3),
                  "node_updates_per_s_real_nodes": tot(real_nodes) * T / (rms * 1e-3),
                  "adjacency_bytes_per_step": int(tot(sum(a.numel() * 4 for a, _ in groups))),
                  "note": "C2 sentence-length distribution, even bucket sizes 4..128, one dense FP32 batch per bucket "
                          "(v < 128: generic gather route, no word-order fold)",
                  "as_one_packed_edge_list": {
                      "value": tot(ng) / (pms * 1e-3), "unit": "graphs/s", "ms_per_step": pms,
                      "node_updates_per_s_padded_rows": tot(padded_nodes) * T / (pms * 1e-3),
                      "node_updates_per_s_real_nodes": tot(real_nodes) * T / (pms * 1e-3),
                      "edges_per_step": int(tot(edges_p.shape[0])),
                      "note": "same length distribution, rows padded to the bucket size like the dense batches, handed over as "
                              "(src, channel, dst) triples in ONE ggnn_propagate_edges call (no dense adjacency, no per-bucket launches)"}}
        del edges_p, hp_, outp
        del groups

    if rank == 0:
        steps_g = stage_steps * graphs_rank                # graph passes in the per-stage pass on this rank
        st_ms = {k: v[0] for k, v in stages.items()}
        st_ln = {k: v[1] for k, v in stages.items()}
        tf32_peak = peaks["bf16_tflops_sustained"] if precision == "f16" else \
            peaks.get("tf32_tflops_sustained", peaks["bf16_tflops_sustained"] / 2.0)            # kind::f16 runs at the BF16 rate
        cand = {
            "compact_mask": ("hbm", steps_g * BYTES_COMPACT_PER_GRAPH / 1e9, "GB/s", peaks["hbm_gbs"]),
            "agg": ("hbm", steps_g * T * BYTES_AGG_PER_GRAPH_STEP / 1e9, "GB/s", peaks["hbm_gbs"]),
            "msg": ("tensor", steps_g * T * FLOP_MSG_PER_GRAPH_STEP / 1e12, "TFLOP/s", tf32_peak),
            "gru_gates": ("tensor", steps_g * T * FLOP_GATES_PER_GRAPH_STEP / 1e12, "TFLOP/s", tf32_peak),
            "gru_cand": ("tensor", steps_g * T * FLOP_CAND_PER_GRAPH_STEP / 1e12, "TFLOP/s", tf32_peak),
        }
        if st_ms.get("gru_cand", 0) == 0 and st_ms.get("gru_gates", 0) > 0:
            # fused aggregation + GRU kernel: one launch does the gate and candidate contractions
            b_, w_, u_, p_ = cand["gru_gates"]
            cand["gru_gates"] = (b_, w_ + cand["gru_cand"][1], u_, p_)
        per_stage = {}
        for k, (bound, work, unit, peak) in cand.items():
            if st_ms.get(k, 0) > 0:
                ach = work / (st_ms[k] * 1e-3)
                per_stage[k] = {"bound": bound, "ms_per_step": st_ms[k] / stage_steps,
                                "launches_per_step": st_ln[k] / stage_steps, "achieved": ach,
                                "unit": unit, "peak": peak, "frac": ach / peak}
        if st_ms.get("agg", 0) == 0 and "gru_gates" in per_stage and st_ms.get("gru_cand", 0) == 0:
            # the aggregation runs inside the fused GRU kernel: report it on SURVEY 8(d)'s algorithmic bytes
            # (FP32 message rows) against the HBM peak, as the survey asks for fused kernels, next to the bytes
            # the kernel really moves (FP16 messages + FP16 state in and out)
            agg_gb = steps_g * T * BYTES_AGG_PER_GRAPH_STEP / 1e9
            moved_gb = steps_g * T * (NNZ_PER_GRAPH * 128 * 2 + 2 * V * 128 * 2) / 1e9
            sec = st_ms["gru_gates"] * 1e-3
            per_stage["gru_gates"]["aggregation_fused"] = {
                "bound": "hbm", "achieved_on_algorithmic_bytes": agg_gb / sec, "achieved_on_moved_bytes": moved_gb / sec,
                "unit": "GB/s", "peak": peaks["hbm_gbs"], "frac_algorithmic": agg_gb / sec / peaks["hbm_gbs"],
                "frac_moved": moved_gb / sec / peaks["hbm_gbs"],
                "note": "algorithmic bytes = nnz*h*4 + nnz*4 + (v+1)*4 + v*h*4 per graph-step (FP32 messages, SURVEY 8d); "
                        "the kernel stores messages and states as FP16, so it moves about half of that"}
        for k in ("compact_rest", "plan", "other"):
            per_stage[k] = {"ms_per_step": st_ms.get(k, 0) / stage_steps,
                            "launches_per_step": st_ln.get(k, 0) / stage_steps}
        dom = max(cand, key=lambda k: st_ms.get(k, 0))
        d = per_stage[dom]
        roofline = {"kernel": dom, "bound": d["bound"], "achieved": d["achieved"], "peak": d["peak"],
                    "unit": d["unit"], "frac": d["frac"],
                    "traffic": (TRAFFIC_COMPACT_PER_GRAPH * steps_g / max(st_ln[dom], 1)) if dom == "compact_mask" else None,
                    "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per graph of one ncu --set full "
                                      "capture (profiles/r1_compact_ncu.md) x graphs per launch" if dom == "compact_mask" else None,
                    "note": ("read-dominated stream (1/32 written): it can exceed the measured COPY bandwidth used as "
                             "peak, which moves read + write; ncu reports 83.8 % of the DRAM peak for this kernel")
                            if dom == "compact_mask" else None,
                    "algorithmic_bytes_per_launch": (BYTES_COMPACT_PER_GRAPH * steps_g / max(st_ln[dom], 1)) if dom == "compact_mask" else None,
                    "avg_launch_ms": st_ms[dom] / max(st_ln[dom], 1),
                    "peak_source": peaks["source"] + ("" if d["bound"] == "hbm" else
                                                      ("; kind::f16 peak = the measured sustained BF16 figure" if precision == "f16" else
                                                       "; TF32 peak taken as half the sustained BF16 figure"))}
        line = {
            "metric": "GGNN propagation graphs/sec (T=8 steps)", "value": value, "unit": "graphs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": {"fp32": "f32", "tf32": "tf32-in/f32-acc", "f16": "f16-in/f32-acc"}[precision], "data": "synthetic",
            "config": {"workload": "C3 synthetic dependency trees v=128 2E=92 h=128 T=8, dense FP32 "
                                   "adjacency -> compaction -> propagation",
                       "global_graphs": args.graphs, "graphs_per_gpu": graphs_rank,
                       "distinct_adjacency_graphs_per_gpu": pool, "precision": precision,
                       "parallelism": f"graph-sharded x{world}, no forward collective",
                       "cache": "inputs (>=24 GB per pass) far exceed the 126 MB L2; no flush needed"},
            "node_updates_per_s": value * V * T,
            "node_updates_per_s_real_nodes": value * V * T,      # every C3 graph has 127 words + ROOT = 128 real nodes
            "ragged_secondary": ragged,
            "clocks": clocks, "e2e": e2e, "e2e_edges": e2e_edges, "other_modes": other_modes, "other_inputs": other_inputs,
            "gpu_launches": int(launches),
            "roofline": roofline, "stages": per_stage,
            "stages_note": "per-stage times come from one extra pass with the stages serialised on the stream (no "
                           "prefetch), each timed alone by CUDA events; the headline is the overlapped pipeline",
            "path_roofline": {"bound": "hbm", "achieved": (TWO_E * V * V * 4 + 2 * V * H * 4) * value / 1e9,
                              "peak": peaks["hbm_gbs"], "unit": "GB/s",
                              "frac": (TWO_E * V * V * 4 + 2 * V * H * 4) * value / 1e9 / peaks["hbm_gbs"],
                              "note": "whole path on its algorithmic bytes (SURVEY 8d): adjacency + h0 + h_T = 6.16 MB per graph"},
            "prefetch": prefetch,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        return line
    return None


# ---------------------------------------------------------------------------------------
# secondary workload: BASELINE config 2 (WSJ-shaped evaluation: embed -> propagate -> heads -> decode -> LAS/UAS)
# ---------------------------------------------------------------------------------------
def measure_c2(args, env):
    import ggnn_b200
    from ggnn_b200 import parser_eval as PE, synth
    world, rank, local, dev = env
    n_sent, Hc, Ec, E_out, Tc = args.c2_sentences, 400, 46, 45, 4
    rng = np.random.default_rng(0)
    r = lambda *s_: (np.sqrt(6.0 / (s_[-2] + s_[-1])) * rng.uniform(-1, 1, s_)).astype(np.float32)
    wd = synth.make_weights(2 * Ec, Hc, 0)
    w = ggnn_b200.from_dict(wd, {"num_timesteps": Tc, "hidden_size": Hc})
    heads = PE.ParserHeads(loc_embeddings=r(150, 80), pos_embeddings=r(48, 50), word_embeddings=r(20000, 100),
                           head_W=r(2 * Hc, 150) * 4, head_b=np.zeros(150, np.float32),
                           label_W=r(2 * Hc, E_out) * 4, label_b=np.zeros(E_out, np.float32))
    recs = synth.wsj_like_corpus(n_sent, Ec, E_out, pos_size=48, vocab=20000, seed=rank)
    ev = PE.ParserEvaluator(w, heads, dev, num_timesteps=Tc)

    def timed(precision, batch):
        ev.evaluate(recs, batch_size=batch, precision=precision)             # warm-up: the same pass (arenas grow to their final size)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res, per = ev.evaluate(recs, batch_size=batch, precision=precision)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0, dev, world)
        return {"value": n_sent * world / dt, "unit": "sentences/s", "seconds": dt, "LAS": res["LAS"], "UAS": res["UAS"]}, per

    runs = {}
    strict20, per_strict = timed("fp32", 20)
    runs["fp32_batch20"] = strict20
    runs["f16_batch20"], per_f16 = timed("f16", 20)
    runs["f16_batch512"], per_f16b = timed("f16", 512)
    runs["fp32_batch512"], per_strictb = timed("fp32", 512)
    if rank != 0:
        return None
    line = {"metric": "WSJ-shaped evaluation sentences/sec (embed -> propagate T=4 -> heads -> decode -> LAS/UAS)",
            "value": runs["f16_batch512"]["value"], "unit": "sentences/s", "n_gpus": world, "higher_is_better": True,
            "scaling": "weak", "data": "synthetic",
            "config": {"workload": "C2 synthetic WSJ twin (SURVEY 8(d)): lognormal lengths, 45 labels, 2E=92, h=400, T=4, "
                                   "buckets of even sizes like the reference; edges fed as triples (N3); host decode + scorer",
                       "sentences_per_gpu": n_sent, "note": "WSJ data / pickle not present (LDC-licensed, SURVEY F6)"},
            "runs": runs,
            "f16_sentences_scoring_differently_from_strict": int((np.abs(per_f16 - per_strict).max(axis=1) > 0).sum()),
            "batch_size_changes_any_score": bool(not np.array_equal(np.sort(per_strictb, 0), np.sort(per_strict, 0)))}
    if world == 1 and not args.no_cpu:
        # the oracle pipeline (CPU, NumPy FP64 propagation) on a small sample: CPU baseline AND the identity check
        from oracle import ggnn_oracle as O
        from oracle import parser_oracle as P
        n = 40
        sample = recs[:n]
        t0 = time.perf_counter()
        per = []
        for v, items in PE.make_batches(sample, 20):
            idx, edges, n_active, _ = PE.batch_arrays(v, items, Ec)
            h0 = P.initial_node_representation(idx, heads.loc_embeddings, heads.pos_embeddings, heads.word_embeddings, Hc)
            adj = np.stack([O.graph_to_adj_mat_bd(rr["graph"], v, Ec) for rr in items])
            hT = O.propagate_dense(h0, adj, wd, Tc)
            ph, pe = P.heads(hT, h0, heads.head_W, heads.head_b, heads.label_W, heads.label_b)
            masks = np.stack([P.get_mask(int(k), v, E_out, PE.OUTPUT_SIZE) for k in n_active])
            masks_e = np.stack([P.get_mask(int(k), v, E_out, PE.OUTPUT_SIZE, is_edge=True) for k in n_active])
            labels = [P.target_to_adj_mat(rr["targets"], v, E_out, v) for rr in items]
            sc, _ = P.score_batch(P.target_values(labels, PE.OUTPUT_SIZE), ph, masks, P.target_values_edges(labels), pe,
                                  masks_e, v, E_out, PE.OUTPUT_SIZE)
            per.append(sc)
        dt = time.perf_counter() - t0
        ref = np.concatenate(per)
        _, got = ev.evaluate(sample, batch_size=20, precision="fp32")
        _, got16 = ev.evaluate(sample, batch_size=20, precision="f16")
        line["cpu_baseline"] = {"value": n / dt, "unit": "sentences/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": f"{n} sentences through the oracle pipeline (NumPy FP64 propagation + the reference's decode / scorer restated)"}
        line["per_sentence_scores_identical_to_oracle"] = {"fp32": bool(np.array_equal(ref, got)), "f16": bool(np.array_equal(ref, got16)),
                                                           "sentences": n}
    return line


# ---------------------------------------------------------------------------------------
# secondary workload: BASELINE config 5 (training step of the path, data parallel)
# ---------------------------------------------------------------------------------------
def c5_inputs(n_graphs, seed, device):
    """C2-shaped synthetic trees: WSJ-like lengths, 2E = 92, rows = bucket size."""
    from ggnn_b200 import synth
    lengths = synth.wsj_like_lengths(n_graphs, seed=seed)
    tri = synth.random_trees(n_graphs, int(lengths.max()), E, seed=seed, lengths=lengths)
    rows = torch.tensor([synth.bucket_size(int(l)) for l in lengths])
    edges, off = synth.packed_edge_list(tri, rows, E)
    N = int(off[-1])
    g = torch.Generator().manual_seed(seed)
    h0 = (torch.rand(N, 400, generator=g) - 0.5) * 0.2
    d = torch.randn(N, 400, generator=g)
    return h0.to(device), edges.to(device), d.to(device), N


def measure_c5(args, env, precision=None):
    import torch.distributed as dist
    import ggnn_b200
    from ggnn_b200 import dist as gdist, synth
    world, rank, local, dev = env
    Hc, Tc, per_gpu = 400, 4, args.c5_graphs
    wd = synth.make_weights(TWO_E, Hc, 0)
    prop = ggnn_b200.Propagator(ggnn_b200.from_dict(wd), dev)
    h0, edges, d, N = c5_inputs(per_gpu, 1000 + rank, dev)
    grads = torch.zeros(prop.grad_floats, device=dev)
    prec = precision or args.precision      # "f16": forward on the FP16 kernels (h = 400 -> wide16.cu) with the FP32 tape, backward TF32

    step_no = [0]

    def step(mode=None):
        # keep-prob 0.9 in the timed run; fresh Philox masks every step, drawn on the device
        drop = None
        if args.c5_keep < 1.0:
            step_no[0] += 1
            drop = {"keep": args.c5_keep, "seed": 1000003 * (rank + 1) + step_no[0]}
        return gdist.train_step(prop, h0, edges, lambda hT: d, Tc, grads=grads, optimizer={}, precision=mode or prec,
                                dropout=drop)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    # The step is ~400 launches in 12 ms: host threads still spinning after a CPU-baseline leg of the previous
    # sub-record (torch's intra-op pool) once stretched this timed region to 15.9 ms (profiles/r2_bench_i.json);
    # the pool goes to sleep within a few hundred milliseconds.  Outside the timed region.
    time.sleep(0.5)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = prop.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    launches = prop.launches - l0
    if world > 1:
        tms = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    ms_per_step = ms / args.steps
    nnz = int(edges.shape[0])
    other = None
    if prec != "fp32" and not args.no_fp32:            # the same training step in the other arithmetic modes, same inputs
        other = {}
        for mode in [m for m in ("tf32", "fp32") if m != prec]:
            step(mode); barrier()
            e0.record()
            for _ in range(max(1, args.steps // 2)):
                step(mode)
            e1.record(); barrier()
            oms = e0.elapsed_time(e1) / max(1, args.steps // 2)
            if world > 1:
                t2 = torch.tensor([oms], device=dev, dtype=torch.float64)
                dist.all_reduce(t2, op=dist.ReduceOp.MAX)
                oms = float(t2.item())
            other[mode] = {"value": per_gpu * world / (oms * 1e-3), "unit": "graphs/s", "ms_per_step": oms}
    # data-parallel replicas must hold bit-identical weights after the timed steps: the all-reduced gradient is the
    # same on every rank and the clip norm is reduced in a fixed order (train_fp32.cu)
    replicas_identical = None
    if world > 1:
        wts = prop.get_weights()
        flat = torch.cat([t.reshape(-1) for t in wts.values() if t is not None]).view(torch.int32).long()
        chk = torch.stack([flat.sum(), (flat * (torch.arange(flat.numel(), device=dev) % 8191 + 1)).sum()])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        replicas_identical = bool(torch.equal(lo, hi))
    if rank == 0:
        flop = 3 * Tc * (nnz * 2 * Hc * Hc + N * 12 * Hc * Hc)       # bwd counted as 2x fwd
        line = {"metric": "GGNN propagation training step graphs/sec (fwd+bwd+clip+Adam, T=4)",
                "value": per_gpu * world / (ms_per_step * 1e-3), "unit": "graphs/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {"fp32": "f32", "tf32": "tf32 operands, f32 accumulate / pointwise / optimiser",
                          "f16": "f16 operands forward, tf32 operands backward, f32 accumulate / pointwise / optimiser"}[prec],
                "data": "synthetic", "other_modes": other,
                "config": {"workload": "C5 synthetic WSJ-shaped trees, h=400 2E=92 T=4, fwd+bwd of the path + "
                                       "one all-reduce of the 62.9 MB flat weight gradient + per-variable clip_by_norm(1.0) + "
                                       "Adam(0.003) on the ctx-owned weights; edge-weight and state dropout, fresh device-generated (Philox) masks per step, "
                                       f"keep-prob {args.c5_keep}",
                           "graphs_per_gpu": per_gpu, "nodes_per_gpu": N, "edges_per_gpu": nnz,
                           "parallelism": f"data parallel x{world}, NCCL all-reduce (sum) of {prop.grad_floats} floats"},
                "achieved_tflops_algorithmic_per_gpu": flop / (ms_per_step * 1e-3) / 1e12,
                "replica_weights_identical_after_timed_steps": replicas_identical,
                "clocks": clocks, "gpu_launches": int(launches)}
        if world == 1 and not args.no_cpu:
            from oracle import grad_oracle as G
            n = 64
            h0c, ec, dc, _ = c5_inputs(n, 7, "cpu")
            torch.set_num_threads(os.cpu_count() or 1)
            t0 = time.perf_counter()
            G.grads(h0c.numpy(), ec.numpy(), wd, Tc, dc.numpy(), dtype=torch.float32)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "graphs/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"{n} C5 graphs, torch-CPU FP32 autograd over the edge-list restatement"}
        return line
    return None


# ---------------------------------------------------------------------------------------
# secondary workload: BASELINE config 4 (large graphs through the edge-list path)
# ---------------------------------------------------------------------------------------
C4_WEIGHT_SCALE = 0.25


def measure_c4(args, env, precision=None):
    import torch.distributed as dist
    import ggnn_b200
    from ggnn_b200 import synth
    world, rank, local, dev = env
    Vn, CH, BASE, Hc, Tc = 4096, 8, 32768, 256, 8
    per_gpu = args.c4_graphs
    prec = precision or args.precision
    # In-degree 16 with Glorot edge weights drives the pre-activations to O(10): the recurrence then amplifies ANY
    # 11-bit operand rounding (emulated on the CPU, pure TF32 rounding alone gives ~1e-2 after 8 steps), so the
    # 2e-3 modes are benchmarked -- and tested, tests/test_gpu_f16.py -- with the edge weights scaled by 0.25
    # (pre-activations O(1), as a degree-normalised aggregation would give).  `parity` below is the error measured
    # on exactly these inputs; the strict-FP32 figure stands beside it.
    wd = synth.make_weights(CH, Hc, 0)
    wd["edge_weights"] = wd["edge_weights"] * np.float32(C4_WEIGHT_SCALE)
    prop = ggnn_b200.Propagator(ggnn_b200.from_dict(wd), dev)
    edges = synth.random_multigraph(per_gpu, Vn, BASE, CH // 2, seed=1000 + rank).to(dev)
    gen = torch.Generator(device=dev).manual_seed(1 + rank)
    h0 = (torch.rand(per_gpu * Vn, Hc, device=dev, generator=gen) - 0.5) * 0.2
    out = torch.empty_like(h0)

    def step(mode=None):
        prop.propagate_edges(h0, edges, num_timesteps=Tc, precision=mode or prec, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    prop.profile(True)
    l0 = prop.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_per_step = max_over_ranks(e0.elapsed_time(e1), dev, world) / args.steps
    clocks = sampler.stop()
    launches = prop.launches - l0
    stages = prop.profile_read()
    prop.profile(False)
    ncheck = 2
    got = {prec: out[:ncheck * Vn].clone()}
    other = {}
    if not args.no_fp32:
        for mode in [m for m in ("f16", "tf32", "fp32") if m != prec]:
            for _ in range(max(1, args.warmup)):
                step(mode)
            barrier()
            e0.record()
            for _ in range(args.steps):
                step(mode)
            e1.record()
            barrier()
            oms = max_over_ranks(e0.elapsed_time(e1), dev, world) / args.steps
            other[mode] = {"value": per_gpu * world / (oms * 1e-3), "unit": "graphs/s", "ms_per_step": oms,
                           "steps": args.steps}
            got[mode] = out[:ncheck * Vn].clone()
    # distinct (source, channel) message rows per graph: what the aggregation has to fetch at least once
    key = edges[:, 0].long() * CH + edges[:, 1].long()
    uniq_rows = int(torch.unique(key).numel()) / per_gpu
    if rank == 0:
        flop_graph = Tc * (min(2 * BASE, Vn * CH) * 2 * Hc * Hc + Vn * 12 * Hc * Hc)     # 60.13 GFLOP
        val = per_gpu * world / (ms_per_step * 1e-3)
        line = {"metric": "GGNN propagation graphs/sec (T=8 steps)", "value": val, "unit": "graphs/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {"fp32": "f32", "tf32": "tf32-in/f32-acc", "f16": "f16-in/f32-acc"}[prec], "data": "synthetic",
                "config": {"workload": "C4 synthetic large graphs v=4096, 8 channels, nnz=65536/graph, h=256, T=8, edge-list path",
                           "graphs_per_gpu": per_gpu, "precision": prec,
                           "edge_weight_scale": C4_WEIGHT_SCALE,
                           "edge_weight_note": "Glorot edge weights x 0.25: with in-degree 16 the unscaled recurrence amplifies "
                                               "any 11-bit operand rounding past 2e-3 (CPU-emulated TF32 rounding alone: ~1e-2); "
                                               "the scaled inputs are the ones the parity tests validate"},
                "node_updates_per_s": val * Vn * Tc,
                "achieved_tflops_algorithmic_per_gpu": flop_graph * per_gpu / (ms_per_step * 1e-3) / 1e12,
                "stages_ms_per_step": {k: v[0] / args.steps for k, v in stages.items() if v[0] > 0},
                "other_modes": other, "clocks": clocks, "gpu_launches": int(launches)}
        # per-stage roofline fractions on SURVEY 8(d)'s algorithmic work (per graph-step): W_e GEMM min(nnz, N*2E)*2h^2
        # flops, GRU N*12h^2 flops; aggregation on the UNIQUE bytes it has to move: distinct message rows * h * s_msg +
        # nnz*4 (index) + (N+1)*4 + N*h*s_out (a row gathered twice is one fetch from HBM and one L2 hit)
        peaks = load_peaks()
        sm = line["stages_ms_per_step"]
        gsteps = per_gpu * Tc
        tpeak = peaks["bf16_tflops_sustained"] if prec != "tf32" else peaks.get("tf32_tflops_sustained", peaks["bf16_tflops_sustained"] / 2.0)
        def tens(ms, flop):
            a = flop * gsteps / (ms * 1e-3) / 1e12
            return {"bound": "tensor", "ms_per_step": ms, "achieved": a, "unit": "TFLOP/s", "peak": tpeak, "frac": a / tpeak}
        st = {}
        if prec != "fp32":
            if "msg" in sm:
                st["msg"] = tens(sm["msg"], min(2 * BASE, Vn * CH) * 2 * Hc * Hc)
            gru_ms = sm.get("gru_gates", 0.0) + sm.get("gru_cand", 0.0)
            if gru_ms > 0:
                st["gru"] = tens(gru_ms, Vn * 12 * Hc * Hc)
        if "agg" in sm:
            s_msg = 4 if prec == "fp32" else 2
            s_out = 2 if prec == "f16" else 4
            by = uniq_rows * Hc * s_msg + 2 * BASE * 4 + (Vn + 1) * 4 + Vn * Hc * s_out
            a = by * gsteps / (sm["agg"] * 1e-3) / 1e9
            st["agg"] = {"bound": "hbm", "ms_per_step": sm["agg"], "achieved": a, "unit": "GB/s", "peak": peaks["hbm_gbs"],
                         "frac": a / peaks["hbm_gbs"], "unique_message_rows_per_graph": uniq_rows,
                         "gathered_message_rows_per_graph": 2 * BASE,
                         "note": "UNIQUE bytes: every distinct (source, channel) message row counted once"}
        line["stages"] = st
        line["peak_source"] = peaks["source"]
        if world == 1 and not args.no_cpu:
            # CPU port on the first graphs of the SAME inputs; its result doubles as the parity check of every mode timed
            from oracle import cpu_reference as R
            torch.set_num_threads(os.cpu_count() or 1)
            n = ncheck
            ec = edges[:n * 2 * BASE].cpu()
            hc = h0[:n * Vn].cpu()
            with torch.no_grad():
                t0 = time.perf_counter()
                ref = R.propagate_edges_torch_cpu(hc, ec, wd, Tc)
                dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "graphs/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"the first {n} C4 graphs of the benched inputs, torch-CPU FP32 edge-list restatement"}
            ref = ref.numpy() if hasattr(ref, "numpy") else np.asarray(ref)
            scale = float(np.abs(ref).max())
            line["parity"] = {m: float(np.abs(g.cpu().numpy() - ref).max() / scale) for m, g in got.items()}
            line["parity"]["note"] = ("norm-wise max|out - ref| / max|ref| of each timed mode on the benched inputs against the "
                                      "FP32 CPU port (tolerances: fp32 1e-5, tf32 / f16 2e-3)")
        return line
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="f16", choices=["fp32", "tf32", "f16"])
    ap.add_argument("--graphs", type=int, default=TOTAL_GRAPHS, help="global batch (all GPUs)")
    ap.add_argument("--pool", type=int, default=4096, help="distinct dense adjacency graphs resident per GPU")
    ap.add_argument("--e2e-graphs", type=int, default=256)
    ap.add_argument("--ragged-graphs", type=int, default=8192, help="graphs per GPU of the ragged secondary C3 run")
    ap.add_argument("--no-secondary", action="store_true", help="c3: skip the c4 / c5 sub-records")
    ap.add_argument("--ref-graphs", type=int, default=64, help="graphs per step of the CPU reference arm")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-prefetch", action="store_true", help="c3: do not announce the next call's adjacency (A-B switch)")
    ap.add_argument("--no-fp32", action="store_true", help="skip the passes in the other arithmetic modes (c3 workload)")
    ap.add_argument("--c2-sentences", type=int, default=2416, help="sentences per GPU of the c2 sub-record (WSJ test set size)")
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5"],
                    help="c3 = the headline forward benchmark; c5 = training step (secondary line)")
    ap.add_argument("--c4-graphs", type=int, default=64, help="graphs per GPU of the c4 workload")
    ap.add_argument("--c5-graphs", type=int, default=2048, help="graphs per GPU of the c5 workload")
    ap.add_argument("--c5-keep", type=float, default=0.9, help="dropout keep-prob of the c5 timed run (1.0 = no masks)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch.distributed as dist
    env = init_env()
    world, rank = env[0], env[1]
    if args.workload == "c2":
        line = measure_c2(args, env)
    elif args.workload == "c5":
        line = measure_c5(args, env)
    elif args.workload == "c4":
        line = measure_c4(args, env)
    else:
        line = measure_c3(args, env)
        if not args.no_secondary:
            # the other two GPU configurations of BASELINE.json ride on the same line (same --steps / --warmup):
            # c4 = large graphs through the edge-list path, c5 = training step (at N > 1: the NCCL all-reduce run)
            torch.cuda.empty_cache()
            c4 = measure_c4(args, env, precision="f16")
            torch.cuda.empty_cache()
            c5 = measure_c5(args, env, precision="f16")
            torch.cuda.empty_cache()
            c2 = measure_c2(args, env)
            if rank == 0:
                line["c2"], line["c4"], line["c5"] = c2, c4, c5
                line["gpu_launches"] += c4["gpu_launches"] + c5["gpu_launches"]
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
