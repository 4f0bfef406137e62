"""Development aid: host-side adjacency bit packing rate on the GPU box's cores (ggnn_host_pack_adjacency)."""
import sys, os, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ggnn_b200 import _lib
l = _lib.lib()
l.ggnn_host_pack_adjacency.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int]
g = 256
n = g * 92 * 128 * 128
a = torch.zeros(n, dtype=torch.float32).pin_memory()
a[::97] = 1.0
bits = torch.zeros(n // 32, dtype=torch.int32).pin_memory()
print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
for th in (1, 2, 4, 8, 16, 32, 0):
    l.ggnn_host_pack_adjacency(a.data_ptr(), 0, n, bits.data_ptr(), th)
    t0 = time.perf_counter()
    for _ in range(4):
        l.ggnn_host_pack_adjacency(a.data_ptr(), 0, n, bits.data_ptr(), th)
    dt = (time.perf_counter() - t0) / 4
    print(f"threads {th:2d}: {n * 4 / dt / 1e9:6.1f} GB/s = {g / dt:8.0f} C3 graphs/s")
d = torch.empty(n, dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(4):
    d.copy_(a, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 4
print(f"H2D of the same tensor: {n * 4 / dt / 1e9:6.1f} GB/s = {g / dt:8.0f} graphs/s")
# both at once: packing while a copy of ANOTHER pinned tensor runs (do they share the memory system gracefully?)
a2 = a.clone().pin_memory()
t0 = time.perf_counter()
for _ in range(4):
    d.copy_(a2, non_blocking=True)
    l.ggnn_host_pack_adjacency(a.data_ptr(), 0, n, bits.data_ptr(), 0)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 4
print(f"pack (all threads) + concurrent H2D: {dt * 1e3:.1f} ms per {g} graphs")
