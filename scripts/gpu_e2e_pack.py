"""Development aid: e2e rate of the packed host route (compute_final_node_representations on pinned host tensors) for a few
chunk sizes / packer thread counts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ggnn_b200
from ggnn_b200 import synth, api
ne = int(sys.argv[1]) if len(sys.argv) > 1 else 256
E, v, h, T = 46, 128, 128, 8
prop = ggnn_b200.Propagator(ggnn_b200.from_dict(synth.make_weights(2 * E, h, 0)), "cuda:0")
tri = synth.random_trees(ne, v - 1, E, seed=1)
adj = synth.dense_adjacency(tri, v, E).pin_memory()
h0 = ((torch.rand(ne, v, h) - 0.5) * 0.2).pin_memory()
per_graph = adj[0].numel() * 4
for graphs_per_chunk in (16, 32, 64, 128):
    for threads in (0, 15, 12, 8):
        api._HOST_PACK_CHUNK_BYTES = graphs_per_chunk * per_graph
        api._HOST_PACK_THREADS = threads
        for _ in range(3):
            r = ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
        t0 = time.perf_counter()
        for _ in range(5):
            r = ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
        dt = (time.perf_counter() - t0) / 5
        print(f"{graphs_per_chunk:4d} graphs per chunk, {threads:2d} packer threads: {dt * 1e3:6.2f} ms per {ne} graphs = {ne / dt:8.0f} graphs/s")

print("hybrid (packer + copy engine at once):")
api._HOST_PACK_THREADS = 15
for hyb, pb, db in ((False, 0, 0), (True, 288, 96), (True, 192, 96), (True, 384, 96), (True, 288, 48), (True, 288, 192)):
    api._HOST_HYBRID = hyb
    api._HOST_PACK_CHUNK_BYTES = 128 * per_graph
    if hyb:
        api._HYBRID_PACK_GRAPH_BYTES, api._HYBRID_DENSE_GRAPH_BYTES = pb << 20, db << 20
    for _ in range(3):
        r = ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
    t0 = time.perf_counter()
    for _ in range(5):
        r = ggnn_b200.compute_final_node_representations(h0, adj, weights=prop, num_timesteps=T, precision="f16")
    dt = (time.perf_counter() - t0) / 5
    print(f"hybrid={hyb} packed unit {pb} MB dense unit {db} MB: {dt * 1e3:6.2f} ms per {ne} graphs = {ne / dt:8.0f} graphs/s")
