"""Host-side multi-process logic (gloo, world_size 2, CPU): graph sharding and the gradient
all-reduce used by the training configuration."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ggnn_b200 import dist as gdist


def test_shard_range_partitions():
    for n in (0, 1, 7, 8, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            blocks = [gdist.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for (a0, a1), (b0, b1) in zip(blocks, blocks[1:]):
                assert a1 == b0 and a0 <= a1
            sizes = [e - b for b, e in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        gdist.shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = gdist.shard_range(n_items, rank, world)
        # "gradient" of this shard: per-item contributions summed locally
        rng = np.random.default_rng(0)
        contrib = torch.from_numpy(rng.standard_normal((n_items, 33)).astype(np.float32))
        flat = contrib[b:e].sum(0)
        gdist.allreduce_grads(flat)
        mean = contrib[b:e].sum(0)
        gdist.allreduce_grads(mean, average=True)
        torch.save({"flat": flat, "mean": mean, "range": (b, e)}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_allreduce_grads_gloo_world2(tmp_path):
    world, n_items = 2, 11
    mp.spawn(_worker, args=(world, _free_port(), n_items, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    rng = np.random.default_rng(0)
    full = torch.from_numpy(rng.standard_normal((n_items, 33)).astype(np.float32)).sum(0)
    for r in res:
        assert torch.allclose(r["flat"], full, atol=1e-5)
        assert torch.allclose(r["mean"], full / world, atol=1e-5)
    assert res[0]["range"][1] == res[1]["range"][0]


def test_allreduce_is_noop_without_process_group():
    t = torch.ones(5)
    assert gdist.allreduce_grads(t) is t and torch.equal(t, torch.ones(5))
