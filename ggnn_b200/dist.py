"""Multi-GPU plumbing for the path: graph sharding (forward needs no collective) and the one
collective of the training configuration -- a sum all-reduce of the flat weight-gradient buffer
(SURVEY.md 8(e)).  torch.distributed is used as-is: NCCL over NVLink on the GPU box, gloo in the
CPU tests.  The reference has no equivalent (its MirroredStrategy is constructed and never used,
chem_tensorflow_dense.py:1327-1328)."""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [begin, end) block of `n_items` graphs for `rank`; sizes differ by at most one,
    earlier ranks take the larger blocks.  Blocks are disjoint and cover range(n_items)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(int(n_items), world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def allreduce_grads(flat: torch.Tensor, average: bool = False, group=None) -> torch.Tensor:
    """In-place sum (or mean) of the flat gradient buffer [dW | dB | dWg | dbg | dWc | dbc] over
    all ranks: ONE collective per training step.  No-op when torch.distributed is not initialised."""
    if not (dist.is_available() and dist.is_initialized()):
        return flat
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat /= dist.get_world_size(group)
    return flat


def train_step(prop, h0: torch.Tensor, edges: torch.Tensor, d_hT_fn, num_timesteps: int,
               grads: torch.Tensor = None, average: bool = False, optimizer: dict = None,
               precision: str = "fp32", dropout=None):
    """One data-parallel step of the path on this rank's shard: forward with the tape, the caller's
    upstream gradient d_hT_fn(h_T) -> dL/dh_T, backward, all-reduce of the weight gradients and -- when
    `optimizer` is a dict of Propagator.optimizer_step keywords ({} = the reference's defaults, lr .003 /
    clip 1.0, chem_tensorflow.py:483-506) -- the clip + Adam update, identical on every rank because the
    all-reduced gradient is.  `precision` "tf32" runs forward and
    backward on the tensor cores (gradients within 2e-3 of FP64).  Returns (h_T, d_h0, flat grads)."""
    if isinstance(dropout, dict):    # {"keep": k, "seed": s}: masks drawn on the device (Philox), A8, this step only
        prop.set_dropout_seed(dropout["keep"], dropout["seed"], dropout.get("drop_weights", True),
                              dropout.get("drop_state", True))
    elif dropout is not None:        # (w_mask [T,2E,h,h] | None, state_mask [T,N,h] | None, keep): supplied masks
        prop.set_dropout(dropout[0], dropout[1], dropout[2], num_timesteps)
    hT = prop.propagate_edges(h0, edges, num_timesteps=num_timesteps, precision=precision, save_tape=True)
    d_hT = d_hT_fn(hT)
    if grads is not None:
        grads.zero_()
    d_h0, grads = prop.backward(d_hT, grads)
    allreduce_grads(grads, average=average)
    if optimizer is not None:
        prop.optimizer_step(grads, **optimizer)
    return hT, d_h0, grads
