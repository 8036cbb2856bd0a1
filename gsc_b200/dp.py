"""Data-parallel plumbing: one process per GPU, states sharded by graph,
replicated parameters, one all-reduce per network per update over its flat gradient buffer.

The reference has no distributed path at all; this is the one strategy north_star names.  Backend
is NCCL on GPUs (NVLink 5 / NVSwitch); the same code runs over gloo on CPU tensors for tests.
"""
from __future__ import annotations

import os
from typing import Tuple

import numpy as np
import torch
import torch.distributed as dist

from .synth import StateBatch


def init_from_env(backend: str = "nccl") -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment; initialises the default
    process group when WORLD_SIZE > 1."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def shard_bounds(num_graphs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split of graph indices: the first ``num_graphs % world`` ranks hold one more."""
    base, rem = divmod(num_graphs, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_states(st: StateBatch, rank: int, world: int) -> StateBatch:
    """Rank's share of a collated batch, re-based so that node ids start at 0."""
    lo, hi = shard_bounds(st.num_graphs, rank, world)
    n0, n1 = int(st.ptr[lo]), int(st.ptr[hi])
    sel = (st.edge_index[0] >= n0) & (st.edge_index[0] < n1)
    return StateBatch(x=st.x[n0:n1], edge_index=np.ascontiguousarray(st.edge_index[:, sel] - n0),
                      batch=st.batch[n0:n1] - lo, ptr=st.ptr[lo:hi + 1] - n0, n=st.n[lo:hi],
                      mask=None if st._mask is None else st._mask[lo:hi], n_max=st.n_max, num_sfcs=st.num_sfcs, num_sfs=st.num_sfs)


def shard_transitions(cur, nxt, actions, rewards, dones, rank: int, world: int):
    lo, hi = shard_bounds(cur.num_graphs, rank, world)
    return (shard_states(cur, rank, world), shard_states(nxt, rank, world), actions[lo:hi], rewards[lo:hi],
            dones[lo:hi])


def allreduce_flat(flat_grad: torch.Tensor, group=None) -> torch.Tensor:
    """Sum a network's flat gradient buffer over all ranks, in place."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


# ---- sharded optimizer (scaling caveat) -----------------------------------
# Optimizer traffic (Adam state + target network, 36 B/param) is replicated work under plain data
# parallelism; at N_max = 1000 it dwarfs the per-sample work.  Instead of all-reduce + full Adam,
# every rank reduce-scatters the flat gradient, runs the fused Adam/soft-update on its 1/W slice of
# the parameters, Adam moments and target network, and all-gathers the updated slices: the same
# bytes on the wire as the all-reduce, optimizer traffic and moment memory divided by W.
def shard_range(numel: int, rank: int, world: int) -> Tuple[int, int]:
    """Equal contiguous slices of a flat buffer (its length is a multiple of the world size)."""
    if numel % world:
        raise ValueError(f"flat buffer of {numel} floats does not split into {world} equal shards")
    chunk = numel // world
    return rank * chunk, (rank + 1) * chunk


def reduce_scatter_flat(flat_grad: torch.Tensor, out: torch.Tensor, group=None) -> torch.Tensor:
    """out = this rank's slice of the sum of ``flat_grad`` over all ranks."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    lo, hi = shard_range(flat_grad.numel(), rank, world)
    if flat_grad.is_cuda:
        dist.reduce_scatter_tensor(out, flat_grad, op=dist.ReduceOp.SUM, group=group)
    else:                       # gloo has no reduce-scatter: same result through an all-reduce
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
        out.copy_(flat_grad[lo:hi])
    return out


def all_gather_flat(flat: torch.Tensor, group=None) -> torch.Tensor:
    """Every rank's slice of ``flat`` (updated in place) is broadcast into the others' copies."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    lo, hi = shard_range(flat.numel(), rank, world)
    if flat.is_cuda:
        dist.all_gather_into_tensor(flat, flat[lo:hi], group=group)     # in place: slice r of the output is the input
    else:
        parts = [torch.empty(hi - lo, dtype=flat.dtype) for _ in range(world)]
        dist.all_gather(parts, flat[lo:hi].clone(), group=group)
        for r, part in enumerate(parts):
            flat[r * (hi - lo):(r + 1) * (hi - lo)].copy_(part)
    return flat
