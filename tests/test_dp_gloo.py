"""world_size-2 gloo tests (CPU) of the data-parallel host logic: graph sharding, the 1/B_global
loss scaling convention and the flat-gradient all-reduce reproduce the single-process gradient."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsc_b200 import dp, synth
from oracle import gsc_oracle as O


def _obs(st):
    return {"x": torch.from_numpy(st.x).double(), "edge_index": torch.from_numpy(st.edge_index),
            "batch": torch.from_numpy(st.batch), "mask": torch.from_numpy(st.mask).double(),
            "num_graphs": st.num_graphs}


def _critic_grad(c_sd, st, actions, y, scale):
    sd = {k: v.clone().requires_grad_(True) for k, v in c_sd.items()}
    q = O.critic_forward(sd, _obs(st), torch.from_numpy(actions).double()).view(-1)
    (scale * ((q - y) ** 2).sum()).backward()
    return torch.cat([v.grad.reshape(-1) for v in sd.values()])


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = dp.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    torch.set_num_threads(1)
    cur, nxt, actions, rewards, dones = synth.make_transitions(7, 8, 3, 8, seed=5, num_topologies=3)
    c_sd = O.init_critic(3, 22, 2, cur.action_dim, seed=9, dtype=torch.float64)
    y_full = torch.from_numpy(rewards).double()
    s_cur, s_nxt, s_act, s_rew, s_don = dp.shard_transitions(cur, nxt, actions, rewards, dones, rank, world)
    lo, hi = dp.shard_bounds(7, rank, world)
    g = _critic_grad(c_sd, s_cur, s_act, y_full[lo:hi], 1.0 / 7)     # scaled by 1/B_global, not 1/B_local
    dp.allreduce_flat(g)
    if rank == 0:
        full = _critic_grad(c_sd, cur, actions, y_full, 1.0 / 7)
        out.put(float((g - full).abs().max() / full.abs().max()))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gradients_allreduce_to_the_full_batch_gradient():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    err = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert err < 1e-12


def test_shard_bounds_and_rebasing():
    assert [dp.shard_bounds(7, r, 2) for r in range(2)] == [(0, 4), (4, 7)]
    assert [dp.shard_bounds(8, r, 4) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
    st = synth.make_states(9, 8, 3, 8, seed=1, num_topologies=4)
    parts = [dp.shard_states(st, r, 4) for r in range(4)]
    assert sum(p.num_graphs for p in parts) == 9
    np.testing.assert_array_equal(np.concatenate([p.x for p in parts]), st.x)
    np.testing.assert_array_equal(np.concatenate([p.mask for p in parts]), st.mask)
    for p in parts:
        if p.num_graphs:
            assert p.ptr[0] == 0 and p.ptr[-1] == p.x.shape[0] and p.batch.min() == 0
            assert p.edge_index.min() >= 0 and p.edge_index.max() < p.x.shape[0]


def _adam(p, g, m, v, step, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    denom = (v.sqrt() / (1 - b2 ** step) ** 0.5).add_(eps)
    p.addcdiv_(m, denom, value=-lr / (1 - b1 ** step))


def _shard_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dp.init_from_env(backend="gloo")
    P = 6720                                                   # one flat-buffer quantum (models._FLAT_MULTIPLE)
    gen = torch.Generator().manual_seed(3)
    p0 = torch.randn(P, generator=gen, dtype=torch.float64)
    grads = [torch.randn(world, P, generator=gen, dtype=torch.float64) for _ in range(3)]   # per step, per rank
    # sharded optimizer: moments exist for this rank's slice only
    p, tgt = p0.clone(), p0.clone()
    lo, hi = dp.shard_range(P, rank, world)
    m, v, gs = torch.zeros(hi - lo, dtype=torch.float64), torch.zeros(hi - lo, dtype=torch.float64), torch.zeros(hi - lo, dtype=torch.float64)
    for step, g in enumerate(grads, 1):
        dp.reduce_scatter_flat(g[rank].clone(), gs)
        _adam(p[lo:hi], gs, m, v, step)
        tgt[lo:hi].mul_(0.99).add_(p[lo:hi], alpha=0.01)
        dp.all_gather_flat(p)
        dp.all_gather_flat(tgt)
    if rank == 0:
        q, qt = p0.clone(), p0.clone()
        mf, vf = torch.zeros(P, dtype=torch.float64), torch.zeros(P, dtype=torch.float64)
        for step, g in enumerate(grads, 1):
            _adam(q, g.sum(0), mf, vf, step)
            qt.mul_(0.99).add_(q, alpha=0.01)
        out.put((float((p - q).abs().max()), float((tgt - qt).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_optimizer_exchange_equals_allreduce_and_full_adam():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    e_p, e_t = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert e_p < 1e-14 and e_t < 1e-14


def test_shard_range_needs_divisible_buffers_and_flat_buffers_are_divisible():
    import pytest
    from gsc_b200.models import _FLAT_MULTIPLE
    assert [dp.shard_range(24, r, 4) for r in range(4)] == [(0, 6), (6, 12), (12, 18), (18, 24)]
    with pytest.raises(ValueError):
        dp.shard_range(25, 0, 4)
    for w in range(1, 9):
        assert _FLAT_MULTIPLE % w == 0 and (_FLAT_MULTIPLE // w) % 4 == 0      # equal, 16-byte aligned shards
