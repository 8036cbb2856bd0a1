"""Two-GPU NCCL tests of the data-parallel DDPG update: sharding the replay
minibatch by graph and (a) all-reducing the flat gradients, (b) reduce-scattering them into a
1/W-sharded fused Adam + soft update followed by an all-gather, or (c) the forked-stream update with
per-layer gradient all-reduces on the side stream must all reproduce the single-GPU
update of the full minibatch.  Skipped on boxes with one GPU (run with ``gpurun --gpus 2``)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _build(spec, dev, seed=11):
    from gsc_b200.models import Actor, QNetwork
    torch.manual_seed(seed)
    nets = [cls(spec) for cls in (Actor, QNetwork, Actor, QNetwork)]
    nets[2].load_state_dict(nets[0].state_dict())
    nets[3].load_state_dict(nets[1].state_dict())
    return [n.to(dev) for n in nets]


def _worker(rank, world, port, n_max, B, steps, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from gsc_b200 import ddpg, dp, synth
    from gsc_b200.graph import GraphBatch
    from gsc_b200.models import AgentSpec
    dp.init_from_env("nccl")
    dev = torch.device("cuda", rank)
    spec = AgentSpec(n_max)
    full = synth.make_transitions(B, n_max, max(2, n_max // 2), n_max, seed=3, num_topologies=5)
    mine = dp.shard_transitions(*full, rank, world)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    results = {}
    for mode in ("allreduce", "sharded", "forked", "bucketed"):
        nets = _build(spec, dev)
        upd = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2, world_size=world, rank=rank,
                               shard_optimizer=mode == "sharded", global_batch=B, fork_streams=mode in ("forked", "bucketed"), bucket_allreduce=mode == "bucketed")
        cur, nxt = GraphBatch.from_states(mine[0], dense_mask=False).to(dev), GraphBatch.from_states(mine[1], dense_mask=False).to(dev)
        for _ in range(steps):
            qv, ql, al = upd.update(cur.view(), nxt.view(), t(mine[2]), t(mine[3]), t(mine[4]))
        losses = torch.stack([ql.reshape(()), al.reshape(())])
        dist.all_reduce(losses)                                  # the ranks' shares of the mean losses
        results[mode] = ([n.flat.clone() for n in nets], losses)
        if mode == "sharded":
            assert upd.m_actor.numel() * world == nets[0].flat.numel(), "Adam moments must be sharded"
    # replicas must hold identical weights after the exchange
    for mode, (flats, _) in results.items():
        for f in flats:
            other = f.clone()
            dist.broadcast(other, src=0)
            assert torch.equal(other, f), f"{mode}: replicas diverged"
    if rank == 0:
        nets = _build(spec, dev)
        upd = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2)
        cur, nxt = GraphBatch.from_states(full[0], dense_mask=False).to(dev), GraphBatch.from_states(full[1], dense_mask=False).to(dev)
        for _ in range(steps):
            qv, ql, al = upd.update(cur.view(), nxt.view(), t(full[2]), t(full[3]), t(full[4]))
        ref_losses = torch.stack([ql.reshape(()), al.reshape(())])
        errs = {}
        for mode, (flats, losses) in results.items():
            e = max(float((a - b.flat).abs().max() / b.flat.abs().max()) for a, b in zip(flats, nets))
            if e > 1e-4:                                        # name the parameter tensors that differ
                for a, b in zip(flats, nets):
                    for (k, _), va, vb in zip(b.named_parameters(), b.param_views(a), b.param_views(b.flat)):
                        d = float((va - vb).abs().max())
                        if d > 1e-5:
                            print(f"[dp parity] {mode} {type(b).__name__}.{k}: max abs diff {d:.3e}, max |w| {float(vb.abs().max()):.3e}", flush=True)
            # Adam moves every weight by ~lr per step whatever the gradient scale: bound the difference of
            # the applied updates as well
            errs[mode] = (e, float((losses - ref_losses).abs().max() / ref_losses.abs().max()))
        out.put(errs)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("n_max,B", [(12, 14), (40, 9)])
def test_two_gpu_update_equals_single_gpu_update(n_max, B):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29700 + (os.getpid() + n_max) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_max, B, 3, out)) for r in range(2)]
    for p in procs:
        p.start()
    errs = out.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for mode, (e_w, e_l) in errs.items():
        assert e_w < 2e-4, (mode, e_w)        # three Adam steps: sign flips of ~0 gradients move a weight by 2*lr
        assert e_l < 1e-4, (mode, e_l)
