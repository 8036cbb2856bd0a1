"""DDPG update on the CUDA path -- the gradient-step block of the reference training loop.

``DDPGUpdater.update`` performs exactly what reference src/rlsp/agents/simple_ddpg.py:312-322 does
per minibatch -- ``_update_critic`` (:204-219), ``_update_actor`` (:221-227),
``_update_target_networks`` (:229-234) -- as a fixed sequence of libgsc_b200 launches with no
autograd graph, no per-tensor optimizer loop and no host synchronisation:

  target actor/critic forward -> critic forward -> TD/MSE -> critic backward -> [all-reduce]
  -> fused Adam+soft-update(critic) -> actor forward -> critic forward (updated weights)
  -> backward through the critic into the action -> actor backward -> [all-reduce]
  -> fused Adam+soft-update(actor).

Data parallelism: every rank holds a replica and ``B/W`` states; losses are
scaled by ``1/B_global`` and the flat gradient buffer of each network is summed with one NCCL
all-reduce before its optimizer step.  The critic gradient that the reference's ``actor_loss.backward()``
also produces is never consumed there (zeroed by the next ``zero_grad``), so it is not computed.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import ops
from .dp import all_gather_flat, allreduce_flat, reduce_scatter_flat, shard_range
from .graph import GraphBatch
from .models import Actor, QNetwork, first_layer_table, net_backward, net_forward


class DDPGUpdater:
    def __init__(self, actor: Actor, qf1: QNetwork, target_actor: Actor, qf1_target: QNetwork,
                 lr: float = 1e-3, gamma: float = 0.99, tau: float = 1e-4, betas=(0.9, 0.999),
                 eps: float = 1e-8, process_group=None, world_size: int = 1, rank: int = 0,
                 shard_optimizer: bool = False, global_batch: Optional[int] = None, fork_streams: bool = False,
                 sm_reserve: int = 8, bucket_allreduce: bool = False, split_optimizer_step: bool = False):
        self.actor, self.qf1, self.target_actor, self.qf1_target = actor, qf1, target_actor, qf1_target
        self.lr, self.gamma, self.tau, self.betas, self.eps = lr, gamma, tau, betas, eps
        self.pg, self.world_size = process_group, int(world_size)
        # Losses are means over the GLOBAL minibatch: with uneven shards the ranks must agree on its size
        # (default: every rank holds the same number of states).
        self.global_batch = None if global_batch is None else int(global_batch)
        for net in (actor, qf1, target_actor, qf1_target):
            net._check_flat()
        # shard_optimizer: reduce-scatter the gradient, Adam + soft update on this rank's 1/W slice
        # (moments exist for that slice only), all-gather the updated online and target weights.
        self.shard = bool(shard_optimizer) and self.world_size > 1
        self.rank = int(rank)
        z = lambda net: torch.zeros_like(net.flat)
        zs = lambda net: torch.zeros(net.flat.numel() // self.world_size, dtype=torch.float32,
                                     device=net.flat.device) if self.shard else z(net)
        self.g_actor, self.m_actor, self.v_actor = z(actor), zs(actor), zs(actor)
        self.g_critic, self.m_critic, self.v_critic = z(qf1), zs(qf1), zs(qf1)
        self.gs_actor = zs(actor) if self.shard else None
        self.gs_critic = zs(qf1) if self.shard else None
        self.step_actor = 0
        self.step_critic = 0
        self.fork_streams = bool(fork_streams)
        self.split_optimizer_step = bool(split_optimizer_step)
        self._side = None
        # Per-layer all-reduces behind the embedder backward (see _bucketed) pay on 2 GPUs (1.92 -> 1.85 ms)
        # but not on 8 (2.04 -> 2.10 ms: six collectives instead of two, and they compete with the
        # backward for SMs), so one all-reduce per network after its backward is the default.
        # "mlp": one all-reduce for the whole dense-layer block (instead of one per layer)
        self.bucket_allreduce = bucket_allreduce if bucket_allreduce == "mlp" else bool(bucket_allreduce)
        if self._bucketed():
            ops.set_sm_reserve(sm_reserve)      # room for the per-layer all-reduces beside the embedder backward
        self._neg = None
        # Adam step counters in device memory: [critic, actor].  Always used, so that an eager update
        # and a CUDA-graph replay of it are the same launches.
        self.steps_dev = torch.zeros(2, dtype=torch.int64, device=actor.flat.device)

    # -- pieces, named after the reference methods ---------------------------------------------
    def _allreduce(self, flat_grad: torch.Tensor) -> None:
        if self.world_size > 1:
            allreduce_flat(flat_grad, self.pg)

    def _bucketed(self) -> bool:
        """Data parallel with the forked backward: every dense layer's gradient block is all-reduced on
        the side stream as soon as it is complete (behind the embedder backward); only the small
        embedder block is left for after it."""
        return self.bucket_allreduce and self.fork_streams and self.world_size > 1 and not self.shard

    def _layer_allreduce(self, net, g):
        if not self._bucketed():
            return None
        _, layers = net.grad_regions()
        if self.bucket_allreduce == "mlp":          # layers are visited last to first: after layer 0 all are complete
            return lambda i: allreduce_flat(g[layers[0][0]:layers[-1][1]], self.pg) if i == 0 else None
        return lambda i: allreduce_flat(g[layers[i][0]:layers[i][1]], self.pg)

    def _split_step(self) -> bool:
        """Step the dense layers on the side stream while the embedder backward runs?  Needs their
        gradients complete there: a single GPU, or the bucketed all-reduce on that stream."""
        return self.fork_streams and self.split_optimizer_step and not self.shard and \
            (self.world_size == 1 or self._bucketed())

    def _mlp_step(self, net, target, g, m, v, step_dev):
        """Adam + soft update of everything behind the embedder block (called on the side stream by
        ``net_backward`` once the dense gradients are complete and the dense weights no longer read)."""
        if not self._split_step():
            return None
        lo = net.grad_regions()[0][1]

        def step():
            ops.adam_soft_update_slice(net.flat[lo:], g[lo:], m[lo:], v[lo:], target.flat[lo:], self.lr,
                                       self.betas[0], self.betas[1], self.eps, step_dev, self.tau, bump_step=False)
        return step

    def _optimizer_step(self, net, target, g, gs, m, v, step_dev) -> None:
        """Gradient exchange + fused Adam + soft target update of one network."""
        if self._split_step():                       # the dense layers were stepped by ``_mlp_step``
            lo, hi = net.grad_regions()[0]
            self._allreduce(g[lo:hi])
            ops.adam_soft_update_slice(net.flat[:hi], g[:hi], m[:hi], v[:hi], target.flat[:hi], self.lr,
                                       self.betas[0], self.betas[1], self.eps, step_dev, self.tau, bump_step=True)
            return
        if not self.shard:
            if self._bucketed():
                lo, hi = net.grad_regions()[0]
                self._allreduce(g[lo:hi])
            else:
                self._allreduce(g)
            ops.adam_soft_update_dev(net.flat, g, m, v, target.flat, self.lr, self.betas[0], self.betas[1],
                                     self.eps, step_dev, self.tau)
            return
        lo, hi = shard_range(net.flat.numel(), self.rank, self.world_size)
        reduce_scatter_flat(g, gs, self.pg)
        ops.adam_soft_update_dev(net.flat[lo:hi], gs, m, v, target.flat[lo:hi], self.lr, self.betas[0],
                                 self.betas[1], self.eps, step_dev, self.tau)
        all_gather_flat(net.flat, self.pg)
        # the other ranks' slices of the target network: soft update applied locally to the online weights
        # just received (12 B/param of HBM traffic) instead of a second all-gather over the wire
        ops.soft_update(net.flat[:lo], target.flat[:lo], self.tau)
        ops.soft_update(net.flat[hi:], target.flat[hi:], self.tau)

    def update_critic(self, nxt: GraphBatch, cur: GraphBatch, dones, rewards, actions):
        """reference _update_critic (simple_ddpg.py:204-219)."""
        B = cur.num_graphs
        scale = 1.0 / (self.global_batch or B * self.world_size)
        ta, tq, q = self.target_actor, self.qf1_target, self.qf1
        a2, _ = net_forward(ta, ta.flat, nxt, None, save=False, mask_output=True, clamp1=True)
        q2, _ = net_forward(tq, tq.flat, nxt, a2, save=False)
        qv, ctx = net_forward(q, q.flat, cur, actions, save=True)
        _, dq, loss = ops.td_mse(rewards, dones, q2, qv, self.gamma, scale)
        net_backward(q, q.flat, cur, ctx, dq.view(B, 1), self.g_critic)
        self.step_critic += 1
        self._optimizer_step(q, tq, self.g_critic, self.gs_critic, self.m_critic, self.v_critic, self.steps_dev[0:1])
        return qv.view(-1), loss

    def update_actor(self, cur: GraphBatch):
        """reference _update_actor (simple_ddpg.py:221-227) + the actor half of
        _update_target_networks (:229-234)."""
        B = cur.num_graphs
        scale = 1.0 / (self.global_batch or B * self.world_size)
        ac, q, ta = self.actor, self.qf1, self.target_actor
        a, ctx_a = net_forward(ac, ac.flat, cur, None, save=True, mask_output=True)
        q_pi, ctx_q = net_forward(q, q.flat, cur, a, save=True, save_gnn=False)
        if self._neg is None or self._neg.shape[0] != B or float(self._neg_scale) != scale:
            self._neg = torch.full((B, 1), -scale, dtype=torch.float32, device=a.device)
            self._neg_scale = scale
        # d(-mean q)/dq = -1/B_global ; back through the critic into the action, times mask
        dz = net_backward(q, q.flat, cur, ctx_q, self._neg, None, want_action_grad=True,
                          action_grad_masked=True)
        net_backward(ac, ac.flat, cur, ctx_a, dz, self.g_actor)
        self.step_actor += 1
        self._optimizer_step(ac, ta, self.g_actor, self.gs_actor, self.m_actor, self.v_actor, self.steps_dev[1:2])
        actor_loss = ops.colsum(q_pi) * (-scale)
        return actor_loss

    def update(self, cur: GraphBatch, nxt: GraphBatch, actions, rewards, dones):
        """One gradient step of reference simple_ddpg.py:312-322.  Returns
        (qf1_a_values [B], qf1_loss [1], actor_loss [1]); the losses are this rank's share."""
        if self.fork_streams:
            return self._update_forked(cur, nxt, actions, rewards, dones)
        cur.csr()
        nxt.share_topology_from(cur)
        qv, q_loss = self.update_critic(nxt, cur, dones, rewards, actions)
        a_loss = self.update_actor(cur)
        return qv, q_loss, a_loss

    def side_streams(self):
        if self._side is None:
            self._side = [torch.cuda.Stream(), torch.cuda.Stream()]
        return self._side

    def _update_forked(self, cur: GraphBatch, nxt: GraphBatch, actions, rewards, dones):
        """The same update with its three independent forward chains on three streams:
             main   : target actor(nxt) -> target critic(nxt, .)
             side 0 : critic(cur, actions)                        (the TD side of the critic loss)
             side 1 : actor(cur)                                  (input of the actor loss; the critic update
                                                                   in between does not touch the actor)
        They join before the TD/MSE kernel; everything after it is one dependent chain.  Same kernels on
        the same data as ``update_critic`` + ``update_actor``, hence bit-identical results; it pays when
        the kernels are too small to fill the GPU (the reference's batch of 100).  Works eagerly and
        under CUDA-graph capture (the fork/join events become graph edges)."""
        B = cur.num_graphs
        scale = 1.0 / (self.global_batch or B * self.world_size)
        ac, q, ta, tq = self.actor, self.qf1, self.target_actor, self.qf1_target
        main = torch.cuda.current_stream()
        self.side_streams()
        # the four first-layer tables depend on the weights only: they are built on a side stream while
        # the main stream packs the batch (seven small launches that leave the GPU mostly idle)
        start = torch.cuda.Event()
        start.record(main)
        with torch.cuda.stream(self._side[1]):
            self._side[1].wait_event(start)
            tab, tab_ready = {}, {}                           # in the order they are needed; one event each
            for net in (ta, tq, q, ac):
                tab[id(net)] = first_layer_table(net, net.flat, cur)
                tab_ready[id(net)] = torch.cuda.Event()
                tab_ready[id(net)].record(self._side[1])
        cur.csr()
        nxt.share_topology_from(cur)
        fork = torch.cuda.Event()
        fork.record(main)
        joins = []
        with torch.cuda.stream(self._side[0]):
            self._side[0].wait_event(fork)
            qv, ctx = net_forward(q, q.flat, cur.view_shared(), actions, save=True, table=tab[id(q)],
                                  table_ready=tab_ready[id(q)])
            ev = torch.cuda.Event(); ev.record(self._side[0]); joins.append(ev)
        with torch.cuda.stream(self._side[1]):
            self._side[1].wait_event(fork)
            # index work of the two embedder backward passes over `cur`: off the critical path, here
            cur.bwd_prep(ac.embedder.hidden_dim, ac.embedder.num_layers)
            a, ctx_a = net_forward(ac, ac.flat, cur.view_shared(), None, save=True, mask_output=True, table=tab[id(ac)])
            ev = torch.cuda.Event(); ev.record(self._side[1]); joins.append(ev)
        a2, _ = net_forward(ta, ta.flat, nxt, None, save=False, mask_output=True, clamp1=True, table=tab[id(ta)],
                            table_ready=tab_ready[id(ta)])
        q2, _ = net_forward(tq, tq.flat, nxt, a2, save=False, table=tab[id(tq)], table_ready=tab_ready[id(tq)])
        for ev in joins:
            main.wait_event(ev)
        # ---- critic update (reference _update_critic, simple_ddpg.py:204-219) ----
        _, dq, q_loss = ops.td_mse(rewards, dones, q2, qv, self.gamma, scale)
        net_backward(q, q.flat, cur, ctx, dq.view(B, 1), self.g_critic, side=self._side[0],
                     after_layer=self._layer_allreduce(q, self.g_critic),
                     mlp_done=self._mlp_step(q, tq, self.g_critic, self.m_critic, self.v_critic, self.steps_dev[0:1]))
        self.step_critic += 1
        self._optimizer_step(q, tq, self.g_critic, self.gs_critic, self.m_critic, self.v_critic, self.steps_dev[0:1])
        # ---- actor update (reference _update_actor, :221-227) ----
        upd_ev = torch.cuda.Event()
        upd_ev.record(main)                                   # critic weights updated
        with torch.cuda.stream(self._side[1]):                # its new table next to its embedder forward
            self._side[1].wait_event(upd_ev)
            tab_q = first_layer_table(q, q.flat, cur)
            tab_q_ready = torch.cuda.Event()
            tab_q_ready.record(self._side[1])
        q_pi, ctx_q = net_forward(q, q.flat, cur, a, save=True, save_gnn=False, table=tab_q, table_ready=tab_q_ready)
        qpi_ev = torch.cuda.Event()
        qpi_ev.record(main)
        with torch.cuda.stream(self._side[1]):                # the reported actor loss: not part of the dependent chain
            self._side[1].wait_event(qpi_ev)
            a_loss = ops.colsum(q_pi) * (-scale)
            loss_ev = torch.cuda.Event()
            loss_ev.record(self._side[1])
        if self._neg is None or self._neg.shape[0] != B or float(self._neg_scale) != scale:
            self._neg = torch.full((B, 1), -scale, dtype=torch.float32, device=a.device)
            self._neg_scale = scale
        dz = net_backward(q, q.flat, cur, ctx_q, self._neg, None, want_action_grad=True, action_grad_masked=True)
        net_backward(ac, ac.flat, cur, ctx_a, dz, self.g_actor, side=self._side[0],
                     after_layer=self._layer_allreduce(ac, self.g_actor),
                     mlp_done=self._mlp_step(ac, ta, self.g_actor, self.m_actor, self.v_actor, self.steps_dev[1:2]))
        self.step_actor += 1
        self._optimizer_step(ac, ta, self.g_actor, self.gs_actor, self.m_actor, self.v_actor, self.steps_dev[1:2])
        main.wait_event(loss_ev)
        return qv.view(-1), q_loss, a_loss


class GraphedUpdate:
    """``DDPGUpdater.update`` captured once into a CUDA graph and replayed per minibatch.

    The whole gradient step (packer, five forward and three backward passes, both fused optimizer
    launches, ~60 kernels) becomes one graph launch, which removes the per-launch host overhead that
    dominates at the reference's batch size of 100.  The batch lives in STATIC device tensors
    (``self.cur``, ``self.nxt``, ``self.actions``, ``self.rewards``, ``self.dones``): write the next
    minibatch into them (``load``) and call ``replay``.  Shapes (B, Nt, E) are frozen at capture.
    Under data parallelism the NCCL collectives of the update are captured too (every rank must
    capture and replay in step); destroy the object before ``destroy_process_group()``."""

    def __init__(self, upd: DDPGUpdater, cur: GraphBatch, nxt: GraphBatch, actions, rewards, dones,
                 warmup: int = 2, replay=None):
        """``replay``: a ``DeviceGraphReplayBuffer``.  The gather of the sampled transitions out of its
        storage is then captured in front of the update (``replay_sampled``): per gradient step the
        host only draws the indices and copies them (and the node / edge offsets) to the device."""
        self.upd = upd
        self.rb = replay
        if replay is not None:
            if not (cur.structured and nxt.structured):
                raise ValueError("GraphedUpdate(replay=...) needs structured batches (what the device buffer samples)")
            B = cur.num_graphs
            dev = cur.x.device
            self.idx = torch.zeros(B, dtype=torch.int64, device=dev)
            self.ptrs = torch.zeros(4, B + 1, dtype=torch.int32, device=dev)
            self.idx_h = [torch.zeros(B, dtype=torch.int64).pin_memory() for _ in range(2)]
            self.ptrs_h = [torch.zeros(4, B + 1, dtype=torch.int32).pin_memory() for _ in range(2)]
            self._copied = [torch.cuda.Event(), torch.cuda.Event()]
            for ev in self._copied:
                ev.record()
            self._flip = 0
            # a valid plan for the warm-up and the capture (drawn without disturbing the caller's RNG stream)
            rng_state = np.random.get_state()
            totals = replay.host_sample_plan(B, self.idx_h[0].numpy(), self.ptrs_h[0].numpy())
            np.random.set_state(rng_state)
            want = (cur.x.shape[0], cur.edge_index.shape[1], nxt.x.shape[0], nxt.edge_index.shape[1])
            if totals != tuple(int(v) for v in want):
                raise ValueError(f"replay minibatches have node/edge totals {totals}, the given batch has {want}: "
                                 "a captured update needs minibatches of constant size")
            self.idx.copy_(self.idx_h[0])
            self.ptrs.copy_(self.ptrs_h[0])
        cl = lambda t: None if t is None else t.clone()
        # The embedder kernels size their shared memory from max_nodes, which is frozen into the graph:
        # a structured batch is captured at the environment limit n_max, so that every later minibatch
        # fits; other batches are checked against the captured value in ``load``.
        cap_nodes = lambda gb: gb.n_max if gb.structured else gb.max_nodes
        mk = lambda gb, share=None: GraphBatch(
            gb.x.clone(), gb.edge_index.clone() if share is None else share.edge_index,
            gb.batch.clone() if share is None else share.batch,
            (cl(gb._mask) if not gb.structured else None) if share is None else share._mask,
            gb.num_graphs, cap_nodes(gb), None,
            cl(gb.n_nodes) if share is None else share.n_nodes, gb.n_max, gb.groups)
        self.cur = mk(cur)
        self.nxt = mk(nxt, self.cur if nxt.edge_index is cur.edge_index else None)
        self.actions, self.rewards, self.dones = actions.clone(), rewards.clone(), dones.clone()
        nets = (upd.actor, upd.qf1, upd.target_actor, upd.qf1_target)
        state = [t.clone() for t in (*[n.flat for n in nets], upd.m_actor, upd.v_actor, upd.m_critic,
                                     upd.v_critic, upd.steps_dev)]
        host_steps = (upd.step_actor, upd.step_critic)

        def restore():
            for dst, src in zip((*[n.flat for n in nets], upd.m_actor, upd.v_actor, upd.m_critic,
                                 upd.v_critic, upd.steps_dev), state):
                dst.copy_(src)
            upd.step_actor, upd.step_critic = host_steps

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):            # sizes every workspace; its weight updates are rolled back
                self._run()
        torch.cuda.current_stream().wait_stream(side)
        restore()
        torch.cuda.synchronize()
        # capture on the stream the warm-up ran on: the workspaces are per stream, so the warm-up has
        # sized exactly the buffers the captured launches use (nothing is allocated during capture)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=side):
            self.out = self._run()
        restore()

    def close(self) -> None:
        """Destroy the captured graph.  A graph that captured NCCL collectives keeps the communicator
        alive: call this (on every rank) before ``torch.distributed.destroy_process_group()``."""
        torch.cuda.synchronize()
        self.graph = None
        self.out = None

    def _run(self):
        if self.rb is not None:
            from .buffer import BufferSample
            aux = self.upd.side_streams()[0] if self.upd.fork_streams else None    # the critic chain's stream
            self.rb.gather_planned(self.idx, self.ptrs, BufferSample(self.cur, self.actions, self.nxt, self.dones,
                                                                    self.rewards), aux_stream=aux)
        return self.upd.update(self.cur.view(), self.nxt.view(), self.actions, self.rewards, self.dones)

    def replay_sampled(self):
        """Draw a minibatch from the replay buffer given at construction and run the update on it: index
        draw on the host (reference RNG), everything else inside the graph."""
        k = self._flip
        self._flip ^= 1
        self._copied[k].synchronize()                      # the pinned pair is free again
        totals = self.rb.host_sample_plan(self.cur.num_graphs, self.idx_h[k].numpy(), self.ptrs_h[k].numpy())
        want = (self.cur.x.shape[0], self.cur.edge_index.shape[1], self.nxt.x.shape[0], self.nxt.edge_index.shape[1])
        if totals != tuple(int(v) for v in want):
            raise ValueError(f"sampled minibatch has node/edge totals {totals}, the captured graph expects {want}")
        self.idx.copy_(self.idx_h[k], non_blocking=True)
        self.ptrs.copy_(self.ptrs_h[k], non_blocking=True)
        self._copied[k].record()
        return self.replay()

    def load(self, cur: GraphBatch, nxt: GraphBatch, actions, rewards, dones, non_blocking: bool = True):
        """Copy a minibatch (host or device tensors of the captured shapes) into the static inputs."""
        nb = non_blocking
        for new, own in ((cur, self.cur), (nxt, self.nxt)):
            if new.max_nodes is not None and own.max_nodes is not None and new.max_nodes > own.max_nodes:
                raise ValueError(f"minibatch holds a graph of {new.max_nodes} nodes; the graph was captured for at "
                                 f"most {own.max_nodes} nodes per graph")
        self.cur.x.copy_(cur.x, non_blocking=nb)
        self.cur.edge_index.copy_(cur.edge_index, non_blocking=nb)
        self.cur.batch.copy_(cur.batch, non_blocking=nb)
        if self.cur.structured:
            self.cur.n_nodes.copy_(cur.n_nodes, non_blocking=nb)
        else:
            self.cur._mask.copy_(cur.mask, non_blocking=nb)
        self.nxt.x.copy_(nxt.x, non_blocking=nb)
        if self.nxt.edge_index is not self.cur.edge_index:
            self.nxt.edge_index.copy_(nxt.edge_index, non_blocking=nb)
            self.nxt.batch.copy_(nxt.batch, non_blocking=nb)
            if self.nxt.structured:
                self.nxt.n_nodes.copy_(nxt.n_nodes, non_blocking=nb)
            else:
                self.nxt._mask.copy_(nxt.mask, non_blocking=nb)
        self.actions.copy_(actions, non_blocking=nb)
        self.rewards.copy_(rewards, non_blocking=nb)
        self.dones.copy_(dones, non_blocking=nb)

    def replay(self):
        """One gradient step on the static inputs; returns (q_values, q_loss, actor_loss) tensors
        that the graph overwrites on every replay."""
        self.graph.replay()
        self.upd.step_actor += 1
        self.upd.step_critic += 1
        return self.out


@torch.no_grad()
def predict(actor: Actor, obs) -> torch.Tensor:
    """reference SimpleDDPG.predict (simple_ddpg.py:331-336) / inference.predict (inference.py:11-15)."""
    gb = GraphBatch.wrap(obs)
    y, _ = net_forward(actor, actor._check_flat(), gb, None, save=False, mask_output=True)
    return y


class GraphedPredict:
    """``predict`` (packer + embedder + actor MLP) for a fixed batch shape, captured once into a CUDA
    graph: batched rollouts / serving call it thousands of times with the same (B, Nt, E), and at
    ~0.2 ms of GPU work the per-launch host overhead of the dozen kernels is what bounds it.
    Write the next batch into the static inputs with ``load`` and call ``replay``; the returned
    tensor is overwritten by the next replay."""

    def __init__(self, actor: Actor, obs, warmup: int = 2):
        gb = GraphBatch.wrap(obs)
        cl = lambda t: None if t is None else t.clone()
        self.actor = actor
        # max_nodes is frozen into the graph (it sizes the embedder's shared memory): a structured batch
        # is captured at the environment limit, other batches are checked in ``load``
        self.gb = GraphBatch(gb.x.clone(), gb.edge_index.clone(), cl(gb.batch),
                             None if gb.structured else cl(gb._mask), gb.num_graphs,
                             gb.n_max if gb.structured else gb.max_nodes, None,
                             cl(gb.n_nodes), gb.n_max, gb.groups)
        if self.gb.max_nodes is None:
            gb.csr()
            self.gb.max_nodes = gb.max_nodes
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                predict(actor, self.gb.view())
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=side):          # same stream as the warm-up (per-stream workspaces)
            self.out = predict(actor, self.gb.view())

    def load(self, obs, non_blocking: bool = True) -> None:
        gb = GraphBatch.wrap(obs)
        nb = non_blocking
        if gb.max_nodes is not None and gb.max_nodes > self.gb.max_nodes:
            raise ValueError(f"batch holds a graph of {gb.max_nodes} nodes; the graph was captured for at most "
                             f"{self.gb.max_nodes} nodes per graph")
        self.gb.x.copy_(gb.x, non_blocking=nb)
        self.gb.edge_index.copy_(gb.edge_index, non_blocking=nb)
        if self.gb.batch is not None:
            self.gb.batch.copy_(gb.batch, non_blocking=nb)
        if self.gb.structured:
            self.gb.n_nodes.copy_(gb.n_nodes, non_blocking=nb)
        else:
            self.gb._mask.copy_(gb.mask, non_blocking=nb)

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.out


def post_process_actions(action, num_nodes: int) -> torch.Tensor:
    """reference SimpleDDPG.post_process_actions (simple_ddpg.py:374-395) on the GPU.

    ``action``: [A] or [R, A] fp32 (tensor or ndarray).  Returns fp64 of the same shape."""
    t = torch.as_tensor(action, dtype=torch.float32)
    if not t.is_cuda:
        t = t.cuda()
    out = ops.actor_head(t, num_nodes)
    return out.view(t.shape) if t.dim() == 1 else out


def choose_action(actor: Actor, obs, noise: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """The non-warm-up branch of reference _choose_action (simple_ddpg.py:187-199) followed by
    post_process_actions, with the Gaussian sample ``noise`` ([R, A] fp64) supplied by the caller."""
    y = predict(actor, obs)
    low, high = float(np.min(actor.low)), float(np.max(actor.high))
    return ops.actor_head(y, num_nodes, noise=noise, low=low, high=high)
