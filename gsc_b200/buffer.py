"""Graph replay buffer with the reference's surface (reference src/rlsp/agents/buffer.py:8-89).

``add`` stores one transition, ``sample`` draws indices with the same numpy global-RNG calls as the
reference (:57-62) and returns a ``BufferSample`` whose observation fields are ``GraphBatch`` objects
in ``Batch.from_data_list`` layout, already on ``device``; their CSR is packed on the GPU the first
time a model touches them.  Host-side storage stays Python, as in the reference.
"""
from __future__ import annotations

from typing import Optional, NamedTuple

import numpy as np
import torch

from .graph import GraphBatch, GraphData


class BufferSample(NamedTuple):
    observations: GraphBatch
    actions: torch.Tensor
    next_observations: GraphBatch
    dones: torch.Tensor
    rewards: torch.Tensor


class GraphReplayBuffer:
    def __init__(self, buffer_size: int, action_space, device="cpu") -> None:
        self.buffer_size = int(buffer_size)
        self.device = device
        self.action_space = int(np.prod(action_space.shape))
        self.dones = np.ones(self.buffer_size, np.float32)
        self.rewards = np.zeros(self.buffer_size, dtype=np.float32)
        self.actions = torch.zeros((self.buffer_size, self.action_space), dtype=torch.float32)
        self.observations = [None] * self.buffer_size
        self.next_observations = [None] * self.buffer_size
        self.pos = 0
        self.full = False

    @staticmethod
    def _as_data(obs) -> GraphData:
        return obs.clone() if isinstance(obs, GraphData) else GraphData(obs.x.clone(), obs.edge_index.clone(),
                                                                        obs.mask.clone())

    def add(self, obs, next_obs, action: np.ndarray, reward, done, infos=None) -> None:
        self.dones[self.pos] = done
        self.rewards[self.pos] = reward
        self.actions[self.pos] = torch.from_numpy(np.asarray(action)).to(torch.float32)
        self.observations[self.pos] = self._as_data(obs)
        self.next_observations[self.pos] = self._as_data(next_obs)
        self.pos += 1
        if self.pos == self.buffer_size:
            self.pos = 0
            self.full = True

    def sample(self, batch_size: int) -> BufferSample:
        if self.full:
            batch_inds = (np.random.randint(1, self.buffer_size, size=batch_size) + self.pos) % self.buffer_size
        else:
            batch_inds = np.random.randint(0, self.pos, size=batch_size)
        cur = GraphBatch.from_data_list([self.observations[i] for i in batch_inds]).to(self.device)
        nxt = GraphBatch.from_data_list([self.next_observations[i] for i in batch_inds]).to(self.device)
        cur.mask = cur.mask.view(-1, self.action_space)
        nxt.mask = nxt.mask.view(-1, self.action_space)
        return BufferSample(
            observations=cur,
            actions=self.actions[batch_inds].to(self.device),
            next_observations=nxt,
            dones=torch.from_numpy(self.dones[batch_inds]).to(self.device),
            rewards=torch.from_numpy(self.rewards[batch_inds]).to(self.device),
        )


class DeviceGraphReplayBuffer:
    """``GraphReplayBuffer`` with the storage in HBM.

    Same ``add`` / ``sample`` surface and the same numpy RNG draws as the reference buffer, but
    transitions are kept on the device in padded struct-of-arrays form -- ``x[cap, N_max, F]``,
    graph-local ``edge_index[cap, 2, E_max]`` (int32), node / edge counts, ``actions[cap, A]`` -- and
    a minibatch is gathered by index with one kernel straight into ``Batch.from_data_list`` layout.
    Masks are not stored: an observation's mask is ``mask[:n, :, :, :n] = 1`` (reference
    src/rlsp/envs/simulator_wrapper.py:302-306), so it is regenerated from ``n`` on the device; ``add``
    verifies that the mask it is given has that form.  Per gradient step the host only sends the
    sampled indices and two small prefix-sum arrays instead of re-collating and copying ~2(4A) bytes
    per state.
    """

    def __init__(self, buffer_size: int, action_space, device="cuda", n_max: int = None, num_feat: int = 3,
                 e_max: int = None, groups_per_node: int = None) -> None:
        from . import ops
        self.ops = ops
        self.buffer_size = int(buffer_size)
        self.device = torch.device(device)
        self.action_space = int(np.prod(action_space.shape))
        if n_max is None:
            raise ValueError("DeviceGraphReplayBuffer needs n_max (EnvironmentLimits.MAX_NODE_COUNT)")
        self.n_max, self.num_feat = int(n_max), int(num_feat)
        self.groups = int(groups_per_node) if groups_per_node else self.action_space // (self.n_max * self.n_max)
        if self.n_max * self.groups * self.n_max != self.action_space:
            raise ValueError("action dimension is not n_max * groups_per_node * n_max")
        self.e_max = int(e_max) if e_max else 8 * self.n_max
        cap, dev = self.buffer_size, self.device
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        self.x, self.next_x = z(cap, self.n_max, self.num_feat), z(cap, self.n_max, self.num_feat)
        self.ei, self.next_ei = z(cap, 2, self.e_max, dt=torch.int32), z(cap, 2, self.e_max, dt=torch.int32)
        self.n, self.next_n = z(cap, dt=torch.int32), z(cap, dt=torch.int32)
        self.ec, self.next_ec = z(cap, dt=torch.int32), z(cap, dt=torch.int32)
        self.actions = z(cap, self.action_space)
        self.rewards, self.dones = z(cap), torch.ones(cap, dtype=torch.float32, device=dev)
        # host mirrors of the counts: minibatch sizes are known without a device round trip
        self.n_h = np.zeros((2, cap), dtype=np.int32)
        self.ec_h = np.zeros((2, cap), dtype=np.int32)
        self.pos, self.full = 0, False

    # ---- add ------------------------------------------------------------------------------------
    def _check_mask(self, obs, n: int) -> None:
        mask = getattr(obs, "mask", None)
        if mask is None:
            return
        m = np.asarray(mask.detach().cpu() if isinstance(mask, torch.Tensor) else mask, dtype=np.float32)
        m = m.reshape(self.n_max, self.groups, self.n_max)
        ok = m[:n, :, :n].all() and m[n:].sum() == 0 and m[:, :, n:].sum() == 0
        if not ok:
            raise ValueError("observation mask is not the structured mask[:n, :, :, :n] = 1; "
                             "use GraphReplayBuffer for arbitrary masks")

    def _store_obs(self, which: int, obs, x_store, ei_store, n_store, ec_store) -> None:
        x = torch.as_tensor(obs.x, dtype=torch.float32)
        ei = torch.as_tensor(obs.edge_index).to(torch.int32)
        n, e = int(x.shape[0]), int(ei.shape[1])
        if n > self.n_max or e > self.e_max:
            raise ValueError(f"observation with {n} nodes / {e} edges exceeds n_max={self.n_max} / e_max={self.e_max}")
        self._check_mask(obs, n)
        p = self.pos
        x_store[p, :n].copy_(x, non_blocking=True)
        ei_store[p, :, :e].copy_(ei, non_blocking=True)
        n_store[p:p + 1].fill_(n)
        ec_store[p:p + 1].fill_(e)
        self.n_h[which, p], self.ec_h[which, p] = n, e

    def add(self, obs, next_obs, action, reward, done, infos=None) -> None:
        p = self.pos
        self._store_obs(0, obs, self.x, self.ei, self.n, self.ec)
        self._store_obs(1, next_obs, self.next_x, self.next_ei, self.next_n, self.next_ec)
        self.actions[p].copy_(torch.as_tensor(np.asarray(action), dtype=torch.float32), non_blocking=True)
        self.rewards[p:p + 1].fill_(float(reward))
        self.dones[p:p + 1].fill_(float(done))
        self.pos += 1
        if self.pos == self.buffer_size:
            self.pos, self.full = 0, True

    def bulk_load(self, cur, nxt, actions, rewards, dones) -> None:
        """Fill the buffer from collated ``gsc_b200.synth.StateBatch`` pairs in a few large copies
        (benchmarks and tests; ``add`` is the per-environment-step path)."""
        B = cur.num_graphs
        if self.pos + B > self.buffer_size:
            raise ValueError("bulk_load would wrap around the buffer")
        for which, st, xs, eis, ns, ecs in ((0, cur, self.x, self.ei, self.n, self.ec),
                                            (1, nxt, self.next_x, self.next_ei, self.next_n, self.next_ec)):
            n = st.n.astype(np.int32)
            xp = np.zeros((B, self.n_max, self.num_feat), dtype=np.float32)
            eip = np.zeros((B, 2, self.e_max), dtype=np.int32)
            ec = np.zeros(B, dtype=np.int32)
            owner = st.batch[st.edge_index[0]]                       # graph of every edge
            starts = np.searchsorted(owner, np.arange(B))
            ends = np.searchsorted(owner, np.arange(B), side="right")
            for g in range(B):
                lo, hi = int(st.ptr[g]), int(st.ptr[g + 1])
                xp[g, :hi - lo] = st.x[lo:hi]
                e = ends[g] - starts[g]
                if e > self.e_max:
                    raise ValueError(f"state with {e} edges exceeds e_max={self.e_max}")
                eip[g, :, :e] = st.edge_index[:, starts[g]:ends[g]] - lo
                ec[g] = e
            sl = slice(self.pos, self.pos + B)
            xs[sl].copy_(torch.from_numpy(xp))
            eis[sl].copy_(torch.from_numpy(eip))
            ns[sl].copy_(torch.from_numpy(n))
            ecs[sl].copy_(torch.from_numpy(ec))
            self.n_h[which, sl], self.ec_h[which, sl] = n, ec
        sl = slice(self.pos, self.pos + B)
        self.actions[sl].copy_(torch.from_numpy(np.ascontiguousarray(actions, dtype=np.float32)))
        self.rewards[sl].copy_(torch.from_numpy(np.ascontiguousarray(rewards, dtype=np.float32)))
        self.dones[sl].copy_(torch.from_numpy(np.ascontiguousarray(dones, dtype=np.float32)))
        self.pos += B
        if self.pos == self.buffer_size:
            self.pos, self.full = 0, True

    # ---- sample ---------------------------------------------------------------------------------
    def sample_indices(self, batch_size: int) -> np.ndarray:
        """The reference's index draw (buffer.py:57-62), same global numpy RNG calls."""
        if self.full:
            return (np.random.randint(1, self.buffer_size, size=batch_size) + self.pos) % self.buffer_size
        return np.random.randint(0, self.pos, size=batch_size)

    def _gather_obs(self, which: int, idx_np: np.ndarray, idx_dev: torch.Tensor, out=None) -> GraphBatch:
        stores = ((self.x, self.ei, self.n, self.ec), (self.next_x, self.next_ei, self.next_n, self.next_ec))[which]
        n_sel, e_sel = self.n_h[which, idx_np], self.ec_h[which, idx_np]
        node_ptr = np.zeros(len(idx_np) + 1, dtype=np.int32)
        edge_ptr = np.zeros(len(idx_np) + 1, dtype=np.int32)
        np.cumsum(n_sel, out=node_ptr[1:])
        np.cumsum(e_sel, out=edge_ptr[1:])
        nt, et = int(node_ptr[-1]), int(edge_ptr[-1])
        ptrs = torch.from_numpy(np.stack([node_ptr, edge_ptr])).to(self.device, non_blocking=True)
        if out is not None and (out.x.shape[0] != nt or out.edge_index.shape[1] != et):
            raise ValueError("static minibatch buffers have a different node / edge count than this sample")
        x, ei, batch, n_dev = self.ops.replay_gather(
            *stores, idx_dev, ptrs[0], ptrs[1], nt, et,
            x_out=None if out is None else out.x, ei_out=None if out is None else out.edge_index,
            batch_out=None if out is None else out.batch, n_out=None if out is None else out.n_nodes)
        if out is not None:
            if out._mask is not None:                 # a dense mask was materialised for this buffer: refresh it
                self.ops.build_mask(n_dev, self.n_max, self.groups, out=out._mask)
            return out
        # structured batch: the [B, A] mask is generated from n only if somebody asks for .mask
        return GraphBatch(x, ei, batch, None, len(idx_np), int(n_sel.max()) if len(n_sel) else 0, None,
                          n_dev, self.n_max, self.groups)

    # ---- sampling inside a CUDA graph -------------------------------------------------------------
    def host_sample_plan(self, batch_size: int, idx_out: np.ndarray, ptrs_out: np.ndarray):
        """Draw the minibatch indices (reference RNG) and write them and the node / edge offsets of
        both states into caller-owned (pinned) host arrays: ``idx_out`` int64 [B], ``ptrs_out`` int32
        [4, B+1] = cur node, cur edge, next node, next edge offsets.  Returns the four totals."""
        idx_np = self.sample_indices(batch_size)
        idx_out[:] = idx_np
        ptrs_out[:, 0] = 0
        np.cumsum(self.n_h[0, idx_np], out=ptrs_out[0, 1:])
        np.cumsum(self.ec_h[0, idx_np], out=ptrs_out[1, 1:])
        np.cumsum(self.n_h[1, idx_np], out=ptrs_out[2, 1:])
        np.cumsum(self.ec_h[1, idx_np], out=ptrs_out[3, 1:])
        return tuple(int(ptrs_out[k, -1]) for k in range(4))

    def gather_planned(self, idx_dev: torch.Tensor, ptrs_dev: torch.Tensor, out: "BufferSample",
                       aux_stream: Optional[torch.cuda.Stream] = None) -> None:
        """The device half of ``sample``: gather the transitions ``idx_dev`` into the static minibatch
        ``out`` using offsets that already sit on the device.  No host work, so it can be captured
        into a CUDA graph in front of the update (``ddpg.GraphedUpdate(..., replay=...)``)."""
        cur, nxt = out.observations, out.next_observations
        self.ops.replay_gather(self.x, self.ei, self.n, self.ec, idx_dev, ptrs_dev[0], ptrs_dev[1],
                               int(cur.x.shape[0]), int(cur.edge_index.shape[1]), x_out=cur.x,
                               ei_out=cur.edge_index, batch_out=cur.batch, n_out=cur.n_nodes)
        self.ops.replay_gather(self.next_x, self.next_ei, self.next_n, self.next_ec, idx_dev, ptrs_dev[2], ptrs_dev[3],
                               int(nxt.x.shape[0]), int(nxt.edge_index.shape[1]), x_out=nxt.x,
                               ei_out=nxt.edge_index, batch_out=nxt.batch, n_out=nxt.n_nodes)
        for gb in (cur, nxt):
            if gb._mask is not None:
                self.ops.build_mask(gb.n_nodes, self.n_max, self.groups, out=gb._mask)
        def rows():
            self.ops.gather_rows(self.actions, idx_dev, out.actions)
            self.ops.gather_rows(self.rewards, idx_dev, out.rewards)
            self.ops.gather_rows(self.dones, idx_dev, out.dones)
        if aux_stream is None:
            rows()
            return
        # the [B, A] action rows are the bulk of the bytes and are first needed by the critic's dense
        # layer: gather them on the stream that consumes them, next to the packer
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        with torch.cuda.stream(aux_stream):
            aux_stream.wait_event(ev)
            rows()

    def sample(self, batch_size: int, out: "BufferSample" = None) -> BufferSample:
        """Minibatch on the device.  ``out`` (a previous BufferSample of identical shapes, e.g. the
        static inputs of a CUDA-graph) is overwritten in place when given."""
        idx_np = self.sample_indices(batch_size)
        idx_dev = torch.from_numpy(idx_np.astype(np.int64)).to(self.device, non_blocking=True)
        cur = self._gather_obs(0, idx_np, idx_dev, None if out is None else out.observations)
        nxt = self._gather_obs(1, idx_np, idx_dev, None if out is None else out.next_observations)
        actions = self.ops.gather_rows(self.actions, idx_dev, None if out is None else out.actions)
        rewards = self.ops.gather_rows(self.rewards, idx_dev, None if out is None else out.rewards)
        dones = self.ops.gather_rows(self.dones, idx_dev, None if out is None else out.dones)
        return BufferSample(observations=cur, actions=actions, next_observations=nxt, dones=dones, rewards=rewards)
