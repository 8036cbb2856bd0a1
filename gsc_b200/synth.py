"""Synthetic network states of the sizes BASELINE.json names.

The reference builds one observation per simulator step in
``SimulatorWrapper._parse_state_as_graph`` (reference src/rlsp/envs/simulator_wrapper.py:237-308):
``x[n, F]`` in [0, 1], ``edge_index[2, 2|E|]`` int64 with both directions listed source-major (the
``from_networkx`` order, :297-299), and ``mask[1, A]`` with ``mask[:n, :, :, :n] = 1`` over the
scheduling shape ``(N_max, C, S, N_max)`` (:302-306).  This module produces batches of such states
with numpy only; the bundled GraphML topologies are not needed at run time.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np


def random_topology(n: int, num_links: int, rng: np.random.Generator) -> np.ndarray:
    """Connected simple graph: random spanning tree plus random extra links.

    Returns ``edge_index[2, 2*num_links]`` int64, both directions, sorted by (src, dst).
    """
    if n <= 1:
        return np.zeros((2, 0), dtype=np.int64)
    num_links = int(min(max(num_links, n - 1), n * (n - 1) // 2))
    perm = rng.permutation(n)
    links = set()
    for k in range(1, n):
        u, v = int(perm[k]), int(perm[rng.integers(0, k)])
        links.add((min(u, v), max(u, v)))
    while len(links) < num_links:
        u, v = int(rng.integers(0, n)), int(rng.integers(0, n))
        if u != v:
            links.add((min(u, v), max(u, v)))
    und = np.array(sorted(links), dtype=np.int64).T
    both = np.concatenate([und, und[::-1]], axis=1)
    order = np.lexsort((both[1], both[0]))
    return np.ascontiguousarray(both[:, order])


def structured_mask(n: np.ndarray, n_max: int, num_sfcs: int = 1, num_sfs: int = 3) -> np.ndarray:
    """``mask[b, i, c, s, j] = (i < n_b) & (j < n_b)`` flattened to ``[B, A]`` float32."""
    n = np.asarray(n).reshape(-1, 1)
    idx = np.arange(n_max).reshape(1, -1)
    valid = (idx < n).astype(np.float32)                       # [B, n_max]
    m = valid[:, :, None, None] * valid[:, None, None, :]      # [B, n_max, 1, n_max]
    m = np.broadcast_to(m, (n.shape[0], n_max, num_sfcs * num_sfs, n_max))
    return np.ascontiguousarray(m.reshape(n.shape[0], -1), dtype=np.float32)


class StateBatch:
    """A collated minibatch in the reference's ``Batch.from_data_list`` layout (buffer.py:65-81).

    ``mask`` ([B, A] f32) is a function of ``n`` (simulator_wrapper.py:302-306); it is materialised on
    first use only, so that very large batches (65 536 states) can be generated without it."""

    def __init__(self, x, edge_index, batch, ptr, n, mask=None, n_max: int = 0, num_sfcs: int = 1,
                 num_sfs: int = 3):
        self.x = x                        # [Nt, F] f32
        self.edge_index = edge_index      # [2, E]  i64, node ids offset per graph
        self.batch = batch                # [Nt]    i64
        self.ptr = ptr                    # [B+1]   i64
        self.n = n                        # [B]     i64 real node count of every state
        self._mask = mask
        self.n_max = int(n_max)
        self.num_sfcs = int(num_sfcs)
        self.num_sfs = int(num_sfs)

    @property
    def mask(self) -> np.ndarray:
        if self._mask is None:
            self._mask = structured_mask(self.n, self.n_max, self.num_sfcs, self.num_sfs)
        return self._mask

    @property
    def num_graphs(self) -> int:
        return int(self.n.shape[0])

    @property
    def action_dim(self) -> int:
        return self.n_max * self.n_max * self.num_sfcs * self.num_sfs

    def graph(self, g: int) -> dict:
        """State ``g`` as the single-``Data`` record the environment emits (batch=None)."""
        lo, hi = int(self.ptr[g]), int(self.ptr[g + 1])
        sel = (self.edge_index[0] >= lo) & (self.edge_index[0] < hi)
        return {"x": self.x[lo:hi].copy(), "edge_index": self.edge_index[:, sel] - lo,
                "mask": self.mask[g:g + 1].copy()}


def make_states(B: int, n_max: int, n_low: int, n_high: int, link_ratio: float = 1.5,
                num_feat: int = 3, seed: int = 1234, num_topologies: Optional[int] = None,
                num_sfcs: int = 1, num_sfs: int = 3, fixed_links: Optional[int] = None) -> StateBatch:
    """``B`` states with ``n ~ U{n_low..n_high}`` nodes and ``round(link_ratio*n)`` links each.

    A pool of ``num_topologies`` distinct topologies is generated and states draw from it (the
    reference rotates through a short list of training topologies: gym_env.py:103-120).
    """
    rng = np.random.default_rng(seed)
    pool_n = int(num_topologies or min(B, 128))
    pool: List[np.ndarray] = []
    pool_sizes = rng.integers(n_low, n_high + 1, size=pool_n)
    for k in range(pool_n):
        nk = int(pool_sizes[k])
        links = fixed_links if fixed_links is not None else int(round(link_ratio * nk))
        pool.append(random_topology(nk, links, rng))
    pick = rng.integers(0, pool_n, size=B)
    n = pool_sizes[pick].astype(np.int64)
    ptr = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(n, out=ptr[1:])
    edges = [pool[int(pick[g])] + ptr[g] for g in range(B)]
    edge_index = np.ascontiguousarray(np.concatenate(edges, axis=1)) if B else np.zeros((2, 0), np.int64)
    batch = np.repeat(np.arange(B, dtype=np.int64), n)
    x = rng.random((int(ptr[-1]), num_feat), dtype=np.float32)
    return StateBatch(x=x, edge_index=edge_index, batch=batch, ptr=ptr, n=n, mask=None, n_max=n_max,
                      num_sfcs=num_sfcs, num_sfs=num_sfs)


def make_actions(states: StateBatch, seed: int = 1234, max_nonzero: int = 3) -> np.ndarray:
    """Replay-buffer style actions: every (src, sfc, sf) group of ``N_max`` destinations holds at
    most ``max_nonzero`` weights >= 0.1 that sum to 1 (what ``post_process_actions`` emits), times mask."""
    rng = np.random.default_rng(seed + 7)
    B, nm = states.num_graphs, states.n_max
    groups = nm * states.num_sfcs * states.num_sfs
    a = np.zeros((B, groups, nm), dtype=np.float32)
    for b in range(B):
        nb = int(states.n[b])
        k = min(max_nonzero, nb)
        dst = np.argsort(rng.random((groups, nb)), axis=1)[:, :k]     # k distinct destinations
        w = rng.random((groups, k)) + 0.2
        w = w / w.sum(1, keepdims=True)
        np.put_along_axis(a[b], dst, w.astype(np.float32), axis=1)
    return a.reshape(B, -1) * states.mask


def make_transitions(B: int, n_max: int, n_low: int, n_high: int, seed: int = 1234, **kw):
    """(cur, nxt, actions, rewards, dones) for one DDPG replay minibatch.

    ``nxt`` shares every state's topology with ``cur`` (one environment step does not change the
    network) and has fresh node features.
    """
    cur = make_states(B, n_max, n_low, n_high, seed=seed, **kw)
    rng = np.random.default_rng(seed + 1)
    nxt = StateBatch(x=rng.random(cur.x.shape, dtype=np.float32), edge_index=cur.edge_index,
                     batch=cur.batch, ptr=cur.ptr, n=cur.n, mask=cur._mask, n_max=cur.n_max,
                     num_sfcs=cur.num_sfcs, num_sfs=cur.num_sfs)
    actions = make_actions(cur, seed)
    rewards = (rng.random(B, dtype=np.float32) * 2 - 1).astype(np.float32)
    dones = (rng.random(B) < 1.0 / 200.0).astype(np.float32)
    return cur, nxt, actions, rewards, dones
