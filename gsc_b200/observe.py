"""Batched construction of graph observations on the device.

The reference emits one observation per environment step on the host: ``deepcopy`` of the networkx
graph, Python loops over its nodes and ``from_networkx`` (reference
src/rlsp/envs/simulator_wrapper.py:237-308).  For many environments stepping in parallel (the
"65 536 parallel network states" configuration) the per-node quantities of all states are written
into three flat arrays and one launch produces the node-feature matrix in the ``Batch`` layout that
the packer and the embedder read; the mask stays a function of the node counts (structured batch).
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import ops
from .graph import GraphBatch


def build_observations(ingress, usage, cap, n_nodes, edge_index, n_max: int, groups: int = 3,
                       device="cuda") -> GraphBatch:
    """``ingress``/``usage``/``cap``: raw per-node values of all states back to back ([Nt], host or
    device); ``n_nodes``: [B] node count of every state; ``edge_index``: [2, E] int64 with node ids
    already offset per state (both directions of every link, as ``from_networkx`` emits them).
    Returns a structured ``GraphBatch`` on ``device`` whose ``x`` is [ingress_traffic, node_load,
    node_cap] normalised per state exactly as ``_parse_state_as_graph`` does."""
    dev = torch.device(device)
    t = lambda a, dt: (a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a))).to(dev, dt, non_blocking=True)
    n = t(n_nodes, torch.int32)
    B = int(n.shape[0])
    ptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    torch.cumsum(n, 0, out=ptr[1:])
    x = ops.node_features(t(ingress, torch.float32), t(usage, torch.float32), t(cap, torch.float32), ptr)
    batch = torch.repeat_interleave(torch.arange(B, dtype=torch.int64, device=dev), n.long(),
                                    output_size=int(x.shape[0]))
    n_host = n_nodes if not isinstance(n_nodes, torch.Tensor) else n_nodes.cpu()
    max_nodes = int(np.asarray(n_host).max()) if B else 0
    return GraphBatch(x, t(edge_index, torch.int64), batch, None, B, max_nodes, None, n, int(n_max), int(groups))
