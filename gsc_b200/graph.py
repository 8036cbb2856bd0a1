"""Graph observation containers with the attribute surface the reference models read.

The reference passes PyG ``Data`` / ``Batch`` objects around and its models touch only ``.x``,
``.edge_index``, ``.batch``, ``.mask`` and ``.to(device)`` (reference src/rlsp/agents/models.py:
88-89,148-149; src/rlsp/agents/simple_ddpg.py:189).  ``GraphData`` / ``GraphBatch`` provide exactly
that surface without PyG, and ``GraphBatch`` additionally carries what the CUDA path needs so that no
device->host sync happens on the hot path: ``num_graphs``, ``max_nodes`` and the packed CSR.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch


def _t(a, dtype):
    if isinstance(a, torch.Tensor):
        return a.to(dtype)
    return torch.as_tensor(np.asarray(a), dtype=dtype)


class GraphData:
    """One network state: what ``_parse_state_as_graph`` emits (reference
    src/rlsp/envs/simulator_wrapper.py:237-308).  ``batch`` is None as for a single PyG ``Data``."""

    batch = None

    def __init__(self, x, edge_index, mask=None, **extra):
        self.x = _t(x, torch.float32)
        self.edge_index = _t(edge_index, torch.int64)
        self.mask = None if mask is None else _t(mask, torch.float32).reshape(1, -1)
        for k, v in extra.items():
            setattr(self, k, v)

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    def clone(self) -> "GraphData":
        return GraphData(self.x.clone(), self.edge_index.clone(),
                         None if self.mask is None else self.mask.clone())

    def to(self, device, non_blocking: bool = False) -> "GraphData":
        g = GraphData.__new__(GraphData)
        g.x = self.x.to(device, non_blocking=non_blocking)
        g.edge_index = self.edge_index.to(device, non_blocking=non_blocking)
        g.mask = None if self.mask is None else self.mask.to(device, non_blocking=non_blocking)
        return g

    def __iter__(self):  # the reference buffer stores rows of (key, tensor) pairs (buffer.py:48-49)
        yield "x", self.x
        yield "edge_index", self.edge_index
        yield "mask", self.mask

    # item access and from_dict: how the reference attaches the mask (simulator_wrapper.py:306) and
    # rebuilds the stored rows when it samples (buffer.py:68,77)
    def __len__(self):  # three (key, tensor) rows: numpy stores a state as one row of an object array
        return 3

    def __setitem__(self, key, value):
        setattr(self, key, value)

    def __getitem__(self, key):
        return getattr(self, key)

    @classmethod
    def from_dict(cls, mapping) -> "GraphData":
        d = dict(mapping)
        return cls(d.pop("x"), d.pop("edge_index"), d.pop("mask", None), **d)


class GraphBatch:
    """A collated minibatch in ``Batch.from_data_list`` layout (reference
    src/rlsp/agents/buffer.py:65-81) plus the packed CSR used by the kernels."""

    def __init__(self, x, edge_index, batch, mask, num_graphs: int, max_nodes: Optional[int] = None,
                 ptr=None, n_nodes=None, n_max: Optional[int] = None, groups: Optional[int] = None):
        self.x, self.edge_index, self.batch = x, edge_index, batch
        self._mask = mask
        self.ptr = ptr
        self.num_graphs = int(num_graphs)
        self.max_nodes = None if max_nodes is None else int(max_nodes)
        # Structured mask: when the per-state node counts are known, mask[b] is mask[:n_b,:,:,:n_b] = 1
        # over the scheduling shape (n_max, groups, n_max) and never has to be stored or read.
        self.n_nodes = n_nodes            # int32 [B] on the same device as x, or None
        self.n_max = None if n_max is None else int(n_max)
        self.groups = None if groups is None else int(groups)
        self._csr = None
        self._prep = {}                   # (hidden, num_layers) -> (buffer, event): see bwd_prep

    @property
    def structured(self) -> bool:
        return self.n_nodes is not None and self.n_max is not None and self.groups is not None

    @property
    def mask(self):
        """Dense [B, A] mask; generated on the device from ``n_nodes`` on first use if not given."""
        if self._mask is None and self.structured and self.n_nodes.is_cuda:
            from . import ops
            self._mask = ops.build_mask(self.n_nodes, self.n_max, self.groups)
        return self._mask

    @mask.setter
    def mask(self, value):
        self._mask = value

    def view(self) -> "GraphBatch":
        """Same tensors, no cached CSR (the packer runs again: used to time it as part of a step)."""
        return GraphBatch(self.x, self.edge_index, self.batch, self._mask, self.num_graphs, self.max_nodes,
                          self.ptr, self.n_nodes, self.n_max, self.groups)

    def view_shared(self) -> "GraphBatch":
        """Same tensors and the same packed CSR, as a separate object (for use on another stream: the
        lazily generated dense mask, if any, is then not shared half-built)."""
        gb = self.view()
        gb._csr = self._csr
        gb._prep = self._prep
        return gb

    # ---- construction -----------------------------------------------------------------------
    @staticmethod
    def from_data_list(data_list: Sequence) -> "GraphBatch":
        """Host-side collate of single observations (any objects with .x/.edge_index/.mask)."""
        xs, eis, masks, sizes = [], [], [], []
        off = 0
        for d in data_list:
            n = int(d.x.shape[0])
            xs.append(d.x)
            eis.append(d.edge_index + off)
            if getattr(d, "mask", None) is not None:
                masks.append(d.mask.reshape(1, -1))
            sizes.append(n)
            off += n
        sizes_t = torch.tensor(sizes, dtype=torch.int64)
        ptr = torch.zeros(len(sizes) + 1, dtype=torch.int64)
        ptr[1:] = torch.cumsum(sizes_t, 0)
        batch = torch.repeat_interleave(torch.arange(len(sizes), dtype=torch.int64), sizes_t)
        return GraphBatch(torch.cat(xs, 0), torch.cat(eis, 1), batch,
                          torch.cat(masks, 0) if masks else None, len(sizes),
                          max(sizes) if sizes else 0, ptr)

    @staticmethod
    def from_arrays(x, edge_index, batch, mask, num_graphs, max_nodes=None, ptr=None, pin=False,
                    n_nodes=None, n_max=None, groups=None):
        def conv(a, dt):
            if a is None:
                return None
            t = _t(a, dt)
            return t.pin_memory() if pin and not t.is_cuda else t
        return GraphBatch(conv(x, torch.float32), conv(edge_index, torch.int64), conv(batch, torch.int64),
                          conv(mask, torch.float32), num_graphs, max_nodes, conv(ptr, torch.int64),
                          conv(n_nodes, torch.int32), n_max, groups)

    @staticmethod
    def from_states(states, pin: bool = False, dense_mask: bool = True) -> "GraphBatch":
        """From a ``gsc_b200.synth.StateBatch`` (carries the node counts, so the mask is structured;
        ``dense_mask=False`` leaves the [B, A] mask to be generated on the device when needed)."""
        return GraphBatch.from_arrays(states.x, states.edge_index, states.batch,
                                      states.mask if dense_mask else None,
                                      states.num_graphs, int(states.n.max()) if len(states.n) else 0,
                                      states.ptr, pin=pin, n_nodes=states.n, n_max=states.n_max,
                                      groups=states.num_sfcs * states.num_sfs)

    @staticmethod
    def wrap(obs) -> "GraphBatch":
        """Adopt any reference-style observation (PyG Data/Batch look-alike).  Objects that do not
        say how many graphs they hold cost one device->host read, like PyG's own
        ``global_mean_pool`` does on ``batch.max()``."""
        if isinstance(obs, GraphBatch):
            return obs
        cached = getattr(obs, "_gsc_batch", None)
        if cached is not None and cached.x is obs.x and cached.mask is getattr(obs, "mask", None):
            return cached
        batch = getattr(obs, "batch", None)
        x = obs.x
        mask = getattr(obs, "mask", None)
        if batch is None:
            gb = GraphBatch(x, obs.edge_index, None, mask, 1, int(x.shape[0]))
        else:
            ng = getattr(obs, "num_graphs", None)
            if ng is None:
                ng = int(batch.max().item()) + 1 if batch.numel() else 1
            gb = GraphBatch(x, obs.edge_index, batch, mask, int(ng), None, getattr(obs, "ptr", None))
        try:
            obs._gsc_batch = gb
        except Exception:
            pass
        return gb

    # ---- reference surface ------------------------------------------------------------------
    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    def to(self, device, non_blocking: bool = True) -> "GraphBatch":
        device = torch.device(device)
        if self.x.device == device:
            return self
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        gb = GraphBatch(mv(self.x), mv(self.edge_index), mv(self.batch), mv(self._mask),
                        self.num_graphs, self.max_nodes, mv(self.ptr), mv(self.n_nodes), self.n_max, self.groups)
        return gb

    # ---- CUDA path --------------------------------------------------------------------------
    def csr(self):
        """(rowptr, col, graph_ptr) int32 on the device; packed once per batch and cached."""
        if self._csr is None:
            from . import ops
            rowptr, col, graph_ptr, stats = ops.pack_csr(self.edge_index, self.batch, self.num_nodes,
                                                         self.num_graphs)
            if self.max_nodes is None:       # one sync, only for foreign observations
                st = stats.tolist()
                if st[2] != 0:
                    raise ValueError(f"gsc_pack_csr: malformed graph batch (flags={st[2]})")
                self.max_nodes = max(1, int(st[1]))
            self._csr = (rowptr, col, graph_ptr, stats)
        return self._csr[:3]

    def bwd_prep(self, hidden: int, num_layers: int):
        """The topology-only index work of the embedder backward (``ops.gnn_bwd_prep``), computed once per
        packed batch on the stream of the first caller and shared by every backward pass over the batch
        (an update runs two); later callers on other streams wait for its event.  None if not applicable."""
        key = (int(hidden), int(num_layers))
        hit = self._prep.get(key)
        if hit is None:
            from . import ops
            rowptr, col, gptr = self.csr()
            buf = ops.gnn_bwd_prep(rowptr, col, gptr, self.num_graphs, self.max_nodes, hidden, num_layers)
            ev = torch.cuda.Event()
            ev.record()
            hit = self._prep[key] = (buf, ev)
        else:
            torch.cuda.current_stream().wait_event(hit[1])
        return hit[0]

    def pack_stats(self):
        """The packer's device-side stats array [Et, max_nodes, error flags, asymmetry flag]."""
        self.csr()
        return self._csr[3]

    def share_topology_from(self, other: "GraphBatch") -> None:
        """Reuse another batch's packed CSR (a transition's next state has the same network)."""
        if other._csr is not None and other.edge_index is self.edge_index:
            self._csr = other._csr
            self._prep = other._prep
            self.max_nodes = other.max_nodes
