"""Drop-in ``GNNEmbedder`` / ``Actor`` / ``QNetwork`` with the reference's call surface.

Mirrors reference src/rlsp/agents/models.py:10-154: same constructor arguments, ``forward``
signatures, attribute names (``.embedder``, ``.actor``, ``.critic``, ``.low``, ``.high``) and
state-dict keys (``embedder.encoder.{att,bias,lin_l.weight,...}``, ``actor.lins.{0,1}.{weight,bias}``
...), so ``SimpleDDPG`` and ``inference.py`` can use them unchanged.  The arithmetic runs in
libgsc_b200 (fused GNN kernel + concat-free GEMMs); autograd is supported through one
``torch.autograd.Function`` per network whose backward calls the hand-written backward kernels.

Every network keeps all of its parameters in ONE flat fp32 buffer (``.flat``); the ``nn.Parameter``s
are views into it.  That is what lets the fused Adam/soft-update kernel and the NCCL all-reduce run
over a single contiguous range.  The first MLP weight is stored with a small front pad and a
row stride that is a multiple of 4 floats so that the ``mask`` and ``action`` column blocks start
16-byte aligned (the logical ``[out, in]`` parameter is a strided view).
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

from . import ops
from .graph import GraphBatch

_ALIGN = 64  # floats; every region of a flat buffer starts 256-byte aligned
# A flat buffer's length is a multiple of this, so that it splits into equal, 16-byte aligned shards
# on any world size from 1 to 8 (sharded optimizer: reduce-scatter / all-gather need equal chunks).
_FLAT_MULTIPLE = 6720   # lcm(64, 2^5 * 3 * 5 * 7)


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


# ------------------------------------------------------------------------------------------------
# Parameter holders (state-dict key compatible with PyG's GATv2Conv / MLP)
# ------------------------------------------------------------------------------------------------
class _Lin(nn.Module):
    def __init__(self, in_dim: int, out_dim: int, glorot: bool):
        super().__init__()
        self.in_channels, self.out_channels = in_dim, out_dim
        w = torch.empty(out_dim, in_dim)
        if glorot:   # PyG Linear(weight_initializer='glorot')
            a = math.sqrt(6.0 / (in_dim + out_dim))
            w.uniform_(-a, a)
        else:        # torch default: kaiming_uniform(a=sqrt(5)) == U(+-1/sqrt(in))
            w.uniform_(-1.0 / math.sqrt(in_dim), 1.0 / math.sqrt(in_dim))
        self.weight = nn.Parameter(w)
        self.bias = nn.Parameter(torch.empty(out_dim).uniform_(-1.0 / math.sqrt(in_dim), 1.0 / math.sqrt(in_dim)))


class GATv2Params(nn.Module):
    """Parameters of one ``GATv2Conv(in, out, aggr='mean')`` (heads=1), PyG registration order."""

    def __init__(self, in_dim: int, out_dim: int):
        super().__init__()
        self.in_channels, self.out_channels = in_dim, out_dim
        a = math.sqrt(6.0 / (1 + out_dim))
        self.att = nn.Parameter(torch.empty(1, 1, out_dim).uniform_(-a, a))
        self.bias = nn.Parameter(torch.zeros(out_dim))
        self.lin_l = _Lin(in_dim, out_dim, glorot=True)
        self.lin_r = _Lin(in_dim, out_dim, glorot=True)

    def ordered(self) -> List[nn.Parameter]:
        return [self.att, self.bias, self.lin_l.weight, self.lin_l.bias, self.lin_r.weight, self.lin_r.bias]


class MLPParams(nn.Module):
    """PyG ``MLP(channel_list, norm=None)``: Linear -> ReLU -> ... -> Linear (plain last)."""

    def __init__(self, channel_list: Sequence[int]):
        super().__init__()
        self.channel_list = [int(c) for c in channel_list]
        self.lins = nn.ModuleList([_Lin(self.channel_list[i], self.channel_list[i + 1], glorot=False)
                                   for i in range(len(self.channel_list) - 1)])


def _flat_view(params: Sequence[torch.Tensor]) -> torch.Tensor:
    """The parameters as one flat tensor: a zero-copy view when they are laid out back to back."""
    first = params[0]
    off, ok = 0, True
    for p in params:
        ok = ok and p.is_contiguous() and p.data_ptr() == first.data_ptr() + 4 * off
        off += p.numel()
    if ok:
        return first.detach().reshape(-1).as_strided((off,), (1,))
    return torch.cat([p.detach().reshape(-1) for p in params])


# ------------------------------------------------------------------------------------------------
# GNN embedder
# ------------------------------------------------------------------------------------------------
class GNNEmbedder(nn.Module):
    """reference models.py:10-53.  ``forward(x, adj_t, batch)`` -> ``[num_graphs, hidden_dim]``."""

    def __init__(self, input_dim: int, hidden_dim: int, num_layers: int, aggr: str, num_iter: int,
                 dropout: float = 0.5):
        super().__init__()
        if aggr != "mean":
            raise NotImplementedError("gsc_b200 implements the reference's configured GNN_aggr='mean' only")
        if hidden_dim > ops._lib.MAX_HIDDEN:
            raise NotImplementedError(f"GNN_features > {ops._lib.MAX_HIDDEN} is not supported")
        self.input_dim, self.hidden_dim = int(input_dim), int(hidden_dim)
        self.num_iter, self.num_layers, self.dropout = int(num_iter), int(num_layers), dropout
        self.encoder = GATv2Params(input_dim, hidden_dim)
        self.process = nn.ModuleList([GATv2Params(hidden_dim, hidden_dim) for _ in range(num_layers - 1)])
        _flatten_into_own_buffer(self, self.ordered_params())

    def ordered_params(self) -> List[nn.Parameter]:
        ps = self.encoder.ordered()
        for conv in self.process:
            ps += conv.ordered()
        return ps

    def param_floats(self) -> int:
        return sum(p.numel() for p in self.ordered_params())

    def flat_params(self) -> torch.Tensor:
        return _flat_view(self.ordered_params())

    def _apply(self, fn, *a, **k):
        super()._apply(fn, *a, **k)
        if not getattr(self, "_adopted", False):
            _flatten_into_own_buffer(self, self.ordered_params())
        return self

    def embed(self, gb: GraphBatch, save: bool = False, params: Optional[torch.Tensor] = None):
        rowptr, col, gptr = gb.csr()
        return ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, gb.num_graphs, gb.max_nodes,
                                 self.flat_params() if params is None else params, self.hidden_dim,
                                 self.num_layers, self.num_iter, save=save)

    def embed_backward(self, gb: GraphBatch, saved, d_emb, d_params=None, params=None):
        rowptr, col, gptr = gb.csr()
        return ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, gb.num_graphs, gb.max_nodes,
                                 self.flat_params() if params is None else params, self.hidden_dim,
                                 self.num_layers, self.num_iter, saved, d_emb, d_params,
                                 pack_stats=gb.pack_stats())

    def forward(self, x, adj_t, batch):
        class _Obs:
            pass
        o = _Obs()
        o.x, o.edge_index, o.batch, o.mask = x, adj_t, batch, None
        gb = GraphBatch.wrap(o)
        return _EmbedFn.apply(self, gb, *self.ordered_params())


def _flatten_into_own_buffer(mod: nn.Module, params: Sequence[nn.Parameter]) -> None:
    total = sum(p.numel() for p in params)
    dev = params[0].device
    flat = torch.empty(_round_up(total, 4), dtype=torch.float32, device=dev)
    off = 0
    for p in params:
        n = p.numel()
        flat[off:off + n].copy_(p.detach().reshape(-1).to(torch.float32))
        p.data = flat[off:off + n].view(p.shape)
        off += n
    mod._own_flat = flat


class _EmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb_mod: GNNEmbedder, gb: GraphBatch, *params):
        emb, saved = emb_mod.embed(gb, save=any(ctx.needs_input_grad))
        ctx.emb_mod, ctx.gb, ctx.saved = emb_mod, gb, saved
        ctx.shapes = [p.shape for p in params]
        return emb

    @staticmethod
    def backward(ctx, d_emb):
        d = ctx.emb_mod.embed_backward(ctx.gb, ctx.saved, d_emb.contiguous())
        grads, off = [], 0
        for s in ctx.shapes:
            n = int(np.prod(s))
            grads.append(d[off:off + n].view(s))
            off += n
        return (None, None, *grads)


# ------------------------------------------------------------------------------------------------
# Flat network layout shared by Actor and QNetwork
# ------------------------------------------------------------------------------------------------
class _FlatNet(nn.Module):
    """Embedder + MLP whose parameters are views into ``self.flat`` (see module docstring)."""

    mlp_attr = "mlp"

    def _mlp(self) -> MLPParams:
        return getattr(self, self.mlp_attr)

    def _layout(self):
        """[(name, param, offset, numel_region, shape, ld, pad)] in state-dict order."""
        items, off = [], 0
        for p in self.embedder.ordered_params():
            items.append((p, off, p.shape, None, 0))
            off += p.numel()
        self._gnn_floats = off
        off = _round_up(off, _ALIGN)
        mlp = self._mlp()
        for i, lin in enumerate(mlp.lins):
            n_out, n_in = lin.out_channels, lin.in_channels
            pad = (-self.embedder.hidden_dim) % 4 if i == 0 else 0
            ld = _round_up(pad + n_in, 4)
            items.append((lin.weight, off, (n_out, n_in), ld, pad))
            off = _round_up(off + n_out * ld, _ALIGN)
            items.append((lin.bias, off, (n_out,), None, 0))
            off = _round_up(off + n_out, _ALIGN)
        return items, _round_up(off, _FLAT_MULTIPLE)

    def _flatten(self) -> None:
        items, total = self._layout()
        dev = items[0][0].device
        flat = torch.zeros(total, dtype=torch.float32, device=dev)
        for p, off, shape, ld, pad in items:
            view = self._view(flat, off, shape, ld, pad)
            view.copy_(p.detach().to(torch.float32))
            p.data = view
        self.flat = flat
        self.embedder._adopted = True
        self._items = [(off, shape, ld, pad) for _, off, shape, ld, pad in items]

    @staticmethod
    def _view(flat, off, shape, ld, pad):
        if ld is None:
            n = int(np.prod(shape))
            return flat[off:off + n].view(shape)
        n_out, n_in = shape
        return flat[off:off + n_out * ld].view(n_out, ld)[:, pad:pad + n_in]

    def views(self, flat: torch.Tensor):
        """Structured views of any buffer with this network's layout (params, grads, Adam state):
        (gnn_flat, [(W_i, b_i, Wp_i, pad_i), ...]).  ``W_i`` is the logical ``[out, in]`` weight,
        ``Wp_i`` the same rows including the ``pad_i`` leading pad columns (16-byte aligned base,
        what the TMA-fed tensor-core GEMM wants)."""
        gnn = flat[:self._gnn_floats]
        n_gnn = len(self.embedder.ordered_params())
        lins = []
        rest = self._items[n_gnn:]
        for i in range(0, len(rest), 2):
            (wo, ws, wld, wpad), (bo, bs, _, _) = rest[i], rest[i + 1]
            n_out, n_in = ws
            wp = flat[wo:wo + n_out * wld].view(n_out, wld)[:, :wpad + n_in]
            lins.append((wp[:, wpad:], self._view(flat, bo, bs, None, 0), wp, wpad))
        return gnn, lins

    def grad_regions(self):
        """(gnn_region, [layer regions]) as (lo, hi) float offsets into a flat buffer: the embedder block
        and, per MLP layer, its weight + bias block.  Everything outside them is alignment padding."""
        n_gnn = len(self.embedder.ordered_params())
        rest = self._items[n_gnn:]
        layers = []
        for i in range(0, len(rest), 2):
            (wo, _, _, _), (bo, bs, _, _) = rest[i], rest[i + 1]
            layers.append((wo, bo + int(bs[0])))
        return (0, rest[0][0]), layers

    def param_views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        """Per-parameter views of ``flat`` in ``self.parameters()`` order."""
        return [self._view(flat, off, shape, ld, pad) for off, shape, ld, pad in self._items]

    def ordered_params(self) -> List[nn.Parameter]:
        ps = self.embedder.ordered_params()
        for lin in self._mlp().lins:
            ps += [lin.weight, lin.bias]
        return ps

    def _apply(self, fn, *a, **k):
        super()._apply(fn, *a, **k)
        self._flatten()
        return self

    def _check_flat(self) -> torch.Tensor:
        p0 = self.embedder.encoder.att
        if p0.data_ptr() != self.flat.data_ptr() or p0.device != self.flat.device:
            self._flatten()   # someone re-assigned .data (e.g. deepcopy/unpickle): restore aliasing
        return self.flat


# ------------------------------------------------------------------------------------------------
# Functional forward / backward shared by the autograd Functions and the fused DDPG updater
# ------------------------------------------------------------------------------------------------
def _structured(gb: GraphBatch) -> bool:
    return ops.use_structured_mask and gb.structured and gb.n_nodes.is_cuda


def first_layer_table(net: _FlatNet, flat: torch.Tensor, gb: GraphBatch) -> Optional[torch.Tensor]:
    """The per-node-count table that replaces the mask column block of the first dense layer for a
    structured batch (None otherwise).  Depends on the weights only, so it can be computed ahead of
    the forward pass (``net_forward(..., table=...)``), e.g. while the batch is still being packed."""
    if not _structured(gb):
        return None
    _, lins = net.views(flat)
    _, _, Wp, pad = lins[0]
    H, A = net.embedder.hidden_dim, net.action_dim
    return ops.mask_table(Wp[:, pad + H:pad + H + A], gb.n_max, gb.groups)


def _single_unit(W: torch.Tensor) -> bool:
    """A dense layer with one output unit and a K the fused head kernels take (ops.linear_n1_*)."""
    return ops.fuse_head and int(W.shape[0]) == 1 and int(W.shape[1]) <= 1024 and W.stride(1) == 1


def net_forward(net: _FlatNet, flat: torch.Tensor, gb: GraphBatch, action: Optional[torch.Tensor],
                save: bool, mask_output: bool = False, clamp1: bool = False,
                save_gnn: Optional[bool] = None, table: Optional[torch.Tensor] = None,
                table_ready: Optional[torch.cuda.Event] = None):
    """emb = GNN(gb); z = MLP(cat[emb, mask(, action)]); returns (z [* mask] [clamped], ctx).

    ``save`` keeps what ``net_backward`` needs; ``save_gnn=False`` skips the embedder activations
    when only the gradient w.r.t. the action input will be asked for.  The embedding is written
    straight into a ``[B, pad + H]`` buffer whose leading ``pad`` columns are zero and line up with
    the pad columns of the first weight, so every K-segment of the first layer is 16-byte aligned.

    When the batch carries its node counts (``gb.structured``) the mask is never read: its column
    block of the first layer becomes a per-``n`` table lookup and the output mask is computed from
    ``n`` in the GEMM epilogue."""
    gnn, lins = net.views(flat)
    H = net.embedder.hidden_dim
    pad0 = lins[0][3]
    st = _structured(gb)
    embp = torch.zeros(gb.num_graphs, pad0 + H, dtype=torch.float32, device=flat.device)
    rowptr, col, gptr = gb.csr()
    _, saved = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, gb.num_graphs, gb.max_nodes, gnn, H,
                                 net.embedder.num_layers, net.embedder.num_iter,
                                 save=save if save_gnn is None else save_gnn, out=embp[:, pad0:])
    A = net.action_dim
    smask = (gb.n_nodes, gb.n_max, gb.groups) if st else None
    inputs = [embp, None if st else gb.mask] + ([action] if action is not None else [])
    hs = []
    h = None
    n = len(lins)
    q_head = None
    for i, (W, b, Wp, pad) in enumerate(lins):
        last = i == n - 1
        if last and q_head is not None:               # the one-unit head came with the layer below
            h = q_head
            break
        # a one-unit last layer (the critic's head) rides on this layer's call
        head = (lins[i + 1][0], lins[i + 1][1]) if (i == n - 2 and _single_unit(lins[i + 1][0]) and not mask_output
                                                    and not clamp1) else None
        out_mul = gb.mask if (mask_output and last and not st) else None
        out_smask = smask if (mask_output and last and st) else None
        if i == 0 and st:
            if table is None:
                table = ops.mask_table(Wp[:, pad + H:pad + H + A], gb.n_max, gb.groups)
            elif table_ready is not None:          # built on another stream: needed only from here on
                torch.cuda.current_stream().wait_event(table_ready)
            ins = [embp] + ([action] if action is not None else [])
            offs = [0] + ([pad + H + A] if action is not None else [])
            h = ops.linear_fwd(ins, Wp, b, relu=not last, mul=out_mul, clamp1=clamp1 and last,
                               col_offsets=offs, row_table=(table, gb.n_nodes), smask=out_smask, head=head)
            if head is not None:
                h, q_head = h
        elif last and i > 0 and _single_unit(W) and out_mul is None and out_smask is None and not clamp1:
            h = ops.linear_n1_fwd(h, W, b)             # the critic's head: one launch, no GEMM machinery
        else:
            h = ops.linear_fwd(inputs if i == 0 else [h], Wp if i == 0 else W, b, relu=not last,
                               mul=out_mul, clamp1=clamp1 and last, smask=out_smask, head=head)
            if head is not None:
                h, q_head = h
        if not last:
            hs.append(h)
    ctx = (embp, saved, inputs, hs, st) if save else None
    return h, ctx


def net_backward(net: _FlatNet, flat: torch.Tensor, gb: GraphBatch, ctx, dz: torch.Tensor,
                 grad_flat: Optional[torch.Tensor], want_action_grad: bool = False,
                 action_grad_masked: bool = False, side: Optional[torch.cuda.Stream] = None,
                 after_layer=None, mlp_done=None):
    """Back-propagate ``dz`` (gradient w.r.t. the pre-mask MLP output).

    Writes parameter gradients into ``grad_flat`` (same layout as ``flat``) unless it is None, and
    returns the gradient w.r.t. the action input (times the mask if ``action_grad_masked``) when
    requested.

    ``side``: a second stream for the weight / bias gradients of the dense layers.  They only read
    the layer's output gradient, while the chain that continues on the current stream (input
    gradients -> embedder backward) is what everything else waits for; the embedder backward is a
    persistent kernel whose last wave leaves more than half of the SMs idle, which the side stream's
    kernels fill.  The streams are joined before returning.  ``after_layer(i)`` is called on that
    stream once layer i's weight and bias gradients are enqueued (data parallelism all-reduces them
    there, behind the embedder backward).  ``mlp_done()`` is called on that stream once every dense
    layer's gradients are enqueued AND the current stream has read the dense weights for the last time
    (the optimizer step of the dense layers then overlaps the embedder backward)."""
    embp, saved, inputs, hs, st = ctx
    main = torch.cuda.current_stream()

    def weight_grads(fn):
        if side is None or grad_flat is None:
            fn()
            return
        ev = torch.cuda.Event()
        ev.record(main)                     # the layer's output gradient is complete here
        with torch.cuda.stream(side):
            side.wait_event(ev)
            fn()

    gnn, lins = net.views(flat)
    g_gnn, g_lins = net.views(grad_flat) if grad_flat is not None else (None, None)
    H = net.embedder.hidden_dim
    A = net.action_dim
    d_action = None
    d = dz
    keep = []          # output gradients that the side stream reads: alive until the streams are joined
    for i in range(len(lins) - 1, -1, -1):
        W, _, Wp, pad = lins[i]
        keep.append(d)
        if i == len(lins) - 1 and i > 0 and _single_unit(W):
            # head with one output unit: input gradient (through the ReLU below), weight and bias gradient
            # in one launch on this stream
            d = ops.linear_n1_bwd(d, hs[i - 1], W, relu_input=True,
                                  dW=g_lins[i][0] if grad_flat is not None else None,
                                  db=g_lins[i][1] if grad_flat is not None else None)
            if grad_flat is not None and after_layer is not None:
                weight_grads(lambda i=i: after_layer(i))
            continue
        if grad_flat is not None:
            def layer_grads(i=i, d=d, pad=pad):
                if i == 0 and st:
                    gWp = g_lins[0][2]
                    ops.linear_wgrad(d, [embp], gWp[:, :pad + H])
                    ops.mask_wgrad(d, gb.n_nodes, gb.n_max, gb.groups, gWp[:, pad + H:pad + H + A])
                    if len(inputs) > 2:
                        ops.linear_wgrad(d, [inputs[2]], gWp[:, pad + H + A:pad + H + 2 * A])
                else:
                    ops.linear_wgrad(d, inputs if i == 0 else [hs[i - 1]], g_lins[i][2] if i == 0 else g_lins[i][0])
                ops.colsum(d, out=g_lins[i][1])
                if after_layer is not None:
                    after_layer(i)
            weight_grads(layer_grads)
        if i > 0:
            d = ops.linear_dgrad(d, W, 0, int(W.shape[1]), gate=hs[i - 1])
        else:
            if want_action_grad:
                mul = gb.mask if (action_grad_masked and not st) else None
                sm = (gb.n_nodes, gb.n_max, gb.groups) if (action_grad_masked and st) else None
                d_action = ops.linear_dgrad(d, Wp, pad + H + A, A, mul=mul, smask=sm)
            if grad_flat is not None:
                d_embp = ops.linear_dgrad(d, Wp, 0, pad + H)
                if mlp_done is not None and side is not None:
                    reads_done = torch.cuda.Event()
                    reads_done.record(main)             # last read of the dense weights on this stream
                    with torch.cuda.stream(side):
                        side.wait_event(reads_done)
                        mlp_done()
                rowptr, col, gptr = gb.csr()
                ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, gb.num_graphs, gb.max_nodes, gnn, H,
                                  net.embedder.num_layers, net.embedder.num_iter, saved, d_embp[:, pad:],
                                  d_params=g_gnn, pack_stats=gb.pack_stats(),
                                  prep=gb.bwd_prep(H, net.embedder.num_layers))
    if side is not None and grad_flat is not None:
        done = torch.cuda.Event()
        done.record(side)
        main.wait_event(done)               # gradients complete, and nothing read on `side` is released early
    del keep
    return d_action


class _NetFn(torch.autograd.Function):
    """autograd bridge: y = net(obs[, action]); backward runs the CUDA backward kernels."""

    @staticmethod
    def forward(ctx, net: _FlatNet, gb: GraphBatch, action, apply_mask: bool, *params):
        flat = net._check_flat()
        need = any(ctx.needs_input_grad)
        z, saved_ctx = net_forward(net, flat, gb, action, save=need, mask_output=apply_mask,
                                   save_gnn=need and any(ctx.needs_input_grad[4:]))
        ctx.net, ctx.gb, ctx.saved_ctx, ctx.apply_mask = net, gb, saved_ctx, apply_mask
        ctx.has_action = action is not None
        return z

    @staticmethod
    def backward(ctx, dy):
        net, gb = ctx.net, ctx.gb
        dy = dy.contiguous()
        dz = ops.mul_gate(dy, mul=gb.mask) if ctx.apply_mask else dy
        want_p = any(ctx.needs_input_grad[4:])
        grad_flat = torch.zeros_like(net.flat) if want_p else None
        want_a = ctx.has_action and ctx.needs_input_grad[2]
        d_action = net_backward(net, net.flat, gb, ctx.saved_ctx, dz, grad_flat, want_action_grad=want_a)
        if not want_p:
            return (None, None, d_action, None, *([None] * len(net._items)))
        return (None, None, d_action, None, *net.param_views(grad_flat))


def _np_prod(shape) -> int:
    return int(np.prod(shape))


class QNetwork(_FlatNet):
    """reference models.py:55-92.  ``forward(obs, a)`` -> ``[B, 1]``."""

    mlp_attr = "critic"

    def __init__(self, agent_helper):
        super().__init__()
        cfg = agent_helper.config
        self.graph_mode = cfg["graph_mode"]
        if self.graph_mode is not True:
            raise NotImplementedError("gsc_b200 implements the reference's graph_mode=True path")
        hidden_layers = list(cfg["critic_hidden_layer_nodes"])
        obs_space = agent_helper.env.observation_space
        action_space = agent_helper.env.action_space
        feature_size = int(cfg["GNN_features"])
        self.embedder = GNNEmbedder(input_dim=obs_space["nodes"].shape[-1], hidden_dim=feature_size,
                                    num_layers=int(cfg["GNN_num_layers"]), num_iter=int(cfg["GNN_num_iter"]),
                                    aggr=cfg["GNN_aggr"])
        self.action_dim = _np_prod(action_space.shape)
        hidden_layers.insert(0, feature_size + 2 * self.action_dim)
        hidden_layers.append(1)
        self.critic = MLPParams(hidden_layers)
        self._flatten()

    def forward(self, x, a):
        gb = GraphBatch.wrap(x)
        return _NetFn.apply(self, gb, a.contiguous(), False, *self.ordered_params())


class Actor(_FlatNet):
    """reference models.py:95-154.  ``forward(obs)`` -> ``[B, A]`` (MLP output times mask)."""

    mlp_attr = "actor"

    def __init__(self, agent_helper):
        super().__init__()
        cfg = agent_helper.config
        self.graph_mode = cfg["graph_mode"]
        if self.graph_mode is not True:
            raise NotImplementedError("gsc_b200 implements the reference's graph_mode=True path")
        hidden_layers = list(cfg["actor_hidden_layer_nodes"])
        obs_space = agent_helper.env.observation_space
        action_space = agent_helper.env.action_space
        feature_size = int(cfg["GNN_features"])
        self.embedder = GNNEmbedder(input_dim=obs_space["nodes"].shape[-1], hidden_dim=feature_size,
                                    num_layers=int(cfg["GNN_num_layers"]), num_iter=int(cfg["GNN_num_iter"]),
                                    aggr=cfg["GNN_aggr"])
        self.action_dim = _np_prod(action_space.shape)
        hidden_layers.insert(0, feature_size + self.action_dim)
        hidden_layers.append(self.action_dim)
        self.actor = MLPParams(hidden_layers)
        self.low = action_space.low
        self.high = action_space.high
        self._flatten()

    def scale_action(self, action: np.ndarray) -> np.ndarray:
        """[low, high] -> [-1, 1]  (reference models.py:127-135)."""
        return 2.0 * ((action - self.low) / (self.high - self.low)) - 1.0

    def unscale_action(self, scaled_action: np.ndarray) -> np.ndarray:
        """[-1, 1] -> [low, high]  (reference models.py:137-144)."""
        return self.low + (0.5 * (scaled_action + 1.0) * (self.high - self.low))

    def forward(self, x):
        gb = GraphBatch.wrap(x)
        return _NetFn.apply(self, gb, None, True, *self.ordered_params())


# ------------------------------------------------------------------------------------------------
# Minimal stand-ins for the two gym objects the constructors read (agent_helper.env.*_space)
# ------------------------------------------------------------------------------------------------
class BoxSpace:
    def __init__(self, low, high, shape, dtype=np.float32):
        self.shape = tuple(shape)
        self.low = np.full(self.shape, low, dtype=dtype)
        self.high = np.full(self.shape, high, dtype=dtype)
        self.dtype = dtype

    def sample(self):
        return np.random.uniform(self.low, self.high).astype(self.dtype)


class EnvSpec:
    """``agent_helper.env`` look-alike: the spaces of reference
    src/rlsp/envs/environment_limits.py:44-90 for ``N_max`` nodes."""

    def __init__(self, n_max: int, num_sfcs: int = 1, num_sfs: int = 3, node_feat: int = 3):
        self.n_max, self.num_sfcs, self.num_sfs = n_max, num_sfcs, num_sfs
        self.action_space = BoxSpace(0.0, 1.0, (n_max * num_sfcs * num_sfs * n_max,))
        self.observation_space = {"nodes": BoxSpace(0.0, 1.0, (n_max, node_feat))}


class AgentSpec:
    """``AgentHelper`` look-alike holding ``config`` and ``env`` (reference agent_helper.py:4-43)."""

    DEFAULTS = dict(graph_mode=True, GNN_features=22, GNN_num_layers=2, GNN_num_iter=2, GNN_aggr="mean",
                    actor_hidden_layer_nodes=[256], critic_hidden_layer_nodes=[64], learning_rate=1e-3,
                    gamma=0.99, target_model_update=1e-4, rand_mu=0.0, rand_sigma=0.3,
                    nb_steps_warmup_critic=200, mem_limit=10000)

    def __init__(self, n_max: int, num_sfcs: int = 1, num_sfs: int = 3, node_feat: int = 3, **overrides):
        self.config = dict(self.DEFAULTS)
        self.config.update(overrides)
        self.env = EnvSpec(n_max, num_sfcs, num_sfs, node_feat)
        self.seed = 1234
        self.episode_steps = 200
