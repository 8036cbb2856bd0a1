"""Tensor-level wrappers over the C-ABI (device pointers of torch CUDA tensors -> libgsc_b200).

torch is used here for device memory and streams only; all arithmetic happens in the library.
Every wrapper enqueues on torch's current stream and never synchronises.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import EPI_BIAS, EPI_CLAMP1, EPI_GATE, EPI_MUL, EPI_RELU, GemmArgs, check

_workspaces = {}
launch_count = 0          # kernels enqueued through this module (bench.py reports it)

# Arithmetic of the dense layers (tolerances of the modes are stated separately in tests/ and DESIGN.md):
#   "tf32x3" (default) tcgen05 tensor cores, fp32-faithful: operands split hi + lo in shared memory,
#                      hi*hi + hi*lo + lo*hi with fp32 accumulation in TMEM -- the parity mode AND the
#                      mode bench.py times;
#   "tf32"             single-pass TF32 operands (10 mantissa bits), the fast mode;
#   "fp32"             exact FMA kernel (also what the tensor-core modes fall back to for operands
#                      that TMA cannot map and for tiny problems).
precision = "tf32x3"
PRECISIONS = ("fp32", "tf32", "tf32x3")
EPI_TF32 = 32
EPI_TF32X3 = 256
# Use the node counts instead of the dense mask whenever a batch carries them (GraphBatch.structured).
use_structured_mask = True
# Layers with one output unit (the critic's head) through the fused one-launch kernels (linear_n1.cu)
# instead of gsc_gemm; False restores the GEMM route (tests compare the two).
fuse_head = True


class use_precision:
    """Context manager: ``with ops.use_precision("tf32"): ...``"""

    def __init__(self, mode: str):
        if mode not in PRECISIONS:
            raise ValueError(mode)
        self.mode = mode

    def __enter__(self):
        global precision
        self.prev, precision = precision, self.mode
        return self

    def __exit__(self, *exc):
        global precision
        precision = self.prev
        return False


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.GscError("gsc_b200 ops need CUDA tensors (there is no CPU fallback)")


def workspace(nbytes: int, device, tag: str = "main", zeroed: bool = False) -> Tuple[Optional[torch.Tensor], int]:
    """Grow-only scratch buffer per (device, tag, stream); the library itself never allocates.
    ``zeroed``: allocated as zeros (for kernels that keep a self-resetting counter in it; the tag must
    then be theirs alone)."""
    if nbytes <= 0:
        return None, 0
    # one buffer per stream: launches on different streams may run concurrently (DDPGUpdater forks the
    # independent forward passes of an update onto side streams)
    key = (torch.device(device).index, tag, _stream())
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = (torch.zeros if zeroed else torch.empty)(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws, ws.numel()


def _pack_launches(num_nodes: int, E: int) -> int:
    """Kernels one gsc_pack_csr call launches (pack_csr.cu: fused scan finish up to 2048 scan tiles)."""
    small = (max(int(num_nodes), 1) + 1023) // 1024 <= 2048
    scan = 2 if small else 4                    # tile scan + (finish | block scan, add, max_nodes)
    if E > 0:
        return 1 + 1 + scan + 3                 # init, count(+graph_ptr), scan, fill, rows, symmetry
    return 1 + 1 + scan + (1 if num_nodes > 0 else 0)


def _count(n: int):
    global launch_count
    launch_count += n


# Optional per-entry-point CUDA-event timing (bench.py uses it for the roofline of the dominant
# kernel).  ``timer`` maps name -> list of (start_event, end_event); events are recorded on the
# launching stream around the C-ABI call, no synchronisation happens here.
timer = None
timeline = None       # optional list of (name, start_event, end_event, stream handle) in launch order


class _Timed:
    __slots__ = ("name", "ev")

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if timer is not None:
            self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            self.ev[0].record(torch.cuda.current_stream())
        return self

    def __exit__(self, *exc):
        if timer is not None:
            self.ev[1].record(torch.cuda.current_stream())
            timer.setdefault(self.name, []).append(self.ev)
            if timeline is not None:
                timeline.append((self.name, self.ev[0], self.ev[1], torch.cuda.current_stream().cuda_stream))
        return False


def _row_major(t: torch.Tensor, what: str) -> torch.Tensor:
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1) or t.dtype != torch.float32:
        raise _lib.GscError(f"{what}: need a 2-D fp32 tensor with unit column stride, got "
                            f"shape={tuple(t.shape)} stride={t.stride()} dtype={t.dtype}")
    return t


def _ld(t: torch.Tensor) -> int:
    return int(t.stride(0)) if t.shape[0] > 1 else max(int(t.stride(0)), int(t.shape[1]))


# ---------------------------------------------------------------------------------------------
def pack_csr(edge_index: torch.Tensor, batch: Optional[torch.Tensor], num_nodes: int, num_graphs: int):
    """COO int64 -> (rowptr, col, graph_ptr, stats) int32 on the device (include/gsc_b200.h)."""
    _need_cuda(edge_index, batch)
    lib = _lib.load()
    dev = edge_index.device
    edge_index = edge_index.contiguous()
    if edge_index.dtype != torch.int64:
        edge_index = edge_index.to(torch.int64)
    if batch is not None:
        batch = batch.contiguous().to(torch.int64)
    E = int(edge_index.shape[1])
    rowptr = torch.empty(num_nodes + 1, dtype=torch.int32, device=dev)
    col = torch.empty(E + num_nodes, dtype=torch.int32, device=dev)
    graph_ptr = torch.empty(num_graphs + 1, dtype=torch.int32, device=dev)
    stats = torch.empty(4, dtype=torch.int32, device=dev)
    ws, wsn = workspace(lib.gsc_pack_csr_workspace_bytes(num_nodes, E), dev, "pack")
    with _Timed("pack_csr"):
        check(lib.gsc_pack_csr(_ptr(edge_index), E, _ptr(batch), num_nodes, num_graphs, _ptr(rowptr),
                               _ptr(col), _ptr(graph_ptr), _ptr(stats), _ptr(ws), wsn, _stream()),
              "gsc_pack_csr")
    _count(_pack_launches(num_nodes, E))
    return rowptr, col, graph_ptr, stats


def pack_csr_noalloc(edge_index: torch.Tensor, batch: Optional[torch.Tensor], num_nodes: int, num_graphs: int,
                     out: Optional[tuple] = None):
    """``pack_csr`` through the entry point that takes no workspace and allocates nothing (its
    scratch is a fixed array inside the library, at most ``_lib.PACK_NOALLOC_MAX_NODES`` nodes).

    For the per-step rollout / serving loop, where a device allocation would stall the training
    stream, and for hosts without an allocator.  Pass ``out=(rowptr, col, graph_ptr, stats)``
    preallocated int32 tensors to keep the call itself allocation-free; calls that overlap on one
    GPU share the scratch and must be serialised."""
    _need_cuda(edge_index, batch)
    lib = _lib.load()
    dev = edge_index.device
    if edge_index.dtype != torch.int64 or not edge_index.is_contiguous():
        raise _lib.GscError("pack_csr_noalloc: edge_index must be a contiguous int64 tensor")
    if batch is not None and (batch.dtype != torch.int64 or not batch.is_contiguous()):
        raise _lib.GscError("pack_csr_noalloc: batch must be a contiguous int64 tensor")
    E = int(edge_index.shape[1])
    if out is None:
        out = (torch.empty(num_nodes + 1, dtype=torch.int32, device=dev),
               torch.empty(E + num_nodes, dtype=torch.int32, device=dev),
               torch.empty(num_graphs + 1, dtype=torch.int32, device=dev),
               torch.empty(4, dtype=torch.int32, device=dev))
    rowptr, col, graph_ptr, stats = out
    with _Timed("pack_csr_noalloc"):
        check(lib.gsc_pack_csr_noalloc(_ptr(edge_index), E, _ptr(batch), num_nodes, num_graphs, _ptr(rowptr),
                                       _ptr(col), _ptr(graph_ptr), _ptr(stats), _stream()),
              "gsc_pack_csr_noalloc")
    _count(_pack_launches(num_nodes, E))
    return rowptr, col, graph_ptr, stats


def build_mask(n: torch.Tensor, n_max: int, groups_per_node: int, out: Optional[torch.Tensor] = None):
    """Structured action mask [B, n_max*groups*n_max] from per-state node counts (int32, device)."""
    _need_cuda(n)
    B = int(n.shape[0])
    A = n_max * groups_per_node * n_max
    if out is None:
        out = torch.empty(B, A, dtype=torch.float32, device=n.device)
    with _Timed(f"build_mask[{B}x{A}]"):
        check(_lib.load().gsc_build_mask(_ptr(n), B, n_max, groups_per_node, _ptr(out), _stream()), "gsc_build_mask")
    _count(1)
    return out


def set_sm_reserve(num_sms: int) -> None:
    """SMs the persistent embedder backward leaves free for concurrent NCCL kernels (data parallel)."""
    check(_lib.load().gsc_set_sm_reserve(int(num_sms)), "gsc_set_sm_reserve")


def node_features(ingress: torch.Tensor, usage: torch.Tensor, cap: torch.Tensor, graph_ptr: torch.Tensor,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x[Nt, 3] = [ingress_traffic, node_load, node_cap] of a batch of states, normalised per state as
    the reference's ``_parse_state_as_graph`` does (include/gsc_b200.h: gsc_node_features)."""
    _need_cuda(ingress, usage, cap, graph_ptr)
    Nt, B = int(ingress.shape[0]), int(graph_ptr.shape[0]) - 1
    if out is None:
        out = torch.empty(Nt, 3, dtype=torch.float32, device=ingress.device)
    f32 = lambda t: t.contiguous().to(torch.float32)
    with _Timed("node_features"):
        check(_lib.load().gsc_node_features(_ptr(f32(ingress)), _ptr(f32(usage)), _ptr(f32(cap)),
                                            _ptr(graph_ptr.contiguous().to(torch.int32)), B, _ptr(out), _stream()),
              "gsc_node_features")
    _count(1)
    return out


def replay_gather(store_x, store_ei, store_n, store_ecount, idx, node_ptr, edge_ptr, num_nodes: int,
                  num_edges: int, x_out=None, ei_out=None, batch_out=None, n_out=None):
    """Gather the transitions ``idx`` from padded device storage into Batch layout (see gsc_b200.h)."""
    lib = _lib.load()
    dev = store_x.device
    B = int(idx.shape[0])
    n_max, F = int(store_x.shape[1]), int(store_x.shape[2])
    e_max = int(store_ei.shape[2])
    if x_out is None:
        x_out = torch.empty(num_nodes, F, dtype=torch.float32, device=dev)
    if ei_out is None:
        ei_out = torch.empty(2, num_edges, dtype=torch.int64, device=dev)
    if batch_out is None:
        batch_out = torch.empty(num_nodes, dtype=torch.int64, device=dev)
    if n_out is None:
        n_out = torch.empty(B, dtype=torch.int32, device=dev)
    with _Timed("replay_gather"):
        check(lib.gsc_replay_gather(_ptr(store_x), _ptr(store_ei), _ptr(store_n), _ptr(store_ecount), _ptr(idx),
                                    _ptr(node_ptr), _ptr(edge_ptr), B, n_max, F, e_max, num_edges, _ptr(x_out),
                                    _ptr(ei_out), _ptr(batch_out), _ptr(n_out), _stream()), "gsc_replay_gather")
    _count(1)
    return x_out, ei_out, batch_out, n_out


def gather_rows(src: torch.Tensor, idx: torch.Tensor, out: Optional[torch.Tensor] = None):
    """out[b] = src[idx[b]] for a [cap, A] (or [cap]) fp32 device tensor."""
    A = int(src.shape[1]) if src.dim() == 2 else 1
    B = int(idx.shape[0])
    if out is None:
        out = torch.empty((B, A) if src.dim() == 2 else (B,), dtype=torch.float32, device=src.device)
    with _Timed(f"gather_rows[{B}x{A}]"):
        check(_lib.load().gsc_gather_rows(_ptr(src), _ptr(idx), B, A, _ptr(out), _stream()), "gsc_gather_rows")
    _count(1)
    return out


def gnn_param_floats(in_dim: int, hidden: int, num_layers: int) -> int:
    return int(_lib.load().gsc_gnn_param_floats(in_dim, hidden, num_layers))


def gnn_embed_fwd(x, rowptr, col, graph_ptr, num_graphs: int, max_nodes: int, params: torch.Tensor,
                  hidden: int, num_layers: int, num_iter: int, save: bool = False,
                  out: Optional[torch.Tensor] = None):
    """``out``: optional [num_graphs, hidden] view with unit column stride (any row stride)."""
    _need_cuda(x, params)
    lib = _lib.load()
    x = x.contiguous()
    Nt, in_dim = int(x.shape[0]), int(x.shape[1])
    emb = out if out is not None else torch.empty(num_graphs, hidden, dtype=torch.float32, device=x.device)
    saved = None
    if save:
        saved = torch.empty(int(lib.gsc_gnn_saved_floats(Nt, hidden, num_layers, num_iter)),
                            dtype=torch.float32, device=x.device)
    with _Timed("gnn_fwd"):
        check(lib.gsc_gnn_embed_fwd(_ptr(x), Nt, in_dim, _ptr(rowptr), _ptr(col), _ptr(graph_ptr),
                                    num_graphs, max_nodes, _ptr(params), hidden, num_layers, num_iter,
                                    _ptr(emb), _ld(emb), _ptr(saved), _stream()), "gsc_gnn_embed_fwd")
    _count(1)
    return emb, saved


def gnn_embed_bwd(x, rowptr, col, graph_ptr, num_graphs: int, max_nodes: int, params, hidden: int,
                  num_layers: int, num_iter: int, saved, d_emb, d_params: Optional[torch.Tensor] = None,
                  pack_stats: Optional[torch.Tensor] = None, prep: Optional[torch.Tensor] = None):
    """``prep``: the buffer of ``gnn_bwd_prep`` for this batch and embedder shape (optional)."""
    _need_cuda(x, params, saved, d_emb)
    lib = _lib.load()
    x = x.contiguous()
    if d_emb.dim() != 2 or (d_emb.shape[1] > 1 and d_emb.stride(1) != 1):
        d_emb = d_emb.contiguous()
    Nt, in_dim = int(x.shape[0]), int(x.shape[1])
    if d_params is None:
        d_params = torch.empty_like(params)
    ws, wsn = workspace(lib.gsc_gnn_bwd_workspace_bytes(Nt, num_graphs, max_nodes, in_dim, hidden, num_layers,
                                                        num_iter), x.device, "gnn_bwd")
    with _Timed("gnn_bwd"):
        check(lib.gsc_gnn_embed_bwd_prepared(_ptr(x), Nt, in_dim, _ptr(rowptr), _ptr(col), _ptr(graph_ptr),
                                             num_graphs, max_nodes, _ptr(params), hidden, num_layers, num_iter,
                                             _ptr(saved), _ptr(d_emb), _ld(d_emb), _ptr(d_params), _ptr(pack_stats),
                                             _ptr(ws), wsn, _ptr(prep), _stream()),
              "gsc_gnn_embed_bwd")
    _count(4 if prep is not None else 5)
    return d_params


def gnn_bwd_prep(rowptr, col, graph_ptr, num_graphs: int, max_nodes: int, hidden: int, num_layers: int):
    """Topology-only index work of the embedder backward for one packed batch (None if the batch has no
    such part: graphs of more than 128 nodes).  Pass the result to every ``gnn_embed_bwd`` over the batch."""
    lib = _lib.load()
    n = int(lib.gsc_gnn_bwd_prep_bytes(num_graphs, max_nodes))
    if n == 0:
        return None
    prep = torch.empty(n, dtype=torch.uint8, device=rowptr.device)
    with _Timed("gnn_bwd_prep"):
        check(lib.gsc_gnn_bwd_prep(_ptr(rowptr), _ptr(col), _ptr(graph_ptr), num_graphs, max_nodes, hidden, num_layers,
                                   _ptr(prep), n, _stream()), "gsc_gnn_bwd_prep")
    _count(1)
    return prep


# ---------------------------------------------------------------------------------------------
def gemm_raw(segs: Sequence[tuple], M: int, N: int, out: torch.Tensor, flags: int = 0,
             bias=None, mul=None, gate=None, tag: str = "gemm", row_table=None, smask=None, dot=None):
    """segs: (A_tensor, a_rs, a_cs, B_tensor, b_rs, b_cs, K) with A/B giving the base pointers.
    ``row_table``: (table [n_max+1, N], row_n int32 [M]) adds table[row_n[i]] to row i;
    ``smask``: (row_n, n_max, groups) zeroes the columns outside the structured mask of row i."""
    lib = _lib.load()
    args = GemmArgs()
    ktot = 0
    for s, (A, ars, acs, B, brs, bcs, K) in enumerate(segs):
        _need_cuda(A, B)
        args.A[s], args.B[s], args.K[s] = A.data_ptr(), B.data_ptr(), K
        args.a_rs[s], args.a_cs[s], args.b_rs[s], args.b_cs[s] = ars, acs, brs, bcs
        ktot += K
    if precision == "tf32":
        flags |= EPI_TF32
    elif precision == "tf32x3":
        flags |= EPI_TF32X3
    elif precision != "fp32":
        raise _lib.GscError(f"ops.precision = {precision!r}: expected one of {PRECISIONS}")
    args.num_seg, args.flags = len(segs), flags
    args.bias = _ptr(bias)
    if mul is not None:
        args.mul, args.ld_mul = mul.data_ptr(), _ld(mul)
    if gate is not None:
        args.gate, args.ld_gate = gate.data_ptr(), _ld(gate)
    if row_table is not None:
        table, row_n = row_table
        args.row_table, args.ld_table, args.row_n = table.data_ptr(), _ld(table), row_n.data_ptr()
        args.flags |= _lib.EPI_ROWTABLE
    if smask is not None:
        row_n, n_max, groups = smask
        args.row_n, args.n_max, args.groups = row_n.data_ptr(), n_max, groups
        args.flags |= _lib.EPI_SMASK
    if dot is not None:
        args.dot_w, args.dot_bias, args.dot_out = dot[0].data_ptr(), _ptr(dot[1]), dot[2].data_ptr()
    need = lib.gsc_gemm_workspace_bytes(M, N, ktot)
    ws, wsn = workspace(need, out.device, "gemm")
    with _Timed(f"{tag}[{M}x{N}x{ktot}{'+m' if mul is not None else ''}{'+g' if gate is not None else ''}]"
                + ("T" if row_table is not None else "") + ("S" if smask is not None else "")):
        check(lib.gsc_gemm(args, M, N, out.data_ptr(), _ld(out), _ptr(ws), wsn, _stream()), "gsc_gemm")
    _count(2 if (need or dot is not None) else 1)
    return out


def linear_fwd(inputs: List[torch.Tensor], weight: torch.Tensor, bias: Optional[torch.Tensor],
               relu: bool = False, mul: Optional[torch.Tensor] = None, clamp1: bool = False,
               out: Optional[torch.Tensor] = None, col_offsets: Optional[Sequence[int]] = None,
               row_table=None, smask=None, head=None):
    """act(cat(inputs, 1) @ weight.T + bias) [* mul] without materialising the concatenation.

    ``head=(w [1, N], b [1] or None)``: a one-unit layer on top of the result, folded into the same
    call (into the split-K reduction when there is one); returns ``(out, q [M, 1])`` then.

    weight: [N, sum K_s] fp32, may be a strided row view (its row stride is used).
    ``col_offsets`` places input s at weight columns [col_offsets[s], +K_s) instead of back to back
    (used when the mask column block is replaced by ``row_table``)."""
    weight = _row_major(weight, "linear_fwd weight")
    M, N = int(inputs[0].shape[0]), int(weight.shape[0])
    ldw = _ld(weight)
    segs, off = [], 0
    for si, x in enumerate(inputs):
        x = _row_major(x, "linear_fwd input")
        K = int(x.shape[1])
        if col_offsets is not None:
            off = int(col_offsets[si])
        segs.append((x, _ld(x), 1, weight[:, off:], 1, ldw, K))
        off += K
    if col_offsets is None and off != weight.shape[1]:
        raise _lib.GscError(f"linear_fwd: inputs have {off} columns, weight expects {weight.shape[1]}")
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=weight.device)
    flags = (EPI_BIAS if bias is not None else 0) | (EPI_RELU if relu else 0) | \
            (EPI_MUL if mul is not None else 0) | (EPI_CLAMP1 if clamp1 else 0)
    if head is None:
        return gemm_raw(segs, M, N, out, flags, bias=bias, mul=mul, tag="linear_fwd", row_table=row_table, smask=smask)
    hw, hb = head
    if tuple(hw.shape) != (1, N) or hw.stride(1) != 1 or N > 1024:
        raise _lib.GscError(f"linear_fwd: head weight must be [1, {N}] with unit column stride, N <= 1024")
    q = torch.empty(M, 1, dtype=torch.float32, device=weight.device)
    gemm_raw(segs, M, N, out, flags, bias=bias, mul=mul, tag="linear_fwd+head", row_table=row_table, smask=smask,
             dot=(hw, hb, q))
    return out, q


def linear_dgrad(dY: torch.Tensor, weight: torch.Tensor, col_off: int, K: int,
                 mul: Optional[torch.Tensor] = None, gate: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None, smask=None) -> torch.Tensor:
    """dX[M, K] = dY[M, N] @ weight[:, col_off:col_off+K]  (* mul) (gated by gate > 0)."""
    dY = _row_major(dY, "linear_dgrad dY")
    weight = _row_major(weight, "linear_dgrad weight")
    M, N = int(dY.shape[0]), int(dY.shape[1])
    if out is None:
        out = torch.empty(M, K, dtype=torch.float32, device=dY.device)
    flags = (EPI_MUL if mul is not None else 0) | (EPI_GATE if gate is not None else 0)
    seg = (dY, _ld(dY), 1, weight[:, col_off:], _ld(weight), 1, N)
    return gemm_raw([seg], M, K, out, flags, mul=mul, gate=gate, tag="linear_dgrad", smask=smask)


def linear_wgrad(dY: torch.Tensor, inputs: List[torch.Tensor], dW: torch.Tensor) -> torch.Tensor:
    """dW[N, sum K_s] = dY^T @ cat(inputs, 1), one GEMM per input segment, written in place."""
    dY = _row_major(dY, "linear_wgrad dY")
    dW = _row_major(dW, "linear_wgrad dW")
    M, N = int(dY.shape[0]), int(dY.shape[1])
    off = 0
    for x in inputs:
        x = _row_major(x, "linear_wgrad input")
        K = int(x.shape[1])
        seg = (dY, 1, _ld(dY), x, _ld(x), 1, M)
        gemm_raw([seg], N, K, dW[:, off:off + K], 0, tag="linear_wgrad")
        off += K
    return dW


def linear_n1_fwd(h: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """h @ weight.T + bias for a layer with ONE output unit (weight [1, K]): one launch, one warp per row."""
    h = _row_major(h, "linear_n1_fwd h")
    M, K = int(h.shape[0]), int(h.shape[1])
    if tuple(weight.shape) != (1, K) or weight.stride(1) != 1:
        raise _lib.GscError(f"linear_n1_fwd: weight must be [1, {K}] with unit column stride")
    out = torch.empty(M, 1, dtype=torch.float32, device=h.device)
    with _Timed(f"linear_n1_fwd[{M}x1x{K}]"):
        check(_lib.load().gsc_linear_n1_fwd(_ptr(h), _ld(h), M, K, _ptr(weight), _ptr(bias), _ptr(out), 1, _stream()),
              "gsc_linear_n1_fwd")
    _count(1)
    return out


def linear_n1_bwd(dq: torch.Tensor, h: torch.Tensor, weight: torch.Tensor, relu_input: bool,
                  want_dh: bool = True, dW: Optional[torch.Tensor] = None, db: Optional[torch.Tensor] = None):
    """Backward of ``linear_n1_fwd`` in one launch: returns d_h [M, K] (gated by h > 0 if ``relu_input``)
    or None, and writes dW [1, K] / db [1] in place when given."""
    h = _row_major(h, "linear_n1_bwd h")
    M, K = int(h.shape[0]), int(h.shape[1])
    if dq.numel() != M:
        raise _lib.GscError("linear_n1_bwd: dq must hold one value per row")
    lib = _lib.load()
    d_h = torch.empty(M, K, dtype=torch.float32, device=h.device) if want_dh else None
    ws, wsn = (None, 0)
    if dW is not None or db is not None:
        ws, wsn = workspace(lib.gsc_linear_n1_bwd_workspace_bytes(K), h.device, "linear_n1", zeroed=True)
    ld_dq = int(dq.stride(0)) if dq.dim() >= 1 and dq.shape[0] > 1 else 1
    with _Timed(f"linear_n1_bwd[{M}x{K}x1]"):
        check(lib.gsc_linear_n1_bwd(_ptr(dq), ld_dq, _ptr(h), _ld(h), M, K, _ptr(weight), 1 if relu_input else 0,
                                    _ptr(d_h), K, _ptr(dW), _ptr(db), _ptr(ws), wsn, _stream()), "gsc_linear_n1_bwd")
    _count(1)
    return d_h


def mul_gate(x: torch.Tensor, mul: Optional[torch.Tensor] = None, gate: Optional[torch.Tensor] = None):
    """x * mul, zeroed where gate <= 0 (standalone form of the GEMM epilogue)."""
    x = _row_major(x, "mul_gate x")
    out = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    check(_lib.load().gsc_mul_gate(_ptr(x), _ld(x), _ptr(mul), _ld(mul) if mul is not None else 0,
                                   _ptr(gate), _ld(gate) if gate is not None else 0, int(x.shape[0]),
                                   int(x.shape[1]), _ptr(out), _ld(out), _stream()), "gsc_mul_gate")
    _count(1)
    return out


def mask_table(w_mask: torch.Tensor, n_max: int, groups: int) -> torch.Tensor:
    """table[n, o] = sum of ``w_mask[o, k]`` over the columns the structured mask of an n-node state
    selects; ``w_mask`` is the mask column block [N1, n_max*groups*n_max] of a first-layer weight."""
    w_mask = _row_major(w_mask, "mask_table weight")
    N1 = int(w_mask.shape[0])
    table = torch.empty(n_max + 1, N1, dtype=torch.float32, device=w_mask.device)
    with _Timed(f"mask_table[{N1}x{int(w_mask.shape[1])}]"):
        check(_lib.load().gsc_mask_table(_ptr(w_mask), _ld(w_mask), N1, n_max, groups, _ptr(table), N1, _stream()),
              "gsc_mask_table")
    _count(1)
    return table


def mask_wgrad(dY: torch.Tensor, n_nodes: torch.Tensor, n_max: int, groups: int, dW: torch.Tensor) -> torch.Tensor:
    """dW[o, k] = sum over {b: k in mask(n_b)} dY[b, o], written into the mask column block ``dW``."""
    dY = _row_major(dY, "mask_wgrad dY")
    dW = _row_major(dW, "mask_wgrad dW")
    lib = _lib.load()
    N1 = int(dY.shape[1])
    ws, wsn = workspace(lib.gsc_mask_wgrad_workspace_bytes(N1, n_max), dY.device, "mask_wgrad")
    with _Timed(f"mask_wgrad[{N1}x{int(dW.shape[1])}]"):
        check(lib.gsc_mask_wgrad(_ptr(dY), _ld(dY), _ptr(n_nodes), int(dY.shape[0]), N1, n_max, groups, _ptr(dW),
                                 _ld(dW), _ptr(ws), wsn, _stream()), "gsc_mask_wgrad")
    _count(2)
    return dW


def colsum(X: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    X = _row_major(X, "colsum")
    if out is None:
        out = torch.empty(X.shape[1], dtype=torch.float32, device=X.device)
    with _Timed(f"colsum[{int(X.shape[0])}x{int(X.shape[1])}]"):
        check(_lib.load().gsc_colsum(_ptr(X), _ld(X), int(X.shape[0]), int(X.shape[1]), _ptr(out), _stream()),
              "gsc_colsum")
    _count(1)
    return out


def td_mse(rewards, dones, q_next, q, gamma: float, loss_scale: float):
    B = int(q.numel())
    y = torch.empty(B, dtype=torch.float32, device=q.device)
    dq = torch.empty(B, dtype=torch.float32, device=q.device)
    loss = torch.empty(1, dtype=torch.float32, device=q.device)
    check(_lib.load().gsc_td_mse(_ptr(rewards.contiguous()), _ptr(dones.contiguous()),
                                 _ptr(q_next.contiguous()), _ptr(q.contiguous()), gamma, loss_scale,
                                 B, _ptr(y), _ptr(dq), _ptr(loss), _stream()), "gsc_td_mse")
    _count(1)
    return y, dq, loss


def adam_soft_update(p, g, m, v, target, lr, beta1, beta2, eps, step: int, tau: float):
    for t in (p, g, m, v, target):
        if t is not None and not t.is_contiguous():
            raise _lib.GscError("adam_soft_update needs flat contiguous buffers")
    with _Timed(f"adam_soft_update[{int(p.numel())}]"):
        check(_lib.load().gsc_adam_soft_update(_ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(target),
                                               int(p.numel()), lr, beta1, beta2, eps, step, tau, _stream()),
              "gsc_adam_soft_update")
    _count(1)


def adam_soft_update_dev(p, g, m, v, target, lr, beta1, beta2, eps, step_dev: torch.Tensor, tau: float):
    """Adam + soft update with the step counter in device memory (int64 scalar tensor, incremented
    by the call) -- the form that can be replayed from a CUDA graph."""
    for t in (p, g, m, v, target):
        if t is not None and not t.is_contiguous():
            raise _lib.GscError("adam_soft_update_dev needs flat contiguous buffers")
    with _Timed(f"adam_soft_update[{int(p.numel())}]"):
        check(_lib.load().gsc_adam_soft_update_dev(_ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(target),
                                                   int(p.numel()), lr, beta1, beta2, eps, _ptr(step_dev), tau,
                                                   _stream()), "gsc_adam_soft_update_dev")
    _count(2)


def adam_soft_update_slice(p, g, m, v, target, lr, beta1, beta2, eps, step_dev: torch.Tensor, tau: float,
                           bump_step: bool):
    """``adam_soft_update_dev`` over a slice of the flat buffers (pass the sliced views); the device
    step counter is incremented only when ``bump_step`` (the last slice of the optimizer step)."""
    for t in (p, g, m, v, target):
        if t is not None and not t.is_contiguous():
            raise _lib.GscError("adam_soft_update_slice needs flat contiguous buffers")
    with _Timed(f"adam_soft_update[{int(p.numel())}]"):
        check(_lib.load().gsc_adam_soft_update_slice(_ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(target),
                                                     int(p.numel()), lr, beta1, beta2, eps, _ptr(step_dev), tau,
                                                     1 if bump_step else 0, _stream()), "gsc_adam_soft_update_slice")
    _count((1 if p.numel() else 0) + (1 if bump_step else 0))


def soft_update(p: torch.Tensor, target: torch.Tensor, tau: float):
    """target <- tau * p + (1 - tau) * target over flat contiguous buffers."""
    if not (p.is_contiguous() and target.is_contiguous()) or p.numel() != target.numel():
        raise _lib.GscError("soft_update needs flat contiguous buffers of equal length")
    if p.numel() == 0:
        return
    with _Timed(f"soft_update[{int(p.numel())}]"):
        check(_lib.load().gsc_soft_update(_ptr(p), _ptr(target), int(p.numel()), tau, _stream()), "gsc_soft_update")
    _count(1)


def actor_head(y: torch.Tensor, group: int, noise: Optional[torch.Tensor] = None,
               low: float = 0.0, high: float = 1.0) -> torch.Tensor:
    """post_process_actions (and, with ``noise``, the exploration step) -> fp64 [R, A]."""
    _need_cuda(y, noise)
    y = y.contiguous().to(torch.float32)
    if y.dim() == 1:
        y = y.unsqueeze(0)
    R, A = int(y.shape[0]), int(y.shape[1])
    if noise is not None:
        noise = noise.contiguous().to(torch.float64)
    out = torch.empty(R, A, dtype=torch.float64, device=y.device)
    check(_lib.load().gsc_actor_head(_ptr(y), _ptr(noise), R, A, group, low, high, _ptr(out), _stream()),
          "gsc_actor_head")
    _count(1)
    return out
