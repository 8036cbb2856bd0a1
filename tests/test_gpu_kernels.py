"""GPU parity tests of the individual kernels, called through the C-ABI (gsc_b200.ops -> ctypes)."""
import numpy as np
import pytest
import torch

from gsc_b200 import ops, synth
from gsc_b200.graph import GraphBatch
from oracle import gsc_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel_err(got, ref):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


# ------------------------------------------------------------------ packer
def _csr_reference(ei, n):
    rows = [[] for _ in range(n)]
    for s, d in zip(ei[0].tolist(), ei[1].tolist()):
        if s != d:
            rows[d].append(s)
    rowptr, col = [0], []
    for i in range(n):
        col += sorted(rows[i]) + [i]
        rowptr.append(len(col))
    return np.array(rowptr), np.array(col)


@pytest.mark.parametrize("B,n_max,lo,hi", [(1, 8, 1, 1), (7, 8, 2, 8), (300, 24, 16, 24), (40, 110, 60, 110)])
def test_pack_csr_matches_numpy(B, n_max, lo, hi):
    st = synth.make_states(B, n_max, lo, hi, seed=B)
    ei = st.edge_index
    if ei.shape[1] > 4:   # sprinkle self loops and duplicate edges (multigraph topologies exist: interroute)
        extra = np.stack([ei[0, :3], ei[0, :3]])
        ei = np.concatenate([ei, extra, ei[:, 1:3]], axis=1)
    rowptr, col, gptr, stats = ops.pack_csr(torch.from_numpy(ei).to(DEV), torch.from_numpy(st.batch).to(DEV),
                                            int(st.ptr[-1]), B)
    want_rp, want_col = _csr_reference(ei, int(st.ptr[-1]))
    st_h = stats.cpu().numpy()
    assert st_h[2] == 0 and st_h[0] == len(want_col) and st_h[1] == int(st.n.max())
    np.testing.assert_array_equal(rowptr.cpu().numpy(), want_rp)
    np.testing.assert_array_equal(col.cpu().numpy()[:len(want_col)], want_col)
    np.testing.assert_array_equal(gptr.cpu().numpy(), st.ptr)


def test_pack_csr_single_graph_no_batch_and_empty_graphs():
    ei = torch.tensor([[0, 1, 1, 2], [1, 0, 2, 1]], dtype=torch.int64, device=DEV)
    rowptr, col, gptr, stats = ops.pack_csr(ei, None, 3, 1)
    assert rowptr.tolist() == [0, 2, 5, 7] and col.tolist()[:7] == [1, 0, 0, 2, 1, 1, 2]
    assert gptr.tolist() == [0, 3] and stats.tolist()[:3] == [7, 3, 0]
    # graphs 1 and 3 hold no nodes
    batch = torch.tensor([0, 0, 2, 4, 4], dtype=torch.int64, device=DEV)
    ei = torch.tensor([[0, 1, 3, 4], [1, 0, 4, 3]], dtype=torch.int64, device=DEV)
    rowptr, col, gptr, stats = ops.pack_csr(ei, batch, 5, 5)
    assert gptr.tolist() == [0, 2, 2, 3, 3, 5] and stats.tolist()[2] == 0
    # an edge across two graphs is flagged
    bad = torch.tensor([[0], [3]], dtype=torch.int64, device=DEV)
    assert ops.pack_csr(bad, batch, 5, 5)[3].tolist()[2] != 0


def test_pack_csr_long_rows_duplicates_and_symmetry_flag():
    """A star with 40 leaves (row longer than the register sort takes), duplicate edges, and the
    symmetry flag: set when a reverse edge (or one copy of a duplicated one) is missing."""
    leaves = np.arange(1, 41)
    und = np.concatenate([np.stack([np.zeros(40, dtype=np.int64), leaves]), np.stack([leaves, np.zeros(40, dtype=np.int64)])], axis=1)
    rng = np.random.default_rng(0)
    und = und[:, rng.permutation(und.shape[1])]
    dup = np.array([[3, 0, 3, 0], [0, 3, 0, 3]])                # edge 0-3 three times in both directions
    for ei, sym in ((und, 0), (np.concatenate([und, dup], axis=1), 0), (np.concatenate([und, dup[:, :3]], axis=1), 1),
                    (und[:, 1:], 1)):
        rowptr, col, gptr, stats = ops.pack_csr(torch.from_numpy(np.ascontiguousarray(ei)).to(DEV), None, 41, 1)
        want_rp, want_col = _csr_reference(ei, 41)
        np.testing.assert_array_equal(rowptr.cpu().numpy(), want_rp)
        np.testing.assert_array_equal(col.cpu().numpy()[:len(want_col)], want_col)
        st_h = stats.tolist()
        assert st_h[1] == 41 and st_h[2] == 0 and st_h[3] == sym, (st_h, sym)


def test_pack_csr_large_scan():
    st = synth.make_states(20000, 24, 16, 24, seed=9, num_topologies=64)      # ~400k rows: multi-block scan
    rowptr, col, gptr, stats = ops.pack_csr(torch.from_numpy(st.edge_index).to(DEV),
                                            torch.from_numpy(st.batch).to(DEV), int(st.ptr[-1]), 20000)
    deg = np.bincount(st.edge_index[1], minlength=int(st.ptr[-1])) + 1
    np.testing.assert_array_equal(rowptr.cpu().numpy(), np.concatenate([[0], np.cumsum(deg)]))
    assert stats.tolist()[0] == int(deg.sum())


# ------------------------------------------------------------------ GNN embedder
def _embed_case(B, n_max, lo, hi, H, F, num_layers, num_iter, seed):
    st = synth.make_states(B, n_max, lo, hi, seed=seed, num_feat=F)
    gen = torch.Generator().manual_seed(seed)
    sd = O.init_embedder(F, H, num_layers, gen, torch.float64, prefix="")
    for k in sd:
        if k.endswith(".bias") and "lin" not in k:
            sd[k] = torch.rand(H, generator=gen, dtype=torch.float64) - 0.5
    return st, sd


def _flat(sd, num_layers):
    names = ["encoder."] + [f"process.{i}." for i in range(num_layers - 1)]
    parts = []
    for nme in names:
        for k in ("att", "bias", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias"):
            parts.append(sd[nme + k].reshape(-1))
    return torch.cat(parts).to(torch.float32).to(DEV)


CASES = [(1, 8, 1, 1, 22, 3, 2, 2), (9, 8, 2, 8, 22, 3, 2, 2), (500, 24, 16, 24, 22, 3, 2, 2),
         (33, 24, 5, 24, 22, 3, 1, 2), (17, 24, 5, 24, 16, 5, 3, 2), (12, 24, 5, 24, 32, 3, 2, 3),
         (21, 24, 5, 24, 7, 4, 2, 1), (6, 300, 100, 300, 22, 3, 2, 2), (3, 1000, 900, 1000, 22, 3, 2, 2)]


@pytest.mark.parametrize("B,n_max,lo,hi,H,F,num_layers,num_iter", CASES)
def test_gnn_embed_fwd_matches_oracle(B, n_max, lo, hi, H, F, num_layers, num_iter):
    st, sd = _embed_case(B, n_max, lo, hi, H, F, num_layers, num_iter, seed=B + H)
    ref = O.gnn_embedder(torch.from_numpy(st.x).double(), torch.from_numpy(st.edge_index),
                         torch.from_numpy(st.batch), sd, num_layers, num_iter, prefix="", num_graphs=B)
    gb = GraphBatch.from_states(st).to(DEV)
    rowptr, col, gptr = gb.csr()
    emb, _ = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, _flat(sd, num_layers), H,
                               num_layers, num_iter)
    assert rel_err(emb.cpu().numpy(), ref.numpy()) < 1e-4     # north_star fp32 tolerance (observed ~1e-6)
    assert rel_err(emb.cpu().numpy(), ref.numpy()) < 2e-5


BWD_CASES = CASES + [(5, 600, 450, 600, 22, 3, 2, 2)]   # > ~450 nodes: node rows in the workspace


@pytest.mark.parametrize("B,n_max,lo,hi,H,F,num_layers,num_iter", BWD_CASES)
def test_gnn_embed_bwd_matches_autograd_of_oracle(B, n_max, lo, hi, H, F, num_layers, num_iter):
    st, sd = _embed_case(B, n_max, lo, hi, H, F, num_layers, num_iter, seed=B + H + 1)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = O.gnn_embedder(torch.from_numpy(st.x).double(), torch.from_numpy(st.edge_index),
                         torch.from_numpy(st.batch), sd, num_layers, num_iter, prefix="", num_graphs=B)
    gen = torch.Generator().manual_seed(1)
    d_emb = torch.randn(B, H, generator=gen, dtype=torch.float64)
    (ref * d_emb).sum().backward()
    gb = GraphBatch.from_states(st).to(DEV)
    rowptr, col, gptr = gb.csr()
    flat = _flat({k: v.detach() for k, v in sd.items()}, num_layers)
    emb, saved = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, num_layers, num_iter, save=True)
    d = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, num_layers, num_iter, saved,
                          d_emb.float().to(DEV), pack_stats=gb.pack_stats())
    assert int(gb.pack_stats()[3]) == 0                       # undirected topology -> gather kernel (if it fits)
    want = _flat({k: v.grad for k, v in sd.items()}, num_layers).cpu().numpy()
    got = d.cpu().numpy()
    assert rel_err(got, want) < 1e-4
    # per-parameter check so that a small tensor cannot hide behind a large one
    off = 0
    names = ["encoder."] + [f"process.{i}." for i in range(num_layers - 1)]
    for nme in names:
        for k in ("att", "bias", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias"):
            n = sd[nme + k].numel()
            # relative to max(own magnitude, 1e-3 of the largest gradient): near-zero gradients
            # (cancellation) hold only fp32 noise and cannot be judged against themselves
            scale = max(np.abs(want[off:off + n]).max(), 1e-3 * np.abs(want).max())
            assert np.abs(got[off:off + n] - want[off:off + n]).max() / scale < 1e-4, nme + k
            off += n
    # the gather kernel has no floating-point atomics on the d_xl path; the small d_att / d_bias /
    # weight-gradient accumulators are shared-memory atomics, so bound the run-to-run jitter
    d2 = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, num_layers, num_iter, saved,
                           d_emb.float().to(DEV), pack_stats=gb.pack_stats())
    assert rel_err(d2.cpu().numpy(), got) < 1e-5
    # without the packer's symmetry flag the scatter (atomic) kernel runs: same gradients
    d3 = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, num_layers, num_iter, saved,
                           d_emb.float().to(DEV))
    assert rel_err(d3.cpu().numpy(), want) < 1e-4


def test_gnn_embed_bwd_prepared_is_bit_identical_and_shared_per_batch():
    """gsc_gnn_bwd_prep + gsc_gnn_embed_bwd_prepared == gsc_gnn_embed_bwd bit for bit (the slot order is a
    stable sort, so the summation orders are fixed), and GraphBatch hands every backward the same buffer."""
    B, H, F = 300, 22, 3
    st, sd = _embed_case(B, 24, 16, 24, H, F, 2, 2, seed=5)
    gb = GraphBatch.from_states(st).to(DEV)
    rowptr, col, gptr = gb.csr()
    flat = _flat(sd, 2)
    emb, saved = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, 2, 2, save=True)
    d_emb = torch.randn(B, H, generator=torch.Generator().manual_seed(3)).to(DEV)
    plain = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, 2, 2, saved, d_emb,
                              pack_stats=gb.pack_stats())
    prep = gb.bwd_prep(H, 2)
    assert prep is not None and gb.bwd_prep(H, 2) is prep and gb.view_shared().bwd_prep(H, 2) is prep
    for _ in range(2):
        got = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, 2, 2, saved, d_emb,
                                pack_stats=gb.pack_stats(), prep=prep)
        assert torch.equal(got, plain)
    big = GraphBatch.from_states(synth.make_states(2, 300, 200, 300, seed=1)).to(DEV)
    assert big.bwd_prep(H, 2) is None                          # graphs beyond 128 nodes have no such part


def test_gnn_embed_bwd_directed_graph_uses_scatter_kernel():
    """Edges without a reverse (stats[3] != 0): the gather kernel must stand down."""
    B, H, F = 7, 22, 3
    st, sd = _embed_case(B, 24, 5, 24, H, F, 2, 2, seed=77)
    keep = np.ones(st.edge_index.shape[1], dtype=bool)
    keep[::5] = False                                         # drop every 5th directed edge
    ei = np.ascontiguousarray(st.edge_index[:, keep])
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = O.gnn_embedder(torch.from_numpy(st.x).double(), torch.from_numpy(ei), torch.from_numpy(st.batch), sd, 2, 2,
                         prefix="", num_graphs=B)
    d_emb = torch.randn(B, H, generator=torch.Generator().manual_seed(2), dtype=torch.float64)
    (ref * d_emb).sum().backward()
    gb = GraphBatch.from_arrays(st.x, ei, st.batch, st.mask, B, int(st.n.max())).to(DEV)
    rowptr, col, gptr = gb.csr()
    assert int(gb.pack_stats()[3]) == 1
    flat = _flat({k: v.detach() for k, v in sd.items()}, 2)
    emb, saved = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, 2, 2, save=True)
    assert rel_err(emb.cpu().numpy(), ref.detach().numpy()) < 2e-5
    d = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, 2, 2, saved, d_emb.float().to(DEV),
                          pack_stats=gb.pack_stats())
    want = _flat({k: v.grad for k, v in sd.items()}, 2).cpu().numpy()
    assert rel_err(d.cpu().numpy(), want) < 1e-4


# ------------------------------------------------------------------ GEMM family
@pytest.fixture(params=["fp32", "tf32x3"])
def dense_precision(request):
    """The rel-2e-5 GEMM tests run on the exact FMA kernel and on the 3xTF32 tensor-core kernel."""
    with ops.use_precision(request.param):
        yield request.param


@pytest.mark.parametrize("M,N,Ks", [(100, 64, (22, 1728, 1728)), (100, 256, (22, 1728)), (37, 5, (3,)),
                                    (4096, 130, (256,)), (100, 1, (64,)), (8, 64, (22, 30000, 30000)),
                                    (1, 256, (22, 1728))])
def test_linear_fwd_segments(M, N, Ks, dense_precision):
    gen = torch.Generator().manual_seed(M + N)
    xs = [torch.randn(M, K, generator=gen, dtype=torch.float64) for K in Ks]
    W = torch.randn(N, sum(Ks), generator=gen, dtype=torch.float64) / np.sqrt(sum(Ks))
    b = torch.randn(N, generator=gen, dtype=torch.float64)
    mul = (torch.rand(M, N, generator=gen) > 0.3).double()
    ref = torch.relu(torch.cat(xs, 1) @ W.t() + b)
    got = ops.linear_fwd([x.float().to(DEV) for x in xs], W.float().to(DEV), b.float().to(DEV), relu=True)
    assert rel_err(got.cpu().numpy(), ref.numpy()) < 2e-5
    ref2 = ((torch.cat(xs, 1) @ W.t() + b) * mul).clamp(-1, 1)
    got2 = ops.linear_fwd([x.float().to(DEV) for x in xs], W.float().to(DEV), b.float().to(DEV), relu=False,
                          mul=mul.float().to(DEV), clamp1=True)
    assert rel_err(got2.cpu().numpy(), ref2.numpy()) < 2e-5


def test_linear_fwd_strided_padded_weight(dense_precision):
    gen = torch.Generator().manual_seed(0)
    M, N, K = 50, 40, 22 + 96
    store = torch.zeros(N, 124, device=DEV)
    W = store[:, 2:2 + K]
    W.copy_(torch.randn(N, K, generator=gen))
    x1, x2 = torch.randn(M, 22, generator=gen).to(DEV), torch.randn(M, 96, generator=gen).to(DEV)
    got = ops.linear_fwd([x1, x2], W, None)
    ref = torch.cat([x1, x2], 1).double() @ W.double().t()
    assert rel_err(got.cpu().numpy(), ref.cpu().numpy()) < 2e-5


@pytest.mark.parametrize("M,N,K", [(100, 64, 1728), (100, 1728, 256), (4096, 256, 22), (64, 30000, 256), (3, 1, 64)])
def test_linear_dgrad_and_wgrad(M, N, K, dense_precision):
    gen = torch.Generator().manual_seed(M + K)
    dY = torch.randn(M, N, generator=gen, dtype=torch.float64)
    W = torch.randn(N, K + 10, generator=gen, dtype=torch.float64)
    X = torch.randn(M, K, generator=gen, dtype=torch.float64)
    gate = torch.randn(M, K, generator=gen, dtype=torch.float64)
    mul = torch.rand(M, K, generator=gen, dtype=torch.float64)
    ref = dY @ W[:, 7:7 + K]
    got = ops.linear_dgrad(dY.float().to(DEV), W.float().to(DEV), 7, K)
    assert rel_err(got.cpu().numpy(), ref.numpy()) < 2e-5
    got = ops.linear_dgrad(dY.float().to(DEV), W.float().to(DEV), 7, K, mul=mul.float().to(DEV), gate=gate.float().to(DEV))
    assert rel_err(got.cpu().numpy(), (ref * mul * (gate > 0)).numpy()) < 2e-5
    dW = torch.zeros(N, K + 10, device=DEV)
    ops.linear_wgrad(dY.float().to(DEV), [X[:, :K // 2].float().to(DEV).contiguous(), X[:, K // 2:].float().to(DEV).contiguous()],
                     dW[:, 3:3 + K])
    assert rel_err(dW[:, 3:3 + K].cpu().numpy(), (dY.t() @ X).numpy()) < 2e-5
    assert float(dW[:, :3].abs().max()) == 0 and float(dW[:, 3 + K:].abs().max()) == 0
    assert rel_err(ops.colsum(dY.float().to(DEV)).cpu().numpy(), dY.sum(0).numpy()) < 2e-5


def test_gemm_is_deterministic_under_split_k(dense_precision):
    gen = torch.Generator().manual_seed(5)
    x = torch.randn(100, 60000, generator=gen).to(DEV)
    W = torch.randn(64, 60000, generator=gen).to(DEV)
    a = ops.linear_fwd([x], W, None).clone()
    b = ops.linear_fwd([x], W, None)
    assert torch.equal(a, b)


def test_mul_gate():
    x = torch.randn(33, 70, device=DEV)
    m = torch.rand(33, 70, device=DEV)
    g = torch.randn(33, 70, device=DEV)
    assert torch.equal(ops.mul_gate(x, mul=m, gate=g), torch.where(g > 0, x * m, torch.zeros_like(x)))


# ------------------------------------------------------------------ optimizer / TD
def test_adam_soft_update_matches_torch_adam():
    gen = torch.Generator().manual_seed(3)
    n = 100003
    p0 = torch.randn(n, generator=gen, dtype=torch.float64)
    t0 = torch.randn(n, generator=gen, dtype=torch.float64)
    p_ref = p0.clone().requires_grad_(True)
    t_ref = t0.clone()
    opt = torch.optim.Adam([p_ref], lr=1e-3, foreach=False)
    p = p0.float().to(DEV)
    tgt = t0.float().to(DEV)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    tau = 1e-2
    for step in range(1, 4):
        g = torch.randn(n, generator=gen, dtype=torch.float64) * (10.0 ** (step - 2))
        p_ref.grad = g.clone()
        opt.step()
        t_ref = tau * p_ref.detach() + (1 - tau) * t_ref
        ops.adam_soft_update(p, g.float().to(DEV), m, v, tgt, 1e-3, 0.9, 0.999, 1e-8, step, tau)
    assert rel_err(p.cpu().numpy(), p_ref.detach().numpy()) < 1e-6
    assert rel_err(tgt.cpu().numpy(), t_ref.numpy()) < 1e-6
    # update itself (not hidden by |p|): compare the applied deltas
    assert rel_err((p.cpu().double() - p0).numpy(), (p_ref.detach() - p0).numpy()) < 1e-3
    # unaligned views take the scalar kernel and give the same numbers
    p2 = torch.zeros(n + 1, device=DEV)[1:]
    p2.copy_(p0.float())
    m2, v2 = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    ops.adam_soft_update(p2, torch.ones(n, device=DEV), m2, v2, None, 1e-3, 0.9, 0.999, 1e-8, 1, 0.0)
    np.testing.assert_allclose(p2.cpu().numpy(), p0.float().numpy() - 1e-3, rtol=0, atol=2e-6)


def test_linear_n1_head_matches_fp64_and_gemm_route():
    """The fused one-output-unit layer (critic head) against an fp64 reference and against the GEMM
    route it replaces; repeated calls reuse the self-resetting counter."""
    gen = torch.Generator().manual_seed(11)
    for M, K in ((1024, 64), (37, 64), (1, 5), (300, 1000)):
        h = torch.relu(torch.randn(M, K, generator=gen)).to(DEV)
        W = (torch.randn(1, K, generator=gen) / 8).to(DEV)
        b = torch.randn(1, generator=gen).to(DEV)
        dq = torch.randn(M, 1, generator=gen).to(DEV)
        q = ops.linear_n1_fwd(h, W, b)
        q_ref = h.double() @ W.double().t() + b.double()
        assert rel_err(q.cpu().numpy(), q_ref.cpu().numpy()) < 2e-6
        q_gemm = ops.linear_fwd([h], W, b)
        assert rel_err(q.cpu().numpy(), q_gemm.cpu().numpy()) < 2e-6
        for rep in range(3):
            dW, db = torch.full((1, K), 7.0, device=DEV), torch.full((1,), 7.0, device=DEV)
            d_h = ops.linear_n1_bwd(dq, h, W, relu_input=True, dW=dW, db=db)
            dh_ref = torch.where(h > 0, dq.double() @ W.double(), torch.zeros((), dtype=torch.float64, device=DEV))
            assert rel_err(d_h.cpu().numpy(), dh_ref.cpu().numpy()) < 1e-6
            assert rel_err(dW.cpu().numpy(), (dq.double().t() @ h.double()).cpu().numpy()) < 2e-6
            assert rel_err(db.cpu().numpy(), dq.double().sum().reshape(1).cpu().numpy()) < 2e-6
        assert torch.equal(d_h, ops.linear_dgrad(dq, W, 0, K, gate=h))       # same products, same gate
        only = ops.linear_n1_bwd(dq, h, W, relu_input=False, want_dh=True)
        assert rel_err(only.cpu().numpy(), (dq.double() @ W.double()).cpu().numpy()) < 1e-6


@pytest.mark.parametrize("M,N,K", [(1024, 64, 30024), (100, 64, 3480), (37, 64, 24), (300, 256, 5000)])
def test_linear_fwd_with_fused_head_equals_two_calls(M, N, K, dense_precision):
    """The one-unit head folded into the layer below (split-K reduction or not) == layer, then head."""
    gen = torch.Generator().manual_seed(M + K)
    x = torch.randn(M, K, generator=gen).to(DEV)
    W = (torch.randn(N, K, generator=gen) / K ** 0.5).to(DEV)
    b = torch.randn(N, generator=gen).to(DEV)
    hw, hb = (torch.randn(1, N, generator=gen) / 8).to(DEV), torch.randn(1, generator=gen).to(DEV)
    h_ref = ops.linear_fwd([x], W, b, relu=True)
    q_ref = ops.linear_n1_fwd(h_ref, hw, hb)
    h, q = ops.linear_fwd([x], W, b, relu=True, head=(hw, hb))
    assert torch.equal(h, h_ref) and torch.equal(q, q_ref)


def test_adam_soft_update_slices_equal_one_launch():
    """Two slices of one optimizer step (the second bumps the device step counter) == one launch over
    the whole buffer, bit for bit, over several steps."""
    gen = torch.Generator().manual_seed(5)
    n, cut = 70001, 1280
    p0, t0 = torch.randn(n, generator=gen).to(DEV), torch.randn(n, generator=gen).to(DEV)
    state = lambda: [p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV), t0.clone(),
                     torch.zeros(1, dtype=torch.int64, device=DEV)]
    pa, ma, va, ta, sa = state()
    pb, mb, vb, tb, sb = state()
    for step in range(3):
        g = torch.randn(n, generator=gen).to(DEV)
        ops.adam_soft_update_dev(pa, g, ma, va, ta, 1e-3, 0.9, 0.999, 1e-8, sa, 1e-2)
        ops.adam_soft_update_slice(pb[cut:], g[cut:], mb[cut:], vb[cut:], tb[cut:], 1e-3, 0.9, 0.999, 1e-8, sb, 1e-2,
                                   bump_step=False)
        ops.adam_soft_update_slice(pb[:cut], g[:cut], mb[:cut], vb[:cut], tb[:cut], 1e-3, 0.9, 0.999, 1e-8, sb, 1e-2,
                                   bump_step=True)
    assert int(sa.item()) == int(sb.item()) == 3
    for a, b in ((pa, pb), (ma, mb), (va, vb), (ta, tb)):
        assert torch.equal(a, b)


def test_td_mse():
    gen = torch.Generator().manual_seed(4)
    B = 1500
    r, d = torch.randn(B, generator=gen), (torch.rand(B, generator=gen) < 0.1).float()
    qn, q = torch.randn(B, generator=gen), torch.randn(B, generator=gen)
    y, dq, loss = ops.td_mse(r.to(DEV), d.to(DEV), qn.to(DEV), q.to(DEV), 0.99, 1.0 / B)
    y_ref = r.double() + (1 - d.double()) * 0.99 * qn.double()
    assert rel_err(y.cpu().numpy(), y_ref.numpy()) < 1e-6
    assert rel_err(loss.cpu().numpy(), ((q.double() - y_ref) ** 2).mean().numpy()) < 1e-5
    assert rel_err(dq.cpu().numpy(), (2 * (q.double() - y_ref) / B).numpy()) < 1e-5


# ------------------------------------------------------------------ actor head (bit exact)
def test_actor_head_golden_bit_exact(golden_head):
    for nn in (3, 8, 24):
        a = torch.from_numpy(golden_head[f"in_{nn}"]).to(DEV)
        out = ops.actor_head(a, nn).cpu().numpy().reshape(-1)
        assert np.array_equal(out, golden_head[f"out_{nn}"]), nn
        noise = torch.from_numpy(golden_head[f"noise_{nn}"]).to(DEV).view(1, -1)
        ex = ops.actor_head(a, nn, noise=noise).cpu().numpy().reshape(-1)
        assert np.array_equal(ex, golden_head[f"explore_{nn}"]), nn


@pytest.mark.parametrize("nn,R", [(24, 40), (100, 3), (7, 50)])
def test_actor_head_random_bit_exact_vs_oracle(nn, R):
    rng = np.random.default_rng(nn)
    A = nn * 3 * nn
    y = (rng.random((R, A)) * (rng.random((R, A)) < 0.15)).astype(np.float32)
    y[0] = rng.random(A).astype(np.float32) * 0.12          # crowd the threshold
    got = ops.actor_head(torch.from_numpy(y).to(DEV), nn).cpu().numpy()
    for r in range(min(R, 6)):
        want = O.post_process_actions(y[r], nn)
        assert np.array_equal(got[r], want)
    sums = got.reshape(R, -1, nn).sum(-1)
    assert np.all(np.abs(sums - 1) < 1e-7)                  # size-independent property for all rows
