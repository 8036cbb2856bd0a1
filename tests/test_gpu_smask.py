"""GPU tests of the structured-mask shortcuts: the mask of an n-node state is mask[:n, :, :, :n] = 1
over (n_max, groups, n_max) (reference src/rlsp/envs/simulator_wrapper.py:298-300), so
  * its column block of the first dense layer is a per-n table lookup (gsc_mask_table + ROWTABLE),
  * the output mask / action-gradient mask is computed from n in the GEMM epilogue (SMASK),
  * the weight gradient of the mask block is a binned column sum (gsc_mask_wgrad).
Every shortcut is compared with the dense-mask path and with fp64 numpy."""
import numpy as np
import pytest
import torch

from gsc_b200 import ddpg, ops, synth
from gsc_b200.graph import GraphBatch
from gsc_b200.models import Actor, AgentSpec, QNetwork

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-4


def rel_err(got, ref):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def dense_masks(n, n_max, groups):
    return synth.structured_mask(np.asarray(n), n_max, groups, 1).astype(np.float64)


@pytest.mark.parametrize("n_max,groups,N1", [(5, 3, 7), (24, 3, 256), (100, 3, 64), (33, 2, 19)])
def test_mask_table_matches_dense_product(n_max, groups, N1):
    A = n_max * groups * n_max
    gen = torch.Generator().manual_seed(n_max + N1)
    W = torch.randn(N1, A + 12, generator=gen)                       # strided view: block inside a wider weight
    tab = ops.mask_table(W.to(DEV)[:, 4:4 + A], n_max, groups).cpu().double().numpy()
    masks = dense_masks(np.arange(n_max + 1), n_max, groups)         # row n = mask of an n-node state
    ref = masks @ W[:, 4:4 + A].double().numpy().T
    assert tab.shape == (n_max + 1, N1)
    assert np.all(tab[0] == 0.0)
    assert rel_err(tab, ref) < 1e-5


@pytest.mark.parametrize("n_max,groups,N1,B", [(5, 3, 7, 9), (24, 3, 256, 300), (100, 3, 64, 128), (33, 2, 19, 1)])
def test_mask_wgrad_matches_dense_product(n_max, groups, N1, B):
    A = n_max * groups * n_max
    rng = np.random.default_rng(B + N1)
    n = rng.integers(0, n_max + 1, size=B).astype(np.int32)
    n[0] = n_max
    dY = rng.standard_normal((B, N1)).astype(np.float32)
    dW = torch.full((N1, A + 8), 7.0, device=DEV)
    ops.mask_wgrad(torch.from_numpy(dY).to(DEV), torch.from_numpy(n).to(DEV), n_max, groups, dW[:, 4:4 + A])
    ref = dY.astype(np.float64).T @ dense_masks(n, n_max, groups)
    got = dW.cpu().double().numpy()
    assert np.all(got[:, :4] == 7.0) and np.all(got[:, 4 + A:] == 7.0), "columns outside the block were touched"
    assert rel_err(got[:, 4:4 + A], ref) < 1e-5


@pytest.mark.parametrize("prec", ["fp32", "tf32x3", "tf32"])
@pytest.mark.parametrize("M,N1,n_max", [(37, 64, 6), (300, 256, 24), (1024, 256, 40)])
def test_row_table_and_smask_epilogues_equal_dense_mask_gemm(prec, M, N1, n_max):
    groups, H = 3, 22
    A = n_max * groups * n_max
    rng = np.random.default_rng(M)
    n = rng.integers(1, n_max + 1, size=M).astype(np.int32)
    n_dev = torch.from_numpy(n).to(DEV)
    mask = torch.from_numpy(dense_masks(n, n_max, groups)).float().to(DEV)
    emb = torch.randn(M, H + 2, device=DEV)
    act = torch.randn(M, A, device=DEV)
    W = torch.randn(N1, H + 2 + 2 * A, device=DEV) / 30
    b = torch.randn(N1, device=DEV)
    tol = 3e-3 if prec == "tf32" else 1e-5
    with ops.use_precision(prec):
        dense = ops.linear_fwd([emb, mask, act], W, b, relu=True)
        table = ops.mask_table(W[:, H + 2:H + 2 + A], n_max, groups)
        fast = ops.linear_fwd([emb, act], W, b, relu=True, col_offsets=[0, H + 2 + A], row_table=(table, n_dev))
        assert rel_err(fast.cpu().numpy(), dense.cpu().numpy()) < tol
        # output mask from n: y = (h @ W2^T + b2) * mask, clamped
        W2 = torch.randn(A, N1, device=DEV) / 16
        b2 = torch.randn(A, device=DEV)
        d_out = ops.linear_fwd([dense], W2, b2, mul=mask, clamp1=True)
        s_out = ops.linear_fwd([dense], W2, b2, smask=(n_dev, n_max, groups), clamp1=True)
        assert torch.equal(d_out, s_out)
        # masked input gradient: dA = (dY @ W[:, block]) * mask
        dY = torch.randn(M, N1, device=DEV)
        d_dg = ops.linear_dgrad(dY, W, H + 2 + A, A, mul=mask)
        s_dg = ops.linear_dgrad(dY, W, H + 2 + A, A, smask=(n_dev, n_max, groups))
        assert torch.equal(d_dg, s_dg)


def _make(n_max, B, seed, min_nodes):
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(B, n_max, min_nodes, n_max, seed=seed, num_topologies=4)
    t = lambda a: torch.from_numpy(a).to(DEV)
    return cur_s, nxt_s, t(actions), t(rewards), t(dones)


def _nets(spec, seed=3):
    torch.manual_seed(seed)
    nets = [cls(spec).to(DEV) for cls in (Actor, QNetwork, Actor, QNetwork)]
    nets[2].load_state_dict(nets[0].state_dict())
    nets[3].load_state_dict(nets[1].state_dict())
    return nets


@pytest.mark.parametrize("prec", ["fp32", "tf32x3", "tf32"])
def test_structured_ddpg_update_equals_dense_mask_update(prec):
    n_max, B = 14, 48
    cur_s, nxt_s, actions, rewards, dones = _make(n_max, B, 21, 5)
    spec = AgentSpec(n_max)
    tol = 5e-3 if prec == "tf32" else 2e-5
    res = {}
    with ops.use_precision(prec):
        for kind in ("dense", "structured"):
            nets = _nets(spec)
            upd = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2)
            cur = GraphBatch.from_states(cur_s, dense_mask=kind == "dense").to(DEV)
            nxt = GraphBatch.from_states(nxt_s, dense_mask=kind == "dense").to(DEV)
            if kind == "dense":
                cur.n_nodes = nxt.n_nodes = None
                assert not cur.structured
            else:
                assert cur.structured and cur._mask is None
            out = upd.update(cur, nxt, actions, rewards, dones)
            if kind == "structured":
                assert cur._mask is None and nxt._mask is None, "the structured path must not materialise the mask"
            res[kind] = (out, upd.g_actor.clone(), upd.g_critic.clone(), nets[0].flat.clone(), nets[3].flat.clone())
    d, s = res["dense"], res["structured"]
    assert rel_err(s[0][0].cpu().numpy(), d[0][0].cpu().numpy()) < tol
    assert rel_err([float(s[0][1]), float(s[0][2])], [float(d[0][1]), float(d[0][2])]) < tol
    for i in (1, 2):
        assert rel_err(s[i].cpu().numpy(), d[i].cpu().numpy()) < tol, "gradients"
    for i in (3, 4):
        assert rel_err(s[i].cpu().numpy(), d[i].cpu().numpy()) < tol, "updated weights"


def test_structured_predict_equals_dense_predict_bitwise_support():
    n_max, B = 24, 64
    st = synth.make_states(B, n_max, 16, 24, seed=8)
    spec = AgentSpec(n_max)
    actor = _nets(spec)[0]
    gs = GraphBatch.from_states(st, dense_mask=False).to(DEV)
    gd = GraphBatch.from_states(st).to(DEV)
    gd.n_nodes = None
    ys, yd = ddpg.predict(actor, gs).cpu().numpy(), ddpg.predict(actor, gd).cpu().numpy()
    assert gs._mask is None
    assert rel_err(ys, yd) < 2e-5
    assert np.array_equal(ys == 0, yd == 0) or np.array_equal((ys != 0) | (st.mask == 0), (yd != 0) | (st.mask == 0))
    assert np.all(ys[st.mask == 0] == 0.0)
    # the lazily generated dense mask is the synthetic generator's mask
    assert np.array_equal(gs.mask.cpu().numpy(), st.mask)


def test_structured_autograd_modules_match_dense_modules():
    n_max, B = 10, 20
    cur_s, _, actions, _, _ = _make(n_max, B, 5, 4)
    spec = AgentSpec(n_max)
    grads = {}
    for kind in ("dense", "structured"):
        actor, qf1 = _nets(spec)[:2]
        gb = GraphBatch.from_states(cur_s, dense_mask=kind == "dense").to(DEV)
        if kind == "dense":
            gb.n_nodes = None
        loss = -qf1(gb, actor(gb)).mean() + qf1(gb, actions).pow(2).mean()
        loss.backward()
        grads[kind] = [p.grad.clone() for p in list(actor.parameters()) + list(qf1.parameters())]
    scale = max(float(g.abs().max()) for g in grads["dense"])
    for gd, gs in zip(grads["dense"], grads["structured"]):
        assert float((gd - gs).abs().max()) <= 2e-5 * max(float(gd.abs().max()), 1e-3 * scale)


def test_device_replay_buffer_returns_structured_batches():
    from gsc_b200.buffer import DeviceGraphReplayBuffer
    n_max = 9
    spec = AgentSpec(n_max)
    rb = DeviceGraphReplayBuffer(64, spec.env.action_space, device=DEV, n_max=n_max)
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(40, n_max, 3, 9, seed=2, num_topologies=5)
    rb.bulk_load(cur_s, nxt_s, actions, rewards, dones)
    np.random.seed(0)
    s = rb.sample(16)
    gb = s.observations
    assert gb.structured and gb._mask is None and gb.n_nodes.dtype == torch.int32
    np.random.seed(0)
    idx = np.random.randint(0, 40, size=16)
    assert np.array_equal(gb.n_nodes.cpu().numpy(), cur_s.n[idx])
    assert np.array_equal(gb.mask.cpu().numpy(), cur_s.mask[idx])
