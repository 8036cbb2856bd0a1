"""GPU tests at the sizes of the large configurations:

  c4  N_max = n = 1000, |E| = 1500 (A = 3 000 000; actor 1.54 G parameters) -- the data-parallel
      training configuration: embedder forward/backward on 1000-node graphs, dense layers with
      K = 22 + 2A, structured-mask table/gradient at n_max = 1000, one full DDPG update;
  c5  inference on 65 536 states of up to 110 nodes (the largest bundled topology, interroute).

The CPU oracle cannot hold these networks in seconds, so the dense layers are checked against plain
fp64 torch matmuls on the GPU (same weights), the embedder against the CPU oracle (it is small), and
the rest through size-independent properties: masked-out actions are exactly zero, whole-batch and
chunked evaluation agree, losses are finite and the update moves every parameter tensor."""
import numpy as np
import pytest
import torch

from gsc_b200 import ddpg, ops, synth
from gsc_b200.graph import GraphBatch
from gsc_b200.models import Actor, AgentSpec, QNetwork, net_backward, net_forward
from oracle import gsc_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel_err(got, ref):
    return float((got.double() - ref.double()).abs().max() / ref.double().abs().max().clamp_min(1e-30))


def _need_gib(gib):
    free, _ = torch.cuda.mem_get_info()
    if free < gib * 2 ** 30:
        pytest.skip(f"needs {gib} GiB of free device memory")


def _embedder_sd(net):
    return {k: v.detach().cpu().double() for k, v in net.state_dict().items() if k.startswith("embedder.")}


def test_c4_shape_networks_match_fp64_reference_and_update_runs():
    _need_gib(120)
    n_max, B = 1000, 4
    spec = AgentSpec(n_max)
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(B, n_max, n_max, n_max, seed=4, fixed_links=1500,
                                                                   num_topologies=2)
    A = cur_s.action_dim
    assert A == 3_000_000
    torch.manual_seed(4)
    with torch.device(DEV):
        actor, qf1 = Actor(spec), QNetwork(spec)
    cur = GraphBatch.from_states(cur_s, dense_mask=False).to(DEV)
    nxt = GraphBatch.from_states(nxt_s, dense_mask=False).to(DEV)
    act = torch.from_numpy(actions).to(DEV)
    mask = cur.mask                                             # [4, 3M] generated from n on the device
    assert int(mask.sum()) == B * A                             # n = n_max: everything valid

    # ---- embedder on 1000-node graphs vs the CPU oracle -------------------------------------
    with torch.no_grad():
        emb = actor.embedder(cur.x, cur.edge_index, cur.batch)
    ref_emb = O.gnn_embedder(torch.from_numpy(cur_s.x).double(), torch.from_numpy(cur_s.edge_index),
                             torch.from_numpy(cur_s.batch), _embedder_sd(actor), num_graphs=B)
    assert rel_err(emb.cpu(), ref_emb) < 2e-5

    # ---- dense layers at K = 22 + A / 22 + 2A vs fp64 torch on the same weights ---------------
    for prec, tol in (("fp32", 2e-5), ("tf32x3", 2e-5), ("tf32", 3e-3)):
        with ops.use_precision(prec), torch.no_grad():
            y = actor(cur)
            q = qf1(cur, act)
        e64 = emb.double()
        W1, b1, W2, b2 = [p.detach() for p in actor.actor.parameters()]
        h = torch.relu(torch.cat([e64, mask.double()], 1) @ W1.double().t() + b1.double())
        y_ref = (h @ W2.double().t() + b2.double()) * mask.double()
        assert rel_err(y, y_ref) < tol, prec
        del h, y_ref
        with torch.no_grad():
            emb_q = qf1.embedder(cur.x, cur.edge_index, cur.batch).double()
        V1, c1, V2, c2 = [p.detach() for p in qf1.critic.parameters()]
        hq = torch.relu(torch.cat([emb_q, mask.double(), act.double()], 1) @ V1.double().t() + c1.double())
        q_ref = hq @ V2.double().t() + c2.double()
        assert rel_err(q, q_ref) < tol, prec
    del W1, W2, V1, V2

    # ---- critic backward at K = 22 + 2A: dense-layer gradients vs fp64 autograd -----------------
    flat = qf1._check_flat()
    g = torch.zeros_like(flat)
    qv, ctx = net_forward(qf1, flat, cur.view(), act, save=True)
    dq = torch.linspace(-1, 1, B, device=DEV).view(B, 1)
    net_backward(qf1, flat, cur, ctx, dq, g)
    _, g_lins = qf1.views(g)
    V1r = qf1.critic.lins[0].weight.detach().double().requires_grad_(True)
    c1r = qf1.critic.lins[0].bias.detach().double().requires_grad_(True)
    V2r = qf1.critic.lins[1].weight.detach().double().requires_grad_(True)
    zin = torch.cat([emb_q, mask.double(), act.double()], 1)
    q64 = torch.relu(zin @ V1r.t() + c1r) @ V2r.t()
    (q64 * dq.double()).sum().backward()
    assert rel_err(g_lins[0][0], V1r.grad) < 2e-5
    assert rel_err(g_lins[0][1], c1r.grad) < 2e-5
    assert rel_err(g_lins[1][0], V2r.grad) < 2e-5
    del V1r, zin, q64, g, ctx

    # ---- one full DDPG update at this size -----------------------------------------------------
    with torch.device(DEV):
        ta, tq = Actor(spec), QNetwork(spec)
    ta.load_state_dict(actor.state_dict())
    tq.load_state_dict(qf1.state_dict())
    upd = ddpg.DDPGUpdater(actor, qf1, ta, tq, lr=1e-3, gamma=0.99, tau=1e-2)
    before = [p.detach().clone() for p in qf1.parameters()] + [next(actor.actor.parameters()).detach()[:, :64].clone()]
    qv, ql, al = upd.update(cur.view(), nxt.view(), act, torch.from_numpy(rewards).to(DEV), torch.from_numpy(dones).to(DEV))
    assert torch.isfinite(qv).all() and torch.isfinite(ql).all() and torch.isfinite(al).all()
    for b, p in zip(before[:-1], qf1.parameters()):
        assert not torch.equal(b, p.detach())
    assert not torch.equal(before[-1], next(actor.actor.parameters()).detach()[:, :64])
    assert not torch.equal(ta.flat, actor.flat) and upd.steps_dev.tolist() == [1, 1]


def test_c5_shape_inference_65536_states_chunk_consistent():
    _need_gib(40)
    n_max, B = 110, 65536
    spec = AgentSpec(n_max)
    st = synth.make_states(B, n_max, 60, 110, seed=5, num_topologies=24)
    torch.manual_seed(5)
    actor = Actor(spec).to(DEV)
    gb = GraphBatch.from_states(st, dense_mask=False).to(DEV)
    y = ddpg.predict(actor, gb)
    assert y.shape == (B, st.action_dim) and gb._mask is None
    assert torch.isfinite(y).all()
    # masked-out entries are exactly zero; valid entries of a trained-from-init actor are not all zero
    n = gb.n_nodes.long()
    idx = torch.arange(n_max, device=DEV)
    valid = (idx[None, :] < n[:, None])
    y4 = y.view(B, n_max, 3, n_max)
    assert float(y4.masked_fill(valid[:, :, None, None] & valid[:, None, None, :], 0.0).abs().max()) == 0.0
    assert int((y != 0).sum()) > 0
    # whole batch == chunks of 8192 states (same kernels, different grids): values to fp32 rounding,
    # decisions (support set and per-group argmax) exactly
    ptr = torch.from_numpy(st.ptr)
    for c0 in range(0, B, 8192 * 3):
        c1 = c0 + 8192
        lo, hi = int(ptr[c0]), int(ptr[c1])
        sel = (st.edge_index[0] >= lo) & (st.edge_index[0] < hi)
        part = synth.StateBatch(x=st.x[lo:hi], edge_index=np.ascontiguousarray(st.edge_index[:, sel] - lo),
                                batch=st.batch[lo:hi] - c0, ptr=st.ptr[c0:c1 + 1] - lo, n=st.n[c0:c1], mask=None,
                                n_max=n_max)
        yp = ddpg.predict(actor, GraphBatch.from_states(part, dense_mask=False).to(DEV))
        assert rel_err(yp, y[c0:c1]) < 1e-5
        assert torch.equal(yp > 0, y[c0:c1] > 0)
        assert torch.equal(yp.view(-1, n_max).argmax(-1), y[c0:c1].reshape(-1, n_max).argmax(-1))
    # against the CPU oracle on the first 64 states
    S = 64
    lo, hi = 0, int(ptr[S])
    sel = st.edge_index[0] < hi
    obs = {"x": torch.from_numpy(st.x[:hi]).double(), "edge_index": torch.from_numpy(np.ascontiguousarray(st.edge_index[:, sel])),
           "batch": torch.from_numpy(st.batch[:hi]), "num_graphs": S,
           "mask": torch.from_numpy(synth.structured_mask(st.n[:S], n_max, 1, 3)).double()}
    sd = {k: v.detach().cpu().double() for k, v in actor.state_dict().items()}
    y_ref = O.actor_forward(sd, obs)
    tol = 3e-3 if ops.precision == "tf32" else 1e-4
    assert rel_err(y[:S].cpu(), y_ref) < tol
