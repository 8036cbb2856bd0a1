"""GPU parity tests of the reference-facing modules and of the fused DDPG step."""
import numpy as np
import pytest
import torch

from gsc_b200 import ddpg, ops, synth
from gsc_b200.buffer import GraphReplayBuffer
from gsc_b200.graph import GraphBatch, GraphData
from gsc_b200.models import Actor, AgentSpec, GNNEmbedder, QNetwork
from oracle import gsc_oracle as O
from conftest import sd_from_golden

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-4          # north_star: embeddings, logits, values, gradients within rel 1e-4 of the fp64 oracle


def rel_err(got, ref):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def grad_err(got, ref, scale):
    """Error of one gradient tensor relative to max(its own magnitude, 1e-3 x the largest gradient of
    the network): tensors whose true gradient is ~0 (e.g. lin_r.bias, to which the segment softmax
    is shift-invariant) hold only fp32 cancellation noise and cannot be judged against themselves."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-3 * scale))


def _gscale(g, prefix):
    return max(float(np.abs(v).max()) for k, v in g.items() if k.startswith(prefix))


def _nets(g, tau=1e-2):
    spec = AgentSpec(int(g["n_max"]), target_model_update=tau)
    nets = []
    for cls, pre in ((Actor, "actor."), (QNetwork, "critic."), (Actor, "actor."), (QNetwork, "critic.")):
        net = cls(spec)
        net.load_state_dict(sd_from_golden(g, pre))
        nets.append(net.to(DEV))
    return spec, nets


def _batches(g):
    B = len(g["n"])
    mk = lambda x: GraphBatch.from_arrays(x, g["edge_index"], g["batch"], g["mask"], B, int(g["n"].max())).to(DEV)
    return mk(g["x"]), mk(g["x_next"])


def test_forward_matches_golden(golden_ddpg):
    g = golden_ddpg
    spec, (actor, critic, _, _) = _nets(g)
    cur, _ = _batches(g)
    with torch.no_grad():
        y = actor(cur)
        q = critic(cur, torch.from_numpy(g["actions"]).to(DEV))
        emb = actor.embedder(cur.x, cur.edge_index, cur.batch)
    assert rel_err(emb.cpu().numpy(), g["emb_f64"]) < TOL
    assert rel_err(y.cpu().numpy(), g["y_f64"]) < TOL
    assert rel_err(q.cpu().numpy(), g["q_f64"]) < TOL
    # decisions: support set and per-group argmax are bit-exact (no ties in the fixture)
    yh, yr = y.cpu().numpy(), g["y_f32"]
    assert np.array_equal(yh > 0, yr > 0)
    nm = int(g["n_max"])
    assert np.array_equal(yh.reshape(len(g["n"]), -1, nm).argmax(-1), yr.reshape(len(g["n"]), -1, nm).argmax(-1))
    # single observation with batch=None, as predict()/inference.py call it
    st0 = GraphData(g["x"][:int(g["n"][0])],
                    g["edge_index"][:, g["edge_index"][0] < int(g["n"][0])], g["mask"][:1]).to(DEV)
    y0 = ddpg.predict(actor, st0)
    assert y0.shape == (1, g["mask"].shape[1])
    assert rel_err(y0.cpu().numpy(), g["y_f64"][:1]) < TOL


def test_autograd_path_matches_golden_gradients(golden_ddpg):
    """The reference's own update code (simple_ddpg.py:204-227) run verbatim on the drop-in modules."""
    g = golden_ddpg
    spec, (actor, qf1, target_actor, qf1_target) = _nets(g)
    cur, nxt = _batches(g)
    actions = torch.from_numpy(g["actions"]).to(DEV)
    rewards, dones = torch.from_numpy(g["rewards"]).to(DEV), torch.from_numpy(g["dones"]).to(DEV)
    q_opt = torch.optim.Adam(list(qf1.parameters()), lr=1e-3)
    with torch.no_grad():
        next_state_actions = target_actor(nxt).clamp(-1, 1)
        qf1_next_target = qf1_target(nxt, next_state_actions)
        next_q_value = rewards.flatten() + (1 - dones.flatten()) * 0.99 * qf1_next_target.view(-1)
    qf1_a_values = qf1(cur, actions).view(-1)
    qf1_loss = torch.nn.functional.mse_loss(qf1_a_values, next_q_value)
    q_opt.zero_grad()
    qf1_loss.backward()
    for k, p in qf1.named_parameters():
        assert grad_err(p.grad.cpu().numpy(), g["gcritic_f64." + k], _gscale(g, "gcritic_f64.")) < TOL, k
    q_opt.step()
    actor_loss = -qf1(cur, actor(cur)).mean()
    actor_loss.backward()
    for k, p in actor.named_parameters():
        assert grad_err(p.grad.cpu().numpy(), g["gactor_f64." + k], _gscale(g, "gactor_f64.")) < TOL, k


def test_fused_ddpg_step_matches_golden(golden_ddpg):
    g = golden_ddpg
    spec, (actor, qf1, target_actor, qf1_target) = _nets(g)
    cur, nxt = _batches(g)
    upd = ddpg.DDPGUpdater(actor, qf1, target_actor, qf1_target, lr=1e-3, gamma=0.99, tau=1e-2)
    actions = torch.from_numpy(g["actions"]).to(DEV)
    rewards, dones = torch.from_numpy(g["rewards"]).to(DEV), torch.from_numpy(g["dones"]).to(DEV)
    for step in range(2):
        qv, ql, al = upd.update(cur, nxt, actions, rewards, dones)
        if step == 0:
            gq = dict(zip([k for k, _ in qf1.named_parameters()], qf1.param_views(upd.g_critic)))
            ga = dict(zip([k for k, _ in actor.named_parameters()], actor.param_views(upd.g_actor)))
            for k, v in gq.items():
                assert grad_err(v.cpu().numpy(), g["gcritic_f64." + k], _gscale(g, "gcritic_f64.")) < TOL, k
            for k, v in ga.items():
                assert grad_err(v.cpu().numpy(), g["gactor_f64." + k], _gscale(g, "gactor_f64.")) < TOL, k
    assert rel_err(np.array([float(ql), float(al)]), g["losses_f64"]) < TOL
    for k, p in actor.named_parameters():
        assert rel_err(p.detach().cpu().numpy(), g["post_actor_f64." + k]) < TOL, k
    for k, p in qf1_target.named_parameters():
        assert rel_err(p.detach().cpu().numpy(), g["post_critic_t_f64." + k]) < TOL, k


def test_fused_step_equals_oracle_on_config3_shape():
    """c3 shape scaled to what the CPU oracle finishes in seconds: N_max = n = 30, B = 12."""
    n_max, B = 30, 12
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(B, n_max, n_max, n_max, seed=77, fixed_links=45)
    A = cur_s.action_dim
    a_sd, c_sd = O.init_actor(3, 22, 2, A, seed=5, dtype=torch.float64), O.init_critic(3, 22, 2, A, seed=6, dtype=torch.float64)
    spec = AgentSpec(n_max)
    nets = []
    for cls, sd in ((Actor, a_sd), (QNetwork, c_sd), (Actor, a_sd), (QNetwork, c_sd)):
        net = cls(spec)
        net.load_state_dict({k: v.float() for k, v in sd.items()})
        nets.append(net.to(DEV))
    upd = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-4)
    ref = O.OracleDDPG(a_sd, c_sd, lr=1e-3, gamma=0.99, tau=1e-4)
    from conftest import obs_from_states
    o_cur, o_nxt = obs_from_states(cur_s, torch.float64), obs_from_states(nxt_s, torch.float64)
    cur, nxt = GraphBatch.from_states(cur_s).to(DEV), GraphBatch.from_states(nxt_s).to(DEV)
    for _ in range(3):
        qv, ql, al = upd.update(cur, nxt, torch.from_numpy(actions).to(DEV), torch.from_numpy(rewards).to(DEV),
                                torch.from_numpy(dones).to(DEV))
        rq, rql, ral = ref.update(o_cur, o_nxt, torch.from_numpy(actions).double(), torch.from_numpy(rewards).double(),
                                  torch.from_numpy(dones).double())
    assert rel_err(qv.cpu().numpy(), rq.numpy()) < TOL
    assert rel_err([float(ql), float(al)], [float(rql), float(ral)]) < TOL
    for k, p in nets[0].named_parameters():
        # Adam's first steps move every weight by ~lr regardless of gradient scale, so compare the
        # applied update, not only the weight
        d_got = p.detach().cpu().double().numpy() - a_sd[k].numpy()
        d_ref = ref.actor[k].detach().numpy() - a_sd[k].numpy()
        assert rel_err(p.detach().cpu().numpy(), ref.actor[k].detach().numpy()) < TOL, k
        assert np.abs(d_got - d_ref).max() < 0.05 * 3e-3 + 1e-7, k


def test_cuda_graph_replay_equals_eager_updates():
    n_max, B = 12, 10
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(B, n_max, 4, 12, seed=5, num_topologies=3)
    spec = AgentSpec(n_max)

    def build():
        torch.manual_seed(7)
        nets = [cls(spec).to(DEV) for cls in (Actor, QNetwork, Actor, QNetwork)]
        nets[2].load_state_dict(nets[0].state_dict())
        nets[3].load_state_dict(nets[1].state_dict())
        return nets, ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2)

    t = lambda a: torch.from_numpy(a).to(DEV)
    cur, nxt = GraphBatch.from_states(cur_s).to(DEV), GraphBatch.from_states(nxt_s).to(DEV)
    nets_e, upd_e = build()
    nets_g, upd_g = build()
    assert torch.equal(nets_e[0].flat, nets_g[0].flat)
    graphed = ddpg.GraphedUpdate(upd_g, cur, nxt, t(actions), t(rewards), t(dones))
    assert torch.equal(nets_e[0].flat, nets_g[0].flat), "capture must not change the weights"
    fresh = lambda gb: gb.view()
    for step in range(3):
        cur_k = cur.view()
        cur_k.x = torch.rand_like(cur.x)                         # a new minibatch each step
        qe, le, ae = upd_e.update(cur_k, fresh(nxt), t(actions), t(rewards) + step, t(dones))
        graphed.load(cur_k, nxt, t(actions), t(rewards) + step, t(dones))
        qg, lg, ag = graphed.replay()
        assert rel_err(qg.cpu().numpy(), qe.cpu().numpy()) < 1e-5
        assert rel_err([float(lg), float(ag)], [float(le), float(ae)]) < 1e-5
    for ne, ng in zip(nets_e, nets_g):
        assert rel_err(ng.flat.cpu().numpy(), ne.flat.cpu().numpy()) < 1e-5
    assert upd_g.steps_dev.tolist() == [3, 3] and upd_e.steps_dev.tolist() == [3, 3]


def test_replay_buffer_to_update_pipeline_runs():
    spec = AgentSpec(8)
    rb = GraphReplayBuffer(32, spec.env.action_space, device=DEV)
    st = synth.make_states(24, 8, 3, 8, seed=3, num_topologies=3)
    acts = synth.make_actions(st)
    for gi in range(24):
        d, nd = st.graph(gi), st.graph((gi + 1) % 24)
        rb.add(GraphData(d["x"], d["edge_index"], d["mask"]), GraphData(nd["x"], nd["edge_index"], nd["mask"]),
               acts[gi], 0.1 * gi, gi % 9 == 0, {})
    nets = [cls(spec).to(DEV) for cls in (Actor, QNetwork, Actor, QNetwork)]
    nets[2].load_state_dict(nets[0].state_dict())
    nets[3].load_state_dict(nets[1].state_dict())
    upd = ddpg.DDPGUpdater(*nets)
    np.random.seed(1)
    before = nets[0].flat.clone()
    for _ in range(3):
        s = rb.sample(10)
        qv, ql, al = upd.update(s.observations, s.next_observations, s.actions, s.rewards, s.dones)
    assert torch.isfinite(qv).all() and torch.isfinite(ql).all() and torch.isfinite(al).all()
    assert not torch.equal(before, nets[0].flat)


def test_device_replay_buffer_samples_what_the_host_buffer_samples():
    from gsc_b200.buffer import DeviceGraphReplayBuffer
    n_max = 8
    spec = AgentSpec(n_max)
    host = GraphReplayBuffer(16, spec.env.action_space, device=DEV)
    dev = DeviceGraphReplayBuffer(16, spec.env.action_space, device=DEV, n_max=n_max)
    st = synth.make_states(21, n_max, 2, 8, seed=4, num_topologies=4)
    acts = synth.make_actions(st)
    for gi in range(21):                                             # wraps around
        d, nd = st.graph(gi), st.graph((gi * 7 + 3) % 21)            # next state on another topology
        o, no = GraphData(d["x"], d["edge_index"], d["mask"]), GraphData(nd["x"], nd["edge_index"], nd["mask"])
        host.add(o, no, acts[gi], 0.25 * gi, gi % 5 == 0, {})
        dev.add(o, no, acts[gi], 0.25 * gi, gi % 5 == 0, {})
    assert dev.full and dev.pos == host.pos
    for B in (1, 6):
        np.random.seed(B)
        a = host.sample(B)
        np.random.seed(B)
        b = dev.sample(B)
        for ga, gb in ((a.observations, b.observations), (a.next_observations, b.next_observations)):
            assert torch.equal(ga.x, gb.x) and torch.equal(ga.batch, gb.batch) and torch.equal(ga.mask, gb.mask)
            assert torch.equal(ga.edge_index, gb.edge_index)
            assert gb.num_graphs == B and gb.max_nodes == ga.max_nodes
        assert torch.equal(a.actions, b.actions) and torch.equal(a.rewards, b.rewards) and torch.equal(a.dones, b.dones)
        # in-place refill of a previous sample (static CUDA-graph inputs) gives the same batch
        np.random.seed(B)
        c = dev.sample(B, out=b)
        assert c.observations is b.observations and torch.equal(c.observations.x, a.observations.x)
    bad = GraphData(st.graph(0)["x"], st.graph(0)["edge_index"], np.ones((1, st.action_dim), np.float32))
    if int(st.n[0]) < n_max:
        with pytest.raises(ValueError):
            dev.add(bad, bad, acts[0], 0.0, False, {})


def test_build_mask_matches_structured_mask():
    n = np.array([1, 5, 8, 3, 8], dtype=np.int32)
    got = ops.build_mask(torch.from_numpy(n).to(DEV), 8, 3).cpu().numpy()
    assert np.array_equal(got, synth.structured_mask(n, 8))
    got = ops.build_mask(torch.from_numpy(n[:2]).to(DEV), 5, 1).cpu().numpy()      # A = 25: scalar kernel
    assert np.array_equal(got, synth.structured_mask(np.minimum(n[:2], 5), 5, 1, 1))


def test_post_process_and_choose_action_entry_points(golden_ddpg):
    g = golden_ddpg
    spec, (actor, _, _, _) = _nets(g)
    cur, _ = _batches(g)
    nm = int(g["n_max"])
    y = ddpg.predict(actor, cur)
    rng = np.random.default_rng(0)
    noise = rng.normal(0, 0.3, size=tuple(y.shape))
    got = ddpg.choose_action(actor, cur, torch.from_numpy(noise).to(DEV), nm).cpu().numpy()
    yh = y.cpu().numpy()
    for r in range(2):
        want = O.post_process_actions(O.explore_action(yh[r], noise[r]), nm)
        assert np.array_equal(got[r], want)
    pp = ddpg.post_process_actions(yh[0], nm).cpu().numpy()
    assert np.array_equal(pp, O.post_process_actions(yh[0], nm))


def test_config2_full_size_properties():
    """BASELINE config 2 at full size (B=4096, N_max=24): spot-check states against the oracle and
    check size-independent properties on all of them."""
    B, n_max = 4096, 24
    st = synth.make_states(B, n_max, 16, 24, seed=1234)
    spec = AgentSpec(n_max)
    sd = O.init_actor(3, 22, 2, st.action_dim, seed=1234, dtype=torch.float64)
    actor = Actor(spec)
    actor.load_state_dict({k: v.float() for k, v in sd.items()})
    actor = actor.to(DEV)
    gb = GraphBatch.from_states(st).to(DEV)
    y = ddpg.predict(actor, gb)
    assert y.shape == (B, st.action_dim)
    yh = y.cpu().numpy()
    assert np.all(yh[st.mask == 0] == 0)                               # masked entries are exactly zero
    for gi in (0, 1, 777, 4095):
        one = st.graph(gi)
        obs = {"x": torch.from_numpy(one["x"]).double(), "edge_index": torch.from_numpy(one["edge_index"]),
               "batch": None, "mask": torch.from_numpy(one["mask"]).double()}
        ref = O.actor_forward(sd, obs).numpy()
        assert rel_err(yh[gi:gi + 1], ref) < TOL
        assert np.array_equal(yh[gi:gi + 1] > 0, ref > 0)
    # batch composition does not matter: a permuted batch gives the permuted result
    perm = np.random.default_rng(0).permutation(B)[:512]
    sub = GraphBatch.from_data_list([GraphData(**st.graph(int(i))) for i in perm]).to(DEV)
    y_sub = ddpg.predict(actor, sub).cpu().numpy()
    assert rel_err(y_sub, yh[perm]) < 1e-5


def test_graphed_predict_equals_eager_predict():
    n_max, B = 24, 96
    st = synth.make_states(B, n_max, 16, 24, seed=12, num_topologies=6)
    st2 = synth.make_states(B, n_max, 16, 24, seed=12, num_topologies=6)
    st2.x = np.random.default_rng(3).random(st.x.shape, dtype=np.float32)     # same topologies, new features
    actor = Actor(AgentSpec(n_max)).to(DEV)
    gb = GraphBatch.from_states(st, dense_mask=False).to(DEV)
    gp = ddpg.GraphedPredict(actor, gb)
    assert torch.equal(gp.replay(), ddpg.predict(actor, gb.view()))
    gb2 = GraphBatch.from_states(st2, dense_mask=False).to(DEV)
    gp.load(gb2)
    assert torch.equal(gp.replay(), ddpg.predict(actor, gb2))


@pytest.mark.parametrize("graphed", [False, True])
def test_forked_stream_update_is_bit_identical(graphed):
    """DDPGUpdater(fork_streams=True) runs the same kernels on the same data, on three streams."""
    n_max, B = 16, 40
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(B, n_max, 6, 16, seed=9, num_topologies=4)
    spec = AgentSpec(n_max)
    t = lambda a: torch.from_numpy(a).to(DEV)
    outs = []
    for fork in (False, True):
        torch.manual_seed(21)
        nets = [cls(spec).to(DEV) for cls in (Actor, QNetwork, Actor, QNetwork)]
        nets[2].load_state_dict(nets[0].state_dict())
        nets[3].load_state_dict(nets[1].state_dict())
        upd = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2, fork_streams=fork)
        cur = GraphBatch.from_states(cur_s, dense_mask=False).to(DEV)
        nxt = GraphBatch.from_states(nxt_s, dense_mask=False).to(DEV)
        if graphed:
            g = ddpg.GraphedUpdate(upd, cur, nxt, t(actions), t(rewards), t(dones))
            for _ in range(3):
                res = g.replay()
        else:
            for _ in range(3):
                res = upd.update(cur.view(), nxt.view(), t(actions), t(rewards), t(dones))
        torch.cuda.synchronize()
        outs.append(([r.clone() for r in res], [n.flat.clone() for n in nets]))
    for a, b in zip(outs[0][0] + outs[0][1], outs[1][0] + outs[1][1]):
        assert torch.equal(a, b)


def test_graphed_update_with_captured_replay_gather_equals_eager_sampling():
    """GraphedUpdate(replay=rb).replay_sampled() == rb.sample(B) + DDPGUpdater.update on the same draws."""
    from gsc_b200.buffer import DeviceGraphReplayBuffer
    n_max, B = 12, 16
    spec = AgentSpec(n_max)
    cur_s, nxt_s, actions, rewards, dones = synth.make_transitions(64, n_max, n_max, n_max, seed=31, fixed_links=18,
                                                                   num_topologies=6)

    def build():
        torch.manual_seed(5)
        nets = [cls(spec).to(DEV) for cls in (Actor, QNetwork, Actor, QNetwork)]
        nets[2].load_state_dict(nets[0].state_dict())
        nets[3].load_state_dict(nets[1].state_dict())
        rb = DeviceGraphReplayBuffer(128, spec.env.action_space, device=DEV, n_max=n_max)
        rb.bulk_load(cur_s, nxt_s, actions, rewards, dones)
        return nets, ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2, fork_streams=True), rb

    nets_e, upd_e, rb_e = build()
    nets_g, upd_g, rb_g = build()
    np.random.seed(3)
    s0 = rb_g.sample(B)
    g = ddpg.GraphedUpdate(upd_g, s0.observations, s0.next_observations, s0.actions, s0.rewards, s0.dones, replay=rb_g)
    np.random.seed(11)
    outs_g = [tuple(t.clone() for t in g.replay_sampled()) for _ in range(3)]
    np.random.seed(11)
    outs_e = []
    for _ in range(3):
        s = rb_e.sample(B)
        outs_e.append(tuple(t.clone() for t in upd_e.update(s.observations, s.next_observations, s.actions, s.rewards, s.dones)))
    torch.cuda.synchronize()
    for a, b in zip(outs_g, outs_e):
        for x, y in zip(a, b):
            assert torch.equal(x.reshape(-1), y.reshape(-1))
    for ng, ne in zip(nets_g, nets_e):
        assert torch.equal(ng.flat, ne.flat)
