"""CPU tests of the oracle: golden-vector regression, an independent loop restatement of GATv2,
fp32-vs-fp64 tolerance anchoring, and the action post-processing semantics."""
import math

import numpy as np
import pytest
import torch

from gsc_b200 import synth
from oracle import gsc_oracle as O
from conftest import sd_from_golden


def _obs(g, dtype, x_key="x"):
    return {"x": torch.from_numpy(g[x_key]).to(dtype), "edge_index": torch.from_numpy(g["edge_index"]),
            "batch": torch.from_numpy(g["batch"]), "mask": torch.from_numpy(g["mask"]).to(dtype),
            "num_graphs": len(g["n"])}


@pytest.mark.parametrize("name,dtype,tol", [("f32", torch.float32, 1e-6), ("f64", torch.float64, 1e-12)])
def test_oracle_matches_golden(golden_ddpg, name, dtype, tol):
    g = golden_ddpg
    a_sd, c_sd = sd_from_golden(g, "actor.", dtype), sd_from_golden(g, "critic.", dtype)
    obs = _obs(g, dtype)
    emb = O.gnn_embedder(obs["x"], obs["edge_index"], obs["batch"], a_sd, 2, 2, num_graphs=len(g["n"]))
    np.testing.assert_allclose(emb.numpy(), g[f"emb_{name}"], rtol=tol, atol=tol)
    np.testing.assert_allclose(O.actor_forward(a_sd, obs).numpy(), g[f"y_{name}"], rtol=tol, atol=tol)
    q = O.critic_forward(c_sd, obs, torch.from_numpy(g["actions"]).to(dtype))
    np.testing.assert_allclose(q.numpy(), g[f"q_{name}"], rtol=tol, atol=tol)


def test_fp32_oracle_within_tolerance_of_fp64(golden_ddpg):
    g = golden_ddpg
    for key in ("emb", "y", "q"):
        ref = g[f"{key}_f64"]
        err = np.abs(g[f"{key}_f32"] - ref).max() / max(np.abs(ref).max(), 1e-30)
        assert err < 1e-5, (key, err)


def _gatv2_loops(x, edge_index, p):
    """Independent restatement with explicit python loops over nodes and edges (float64)."""
    n, H = x.shape[0], p["bias"].shape[0]
    Wl, bl, Wr, br = p["lin_l.weight"], p["lin_l.bias"], p["lin_r.weight"], p["lin_r.bias"]
    att, bias = p["att"].reshape(-1), p["bias"]
    xl = [[sum(Wl[c][k] * x[i][k] for k in range(x.shape[1])) + bl[c] for c in range(H)] for i in range(n)]
    xr = [[sum(Wr[c][k] * x[i][k] for k in range(x.shape[1])) + br[c] for c in range(H)] for i in range(n)]
    nbrs = [[] for _ in range(n)]
    for s, d in zip(edge_index[0].tolist(), edge_index[1].tolist()):
        if s != d:
            nbrs[d].append(s)
    out = np.zeros((n, H))
    for i in range(n):
        js = nbrs[i] + [i]
        scores = []
        for j in js:
            sc = 0.0
            for c in range(H):
                u = xl[j][c] + xr[i][c]
                sc += att[c] * (u if u > 0 else 0.2 * u)
            scores.append(sc)
        m = max(scores)
        ex = [math.exp(s - m) for s in scores]
        den = sum(ex) + 1e-16
        for c in range(H):
            out[i][c] = sum((e / den) * xl[j][c] for e, j in zip(ex, js)) / len(js) + bias[c]
    return out


def test_gatv2_against_loop_restatement():
    rng = np.random.default_rng(3)
    n, F, H = 7, 3, 5
    ei = synth.random_topology(n, 9, rng)
    ei = np.concatenate([ei, np.array([[2, 4, 4], [2, 1, 1]])], axis=1)   # a self loop and a duplicate edge
    gen = torch.Generator().manual_seed(1)
    p = O.init_gatv2(F, H, gen, torch.float64)
    p["bias"] = torch.linspace(-1, 1, H, dtype=torch.float64)
    x = torch.from_numpy(rng.random((n, F)))
    got = O.gatv2_conv(x, torch.from_numpy(ei), p).numpy()
    want = _gatv2_loops(x.numpy(), torch.from_numpy(ei), {k: v.numpy() for k, v in p.items()})
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)


def test_isolated_node_and_single_graph_pool():
    gen = torch.Generator().manual_seed(2)
    p = O.init_gatv2(3, 4, gen, torch.float64)
    x = torch.rand(1, 3, dtype=torch.float64)
    out = O.gatv2_conv(x, torch.zeros(2, 0, dtype=torch.int64), p)
    xl = x @ p["lin_l.weight"].t() + p["lin_l.bias"]
    np.testing.assert_allclose(out.numpy(), (xl + p["bias"]).numpy(), rtol=1e-12)   # alpha == 1, deg == 1
    h = torch.rand(5, 4, dtype=torch.float64)
    np.testing.assert_allclose(O.global_mean_pool(h, None).numpy(), h.mean(0, keepdim=True).numpy())


def test_conv_schedule_follows_reference_control_flow():
    assert O.conv_schedule(2, 2) == [(0, True), (1, True), (1, False)]
    assert O.conv_schedule(1, 2) == [(0, True)]
    assert O.conv_schedule(3, 2) == [(0, True), (1, True), (2, True), (1, True), (2, False)]


def test_embedder_is_permutation_equivariant_over_graph_order():
    st = synth.make_states(5, 8, 3, 8, seed=7, num_topologies=5)
    sd = O.init_actor(3, 22, 2, st.action_dim, seed=3, dtype=torch.float64)
    x = torch.from_numpy(st.x).double()
    emb = O.gnn_embedder(x, torch.from_numpy(st.edge_index), torch.from_numpy(st.batch), sd, 2, 2, num_graphs=5)
    for gidx in range(5):
        one = st.graph(gidx)
        e1 = O.gnn_embedder(torch.from_numpy(one["x"]).double(), torch.from_numpy(one["edge_index"]), None, sd, 2, 2)
        np.testing.assert_allclose(e1.numpy()[0], emb.numpy()[gidx], rtol=1e-12, atol=1e-14)


# ---- action post-processing ---------------------------------------------------------------------
def test_post_process_matches_golden(golden_head):
    for nn in (3, 8, 24):
        out = O.post_process_actions(golden_head[f"in_{nn}"], nn)
        assert out.dtype == np.float64
        assert np.array_equal(out, golden_head[f"out_{nn}"])
        ex = O.post_process_actions(O.explore_action(golden_head[f"in_{nn}"], golden_head[f"noise_{nn}"]), nn)
        assert np.array_equal(ex, golden_head[f"explore_{nn}"])


def test_post_process_properties():
    rng = np.random.default_rng(0)
    for nn in (2, 5, 24):
        a = (rng.random(6 * nn) * rng.integers(0, 2, 6 * nn)).astype(np.float32)
        out = O.post_process_actions(a, nn).reshape(-1, nn)
        assert np.all(np.abs(out.sum(1) - 1) < O.SCHED_ACCURACY)
        assert np.all(out >= 0)
        # thresholded entries stay (numerically) empty unless the whole group was empty; the
        # reference adds the rounding offset to the FIRST entry that stays >= 0, zero or not
        inp = a.reshape(-1, nn)
        for r in range(inp.shape[0]):
            if (inp[r] >= 0.1).any():
                assert np.all(out[r][inp[r] < 0.1] < 1e-8)
                assert np.all(out[r][1:][inp[r][1:] < 0.1] == 0) or out[r][0] + 0 >= 0
            else:
                assert np.allclose(out[r], 1.0 / nn, atol=1e-9)


def test_normalize_uses_plain_left_to_right_sum():
    # 0.1 * 10 summed left to right in fp64 is 0.9999999999999999, not 1.0 (what a compensated sum gives)
    lst = [np.float64(0.1)] * 10
    assert O._seq_sum(lst) == 0.9999999999999999
    out = O.normalize_scheduling_probabilities(lst)
    assert out[0] == np.float64(0.1) + (1 - 0.9999999999999999)


def test_node_features_follow_the_reference_lines():
    """simulator_wrapper.py:255-294 for one state, written with explicit float32/float64 steps."""
    ingress = np.array([0.0, 12.5, 0.0, 3.25], dtype=np.float32)
    usage = np.array([1.0, 0.0, 7.0, 2.0], dtype=np.float32)
    cap = np.array([4.0, 0.0, 2.0, 8.0], dtype=np.float32)
    x = O.node_features(ingress, usage, cap, np.array([0, 4]))
    assert x.dtype == np.float32 and x.shape == (4, 3)
    d_in = np.float32(np.float64(np.float32(12.5)) + 1e-3)
    assert np.array_equal(x[:, 0], np.clip(ingress / d_in, 0, 1).astype(np.float32))
    util = np.array([0.25, 1.0, 3.5, 0.25], dtype=np.float32)        # cap == 0 -> 1
    d_ut = np.float32(np.float64(np.float32(3.5)) + 1e-3)
    assert np.array_equal(x[:, 1], np.clip(util / d_ut, 0, 1).astype(np.float32))
    assert x[1, 2] == 0.0 and abs(float(x[3, 2]) - 8.0 / 8.001) < 1e-6
    # two states are normalised independently
    x2 = O.node_features(np.tile(ingress, 2) * np.repeat([1, 3], 4).astype(np.float32), np.tile(usage, 2),
                         np.tile(cap, 2), np.array([0, 4, 8]))
    assert np.array_equal(x2[:4, 1:], x[:, 1:]) and np.array_equal(x2[4:, 1:], x[:, 1:])
