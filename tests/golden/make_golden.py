"""Generates tests/golden/*.npz from the CPU oracle (oracle/gsc_oracle.py).

The reference cannot be imported in the authoring container (torch-geometric 2.3.0, gym, simpy,
stable-baselines3 absent) and it ships no golden vectors for this path, so
these fixtures pin the ORACLE (a regression anchor for it and the comparison target of the CUDA
path), not the reference itself -- parity stays "unpinned" in the sense of DESIGN.md.

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gsc_b200 import synth  # noqa: E402
from oracle import gsc_oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def obs_of(st, dtype):
    return {"x": torch.from_numpy(st.x).to(dtype), "edge_index": torch.from_numpy(st.edge_index),
            "batch": torch.from_numpy(st.batch), "mask": torch.from_numpy(st.mask).to(dtype),
            "num_graphs": st.num_graphs}


def main():
    torch.manual_seed(0)
    n_max, B, H, F = 8, 6, 22, 3
    cur, nxt, actions, rewards, dones = synth.make_transitions(B, n_max, 3, 8, seed=1234, num_topologies=4)
    dones[1] = 1.0
    A = cur.action_dim
    actor_sd = O.init_actor(F, H, 2, A, seed=11)
    critic_sd = O.init_critic(F, H, 2, A, seed=12)
    for sd in (actor_sd, critic_sd):          # GATv2 output bias is zero-initialised: make it count
        for k in sd:
            if k.endswith("encoder.bias") or k.endswith("process.0.bias"):
                sd[k] = torch.linspace(-0.1, 0.1, H)
    out = {"n_max": n_max, "x": cur.x, "x_next": nxt.x, "edge_index": cur.edge_index, "batch": cur.batch,
           "n": cur.n, "mask": cur.mask, "actions": actions, "rewards": rewards, "dones": dones}
    for k, v in actor_sd.items():
        out["actor." + k] = v.numpy()
    for k, v in critic_sd.items():
        out["critic." + k] = v.numpy()
    for name, dt in (("f32", torch.float32), ("f64", torch.float64)):
        a_sd = {k: v.to(dt) for k, v in actor_sd.items()}
        c_sd = {k: v.to(dt) for k, v in critic_sd.items()}
        o_cur, o_nxt = obs_of(cur, dt), obs_of(nxt, dt)
        emb = O.gnn_embedder(o_cur["x"], o_cur["edge_index"], o_cur["batch"], a_sd, 2, 2, num_graphs=B)
        y = O.actor_forward(a_sd, o_cur)
        q = O.critic_forward(c_sd, o_cur, torch.from_numpy(actions).to(dt))
        out[f"emb_{name}"], out[f"y_{name}"], out[f"q_{name}"] = emb.numpy(), y.numpy(), q.numpy()
        ddpg = O.OracleDDPG(a_sd, c_sd, lr=1e-3, gamma=0.99, tau=1e-2)
        for step in range(2):
            qv, ql, al = ddpg.update(o_cur, o_nxt, torch.from_numpy(actions).to(dt),
                                     torch.from_numpy(rewards).to(dt), torch.from_numpy(dones).to(dt))
            if step == 0:
                for k, v in ddpg.last_critic_grads.items():
                    out[f"gcritic_{name}.{k}"] = v.numpy()
                for k, v in ddpg.last_actor_grads.items():
                    out[f"gactor_{name}.{k}"] = v.numpy()
        out[f"losses_{name}"] = np.array([float(ql), float(al)])
        for k, v in ddpg.actor.items():
            out[f"post_actor_{name}.{k}"] = v.detach().numpy()
        for k, v in ddpg.critic_t.items():
            out[f"post_critic_t_{name}.{k}"] = v.detach().numpy()
    np.savez_compressed(os.path.join(HERE, "ddpg_small.npz"), **out)

    # action post-processing cases (fp32 inputs, fp64 outputs)
    rng = np.random.default_rng(5)
    cases = []
    for nn in (3, 8, 24):
        g = 3 * nn
        a = rng.random((g, nn)).astype(np.float32)
        a[a < 0.55] = 0                                   # sparse-ish rows
        a[0] = 0                                          # all below threshold -> uniform
        a[1] = np.float32(1.0 / nn)                       # equal split
        a[2, :] = 0
        a[2, 0] = np.float32(0.6)
        a[2, -1] = np.float32(0.4)                        # already normalised
        if nn >= 8:
            a[3] = rng.random(nn).astype(np.float32) * 0.2  # many just around the threshold
            a[4] = np.float32(0.1)
        cases.append((nn, a.reshape(-1)))
    pp = {}
    for nn, a in cases:
        pp[f"in_{nn}"] = a
        pp[f"out_{nn}"] = O.post_process_actions(a, nn)
        noise = rng.normal(0.0, 0.3, size=a.shape)
        pp[f"noise_{nn}"] = noise
        pp[f"explore_{nn}"] = O.post_process_actions(O.explore_action(a, noise), nn)
    np.savez_compressed(os.path.join(HERE, "actor_head.npz"), **pp)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
