"""CPU oracle for the GSC hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  Nothing under ``gsc_b200/`` imports it.

PARITY UNPINNED.  The reference (farzad1132/GSC) delegates all of this path's arithmetic to
``torch-geometric==2.3.0`` on ``torch==1.13.1`` (reference ``requirements.txt:59-60``).  PyG is not
vendored under /root/reference and is not installed here, and the reference holds no test or golden
vector for ``models.py`` / ``simple_ddpg.py`` / ``buffer.py``.  This file
therefore restates the *published* PyG 2.3.0 algorithms in plain PyTorch index ops and anchors on the
reference's own call sites:

  * ``GATv2Conv(in, out, aggr='mean')``      reference src/rlsp/agents/models.py:22,26
  * ``GNNEmbedder.forward``                  reference src/rlsp/agents/models.py:37-53
  * PyG ``MLP(channel_list, norm=None)``     reference src/rlsp/agents/models.py:80-83,119-122
  * ``Actor.forward`` / ``QNetwork.forward`` reference src/rlsp/agents/models.py:146-154 / 86-92
  * DDPG update order                        reference src/rlsp/agents/simple_ddpg.py:204-234
  * action post-processing                   reference src/rlsp/agents/simple_ddpg.py:338-395
  * replay-batch layout                      reference src/rlsp/agents/buffer.py:56-89

All functions work in the dtype of their inputs (fp32 for the timed baseline, fp64 to anchor
tolerances).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

NEG_SLOPE = 0.2  # GATv2Conv default negative_slope [PyG 2.3.0]


# --------------------------------------------------------------------------------------------
# Parameter initialisation (PyG 2.3.0 defaults; only matters for from-scratch training)
# --------------------------------------------------------------------------------------------
def _glorot(shape, gen):
    # torch_geometric.nn.inits.glorot: U(+-sqrt(6 / (size(-2) + size(-1))))
    a = math.sqrt(6.0 / (shape[-2] + shape[-1]))
    return (torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * a


def _uniform(shape, bound, gen):
    return (torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * bound


def init_gatv2(in_dim: int, out_dim: int, gen, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """State-dict entries of one GATv2Conv(in_dim, out_dim) (heads=1), PyG registration order."""
    p = {
        "att": _glorot((1, 1, out_dim), gen),
        "bias": torch.zeros(out_dim, dtype=torch.float64),
        "lin_l.weight": _glorot((out_dim, in_dim), gen),
        "lin_l.bias": _uniform((out_dim,), 1.0 / math.sqrt(in_dim), gen),
        "lin_r.weight": _glorot((out_dim, in_dim), gen),
        "lin_r.bias": _uniform((out_dim,), 1.0 / math.sqrt(in_dim), gen),
    }
    return {k: v.to(dtype) for k, v in p.items()}


def init_linear(in_dim: int, out_dim: int, gen, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    # kaiming_uniform(a=sqrt(5)) == U(+-1/sqrt(fan_in)); bias U(+-1/sqrt(fan_in))
    b = 1.0 / math.sqrt(in_dim)
    return {"weight": _uniform((out_dim, in_dim), b, gen).to(dtype),
            "bias": _uniform((out_dim,), b, gen).to(dtype)}


def init_embedder(in_dim, hidden, num_layers, gen, dtype=torch.float32, prefix="embedder."):
    sd = {}
    for k, v in init_gatv2(in_dim, hidden, gen, dtype).items():
        sd[f"{prefix}encoder.{k}"] = v
    for i in range(num_layers - 1):
        for k, v in init_gatv2(hidden, hidden, gen, dtype).items():
            sd[f"{prefix}process.{i}.{k}"] = v
    return sd


def init_actor(node_feat, hidden, num_layers, action_dim, mlp_hidden=(256,), seed=1234,
               dtype=torch.float32):
    """State dict with the keys of reference ``Actor`` (models.py:95-125)."""
    gen = torch.Generator().manual_seed(seed)
    sd = init_embedder(node_feat, hidden, num_layers, gen, dtype)
    chans = [hidden + action_dim, *mlp_hidden, action_dim]
    for i in range(len(chans) - 1):
        for k, v in init_linear(chans[i], chans[i + 1], gen, dtype).items():
            sd[f"actor.lins.{i}.{k}"] = v
    return sd


def init_critic(node_feat, hidden, num_layers, action_dim, mlp_hidden=(64,), seed=4321,
                dtype=torch.float32):
    """State dict with the keys of reference ``QNetwork`` (models.py:55-83)."""
    gen = torch.Generator().manual_seed(seed)
    sd = init_embedder(node_feat, hidden, num_layers, gen, dtype)
    chans = [hidden + 2 * action_dim, *mlp_hidden, 1]
    for i in range(len(chans) - 1):
        for k, v in init_linear(chans[i], chans[i + 1], gen, dtype).items():
            sd[f"critic.lins.{i}.{k}"] = v
    return sd


# --------------------------------------------------------------------------------------------
# GATv2Conv(aggr='mean'), global_mean_pool, GNNEmbedder   
# --------------------------------------------------------------------------------------------
def with_self_loops(edge_index: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """remove_self_loops + add_self_loops, as GATv2Conv.forward does on every call [PyG 2.3.0]."""
    keep = edge_index[0] != edge_index[1]
    ei = edge_index[:, keep]
    loop = torch.arange(num_nodes, dtype=edge_index.dtype)
    return torch.cat([ei, torch.stack([loop, loop])], dim=1)


def gatv2_conv(x, edge_index, p: Dict[str, torch.Tensor], prefix: str = ""):
    """One GATv2Conv(heads=1, concat=True, add_self_loops=True, aggr='mean') application."""
    n = x.shape[0]
    Wl, bl = p[prefix + "lin_l.weight"], p[prefix + "lin_l.bias"]
    Wr, br = p[prefix + "lin_r.weight"], p[prefix + "lin_r.bias"]
    att, bias = p[prefix + "att"].reshape(-1), p[prefix + "bias"]
    xl = x @ Wl.t() + bl
    xr = x @ Wr.t() + br
    ei = with_self_loops(edge_index, n)
    src, dst = ei[0], ei[1]
    u = F.leaky_relu(xl[src] + xr[dst], NEG_SLOPE)
    s = (u * att).sum(-1)                                            # [Et]
    # torch_geometric.utils.softmax: max detached, eps 1e-16 on the sum
    smax = torch.full((n,), -float("inf"), dtype=s.dtype)
    smax = smax.scatter_reduce(0, dst, s.detach(), reduce="amax", include_self=True)
    ex = (s - smax[dst]).exp()
    den = torch.zeros(n, dtype=s.dtype).index_add(0, dst, ex) + 1e-16
    alpha = ex / den[dst]
    msg = xl[src] * alpha.unsqueeze(-1)
    # aggr='mean': scatter-sum divided by the in-degree (self-loop included), clamp(min=1)
    out = torch.zeros(n, xl.shape[1], dtype=x.dtype).index_add(0, dst, msg)
    deg = torch.zeros(n, dtype=x.dtype).index_add(0, dst, torch.ones_like(s)).clamp(min=1)
    out = out / deg.unsqueeze(-1)
    return out + bias


def global_mean_pool(h, batch: Optional[torch.Tensor], num_graphs: Optional[int] = None):
    if batch is None:
        return h.mean(0, keepdim=True)
    if num_graphs is None:
        num_graphs = int(batch.max()) + 1
    out = torch.zeros(num_graphs, h.shape[1], dtype=h.dtype).index_add(0, batch, h)
    cnt = torch.zeros(num_graphs, dtype=h.dtype).index_add(0, batch, torch.ones(h.shape[0], dtype=h.dtype))
    return out / cnt.clamp(min=1).unsqueeze(-1)


def conv_schedule(num_layers: int, num_iter: int):
    """Sequence of (conv_index, relu_after) that GNNEmbedder.forward executes (models.py:37-53).

    conv_index 0 is ``encoder``; i>=1 is ``process[i-1]``.  The last application feeds the pool and
    has no ReLU, except for num_layers == 1 where the encoder output is ReLU'd and pooled.
    """
    sched = [(0, True)]
    if num_layers == 1:
        return sched
    for it in range(num_iter):
        for i in range(num_layers - 1):
            last = (i == num_layers - 2) and (it == num_iter - 1)
            sched.append((i + 1, not last))
    return sched


def gnn_embedder(x, edge_index, batch, sd, num_layers=2, num_iter=2, prefix="embedder.",
                 num_graphs=None):
    h = x
    for ci, relu in conv_schedule(num_layers, num_iter):
        pfx = prefix + ("encoder." if ci == 0 else f"process.{ci - 1}.")
        h = gatv2_conv(h, edge_index, sd, pfx)
        if relu:
            h = F.relu(h)
    return global_mean_pool(h, batch, num_graphs)


def mlp(z, sd, prefix: str):
    """PyG MLP(norm=None): Linear -> ReLU -> ... -> Linear (plain last) [PyG 2.3.0]."""
    n = len([k for k in sd if k.startswith(prefix + "lins.") and k.endswith(".weight")])
    for i in range(n):
        z = z @ sd[f"{prefix}lins.{i}.weight"].t() + sd[f"{prefix}lins.{i}.bias"]
        if i < n - 1:
            z = F.relu(z)
    return z


def actor_forward(sd, obs, num_layers=2, num_iter=2):
    """reference Actor.forward (models.py:146-154): MLP(cat[emb, mask]) * mask."""
    emb = gnn_embedder(obs["x"], obs["edge_index"], obs.get("batch"), sd, num_layers, num_iter,
                       num_graphs=obs.get("num_graphs"))
    mask = obs["mask"]
    y = mlp(torch.cat([emb, mask], 1), sd, "actor.")
    return y * mask


def critic_forward(sd, obs, a, num_layers=2, num_iter=2):
    """reference QNetwork.forward (models.py:86-92): MLP(cat[emb, mask, a])."""
    emb = gnn_embedder(obs["x"], obs["edge_index"], obs.get("batch"), sd, num_layers, num_iter,
                       num_graphs=obs.get("num_graphs"))
    return mlp(torch.cat([emb, obs["mask"], a], 1), sd, "critic.")


# --------------------------------------------------------------------------------------------
# DDPG update (simple_ddpg.py:204-234), torch autograd + torch.optim.Adam like the reference
# --------------------------------------------------------------------------------------------
class OracleDDPG:
    def __init__(self, actor_sd, critic_sd, lr=1e-3, gamma=0.99, tau=1e-4, num_layers=2, num_iter=2):
        cl = lambda sd, g: {k: v.clone().requires_grad_(g) for k, v in sd.items()}
        self.actor, self.critic = cl(actor_sd, True), cl(critic_sd, True)
        self.actor_t, self.critic_t = cl(actor_sd, False), cl(critic_sd, False)
        self.q_opt = torch.optim.Adam(list(self.critic.values()), lr=lr, foreach=False)
        self.a_opt = torch.optim.Adam(list(self.actor.values()), lr=lr, foreach=False)
        self.gamma, self.tau, self.nl, self.ni = gamma, tau, num_layers, num_iter

    def update(self, cur, nxt, actions, rewards, dones, world_grads=None):
        nl, ni = self.nl, self.ni
        with torch.no_grad():                                        # simple_ddpg.py:205-209
            a2 = actor_forward(self.actor_t, nxt, nl, ni).clamp(-1, 1)
            q2 = critic_forward(self.critic_t, nxt, a2, nl, ni)
            y = rewards.flatten() + (1 - dones.flatten()) * self.gamma * q2.view(-1)
        q = critic_forward(self.critic, cur, actions, nl, ni).view(-1)  # :211
        q_loss = F.mse_loss(q, y)
        self.q_opt.zero_grad()
        q_loss.backward()
        self.last_critic_grads = {k: v.grad.clone() for k, v in self.critic.items()}
        self.q_opt.step()
        a_loss = -critic_forward(self.critic, cur, actor_forward(self.actor, cur, nl, ni), nl, ni).mean()
        self.a_opt.zero_grad()                                       # :222-225
        a_loss.backward()
        self.last_actor_grads = {k: v.grad.clone() for k, v in self.actor.items()}
        self.a_opt.step()
        with torch.no_grad():                                        # :229-234
            for sd, sdt in ((self.actor, self.actor_t), (self.critic, self.critic_t)):
                for k in sd:
                    sdt[k].copy_(self.tau * sd[k] + (1 - self.tau) * sdt[k])
        return q.detach(), q_loss.detach(), a_loss.detach()


# --------------------------------------------------------------------------------------------
# Action post-processing (simple_ddpg.py:338-395)
#
# The reference runs these loops on np.float32 scalars under its pinned numpy 1.23.4 and
# Python <= 3.11 (torch 1.13.1 has no 3.12 wheels).  There, a np.float32 scalar combined with a
# Python int/float or a np.float64 scalar promotes to np.float64 (pre-NEP-50 scalar promotion),
# and builtin ``sum`` adds left to right without compensation.  numpy 2.x and Python 3.12 changed
# both (weak Python scalars; Neumaier ``sum`` for exact floats), so this restatement widens the
# inputs to np.float64 explicitly and sums with a plain loop: the arithmetic the reference
# executes, independent of the interpreter/numpy this oracle runs on.
# --------------------------------------------------------------------------------------------
SCHED_ACCURACY = np.sqrt(np.finfo(np.float64).eps)                   # simple_ddpg.py:131
SCHED_THRESHOLD = 0.1                                                # simple_ddpg.py:130


def _seq_sum(values):
    """Python <= 3.11 ``sum``: start at int 0, add left to right."""
    total = 0
    for v in values:
        total = total + v
    return total


def normalize_scheduling_probabilities(input_list: list) -> list:
    output_list: list = []
    if len(input_list) == 0:
        return output_list
    total = _seq_sum(input_list)
    offset = 1 - total
    if total == 0:
        output_list = [round(1 / len(input_list), 10)] * len(input_list)
    elif abs(offset) > SCHED_ACCURACY:
        output_list = [round(prob / total, 10) for prob in input_list]
    else:
        output_list = list(input_list)
    new_offset = 1 - _seq_sum(output_list)
    if new_offset != 0:
        i = 0
        while output_list[i] + new_offset < 0:
            i += 1
        output_list[i] = output_list[i] + new_offset
    return output_list


def post_process_actions(action: np.ndarray, num_nodes: int) -> np.ndarray:
    """action: flat fp32 [G * num_nodes]; groups of ``num_nodes`` destinations (simple_ddpg.py:374-395).

    Returns fp64 (what ``np.array(processed_action)`` yields in the reference whenever any group
    was touched)."""
    assert action.shape[0] % num_nodes == 0
    out: list = []
    for g in range(action.shape[0] // num_nodes):
        probs = [np.float64(p) for p in action[g * num_nodes:(g + 1) * num_nodes]]
        for _ in range(2):
            rounded = [p if p >= SCHED_THRESHOLD else 0 for p in probs]
            probs = [np.float64(p) for p in normalize_scheduling_probabilities(rounded)]
        out.extend(probs)
    return np.array(out, dtype=np.float64)


def explore_action(y: np.ndarray, noise: np.ndarray, low=0.0, high=1.0) -> np.ndarray:
    """simple_ddpg.py:190-199 with the Gaussian sample passed in.  fp32 array math; the fp64 noise
    is added in fp64 and the in-place ``+=`` rounds back to fp32.  The ``.clip(-1, 1)`` at :197 is
    a no-op in the reference (its result is discarded)."""
    y = y.astype(np.float32)
    low32, high32 = np.float32(low), np.float32(high)
    scaled = np.float32(2.0) * ((y - low32) / (high32 - low32)) - np.float32(1.0)
    scaled = (scaled.astype(np.float64) + noise.astype(np.float64)).astype(np.float32)
    a = low32 + (np.float32(0.5) * (scaled + np.float32(1.0)) * (high32 - low32))
    return a.clip(low32, high32).astype(np.float32)


def node_features(ingress, usage, cap, ptr) -> np.ndarray:
    """Node features of a batch of states: reference simulator_wrapper.py:255-294 per state.

    The reference computes with float32 arrays under its pinned numpy 1.23.4: ``np.max(arr) + 1e-3``
    promotes the float32 scalar to float64, and ``arr / that`` casts the scalar back to float32
    (value-based casting); ``usage / cap`` is a Python-float division stored into a float32 array.
    Written out explicitly so that the result does not depend on the numpy this file runs on."""
    ingress, usage, cap = (np.asarray(a, dtype=np.float32) for a in (ingress, usage, cap))
    x = np.zeros((ingress.shape[0], 3), dtype=np.float32)
    for g in range(len(ptr) - 1):
        lo, hi = int(ptr[g]), int(ptr[g + 1])
        if hi <= lo:
            continue
        util = np.ones(hi - lo, dtype=np.float32)
        nz = cap[lo:hi] != 0
        util[nz] = (usage[lo:hi][nz].astype(np.float64) / cap[lo:hi][nz].astype(np.float64)).astype(np.float32)
        for k, arr in enumerate((ingress[lo:hi], util, cap[lo:hi])):
            denom = np.float32(np.float64(arr.max()) + 1e-3)
            x[lo:hi, k] = np.clip(arr / denom, np.float32(0), np.float32(1))
    return x


def obs_to(obs: dict, dtype) -> dict:
    out = dict(obs)
    for k in ("x", "mask"):
        out[k] = obs[k].to(dtype)
    return out
