#!/usr/bin/env python
"""Benchmark of the GSC hot path on B200 (contract: see the task statement / DESIGN.md section 5).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5] [--batch B]
  python bench.py --impl reference ...     # the CPU oracle on the box's host cores, same workload

A "step" is one pass of the hot path over one batch of synthetic network states:
  c2  batch packer + GNN embedder + actor MLP inference  (B = 4096 states, N_max = 24, n ~ U{16..24})
  c3  full DDPG update: packer, 5 forward + 3 backward passes, 2 fused Adam/soft-update launches
      (N_max = n = 100, |E| = 150, replay minibatch of B = 1024 states per GPU)      [default]
  c4  the same update on 1000-node topologies (A = 3 000 000, 1.92 G parameters), B = 32 states per
      GPU; for N > 1 the optimizer is sharded (reduce-scatter, Adam on 1/N, all-gather)
  c5  inference on 65 536 states of the largest bundled topology size (N_max = n = 110, |E| = 148)
For N > 1 every rank works on its own B states (weak scaling); the training workloads exchange the
flat gradient buffers over NCCL, the inference workloads have no collective (replicas).  One JSON
line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "graph-states/sec GNN+actor/critic fwd+bwd at 1/2/4/8 B200; % HBM roofline"
UNIT = "graph-states/s"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------
# name -> (kind, N_max, n_low, n_high, links per state (None: round(1.5 n)), default states per GPU)
# kind "c2" = inference step, "c3" = DDPG update step (the two step shapes of the hot path).
WORKLOADS = {"c2": ("c2", 24, 16, 24, None, 4096), "c3": ("c3", 100, 100, 100, 150, 1024),
             "c4": ("c3", 1000, 1000, 1000, 1500, 32), "c5": ("c2", 110, 110, 110, 148, 65536)}


def make_workload(args, rank: int, max_states: int = 0):
    """``max_states``: generate only that many states (the CPU arm's bounded sample of a large workload)."""
    from gsc_b200 import synth
    seed = 1234 + 1000 * rank
    kind, n_max, lo, hi, links, B = WORKLOADS[args.workload]
    B = args.batch or B
    G = min(B, max_states) if max_states else B
    base = dict(kind=kind, name=args.workload, B=B, n_max=n_max, n_low=lo, n_high=hi, links=links)
    if kind == "c2":
        st = synth.make_states(G, n_max, lo, hi, seed=seed, fixed_links=links,
                               num_topologies=None if args.workload == "c2" else 24)
        return dict(base, cur=st)
    cur, nxt, actions, rewards, dones = synth.make_transitions(G, n_max, lo, hi, seed=seed, fixed_links=links,
                                                               num_topologies=32 if n_max <= 100 else 4)
    return dict(base, cur=cur, nxt=nxt, actions=actions, rewards=rewards, dones=dones)


def workload_name(w):
    size = (f"N_max={w['n_max']}, n~U{{{w['n_low']}..{w['n_high']}}}" if w["n_low"] != w["n_high"]
            else f"N_max=n={w['n_max']}")
    links = "|E|=round(1.5n)" if w["links"] is None else f"|E|={w['links']}"
    what = "batch packer + GNN embedder + actor inference" if w["kind"] == "c2" else "full DDPG update (fwd+bwd+Adam+target)"
    return f"{w['name']}: {what}, B={w['B']}/GPU, {size}, {links}"


def make_config(w):
    """The workload description; identical in the GPU arm and in the --impl reference arm."""
    return {"workload": workload_name(w), "states_per_gpu": w["B"], "n_max": w["n_max"],
            "action_dim": w["n_max"] * w["n_max"] * 3, "gnn": "3 features -> 22 hidden, 2 layers x 2 iterations, mean aggregation",
            "weights": "reference init (PyG glorot / torch Linear), seeded", "seed": 1234}


def algorithmic_bytes(w, Pa: int, Pc: int, opt_div: int = 1) -> float:
    """SURVEY.md section 8d formulas with the actual n, e, A, B of this batch (``opt_div``: the
    optimizer's share of the 48 B/param is divided by the world size when it is sharded)."""
    st = w["cur"]
    Nt, E, B, A = st.x.shape[0], st.edge_index.shape[1], st.num_graphs, st.action_dim
    graph = 4.0 * Nt * 3 + 16.0 * E + 8.0 * Nt          # 4nF + 16e (two int64 per directed edge) + 8n (batch)
    if w["kind"] == "c2":
        return graph + 4.0 * A * B * 2 + 4.0 * Pa
    return 2 * (graph + 4.0 * A * B) + (4.0 * A + 8) * B + (16.0 + 32.0 / opt_div) * (Pa + Pc)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU through NVML on a background thread while
    the timed region runs (B200_PROFILING.md: the clocks line)."""

    def __init__(self, index: int, period_s: float = 0.005):
        import threading
        self.samples, self.max_mhz, self.reasons, self.err = [], None, set(), None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:          # no NVML: report it, never fail the bench
            self.nv, self.err = None, repr(e)
            return
        self.period = period_s
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def _run(self):
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nme in names.items():
                    if r & bit:
                        self.reasons.add(nme)
            except Exception as e:
                self.err = repr(e)
                break
            self._stop.wait(self.period)

    def stop(self):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [f"nvml unavailable: {self.err}"], "samples": 0}
        self._stop.set()
        self.th.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    from gsc_b200 import ddpg, ops
    from gsc_b200.graph import GraphBatch
    from gsc_b200.models import Actor, AgentSpec, QNetwork

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gsc_b200 has no CPU fallback (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    shard_opt = (args.shard_optimizer if args.shard_optimizer is not None else args.workload == "c4") and world > 1
    w = make_workload(args, rank)
    B, n_max = w["B"], w["n_max"]
    ops.precision = args.precision
    spec = AgentSpec(n_max)
    torch.manual_seed(1234)
    actor = Actor(spec).to(dev)
    Pa = sum(p.numel() for p in actor.parameters())
    Pc = 0
    upd = None
    if w["kind"] == "c3":
        qf1 = QNetwork(spec).to(dev)
        Pc = sum(p.numel() for p in qf1.parameters())
        ta, tq = Actor(spec).to(dev), QNetwork(spec).to(dev)
        ta.load_state_dict(actor.state_dict())
        tq.load_state_dict(qf1.state_dict())
        upd = ddpg.DDPGUpdater(actor, qf1, ta, tq, lr=1e-3, gamma=0.99, tau=1e-4,
                               world_size=world, rank=rank, shard_optimizer=shard_opt,
                               fork_streams=args.fork, sm_reserve=args.sm_reserve,
                               bucket_allreduce=args.bucket_allreduce, split_optimizer_step=args.split_step)

    def host_batch(st):
        return GraphBatch.from_states(st, pin=True, dense_mask=args.dense_mask)

    h_cur = host_batch(w["cur"])
    host = {"cur": h_cur}
    if w["kind"] == "c3":
        host["nxt"] = host_batch(w["nxt"])
        host["nxt"].edge_index, host["nxt"].batch, host["nxt"]._mask = h_cur.edge_index, h_cur.batch, h_cur._mask
        for k in ("actions", "rewards", "dones"):
            host[k] = torch.from_numpy(w[k]).pin_memory()

    def to_dev(hb, share=None):
        mv = lambda t: t.to(dev, non_blocking=True)
        mo = lambda t: None if t is None else mv(t)
        if share is None:
            return GraphBatch(mv(hb.x), mv(hb.edge_index), mv(hb.batch), mo(hb._mask), hb.num_graphs, hb.max_nodes,
                              None, None if args.dense_mask else mv(hb.n_nodes), hb.n_max, hb.groups)
        return GraphBatch(mv(hb.x), share.edge_index, share.batch, share._mask, hb.num_graphs, hb.max_nodes,
                          None, share.n_nodes, hb.n_max, hb.groups)

    def h2d_all():
        d = {"cur": to_dev(host["cur"])}
        if w["kind"] == "c3":
            d["nxt"] = to_dev(host["nxt"], share=d["cur"])
            for k in ("actions", "rewards", "dones"):
                d[k] = host[k].to(dev, non_blocking=True)
        return d

    def fresh(gb):   # same device tensors, CSR not packed yet: the packer is part of every step
        return gb.view()

    def step(d):
        if w["kind"] == "c2":
            if gpredict is not None:            # static inputs already hold the batch
                return gpredict.replay()
            return ddpg.predict(actor, fresh(d["cur"]))
        if graphed is not None:                 # static inputs already hold the batch
            return graphed.replay()
        cur, nxt = fresh(d["cur"]), fresh(d["nxt"])
        return upd.update(cur, nxt, d["actions"], d["rewards"], d["dones"])

    graphed = None
    gpredict = None
    resident = h2d_all()
    torch.cuda.synchronize()
    if w["kind"] == "c2" and args.graph:
        gpredict = ddpg.GraphedPredict(actor, resident["cur"])
    use_graph = w["kind"] == "c3" and args.graph
    if use_graph:
        graphed = ddpg.GraphedUpdate(upd, resident["cur"], resident["nxt"], resident["actions"],
                                     resident["rewards"], resident["dones"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    out_host = torch.empty((B, w["cur"].action_dim), dtype=torch.float32).pin_memory() if w["kind"] == "c2" else \
        torch.empty(2, dtype=torch.float32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) ----
    for _ in range(args.warmup):
        flush.zero_()
        step(resident)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ops.timer = {}
    launches0 = ops.launch_count
    evs = []
    barrier()
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()                                               # L2 flush, outside the event pair
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step(resident)
        e1.record()
        evs.append((e0, e1))
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = (ops.launch_count - launches0) // max(args.steps, 1)
    timer, ops.timer = ops.timer, None
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    clocks = sampler.stop() if sampler else None
    op_steps = args.steps
    if graphed is not None or gpredict is not None:
        # A graph replay has no host code between kernels, so per-kernel CUDA events are taken on a
        # kernel-by-kernel pass of the same update right after the timed region (same inputs, L2
        # flushed); the kernel count is what the capture enqueued.
        op_steps = min(args.steps, 5)
        ops.timer = {}
        launches0 = ops.launch_count
        g_keep, graphed, p_keep, gpredict = graphed, None, gpredict, None
        for i_op in range(op_steps):
            flush.zero_()
            if args.timeline and i_op == op_steps - 1:
                ops.timeline = []
                tl0 = torch.cuda.Event(enable_timing=True)
                tl0.record()
            step(resident)
        graphed, gpredict = g_keep, p_keep
        barrier()
        launches = (ops.launch_count - launches0) // op_steps
        timer, ops.timer = ops.timer, None
    if args.timeline and ops.timeline and rank == 0:
        # where every entry point of the last kernel-by-kernel step ran: stream, start, end (ms from step start)
        torch.cuda.synchronize()
        streams = {}
        for name, a, b, sh in sorted(ops.timeline, key=lambda r: tl0.elapsed_time(r[1])):
            sid = streams.setdefault(sh, len(streams))
            print(f"[timeline] s{sid} {tl0.elapsed_time(a):8.3f} -> {tl0.elapsed_time(b):8.3f}  {name}", file=sys.stderr)
        ops.timeline = None
    per_op = {k: sum(a.elapsed_time(b) for a, b in v) / op_steps for k, v in timer.items()}
    per_op_n = {k: len(v) / op_steps for k, v in timer.items()}

    # ---- end-to-end timing through the public API with host buffers (e2e) ----
    # c2: the states (x, edge_index, batch, n) come from pinned host memory every step, the mask is
    #     generated on the device from n, the action matrix goes back to pinned host memory.
    # c3: the replay buffer lives in HBM (DeviceGraphReplayBuffer).  Every step one new transition
    #     arrives from pinned host memory (what an environment step produces), a minibatch is drawn
    #     with the reference's host RNG and gathered on the device, the update runs, and both losses
    #     are read back.
    from gsc_b200.buffer import BufferSample, DeviceGraphReplayBuffer
    from gsc_b200.graph import GraphData
    st0 = w["cur"]
    g_e2e_ref = []
    if w["kind"] == "c2":
        h_n = torch.from_numpy(st0.n.astype(np.int32)).pin_memory()
        e2e_in = [host["cur"].x, host["cur"].edge_index, host["cur"].batch, h_n]

        def e2e_step():
            if gpredict is not None and not args.dense_mask:   # H2D straight into the graph's static inputs
                s_gb = gpredict.gb
                s_gb.x.copy_(host["cur"].x, non_blocking=True)
                s_gb.edge_index.copy_(host["cur"].edge_index, non_blocking=True)
                s_gb.batch.copy_(host["cur"].batch, non_blocking=True)
                s_gb.n_nodes.copy_(h_n, non_blocking=True)
                out_host.copy_(gpredict.replay(), non_blocking=True)
                return
            mv = lambda t_: t_.to(dev, non_blocking=True)
            n_dev = mv(h_n)
            gb = GraphBatch(mv(host["cur"].x), mv(host["cur"].edge_index), mv(host["cur"].batch),
                            ops.build_mask(n_dev, n_max, 3) if args.dense_mask else None, B,
                            host["cur"].max_nodes, None, n_dev, n_max, 3)
            out_host.copy_(ddpg.predict(actor, gb), non_blocking=True)
        h2d = sum(t_.numel() * t_.element_size() for t_ in e2e_in)
    else:
        cap = max(2 * B, 2048 if n_max <= 100 else 64)      # the action rows are 4A bytes each (12 MB at N_max = 1000)
        rb = DeviceGraphReplayBuffer(cap, spec.env.action_space, device=dev, n_max=n_max, e_max=2 * (w["links"] or 2 * n_max) + 4)
        for _ in range(cap // B - 1):                                   # leave room for the streamed adds
            rb.bulk_load(w["cur"], w["nxt"], w["actions"], w["rewards"], w["dones"])
        g0 = st0.graph(0)
        new_obs = GraphData(torch.from_numpy(g0["x"]).pin_memory(), torch.from_numpy(g0["edge_index"]).pin_memory(), None)
        new_act = torch.from_numpy(w["actions"][0]).pin_memory()
        static = None
        g_e2e = None
        if graphed is not None:
            static = BufferSample(graphed.cur, graphed.actions, graphed.nxt, graphed.dones, graphed.rewards)
            if not args.dense_mask:
                np.random.seed(99 + rank)
                smp0 = rb.sample(B)
                c0, n0 = smp0.observations, smp0.next_observations
                # an environment step does not change the network: the next state reuses the topology
                # (and so the packed CSR) of the current one, as in the device-resident timing above
                n0 = GraphBatch(n0.x, c0.edge_index, c0.batch, None, n0.num_graphs, c0.max_nodes, None,
                                c0.n_nodes, c0.n_max, c0.groups)
                g_e2e = ddpg.GraphedUpdate(upd, c0, n0, smp0.actions, smp0.rewards, smp0.dones, replay=rb)
                g_e2e_ref.append(g_e2e)

        def e2e_step():
            if rb.pos + 1 >= rb.buffer_size:
                rb.pos = rb.buffer_size - B                              # keep cycling inside the buffer
            rb.add(new_obs, new_obs, new_act, 0.5, False)
            if g_e2e is not None:            # index draw on the host, gather + update in one graph launch
                r = g_e2e.replay_sampled()
            elif static is not None:
                rb.sample(B, out=static)
                r = graphed.replay()
            else:
                smp = rb.sample(B)
                r = upd.update(smp.observations, smp.next_observations, smp.actions, smp.rewards, smp.dones)
            out_host[0:1].copy_(r[1], non_blocking=True)
            out_host[1:2].copy_(r[2], non_blocking=True)
        h2d = 2 * (new_obs.x.numel() * 4 + new_obs.edge_index.numel() * 4) + new_act.numel() * 4 + 8 * B + 2 * 2 * (B + 1) * 4

    np.random.seed(1234 + rank)
    for _ in range(min(args.warmup, 3)):
        e2e_step()
    barrier()
    e2e_evs = []
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        e2e_step()
        e1.record()
        e2e_evs.append((e0, e1))
    barrier()
    e2e_ms = sum(a.elapsed_time(b) for a, b in e2e_evs)

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])

    parity = dp_parity(upd, dev, rank, world, args, shard_opt) if (world > 1 and upd is not None) else None

    if rank == 0:
        hbm, which = measured_peaks()
        total_states = B * world * args.steps
        value = total_states / (dev_ms / 1e3)
        e2e_value = total_states / (e2e_ms / 1e3)
        d2h = out_host.numel() * 4
        # Dominant kernel + roofline.  Every timed entry point (name[shape]) launches one kernel (plus, for
        # split-K GEMMs, a small finalize); the entry point with the largest total time in the step is
        # reported under its kernel's name: algorithmic bytes of its launches / its device time.  The
        # per-kernel times come from a kernel-by-kernel pass of the same step (serialised: no overlap
        # between streams), so their sum exceeds the graph-timed step; shares, not absolutes, compare.
        def kernel_of(name):
            base = name.split("[")[0]
            if base.startswith(("linear_", "gemm")):
                role = {"linear_fwd": "<K-major,K-major>", "linear_dgrad": "<K-major,MN-major>",
                        "linear_wgrad": "<MN-major,MN-major>"}.get(base, "")
                kern = {"tf32": "gemm_tc_kernel", "tf32x3": "gemm_tc_kernel", "fp32": "gemm_simt_kernel"}[args.precision]
                mode = {"tf32": "tcgen05 TF32", "tf32x3": "tcgen05 3xTF32", "fp32": "fp32 FMA"}[args.precision]
                return f"{kern}{role} {name[name.index('['):]} ({mode})"
            return {"gnn_bwd": "gnn_bwd_gather_kernel + gnn_wgrad_kernel" if w["n_max"] <= 128 else "gnn_bwd_tc_kernel / gnn_bwd_kernel",
                    "gnn_fwd": "gnn_fwd_tc_kernel", "adam_soft_update": "adam_soft_dev_kernel",
                    "pack_csr": "pack_csr kernels"}.get(base, base + "_kernel") + (name[name.index("["):] if "[" in name else "")
        top_op = max(per_op, key=per_op.get)
        top_ms = per_op[top_op]
        alg_b = op_bytes(top_op, w, Pa, Pc) * per_op_n[top_op]
        step_bytes = algorithmic_bytes(w, Pa, Pc, world if shard_opt else 1)
        fam = {}
        for k, ms in per_op.items():
            f = k.split("[")[0]
            fam[f] = fam.get(f, 0.0) + ms
        roof = {"bound": "hbm", "kernel": kernel_of(top_op), "achieved": alg_b / (top_ms / 1e3) / 1e9, "peak": hbm,
                "unit": "GB/s", "frac": alg_b / (top_ms / 1e3) / 1e9 / hbm,
                "traffic": profiled_traffic(w, args, top_op),
                "peak_source": which, "kernel_ms_per_step": top_ms, "kernel_launches_per_step": per_op_n[top_op],
                "kernel_share_of_step": top_ms / sum(per_op.values()),
                "algorithmic_bytes_per_launch": alg_b / per_op_n[top_op],
                "timing": "CUDA events around each entry point, kernel-by-kernel pass of the same step (serialised)",
                "step_algorithmic_bytes": step_bytes,
                "step_frac_of_hbm_roofline": step_bytes / (dev_ms / args.steps / 1e3) / 1e9 / hbm,
                "per_entry_point_family_ms": {k: round(v, 4) for k, v in sorted(fam.items(), key=lambda kv: -kv[1])}}
        # the CPU oracle is timed on rank 0 at N = 1 only (other ranks would spin in a barrier meanwhile)
        cpu = cpu_baseline(args, w, sample_states=min(B, {"c2": 4096, "c3": 256, "c4": 2, "c5": 1024}[w["name"]])) \
            if world == 1 else None
        par = (f"dp{world}" + ("+sharded-optimizer" if shard_opt else "")) if w["kind"] == "c3" else f"replicas{world}"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "f32", "tf32": "tf32 tensor-core GEMMs (fp32 accumulate), f32 elsewhere",
                      "tf32x3": "f32 (dense layers: 3xTF32 split on tcgen05, fp32 accumulate; fp32-faithful)"}[args.precision],
            "data": "synthetic",
            "config": dict(make_config(w), parallelism=par),
            "impl_config": {"precision": args.precision,
                            "streams": "3 forward chains forked" if (args.fork and w["kind"] == "c3") else "single",
                            "mask": "dense [B, A] tensor" if args.dense_mask else "structured: mask[:n,:,:,:n] from node counts",
                            "cuda_graph": bool(graphed is not None or gpredict is not None),
                            "l2": "flushed between timed steps (256 MiB write)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_ms / args.steps,
                    "path": ("pinned host states -> H2D -> mask from n on device -> packer+embedder+actor -> D2H actions"
                             if w["kind"] == "c2" else
                             "DeviceGraphReplayBuffer.add(1 transition from pinned host) -> sample(B): index draw with the "
                             "reference's host RNG, indices H2D, gather from device storage + DDPG update as one "
                             "CUDA-graph launch -> D2H losses")},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
            "per_op_ms": {k: round(v, 4) for k, v in sorted(per_op.items(), key=lambda kv: -kv[1])[:14]},
            "wall_ms_per_step_incl_flush": t_wall / args.steps * 1e3,
        }
        if parity is not None:
            line["dp_parity"] = parity
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        for g in (graphed, g_e2e_ref[0] if g_e2e_ref else None):
            if g is not None:
                g.close()            # a captured graph keeps the communicator alive: destroy it first
        dist.destroy_process_group()


def dp_parity(upd, dev, rank, world, args, shard_opt):
    """N > 1: (a) the replicas of the benchmarked networks are bit-identical after the run; (b) on a small
    problem (N_max = 24, 8 states per rank) three data-parallel updates in the benchmarked mode equal three
    single-GPU updates of the full minibatch (weights rel 2e-4, losses rel 1e-4: what
    tests/test_gpu_dp.py checks on a 2-GPU box)."""
    import torch.distributed as dist
    from gsc_b200 import ddpg, dp, synth
    from gsc_b200.graph import GraphBatch
    from gsc_b200.models import Actor, AgentSpec, QNetwork
    same = True
    for net in (upd.actor, upd.qf1, upd.target_actor, upd.qf1_target):
        other = net.flat.clone()
        dist.broadcast(other, src=0)
        same = same and bool(torch.equal(other, net.flat))
    flag = torch.tensor([1 if same else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    n_max, Bg = 24, 8 * world
    spec = AgentSpec(n_max)

    def build():
        torch.manual_seed(11)
        nets = [cls(spec) for cls in (Actor, QNetwork, Actor, QNetwork)]
        nets[2].load_state_dict(nets[0].state_dict())
        nets[3].load_state_dict(nets[1].state_dict())
        return [n.to(dev) for n in nets]
    full = synth.make_transitions(Bg, n_max, 12, n_max, seed=3, num_topologies=5)
    mine = dp.shard_transitions(*full, rank, world)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    def run(nets, data, **kw):
        u = ddpg.DDPGUpdater(*nets, lr=1e-3, gamma=0.99, tau=1e-2, fork_streams=args.fork, **kw)
        cur = GraphBatch.from_states(data[0], dense_mask=False).to(dev)
        nxt = GraphBatch.from_states(data[1], dense_mask=False).to(dev)
        for _ in range(3):
            _, ql, al = u.update(cur.view(), nxt.view(), t(data[2]), t(data[3]), t(data[4]))
        return torch.stack([ql.reshape(()), al.reshape(())])
    nets_dp = build()
    losses = run(nets_dp, mine, world_size=world, rank=rank, shard_optimizer=shard_opt, global_batch=Bg,
                 bucket_allreduce=args.bucket_allreduce)
    dist.all_reduce(losses)
    out = {"replicas_bit_identical": bool(int(flag[0]))}
    if rank == 0:
        nets_1 = build()
        ref = run(nets_1, full)
        out["weights_rel_err_vs_single_gpu"] = max(float((a.flat - b.flat).abs().max() / b.flat.abs().max())
                                                   for a, b in zip(nets_dp, nets_1))
        out["losses_rel_err_vs_single_gpu"] = float((losses - ref).abs().max() / ref.abs().max())
        out["ok"] = bool(out["replicas_bit_identical"] and out["weights_rel_err_vs_single_gpu"] < 2e-4
                         and out["losses_rel_err_vs_single_gpu"] < 1e-4)
        out["what"] = f"3 updates, N_max=24, {Bg} states sharded over {world} ranks vs 1 GPU on the full minibatch"
    torch.cuda.synchronize()
    dist.barrier()
    return out


def profiled_traffic(w, args, op_name: str):
    """DRAM bytes (read + written) per launch of the kernel behind the dominant entry point, from the committed
    ncu --set full capture of this workload (profiles/r2_traffic.json: {workload_bB_precision: {kernel name:
    {"dram_bytes_per_launch": ...}}}, written from the raw page by tools/ncu_traffic.py; for the GEMM kernel
    the figure is the mean over the shapes of the step).  A pointer to the capture, not a measurement of this
    run; None when no capture of this workload/precision is committed."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r2_traffic.json")
    try:
        with open(path) as f:
            table = json.load(f)
    except (OSError, ValueError):
        return None
    base = op_name.split("[")[0]
    kernels = {"gnn_bwd": ["gnn_bwd_gather_kernel", "gnn_bwd_tc_kernel", "gnn_bwd_kernel"],
               "gnn_fwd": ["gnn_fwd_tc_kernel", "gnn_fwd_kernel"], "adam_soft_update": ["adam_soft_dev_kernel"],
               "linear_fwd": ["gemm_tc_kernel"], "linear_dgrad": ["gemm_tc_kernel"], "linear_wgrad": ["gemm_tc_kernel"]}
    entry = table.get(f"{w['name']}_b{w['B']}_{args.precision}", {})
    for k in kernels.get(base, [base + "_kernel"]):
        if k in entry:
            return entry[k].get("dram_bytes_per_launch")
    return None


def op_bytes(name: str, w, Pa, Pc) -> float:
    """Algorithmic bytes of one launch of the named entry point."""
    st = w["cur"]
    Nt, E, B = st.x.shape[0], st.edge_index.shape[1], st.num_graphs
    if name.startswith(("linear_", "gemm")):
        dims = name[name.index("[") + 1:name.index("]")]
        extra = dims.count("+")
        M, N, K = (int(v) for v in dims.split("+")[0].split("x"))
        return 4.0 * (M * K + K * N + M * N * (1 + extra))
    if name.startswith("gnn_fwd"):
        return 4.0 * Nt * 3 + 4.0 * (E + Nt) + 4.0 * (Nt + 1) + 4.0 * (B + 1) + 4.0 * B * 22
    if name.startswith("gnn_bwd"):
        return 4.0 * Nt * 3 + 4.0 * (E + Nt) + 4.0 * (Nt + 1) + 4.0 * (B + 1) + 4.0 * B * 22 + 3 * 4.0 * Nt * 24
    if name.startswith("adam"):
        n = int(name[name.index("[") + 1:name.index("]")])
        return 36.0 * n
    if name.startswith("pack_csr"):
        return 2 * 16.0 * E + 8.0 * Nt + 4.0 * (E + Nt) + 4.0 * Nt
    if name.startswith("mask_table"):      # reads the [N1, A] mask block of the weight, writes the table
        N1, A = (int(v) for v in name[name.index("[") + 1:name.index("]")].split("x"))
        return 4.0 * N1 * A + 4.0 * N1 * (w["n_max"] + 1)
    if name.startswith("mask_wgrad"):      # reads dY [B, N1] and n, writes the [N1, A] gradient block
        N1, A = (int(v) for v in name[name.index("[") + 1:name.index("]")].split("x"))
        return 4.0 * N1 * A + 4.0 * B * N1 + 4.0 * B
    if name.startswith("colsum"):
        M, N = (int(v) for v in name[name.index("[") + 1:name.index("]")].split("x"))
        return 4.0 * M * N
    return 0.0


# ------------------------------------------------------------------------------------------------
def oracle_step_fn(w, sample_states: int):
    """The CPU oracle on the first ``sample_states`` states of the workload (fp32, all host threads)."""
    from oracle import gsc_oracle as O
    from gsc_b200 import synth
    st = w["cur"]
    S = min(sample_states, st.num_graphs)
    nt = int(st.ptr[S])
    e_sel = st.edge_index[0] < nt

    def obs_of(s):
        return {"x": torch.from_numpy(s.x[:nt]), "edge_index": torch.from_numpy(np.ascontiguousarray(s.edge_index[:, e_sel])),
                "batch": torch.from_numpy(s.batch[:nt]), "mask": torch.from_numpy(s.mask[:S]), "num_graphs": S}
    A = st.action_dim
    a_sd = O.init_actor(3, 22, 2, A, seed=1234)
    if w["kind"] == "c2":
        o = obs_of(st)

        def fn():
            with torch.no_grad():
                return O.actor_forward(a_sd, o)
        return fn, S
    c_sd = O.init_critic(3, 22, 2, A, seed=4321)
    dd = O.OracleDDPG(a_sd, c_sd)
    o_cur, o_nxt = obs_of(w["cur"]), obs_of(w["nxt"])
    acts, rew, don = (torch.from_numpy(w[k][:S]) for k in ("actions", "rewards", "dones"))

    def fn():
        return dd.update(o_cur, o_nxt, acts, rew, don)
    return fn, S


def cpu_baseline(args, w, sample_states: int, reps: int = 3):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    fn, S = oracle_step_fn(w, sample_states)
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return {"value": S / statistics.median(ts), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{S} states of the same workload, fp32 oracle (plain PyTorch restatement of the PyG path), "
                      f"median of {reps} after 1 warm-up"}


def run_reference(args):
    """The reference's CPU path (the oracle port: PyG is not installable here) on the same workload, with
    exactly --steps timed steps after --warmup untimed ones.  c2/c3 run the full batch of one GPU per step;
    c4/c5 a bounded sample of it (stated in cpu_baseline.sample) so that the run ends within minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cap = {"c2": 0, "c3": 0, "c4": 2, "c5": 1024}[args.workload]
    w = make_workload(args, 0, max_states=cap)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    fn, S = oracle_step_fn(w, cap or w["B"])
    warm = max(0, args.warmup)
    for _ in range(warm):
        fn()
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = time.perf_counter() - t0
    value = S * steps / dt
    world = int(os.environ.get("WORLD_SIZE", "1"))
    par = f"dp{world}" if w["kind"] == "c3" else f"replicas{world}"
    if w["name"] == "c4" and world > 1:
        par += "+sharded-optimizer"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": dict(make_config(w), parallelism=par),
            "impl_config": {"note": "reference CPU path = oracle port (PyG 2.3.0 is not installable here); one process, "
                                    "all host threads; no GPU"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{S} of the {w['B']} states of one GPU's batch per step x {steps} steps"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("GSC_BENCH_WORKLOAD", "c3"), choices=sorted(WORKLOADS),
                    help="c3 (default): full DDPG update, the fwd+bwd+Adam step BASELINE.json's metric names; "
                         "c4: the same update on 1000-node topologies; c2 / c5: embedder + actor inference")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--bucket-allreduce", nargs="?", const=True, default=False, choices=[True, "mlp", "layer"],
                    help="N > 1: all-reduce the dense-layer gradient blocks on the side stream behind the embedder "
                         "backward: per layer (default when given) or as one block ('mlp')")
    ap.add_argument("--sm-reserve", type=int, default=8,
                    help="N > 1: SMs the embedder backward leaves free for the concurrent NCCL kernels")
    ap.add_argument("--split-step", dest="split_step", action="store_true",
                    help="step the dense layers on the side stream while the embedder backward runs (measured: "
                         "no gain, the embedder backward leaves no SM resources for it; default: one optimizer "
                         "launch per network after its backward)")
    ap.add_argument("--no-fork", dest="fork", action="store_false",
                    help="run the three independent forward chains of the update on one stream")
    ap.add_argument("--shard-optimizer", action=argparse.BooleanOptionalAction, default=None,
                    help="N > 1: reduce-scatter gradients, Adam + soft update on a 1/N parameter shard, all-gather "
                         "(default: on for c4, off otherwise)")
    ap.add_argument("--dense-mask", action="store_true",
                    help="carry the dense [B, A] mask tensor instead of the per-state node counts")
    ap.add_argument("--timeline", action="store_true",
                    help="print, per entry point of one kernel-by-kernel step, its stream and start/end time (stderr)")
    ap.add_argument("--no-graph", dest="graph", action="store_false",
                    help="c3: launch the update kernel by kernel instead of replaying its CUDA graph")
    ap.add_argument("--precision", default=os.environ.get("GSC_BENCH_PRECISION", "tf32x3"),
                    choices=["fp32", "tf32", "tf32x3"],
                    help="dense layers: tf32x3 (default) = tcgen05 tensor cores, fp32-faithful 3xTF32 split -- the mode "
                         "the parity tests run; tf32 = single-pass TF32 operands; fp32 = exact FMA kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
