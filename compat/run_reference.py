#!/usr/bin/env python
"""Run the reference's own ``main.py`` (training + appended test rollout) and ``inference.py`` on gsc_b200.

  python compat/run_reference.py --reference /root/reference [--episodes 2] [--network configs/networks/2node/...]

The reference tree is used as it is.  What this script does around it:
  * puts ``compat/shims`` (stand-ins for simpy, gym, geopy, stable-baselines3 and the torch-geometric
    containers, none of which is installable here) and the reference root on ``sys.path``;
  * registers ``compat.reference_models`` as ``src.rlsp.agents.models`` -- the effect of replacing that
    file by the one-line import of INTEGRATION.md section 3;
  * works in a scratch copy of the reference's ``configs/`` (the reference writes ``results/`` and ``runs/``
    into its working directory), with a scheduler yaml that points at the chosen topology and, when asked,
    the agent yaml's ``episode_steps`` shortened;
  * closes interpreter / library version gaps (``compat/pycompat.py``): ``torch.load`` unpickles whole
    modules as torch 1.13 did (``inference.py`` loads ``th.save(actor)``), Python 3.12 and numpy 2 details;
  * links the fresh result directory to the path hard-coded in ``inference.py:17`` before executing it.
Needs a CUDA device (gsc_b200 has no CPU path)."""
import argparse
import glob
import os
import shutil
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def prepare(reference: str):
    sys.path.insert(0, ROOT)
    from compat import pycompat
    pycompat.install(reference, os.path.join(HERE, "shims"), ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default=os.environ.get("GSC_REFERENCE", "/root/reference"))
    ap.add_argument("--workdir", default=None)
    ap.add_argument("--episodes", type=int, default=2)
    ap.add_argument("--episode-steps", type=int, default=0, help="override episode_steps of the agent yaml (0: keep 200)")
    ap.add_argument("--network", default="configs/networks/2node/2node-in1-cap1-delay10.graphml")
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--skip-inference", action="store_true")
    ap.add_argument("--inference-only", metavar="RESULT_DIR", default=None,
                    help="only run the reference's inference.py on an existing result directory")
    args = ap.parse_args()
    reference = os.path.abspath(args.reference)
    if args.inference_only:
        return run_inference(reference, os.path.abspath(args.workdir), os.path.abspath(args.inference_only))
    if not os.path.exists(os.path.join(reference, "main.py")):
        raise SystemExit(f"{reference} does not hold the reference tree (main.py not found)")
    work = os.path.abspath(args.workdir or os.path.join(ROOT, "gpurun_out", "reference_run"))
    shutil.rmtree(work, ignore_errors=True)
    os.makedirs(work)
    shutil.copytree(os.path.join(reference, "configs"), os.path.join(work, "configs"))
    shutil.copy(os.path.join(reference, "logging.conf"), work)
    os.chdir(work)
    import yaml
    sched = "configs/config/scheduler_compat.yaml"
    with open(sched, "w") as f:
        yaml.safe_dump({"training_network_files": [args.network], "inference_network": args.network, "period": 10}, f)
    agent_yaml = "configs/config/agent/sample_agent.yaml"
    if args.episode_steps:
        with open(agent_yaml) as f:
            cfg = yaml.safe_load(f)
        cfg["episode_steps"] = args.episode_steps
        cfg["nb_steps_warmup_critic"] = min(cfg.get("nb_steps_warmup_critic", 200), args.episode_steps)
        agent_yaml = "configs/config/agent/sample_agent_compat.yaml"
        with open(agent_yaml, "w") as f:
            yaml.safe_dump(cfg, f)
    prepare(reference)
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("compat/run_reference.py needs a CUDA device: gsc_b200 has no CPU path")
    from gsc_b200 import ops
    import main as ref_main                        # the reference's root main.py
    from compat import pycompat
    pycompat.restore_numpy1_scalar_promotion()
    t0 = time.time()
    launches0 = ops.launch_count
    ref_main.cli.main([agent_yaml, args.network, "configs/service_functions/abc.yaml",
                       "configs/config/simulator/sample_config.yaml", sched, str(args.episodes),
                       "--seed", str(args.seed), "--append-test"], standalone_mode=False)
    torch.cuda.synchronize()
    print(f"[compat] reference main.py finished: {args.episodes} episodes in {time.time() - t0:.1f} s, "
          f"{ops.launch_count - launches0} gsc_b200 kernel launches", flush=True)
    runs = sorted(glob.glob(os.path.join(work, "results", "**", "trained_actor.pt"), recursive=True))
    assert runs, "main.py did not save trained_actor.pt"
    result_dir = os.path.dirname(runs[-1])
    print(f"[compat] saved actor: {os.path.relpath(runs[-1], work)} ({os.path.getsize(runs[-1])} bytes)", flush=True)
    if args.skip_inference:
        return
    # inference.py in a process of its own, as a user would run it (the training process has hidden the GPUs
    # from later CUDA queries: reference src/rlsp/agents/main.py:242-243 sets CUDA_VISIBLE_DEVICES="")
    import subprocess
    env = {k: v for k, v in os.environ.items() if k != "CUDA_VISIBLE_DEVICES"}
    rc = subprocess.call([sys.executable, os.path.abspath(__file__), "--reference", reference, "--workdir", work,
                          "--inference-only", result_dir], env=env)
    if rc:
        raise SystemExit(rc)


def run_inference(reference: str, work: str, result_dir: str):
    """Execute the reference's inference.py against ``result_dir`` (the directory main.py wrote)."""
    import re
    import runpy
    os.chdir(work)
    prepare(reference)
    import torch
    from gsc_b200 import ops
    # inference.py reads a directory name that is hard-coded in the script (inference.py:17)
    src = open(os.path.join(reference, "inference.py")).read()
    hard = re.search(r'^dir = "([^"]+)"', src, re.M).group(1)
    os.makedirs(os.path.dirname(hard), exist_ok=True)
    if os.path.lexists(hard):
        os.remove(hard)
    os.symlink(result_dir, hard)
    t0 = time.time()
    launches0 = ops.launch_count
    runpy.run_path(os.path.join(reference, "inference.py"), run_name="__main__")
    torch.cuda.synchronize()
    print(f"[compat] reference inference.py finished in {time.time() - t0:.1f} s, "
          f"{ops.launch_count - launches0} gsc_b200 kernel launches", flush=True)


if __name__ == "__main__":
    main()
