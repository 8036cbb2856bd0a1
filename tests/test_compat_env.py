"""The reference's own environment (gym env, coordsim SimPy simulator, observation builder) on the import
stand-ins of ``compat/shims`` -- simpy core, gym spaces/wrappers, geopy distance, the torch-geometric
containers -- with no agent in the loop.  CPU only; reads the reference tree, so it is skipped where
that does not exist (the GPU box)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("GSC_REFERENCE", "/root/reference")

pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "main.py")), reason="reference tree not present")

SCRIPT = r"""
import os, sys, shutil, tempfile
root, ref = sys.argv[1], sys.argv[2]
work = tempfile.mkdtemp()
shutil.copytree(os.path.join(ref, "configs"), os.path.join(work, "configs"))
shutil.copy(os.path.join(ref, "logging.conf"), work)
os.chdir(work)
net = "configs/networks/2node/2node-in1-cap1-delay10.graphml"
open("sched.yaml", "w").write(f"training_network_files:\n  - {net}\ninference_network: {net}\nperiod: 10\n")
sys.path.insert(0, root)
from compat import pycompat
pycompat.install(ref, os.path.join(root, "compat", "shims"), root)
import numpy as np, torch
from src.rlsp.agents.main import DATETIME, setup
from src.rlsp.agents import simple_ddpg as sd
from oracle import gsc_oracle as O
ah = setup("configs/config/agent/sample_agent.yaml", net, "configs/service_functions/abc.yaml",
           "configs/config/simulator/sample_config.yaml", "sched.yaml", 1234, 2, None, False, DATETIME,
           None, True, None, None, None)
env = sd.simple_make_env(ah)
env.action_space.seed(7)
obs, _ = env.reset(seed=1234)
n_max = ah.env.env_limits.MAX_NODE_COUNT
assert type(obs).__name__ == "GraphData" and tuple(obs.x.shape) == (2, 3)
assert obs.edge_index.tolist() == [[0, 1], [1, 0]]                 # both directions, source-major
assert tuple(obs.mask.shape) == (1, n_max * 3 * n_max) and float(obs.mask.sum()) == 2 * 3 * 2
assert env.action_space.shape == (n_max * 3 * n_max,)
rows = list(obs)                                                   # what the reference buffer stores per state
assert [k for k, _ in rows] == ["x", "edge_index", "mask"]
total, ends = 0.0, []
for t in range(205):
    a = env.action_space.sample() * obs.mask.numpy().squeeze()
    a = O.post_process_actions(a.astype(np.float32), n_max)
    obs, r, done, trunc, info = env.step(a)
    total += float(r)
    if done:
        ends.append((t, sorted(info), int(info["episode"]["l"])))
assert ends == [(199, ["episode", "final_info", "final_observation"], 200)], ends
assert float(obs.x.min()) >= 0.0 and float(obs.x.max()) <= 1.0
print("RESULT", repr(total))
shutil.rmtree(work, ignore_errors=True)
"""


def _run():
    out = subprocess.run([sys.executable, "-c", SCRIPT, ROOT, REF], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("RESULT")]
    assert line, out.stdout[-2000:]
    return float(line[-1].split(" ", 1)[1])


def test_reference_environment_runs_on_the_shims_and_is_seed_deterministic():
    a = _run()
    b = _run()
    assert a == b and a < 0
