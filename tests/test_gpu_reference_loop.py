"""The reference's own training loop and inference script, unmodified, on gsc_b200 (SURVEY.md section 8,
rows b and f-4): ``main.py`` -> ``SimpleDDPG.train`` (its ``_choose_action``, ``post_process_actions``,
``GraphReplayBuffer``, ``_update_critic`` / ``_update_actor`` / ``_update_target_networks`` and two
``optim.Adam``) with ``src/rlsp/agents/models.py`` replaced by gsc_b200's modules, then ``th.save`` /
``inference.py``.  Runs from the staged copy of the reference (``baseline/_ref/GSC``, see
``__graft_entry__.stage_reference``); skipped where no copy exists."""
import os
import re
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_root():
    for cand in (os.environ.get("GSC_REFERENCE"), os.path.join(ROOT, "baseline", "_ref", "GSC"), "/root/reference"):
        if cand and os.path.exists(os.path.join(cand, "main.py")):
            return cand
    return None


@pytest.mark.skipif(_reference_root() is None, reason="no copy of the reference tree")
def test_reference_main_and_inference_run_on_gsc_b200(tmp_path):
    # two episodes of 40 steps: one warm-up episode, then an episode followed by 40 gradient steps at the
    # reference's minibatch of 100; the appended test rollout and inference.py use the trained actor
    cmd = [sys.executable, os.path.join(ROOT, "compat", "run_reference.py"), "--reference", _reference_root(),
           "--workdir", str(tmp_path / "run"), "--episodes", "2", "--episode-steps", "40"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=1200)
    log = out.stdout + out.stderr
    assert out.returncode == 0, log[-4000:]
    m = re.search(r"reference main.py finished: 2 episodes in [\d.]+ s, (\d+) gsc_b200 kernel launches", log)
    assert m and int(m.group(1)) > 40 * 30, log[-2000:]          # every gradient step runs tens of our kernels
    assert "saved actor:" in log
    m = re.search(r"reference inference.py finished in [\d.]+ s, (\d+) gsc_b200 kernel launches", log)
    assert m and int(m.group(1)) >= 40, log[-2000:]               # one actor forward per environment step
