"""Interpreter / library version gaps between the reference's pinned stack (Python <= 3.11, numpy 1.23,
torch 1.13) and this image (Python 3.12, numpy 2.3, torch 2.11), closed without touching the reference."""
import functools
import sys


def install(reference_root: str, shims_dir: str, repo_root: str) -> None:
    for p in (repo_root, reference_root, shims_dir):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    # coordsim/*/__init__.py import their sibling modules through FileFinder.find_module(...).load_module(...),
    # which Python 3.12 removed
    from importlib.machinery import FileFinder
    if not hasattr(FileFinder, "find_module"):
        def find_module(self, fullname, path=None):
            spec = self.find_spec(fullname)
            return None if spec is None else spec.loader
        FileFinder.find_module = find_module
    # random.seed(numpy integer): hashed to the same int by Python <= 3.10, a TypeError since 3.11
    # (reference siminterface/simulator.py:112 seeds with the np.int32 drawn in gym_env.py:142)
    import random
    if not getattr(random.seed, "_gsc_compat", False):
        orig_seed = random.seed

        def seed(a=None, version=2):
            if a is not None and not isinstance(a, (int, float, str, bytes, bytearray)) and hasattr(a, "__index__"):
                a = int(a)
            return orig_seed(a, version)
        seed._gsc_compat = True
        random.seed = seed
    # inference.py loads a whole pickled module (th.save(actor)); torch >= 2.6 refuses that by default
    import torch
    if not getattr(torch.load, "_gsc_compat", False):
        orig = torch.load

        @functools.wraps(orig)
        def load(*a, **k):
            k.setdefault("weights_only", False)
            return orig(*a, **k)
        load._gsc_compat = True
        torch.load = load
    # the module that takes the place of src/rlsp/agents/models.py
    import compat.reference_models as rm
    sys.modules["src.rlsp.agents.models"] = rm


def restore_numpy1_scalar_promotion() -> None:
    """``SimpleDDPG.post_process_actions`` (reference src/rlsp/agents/simple_ddpg.py:338-395) walks a float32
    action vector element by element.  Under the reference's pinned numpy 1.23 every such float32 scalar is
    widened to float64 by the first Python number it meets (``0 + p``, ``p >= 0.1``, ``1 - sum``), and the
    function's own assertion ``abs(1 - sum) < 1.5e-8`` relies on that.  numpy >= 2 (NEP 50) keeps float32 and
    the assertion fails inside the reference.  Handing the function the same values as float64 reproduces
    the pinned behaviour exactly (widening float32 -> float64 is exact, and every later operation is then
    the float64 operation numpy 1.23 performed).  Call after ``src.rlsp.agents.simple_ddpg`` is importable."""
    import numpy as np
    if int(np.__version__.split(".")[0]) < 2:
        return
    from src.rlsp.agents import simple_ddpg as sd
    orig = sd.SimpleDDPG.post_process_actions
    if getattr(orig, "_gsc_compat", False):
        return

    def post_process_actions(self, action):
        return orig(self, np.asarray(action, dtype=np.float64))
    post_process_actions._gsc_compat = True
    sd.SimpleDDPG.post_process_actions = post_process_actions
