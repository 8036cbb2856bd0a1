import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def golden_ddpg():
    return dict(np.load(os.path.join(GOLDEN, "ddpg_small.npz")))


@pytest.fixture(scope="session")
def golden_head():
    return dict(np.load(os.path.join(GOLDEN, "actor_head.npz")))


def sd_from_golden(g, prefix, dtype=torch.float32):
    return {k[len(prefix):]: torch.from_numpy(v).to(dtype) for k, v in g.items() if k.startswith(prefix)}


def obs_from_states(st, dtype=torch.float32, x=None):
    return {"x": torch.from_numpy(st.x if x is None else x).to(dtype),
            "edge_index": torch.from_numpy(st.edge_index), "batch": torch.from_numpy(st.batch),
            "mask": torch.from_numpy(st.mask).to(dtype), "num_graphs": st.num_graphs}
