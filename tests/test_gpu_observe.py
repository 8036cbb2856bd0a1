"""GPU test of the batched observation builder (gsc_node_features) against the oracle restatement
of reference simulator_wrapper.py:255-294: float32 results bit-exact."""
import numpy as np
import pytest
import torch

from gsc_b200 import ddpg, observe, ops, synth
from gsc_b200.models import Actor, AgentSpec
from oracle import gsc_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _raw(n, seed):
    rng = np.random.default_rng(seed)
    Nt = int(n.sum())
    ingress = np.where(rng.random(Nt) < 0.3, rng.random(Nt) * 40, 0).astype(np.float32)
    cap = rng.integers(0, 6, size=Nt).astype(np.float32) * np.float32(2.5)          # some nodes with cap == 0
    usage = (rng.random(Nt) * 12).astype(np.float32)
    return ingress, usage, cap


@pytest.mark.parametrize("B,lo,hi", [(1, 1, 1), (7, 2, 11), (300, 16, 24), (33, 900, 1000)])
def test_node_features_bit_exact(B, lo, hi):
    rng = np.random.default_rng(B)
    n = rng.integers(lo, hi + 1, size=B).astype(np.int64)
    ptr = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(n, out=ptr[1:])
    ingress, usage, cap = _raw(n, B + 1)
    if B > 1:                                                   # an all-zero state: 0 / 1e-3
        ingress[ptr[1]:ptr[2]] = 0
        usage[ptr[1]:ptr[2]] = 0
    want = O.node_features(ingress, usage, cap, ptr)
    t = lambda a: torch.from_numpy(a).to(DEV)
    got = ops.node_features(t(ingress), t(usage), t(cap), t(ptr.astype(np.int32))).cpu().numpy()
    assert got.dtype == np.float32 and np.array_equal(got, want)
    assert got.min() >= 0.0 and got.max() <= 1.0


def test_build_observations_feeds_the_actor():
    n_max, B = 24, 50
    st = synth.make_states(B, n_max, 10, 24, seed=2, num_topologies=5)
    ingress, usage, cap = _raw(st.n, 9)
    gb = observe.build_observations(ingress, usage, cap, st.n, st.edge_index, n_max)
    assert gb.structured and gb._mask is None and gb.x.shape == (int(st.n.sum()), 3)
    assert np.array_equal(gb.x.cpu().numpy(), O.node_features(ingress, usage, cap, st.ptr))
    assert np.array_equal(gb.batch.cpu().numpy(), st.batch)
    actor = Actor(AgentSpec(n_max)).to(DEV)
    y = ddpg.predict(actor, gb)
    assert y.shape == (B, st.action_dim) and torch.isfinite(y).all()
    assert np.all(y.cpu().numpy()[st.mask == 0] == 0.0)
