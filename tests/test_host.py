"""CPU tests of the host side: C-ABI exports, containers, replay buffer, module surface."""
import io
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from gsc_b200 import _lib, synth
from gsc_b200.buffer import GraphReplayBuffer
from gsc_b200.graph import GraphBatch, GraphData
from gsc_b200.models import Actor, AgentSpec, GNNEmbedder, QNetwork

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    text = open(os.path.join(ROOT, "include", "gsc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gsc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _header_functions()
    assert len(names) >= 15
    lib = _lib.load()                      # raises if the .so was not built
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gsc_b200.h but not exported"
    assert sorted(_lib.exported_symbols()) == names
    assert lib.gsc_version() >= 100


def test_ctypes_mirror_of_gsc_gemm_args_matches_the_header(tmp_path):
    """The struct crosses the C-ABI by pointer: size and field offsets of the ctypes mirror must be what a
    C compiler makes of include/gsc_b200.h."""
    import ctypes
    fields = [n for n, _ in _lib.GemmArgs._fields_]
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "gsc_b200.h"\nint main(void) {\n'
                   '  printf("%zu\\n", sizeof(gsc_gemm_args));\n' +
                   "".join(f'  printf("%zu\\n", offsetof(gsc_gemm_args, {f}));\n' for f in fields) +
                   "  return 0;\n}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert out[0] == ctypes.sizeof(_lib.GemmArgs)
    assert out[1:] == [getattr(_lib.GemmArgs, f).offset for f in fields]
    # the size queries added with the round-2 entry points answer without a GPU
    lib = _lib.load()
    assert lib.gsc_linear_n1_bwd_workspace_bytes(64) >= 256 + 65 * 4
    assert lib.gsc_gnn_bwd_prep_bytes(1024, 100) > 1024 * 128 * 2 and lib.gsc_gnn_bwd_prep_bytes(4, 300) == 0
    assert lib.gsc_version() >= 200


def test_size_queries_need_no_gpu():
    lib = _lib.load()
    assert lib.gsc_gnn_param_floats(3, 22, 2) == 220 + 1056
    assert lib.gsc_gnn_saved_floats(100, 22, 2, 2) == 3 * 100 * 24
    assert lib.gsc_pack_csr_workspace_bytes(1000, 5000) >= 4000
    assert lib.gsc_gemm_workspace_bytes(100, 64, 60022) > 0           # Linear(22+2A -> 64) is split along K
    assert lib.gsc_gemm_workspace_bytes(4096, 1728, 256) == 0


def test_product_code_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gsc_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


def test_ops_refuse_cpu_tensors():
    from gsc_b200 import ops
    with pytest.raises(_lib.GscError):
        ops.pack_csr(torch.zeros(2, 3, dtype=torch.int64), None, 4, 1)


def test_synth_states_layout():
    st = synth.make_states(16, 24, 16, 24, seed=1)
    assert st.x.shape == (st.ptr[-1], 3) and st.x.dtype == np.float32
    assert st.edge_index.dtype == np.int64 and st.mask.shape == (16, 24 * 24 * 3)
    for g in range(16):
        lo, hi, n = st.ptr[g], st.ptr[g + 1], st.n[g]
        sel = (st.edge_index[1] >= lo) & (st.edge_index[1] < hi)
        src = st.edge_index[0][sel]
        assert np.all((src >= lo) & (src < hi))                       # block diagonal
        assert sel.sum() == 2 * round(1.5 * n)                        # both directions
        m = st.mask[g].reshape(24, 3, 24)
        assert m[:n, :, :n].all() and m[n:].sum() == 0 and m[:, :, n:].sum() == 0
    # symmetric and sorted source-major like from_networkx
    e = st.edge_index
    assert set(map(tuple, e.T.tolist())) == set(map(tuple, e[::-1].T.tolist()))
    a = synth.make_actions(st)
    groups = a.reshape(16, 72, 24)
    s = groups.sum(-1)
    assert np.all((np.abs(s - 1) < 1e-5) | (s == 0))


def test_graph_batch_collate_matches_from_data_list_layout():
    st = synth.make_states(5, 8, 3, 8, seed=2, num_topologies=5)
    datas = []
    for g in range(5):
        d = st.graph(g)
        datas.append(GraphData(d["x"], d["edge_index"], d["mask"]))
    gb = GraphBatch.from_data_list(datas)
    assert gb.num_graphs == 5 and gb.max_nodes == int(st.n.max())
    np.testing.assert_array_equal(gb.x.numpy(), st.x)
    np.testing.assert_array_equal(gb.batch.numpy(), st.batch)
    np.testing.assert_array_equal(gb.ptr.numpy(), st.ptr)
    np.testing.assert_array_equal(gb.mask.numpy(), st.mask)
    got = set(map(tuple, gb.edge_index.numpy().T.tolist()))
    assert got == set(map(tuple, st.edge_index.T.tolist()))


def test_replay_buffer_surface_and_sampling():
    spec = AgentSpec(8)
    rb = GraphReplayBuffer(16, spec.env.action_space, device="cpu")
    st = synth.make_states(20, 8, 3, 8, seed=3, num_topologies=3)
    acts = synth.make_actions(st)
    for g in range(20):                                   # wraps around once
        d = st.graph(g)
        nd = st.graph((g + 1) % 20)
        rb.add(GraphData(d["x"], d["edge_index"], d["mask"]), GraphData(nd["x"], nd["edge_index"], nd["mask"]),
               acts[g], 0.5 * g, g % 7 == 0, {})
    assert rb.full and rb.pos == 4
    np.random.seed(0)
    expect = (np.random.randint(1, 16, size=6) + rb.pos) % 16
    np.random.seed(0)
    s = rb.sample(6)
    assert s.observations.num_graphs == 6 and s.observations.mask.shape == (6, 192)
    assert s.actions.shape == (6, 192) and s.dones.shape == (6,) and s.rewards.shape == (6,)
    np.testing.assert_allclose(s.rewards.numpy(), rb.rewards[expect])
    assert s.observations.x.shape[0] == int(s.observations.ptr[-1])
    assert int(s.observations.batch.max()) == 5


def test_module_surface_matches_reference():
    spec = AgentSpec(24)
    actor, critic = Actor(spec), QNetwork(spec)
    keys = ["embedder.encoder." + k for k in ("att", "bias", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias")]
    keys += ["embedder.process.0." + k for k in ("att", "bias", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias")]
    assert list(actor.state_dict().keys()) == keys + [f"actor.lins.{i}.{k}" for i in (0, 1) for k in ("weight", "bias")]
    assert list(critic.state_dict().keys()) == keys + [f"critic.lins.{i}.{k}" for i in (0, 1) for k in ("weight", "bias")]
    assert sum(p.numel() for p in actor.parameters()) == 893628
    assert sum(p.numel() for p in critic.parameters()) == 223997
    assert actor.actor.lins[0].weight.shape == (256, 22 + 1728)
    assert critic.critic.lins[0].weight.shape == (64, 22 + 2 * 1728)
    assert actor.embedder.encoder.att.shape == (1, 1, 22)
    assert actor.low.shape == (1728,) and actor.high.max() == 1.0
    a = np.array([0.0, 0.25, 1.0], dtype=np.float32)
    np.testing.assert_allclose(actor.unscale_action(actor.scale_action(np.resize(a, 1728))), np.resize(a, 1728), atol=1e-6)


def test_flat_parameter_buffer_aliases_parameters_and_survives_io():
    spec = AgentSpec(8)
    actor = Actor(spec)
    for p, v in zip(actor.ordered_params(), actor.param_views(actor.flat)):
        assert p.data_ptr() == v.data_ptr() and p.shape == v.shape
    w1 = actor.actor.lins[0].weight
    assert (w1.data_ptr() + 22 * 4) % 16 == 0 and w1.stride(0) % 4 == 0   # mask block 16-byte aligned
    with torch.no_grad():
        actor.actor.lins[1].bias.fill_(3.0)
    _, lins = actor.views(actor.flat)
    assert float(lins[1][1].min()) == 3.0
    other = Actor(spec)
    other.load_state_dict(actor.state_dict())
    assert torch.equal(other.flat, actor.flat)
    buf = io.BytesIO()
    torch.save(actor, buf)                                  # whole-module pickle, as reference main.py:47
    buf.seek(0)
    back = torch.load(buf, weights_only=False)
    assert torch.equal(back.flat, actor.flat)
    # optimizers that the reference constructs work on the view parameters
    opt = torch.optim.Adam(list(actor.parameters()), lr=1e-3)
    for p in actor.parameters():
        p.grad = torch.ones_like(p)
    opt.step()
    assert float(lins[1][1].max()) < 3.0


def test_embedder_constructor_signature():
    e = GNNEmbedder(input_dim=3, hidden_dim=22, num_layers=2, aggr="mean", num_iter=2)
    assert e.param_floats() == 1276 and e.flat_params().numel() == 1276
    assert e.flat_params().data_ptr() == e.encoder.att.data_ptr()
    with pytest.raises(NotImplementedError):
        GNNEmbedder(3, 22, 2, "add", 2)


def test_grad_regions_cover_every_parameter_and_nothing_else():
    """DDPGUpdater all-reduces the embedder block and one block per dense layer: together they must
    contain every parameter of the flat buffer, disjointly."""
    from gsc_b200.models import Actor, AgentSpec, QNetwork
    for cls in (Actor, QNetwork):
        net = cls(AgentSpec(8))
        gnn, layers = net.grad_regions()
        marks = torch.zeros(net.flat.numel(), dtype=torch.int32)
        for lo, hi in [gnn] + layers:
            assert 0 <= lo < hi <= net.flat.numel()
            marks[lo:hi] += 1
        assert int(marks.max()) == 1
        probe = torch.zeros_like(net.flat)
        for v in net.param_views(probe):
            v.fill_(1.0)
        assert bool(((probe > 0) <= (marks > 0)).all()), "a parameter lies outside the all-reduced blocks"
        assert gnn == (0, sum(p.numel() for p in net.embedder.ordered_params()) + (-sum(p.numel() for p in net.embedder.ordered_params())) % 64)
