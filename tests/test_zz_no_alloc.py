"""GPU test of gsc_pack_csr_noalloc, the packer entry point that takes no workspace.

Why it exists: the rollout / serving loop packs one small graph per environment step while a
training stream may be busy on the same GPU, and some hosts bind the C-ABI with a fixed memory
budget and no allocator; neither can afford (or perform) a device allocation per call.  The entry
point therefore keeps its scratch in a fixed array inside the library.  Collected last on purpose
(file name), so a run that stops at the first failure has already executed every other test.

Checks, through the C-ABI: two calls in a row (the second on a different, smaller batch, so it sees
the first call's leftovers in the shared scratch) give exactly the CSR that gsc_pack_csr gives with
a caller workspace -- integer outputs, so the tolerance is bit-exact -- and a batch beyond the
built-in limit is refused with a clear error instead of being run.
"""
import ctypes

import numpy as np
import pytest
import torch

from gsc_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _call_noalloc(ei, batch, num_nodes, num_graphs):
    lib = _lib.load()
    E = int(ei.shape[1])
    rowptr = torch.full((num_nodes + 1,), -7, dtype=torch.int32, device=DEV)
    col = torch.full((E + num_nodes,), -7, dtype=torch.int32, device=DEV)
    gptr = torch.full((num_graphs + 1,), -7, dtype=torch.int32, device=DEV)
    stats = torch.full((4,), -7, dtype=torch.int32, device=DEV)
    rc = lib.gsc_pack_csr_noalloc(ei.data_ptr(), E, batch.data_ptr(), num_nodes, num_graphs, rowptr.data_ptr(),
                                  col.data_ptr(), gptr.data_ptr(), stats.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    return rc, rowptr, col, gptr, stats


def _same_as_workspace_entry_point(st):
    ei = torch.from_numpy(st.edge_index).to(DEV)
    batch = torch.from_numpy(st.batch).to(DEV)
    nt, B = int(st.ptr[-1]), st.num_graphs
    rc, rowptr, col, gptr, stats = _call_noalloc(ei, batch, nt, B)
    assert rc == 0, _lib.load().gsc_last_error().decode()
    want = ops.pack_csr(ei, batch, nt, B)                      # existing entry point, caller workspace
    torch.cuda.synchronize()
    et = int(want[3][0])
    assert torch.equal(rowptr, want[0])
    assert torch.equal(col[:et], want[1][:et])
    assert torch.equal(gptr, want[2])
    assert torch.equal(stats, want[3])
    assert int(stats[2]) == 0 and int(stats[0]) == st.edge_index.shape[1] + nt


def test_noalloc_packer_matches_workspace_packer_on_two_calls_in_a_row():
    big = synth.make_states(3000, 24, 16, 24, seed=21, num_topologies=32)      # several scan blocks
    small = synth.make_states(5, 24, 2, 9, seed=22, num_topologies=5)          # reuses a dirty scratch
    _same_as_workspace_entry_point(big)
    _same_as_workspace_entry_point(small)
    _same_as_workspace_entry_point(big)


def test_noalloc_packer_rejects_batches_beyond_its_scratch():
    lib = _lib.load()
    dummy = torch.zeros(8, dtype=torch.int64, device=DEV)
    out = torch.zeros(8, dtype=torch.int32, device=DEV)
    too_many = _lib.PACK_NOALLOC_MAX_NODES + 1
    rc = lib.gsc_pack_csr_noalloc(dummy.data_ptr(), 0, dummy.data_ptr(), too_many, 1, out.data_ptr(), out.data_ptr(),
                                  out.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -1                                                            # GSC_EINVAL, nothing launched
    msg = lib.gsc_last_error().decode()
    assert "exceed" in msg and str(_lib.PACK_NOALLOC_MAX_NODES) in msg
    torch.cuda.synchronize()
    assert int(out.abs().sum()) == 0
