"""ctypes binding of libgsc_b200.so (the C-ABI declared in include/gsc_b200.h).

There is no CPU fallback: if the library is missing every op raises.  ``load()`` is lazy so that
pure-host modules (synth, graph containers, the replay buffer) import without the CUDA build.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64,
                    c_size_t, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgsc_b200.so")
_lib = None

EPI_BIAS, EPI_RELU, EPI_MUL, EPI_GATE, EPI_CLAMP1 = 1, 2, 4, 8, 16
EPI_ROWTABLE, EPI_SMASK = 64, 128
MAX_HIDDEN = 32
PACK_NOALLOC_MAX_NODES = 1 << 20


class GemmArgs(Structure):
    _fields_ = [("A", c_void_p * 3), ("B", c_void_p * 3), ("K", c_int64 * 3),
                ("a_rs", c_int64 * 3), ("a_cs", c_int64 * 3), ("b_rs", c_int64 * 3),
                ("b_cs", c_int64 * 3), ("num_seg", c_int), ("flags", c_int), ("bias", c_void_p),
                ("mul", c_void_p), ("ld_mul", c_int64), ("gate", c_void_p), ("ld_gate", c_int64),
                ("row_table", c_void_p), ("ld_table", c_int64), ("row_n", c_void_p), ("n_max", c_int32),
                ("groups", c_int32), ("dot_w", c_void_p), ("dot_bias", c_void_p), ("dot_out", c_void_p)]


class GscError(RuntimeError):
    pass


_PROTOS = {
    "gsc_version": (c_int, []),
    "gsc_set_sm_reserve": (c_int, [c_int]),
    "gsc_last_error": (c_char_p, []),
    "gsc_pack_csr_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "gsc_pack_csr": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_void_p,
                             c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "gsc_pack_csr_noalloc": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p]),
    "gsc_build_mask": (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "gsc_replay_gather": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                  c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "gsc_gather_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "gsc_gnn_param_floats": (c_int64, [c_int, c_int, c_int]),
    "gsc_gnn_saved_floats": (c_int64, [c_int64, c_int, c_int, c_int]),
    "gsc_gnn_embed_fwd": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64,
                                  c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "gsc_node_features": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "gsc_gnn_bwd_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int, c_int, c_int, c_int, c_int]),
    "gsc_gnn_embed_bwd": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64,
                                  c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int64, c_void_p,
                                  c_void_p, c_void_p, c_size_t, c_void_p]),
    "gsc_gnn_embed_bwd_prepared": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64,
                                           c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int64, c_void_p,
                                           c_void_p, c_void_p, c_size_t, c_void_p, c_void_p]),
    "gsc_gnn_bwd_prep_bytes": (c_size_t, [c_int64, c_int]),
    "gsc_gnn_bwd_prep": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_size_t,
                                 c_void_p]),
    "gsc_gemm_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64]),
    "gsc_gemm": (c_int, [POINTER(GemmArgs), c_int64, c_int64, c_void_p, c_int64, c_void_p, c_size_t,
                         c_void_p]),
    "gsc_mask_table": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "gsc_mask_wgrad_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "gsc_mask_wgrad": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int64,
                               c_void_p, c_size_t, c_void_p]),
    "gsc_mul_gate": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64,
                             c_void_p, c_int64, c_void_p]),
    "gsc_colsum": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "gsc_td_mse": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_int64,
                           c_void_p, c_void_p, c_void_p, c_void_p]),
    "gsc_adam_soft_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                     c_float, c_float, c_float, c_float, c_int64, c_float, c_void_p]),
    "gsc_adam_soft_update_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                         c_float, c_float, c_float, c_float, c_void_p, c_float, c_void_p]),
    "gsc_adam_soft_update_slice": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                           c_float, c_float, c_float, c_float, c_void_p, c_float, c_int, c_void_p]),
    "gsc_soft_update": (c_int, [c_void_p, c_void_p, c_int64, c_float, c_void_p]),
    "gsc_linear_n1_fwd": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "gsc_linear_n1_bwd_workspace_bytes": (c_size_t, [c_int]),
    "gsc_linear_n1_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p, c_int, c_void_p,
                                  c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "gsc_actor_head": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_float, c_float,
                               c_void_p, c_void_p]),
}


def exported_symbols():
    return sorted(_PROTOS)


def load() -> ctypes.CDLL:
    """Load the C-ABI library, failing loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GscError(
                f"{LIB_PATH} is missing: build it with `python -m gsc_b200.build` "
                "(or __graft_entry__.build()).  gsc_b200 has no CPU or PyTorch fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().gsc_last_error().decode(errors="replace")
        raise GscError(f"{what} failed (code {rc}): {msg}")
