"""ctypes binding of include/posteriflow_b200.h (the C-ABI drop-in boundary).

There is deliberately NO fallback: if the CUDA library is missing or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "lib", "libposteriflow_b200.so")

PF_OK = 0
PF_TIER_FP32, PF_TIER_BF16, PF_TIER_BF16X3 = 0, 1, 2
PF_FLAG_SANITIZE_CTX = 1 << 0
PF_FLAG_CLAMP = 1 << 1
PF_FLAG_WRAP = 1 << 2
PF_FLAG_DENORMALIZE = 1 << 3
PF_FLAG_CTX_IS_PROJ = 1 << 4
PF_FLAG_SORT_MASSES = 1 << 5
PF_OP_CTXPROJ, PF_OP_INVERSE, PF_OP_FORWARD, PF_OP_BACKWARD, PF_OP_SAMPLE = 0, 1, 2, 3, 4

EXPORTS = [
    "pf_flow_create", "pf_flow_destroy", "pf_flow_pack", "pf_flow_ctxproj", "pf_flow_inverse",
    "pf_sample_fused", "pf_flow_forward", "pf_flow_backward", "pf_flow_workspace_bytes",
    "pf_flow_saved_bytes", "pf_flow_last_launch_count", "pf_flow_hidden_perm", "pf_status_str",
    "pf_last_cuda_error", "pf_abi_version", "pf_physical_logq", "pf_posterior_summary", "pf_flow_check", "pf_flow_fused_training",
    "pf_sample_fused_static", "pf_slot_prepare", "pf_slot_loss", "pf_slot_ctx_grad",
    "pf_rows_prepare", "pf_rows_ctx_grad",
]

_fp = C.c_void_p  # every device pointer is passed as an integer address


class FlowDesc(C.Structure):
    _fields_ = [("features", C.c_int32), ("context_features", C.c_int32), ("hidden_features", C.c_int32),
                ("num_layers", C.c_int32), ("num_bins", C.c_int32), ("num_blocks", C.c_int32),
                ("tail_bound", C.c_float), ("clamp_bound", C.c_float), ("tier", C.c_int32),
                ("device", C.c_int32)]


class LayerPtrs(C.Structure):
    _fields_ = [("initial_w", _fp), ("initial_b", _fp), ("initial_mask", _fp), ("ctx_w", _fp), ("ctx_b", _fp),
                ("blk_ctx_w", _fp * 2), ("blk_ctx_b", _fp * 2), ("blk_lin_w", (_fp * 2) * 2),
                ("blk_lin_b", (_fp * 2) * 2), ("blk_lin_mask", (_fp * 2) * 2), ("final_w", _fp),
                ("final_b", _fp), ("final_mask", _fp)]


class LayerGrads(C.Structure):
    _fields_ = [("initial_w", _fp), ("initial_b", _fp), ("ctx_w", _fp), ("ctx_b", _fp),
                ("blk_ctx_w", _fp * 2), ("blk_ctx_b", _fp * 2), ("blk_lin_w", (_fp * 2) * 2),
                ("blk_lin_b", (_fp * 2) * 2), ("final_w", _fp), ("final_b", _fp)]


class Scaler(C.Structure):
    _fields_ = [("lo", _fp), ("hi", _fp), ("log_mask", _fp), ("circ_mask", _fp)]


_lib = None


def load_library() -> C.CDLL:
    """Load the in-tree CUDA library; raise if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"posteriflow_b200: CUDA library not built ({LIB_PATH}). Run `python -m posteriflow_b200.build` "
            "(or __graft_entry__.build()). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    i64, u64, u32, i32, sz = C.c_int64, C.c_uint64, C.c_uint32, C.c_int32, C.c_size_t
    lib.pf_flow_create.argtypes = [C.POINTER(FlowDesc), C.POINTER(C.c_void_p)]
    lib.pf_flow_create.restype = i32
    lib.pf_flow_destroy.argtypes = [C.c_void_p]
    lib.pf_flow_destroy.restype = None
    lib.pf_flow_pack.argtypes = [C.c_void_p, C.POINTER(LayerPtrs), _fp]
    lib.pf_flow_pack.restype = i32
    lib.pf_flow_fused_training.argtypes = [C.c_void_p]
    lib.pf_flow_fused_training.restype = i32
    lib.pf_flow_check.argtypes = [C.c_void_p, _fp]
    lib.pf_flow_check.restype = i32
    lib.pf_flow_ctxproj.argtypes = [C.c_void_p, _fp, i64, u32, _fp, _fp]
    lib.pf_flow_ctxproj.restype = i32
    lib.pf_flow_inverse.argtypes = [C.c_void_p, _fp, _fp, _fp, i64, i64, i64, _fp, _fp, u32, C.POINTER(Scaler),
                                    _fp, sz, _fp]
    lib.pf_flow_inverse.restype = i32
    lib.pf_sample_fused.argtypes = [C.c_void_p, _fp, C.POINTER(Scaler), _fp, u64, i64, i64, _fp, _fp, _fp, u32,
                                    _fp, sz, _fp]
    lib.pf_sample_fused.restype = i32
    lib.pf_sample_fused_static.argtypes = [C.c_void_p, _fp, C.POINTER(Scaler), _fp, u64, i64, i64, _fp, _fp, _fp, u32, _fp]
    lib.pf_sample_fused_static.restype = i32
    lib.pf_slot_prepare.argtypes = [_fp, _fp, _fp, _fp, C.POINTER(Scaler), _fp, _fp, i32, i32, i32, i32, i32, _fp, _fp, _fp,
                                    i32, _fp]
    lib.pf_slot_prepare.restype = i32
    lib.pf_rows_prepare.argtypes = [_fp, _fp, _fp, _fp, C.POINTER(Scaler), _fp, _fp, i32, i32, i32, i32, i32, i32, _fp, _fp, _fp,
                                    _fp, i32, _fp]
    lib.pf_rows_prepare.restype = i32
    lib.pf_rows_ctx_grad.argtypes = [_fp, _fp, i32, i32, i32, i32, _fp, _fp, i32, _fp]
    lib.pf_rows_ctx_grad.restype = i32
    lib.pf_slot_loss.argtypes = [_fp, _fp, _fp, i64, i32, _fp, _fp, _fp, i32, _fp]
    lib.pf_slot_loss.restype = i32
    lib.pf_slot_ctx_grad.argtypes = [_fp, i32, i32, i32, i32, _fp, _fp, i32, _fp]
    lib.pf_slot_ctx_grad.restype = i32
    lib.pf_flow_forward.argtypes = [C.c_void_p, _fp, _fp, _fp, i64, i64, i64, _fp, _fp, _fp, _fp, u32, _fp, _fp,
                                    sz, _fp]
    lib.pf_flow_forward.restype = i32
    lib.pf_flow_backward.argtypes = [C.c_void_p, _fp, _fp, _fp, i64, i64, i64, _fp, _fp, _fp, _fp,
                                     C.POINTER(LayerGrads), _fp, sz, _fp]
    lib.pf_flow_backward.restype = i32
    lib.pf_physical_logq.argtypes = [_fp, _fp, C.POINTER(Scaler), i32, i64, _fp, _fp, u32, i32, _fp]
    lib.pf_physical_logq.restype = i32
    lib.pf_posterior_summary.argtypes = [_fp, i64, i64, i32, C.POINTER(C.c_double), i32, _fp, _fp, _fp, _fp, _fp, _fp,
                                         i32, _fp]
    lib.pf_posterior_summary.restype = i32
    lib.pf_flow_workspace_bytes.argtypes = [C.c_void_p, i32, i64, i64]
    lib.pf_flow_workspace_bytes.restype = sz
    lib.pf_flow_saved_bytes.argtypes = [C.c_void_p, i64, i64]
    lib.pf_flow_saved_bytes.restype = sz
    lib.pf_flow_last_launch_count.argtypes = [C.c_void_p]
    lib.pf_flow_last_launch_count.restype = i64
    lib.pf_flow_hidden_perm.argtypes = [C.c_void_p, C.POINTER(i32)]
    lib.pf_flow_hidden_perm.restype = i32
    lib.pf_status_str.argtypes = [i32]
    lib.pf_status_str.restype = C.c_char_p
    lib.pf_last_cuda_error.argtypes = []
    lib.pf_last_cuda_error.restype = C.c_char_p
    lib.pf_abi_version.argtypes = []
    lib.pf_abi_version.restype = i32
    _lib = lib
    return lib


class PfError(RuntimeError):
    pass


def check(status: int, what: str) -> None:
    if status != PF_OK:
        lib = load_library()
        msg = lib.pf_status_str(status).decode()
        if status == -2:
            msg += ": " + lib.pf_last_cuda_error().decode()
        raise PfError(f"{what} failed: {msg} (status {status})")
