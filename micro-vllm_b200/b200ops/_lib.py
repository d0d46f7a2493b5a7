"""ctypes binding of libb200ops.so (include/b200_abi.h).

There is deliberately no fallback: if the shared library cannot be found or
built, importing this module raises, and every operator in core/ and kernel/
fails with it.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_i32, _i64, _f32, _vp = C.c_int32, C.c_int64, C.c_float, C.c_void_p

E_NULL, E_SHAPE, E_ALIGN, E_RANGE, E_UNSUPPORTED = -1, -2, -3, -4, -5
DT_BF16, DT_F32 = 0, 1
ABI_VERSION = 17


class DecodeArgs(C.Structure):
    """struct b200_decode_args (include/b200_abi.h)."""
    _fields_ = [
        ("q", _vp), ("k_new", _vp), ("v_new", _vp), ("k_cache", _vp), ("v_cache", _vp),
        ("rotary_cos", _vp), ("rotary_sin", _vp), ("cache_seqlens", _vp), ("block_table", _vp),
        ("out", _vp), ("softmax_lse", _vp), ("workspace", _vp), ("workspace_bytes", _i64),
        ("B", _i32), ("Hq", _i32), ("Hkv", _i32), ("D", _i32),
        ("rotary_dim", _i32), ("rotary_interleaved", _i32), ("rot_rows", _i32),
        ("n_blocks", _i32), ("block_size", _i32), ("max_blocks_per_seq", _i32),
        ("q_stride_b", _i64), ("q_stride_h", _i64),
        ("k_stride_b", _i64), ("k_stride_h", _i64),
        ("v_stride_b", _i64), ("v_stride_h", _i64),
        ("o_stride_b", _i64), ("o_stride_h", _i64),
        ("kc_stride_blk", _i64), ("kc_stride_tok", _i64), ("kc_stride_h", _i64),
        ("vc_stride_blk", _i64), ("vc_stride_tok", _i64), ("vc_stride_h", _i64),
        ("bt_stride_b", _i64),
        ("num_splits", _i32), ("variant", _i32), ("softmax_scale", _f32), ("flags", _i32),
    ]


DECODE_FLAG_PDL = 1
DECODE_FLAG_PDL_LATE = 2
DECODE_FLAG_BALANCED = 4
DECODE_FLAG_NO_CLUSTER = 8
DECODE_BALANCED_MAX_B = 512


class PrefillArgs(C.Structure):
    """struct b200_prefill_args (include/b200_abi.h)."""
    _fields_ = [
        ("q", _vp), ("k", _vp), ("v", _vp), ("k_cache", _vp), ("v_cache", _vp),
        ("rotary_cos", _vp), ("rotary_sin", _vp), ("cache_seqlens", _vp), ("block_table", _vp),
        ("out", _vp), ("workspace", _vp), ("workspace_bytes", _i64),
        ("B", _i32), ("S", _i32), ("Hq", _i32), ("Hkv", _i32), ("D", _i32),
        ("rotary_dim", _i32), ("rotary_interleaved", _i32), ("rot_rows", _i32),
        ("n_blocks", _i32), ("block_size", _i32), ("max_blocks_per_seq", _i32),
        ("max_prefix", _i32), ("causal", _i32),
        ("q_stride_b", _i64), ("q_stride_s", _i64), ("q_stride_h", _i64),
        ("k_stride_b", _i64), ("k_stride_s", _i64), ("k_stride_h", _i64),
        ("v_stride_b", _i64), ("v_stride_s", _i64), ("v_stride_h", _i64),
        ("o_stride_b", _i64), ("o_stride_s", _i64), ("o_stride_h", _i64),
        ("kc_stride_blk", _i64), ("kc_stride_tok", _i64), ("kc_stride_h", _i64),
        ("vc_stride_blk", _i64), ("vc_stride_tok", _i64), ("vc_stride_h", _i64),
        ("bt_stride_b", _i64),
        ("variant", _i32), ("softmax_scale", _f32),
    ]


GEMM_EPI_NONE, GEMM_EPI_BIAS, GEMM_EPI_SWIGLU = 0, 1, 2
GEMM_FLAG_PDL = 1
AR_MAX_WORLD, AR_MAX_CTAS, IPC_HANDLE_BYTES = 8, 1024, 64
AR_MODE_ROW_SHARDED = 1
AR_MODE_LL = 2
AR_MODE_NVLS = 4
AR_MODE_NVLS_ONEHOP = 8
AR_FLAG_PDL = 0x100
AR_FLAG_PACKET_TAIL = 0x200
AR_NVLS_READY_WORD = AR_MAX_CTAS * (AR_MAX_WORLD + 1)
AR_NVLS_DONE_WORD = AR_NVLS_READY_WORD + 32


class ArArgs(C.Structure):
    """struct b200_ar_args (include/b200_abi.h)."""
    _fields_ = [
        ("x_peers", _vp * AR_MAX_WORLD), ("flag_peers", _vp * AR_MAX_WORLD),
        ("epoch", _vp), ("residual", _vp), ("weight", _vp), ("y", _vp), ("res_out", _vp),
        ("rows", _i64), ("n", _i64),
        ("stride_x", _i64), ("stride_r", _i64), ("stride_y", _i64), ("stride_res", _i64),
        ("eps", _f32), ("w_dtype", _i32), ("rank", _i32), ("world", _i32),
        ("ss_rounded", _i32), ("timeout_ms", _i32), ("mode", _i32),
        ("y_peers", _vp * AR_MAX_WORLD),
        ("xstage_peers", _vp * AR_MAX_WORLD), ("ystage_peers", _vp * AR_MAX_WORLD), ("trace", _vp),
        ("x_mc", _vp), ("y_mc", _vp), ("flags_mc", _vp), ("ystage_mc", _vp),
    ]


class GatherArgs(C.Structure):
    """struct b200_gather_args (include/b200_abi.h)."""
    _fields_ = [
        ("src", _vp), ("dst_peers", _vp * AR_MAX_WORLD), ("dst_mc", _vp),
        ("flag_peers", _vp * AR_MAX_WORLD), ("flags_mc", _vp), ("epoch", _vp),
        ("rows", _i64), ("cols_local", _i64), ("stride_src", _i64), ("stride_dst", _i64),
        ("rank", _i32), ("world", _i32), ("timeout_ms", _i32), ("flags", _i32),
    ]


# name -> (restype, argtypes); every symbol include/b200_abi.h declares
SIGNATURES = {
    "b200_version": (C.c_int, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_sm_count": (C.c_int, []),
    "b200_swiglu": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "b200_rmsnorm": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _f32, C.c_int, _vp]),
    "b200_rmsnorm_residual": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64,
                                         _f32, C.c_int, _vp]),
    "b200_store_kvcache": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64, _i64,
                                      _i64, _i64, _i64, _i64, _i64, _vp]),
    "b200_greedy_argmax_workspace_bytes": (_i64, [_i64]),
    "b200_greedy_argmax": (C.c_int, [_vp, C.c_int, _i64, _i64, _i64, _vp, _vp, _i64, _vp]),
    "b200_greedy_argmax_noalloc": (C.c_int, [_vp, C.c_int, _i64, _i64, _i64, _vp, _vp]),
    "b200_topk_topp_sample": (C.c_int, [_vp, C.c_int, _i64, _i64, _i64, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200_table_advance": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _i64, C.c_int, C.c_int, _vp, C.c_int, _vp]),
    "b200_l2_prefetch": (C.c_int, [_vp, _i64, _i64, _i64, _vp]),
    "b200_stream_gemm_workspace_bytes": (_i64, [C.c_int, _i64, C.c_int]),
    "b200_debug_stream_gemm_trace": (None, [_vp]),
    "b200_stream_gemm": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, C.c_int, _i64, _i64, _i64, _i64, _i64,
                                    C.c_int, C.c_int, C.c_int, _vp]),
    "b200_skinny_gemm_splits": (C.c_int, [C.c_int, _i64, _i64, C.c_int]),
    "b200_skinny_gemm_workspace_bytes": (_i64, [C.c_int, _i64, C.c_int]),
    "b200_skinny_gemm": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, C.c_int, _i64, _i64, _i64, _i64, _i64,
                                    C.c_int, C.c_int, _vp]),
    "b200_decode_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32, _i32]),
    "b200_decode_balanced_workspace_bytes": (_i64, [_i32]),
    "b200_paged_attn_decode": (C.c_int, [C.POINTER(DecodeArgs), _vp]),
    "b200_lse_combine": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _i64, _vp]),
    "b200_prefill_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32]),
    "b200_paged_attn_prefill": (C.c_int, [C.POINTER(PrefillArgs), _vp]),
    "b200_debug_prefill_trace": (None, [_vp]),
    "b200_debug_decode_trace": (None, [_vp]),
    "b200_allreduce_residual_rmsnorm": (C.c_int, [C.POINTER(ArArgs), _vp]),
    "b200_allgather_cols": (C.c_int, [C.POINTER(GatherArgs), _vp]),
    "b200_comm_alloc": (C.c_int, [_i64, C.POINTER(_vp)]),
    "b200_comm_free": (C.c_int, [_vp]),
    "b200_ipc_export": (C.c_int, [_vp, _vp]),
    "b200_ipc_import": (C.c_int, [_vp, C.POINTER(_vp)]),
    "b200_ipc_close": (C.c_int, [_vp]),
}


class B200Error(RuntimeError):
    def __init__(self, code: int, where: str, msg: str):
        kind = "argument error" if code < 0 else "cudaError"
        super().__init__(f"{where}: {kind} {code}: {msg}")
        self.code = code


def _load() -> C.CDLL:
    path = os.environ.get("B200OPS_LIB") or _build.LIB_PATH
    if not os.path.exists(path) or (path == _build.LIB_PATH and not _build.is_fresh()):
        try:
            path = _build.build()
        except Exception as e:   # noqa: BLE001 - re-raised as a hard import error
            if not os.path.exists(path):
                raise ImportError(
                    f"libb200ops.so is missing and could not be built ({e}); "
                    "the B200 operators have no fallback path") from e
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError here = ABI mismatch, fail loudly
        fn.restype = res
        fn.argtypes = args
    ver = lib.b200_version()
    if ver != ABI_VERSION:
        raise ImportError(f"libb200ops.so ABI version {ver} != binding version {ABI_VERSION}; rebuild")
    return lib


lib = _load()
LIB_PATH = lib._name


def check(code: int, where: str) -> None:
    if code != 0:
        raise B200Error(code, where, lib.b200_last_error().decode(errors="replace"))
