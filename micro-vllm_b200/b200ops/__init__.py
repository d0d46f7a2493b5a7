"""b200ops — thin torch-tensor front end over the libb200ops C-ABI.

Tensors are only used for their device pointers, shapes and strides; every
function launches on torch's *current* CUDA stream, allocates nothing unless
stated, and never synchronises, so all of them can be captured in a CUDA graph
(reference contract: core/layer/model_graph.py:240-250).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch

from . import _lib
from ._lib import B200Error, DecodeArgs, PrefillArgs, lib, check, LIB_PATH  # noqa: F401

__all__ = [
    "swiglu", "rmsnorm", "rmsnorm_residual", "store_kvcache", "paged_attn_decode",
    "paged_attn_prefill", "lse_combine", "greedy_argmax", "topk_topp_sample", "skinny_gemm", "skinny_gemm_ok", "stream_gemm", "stream_gemm_ok", "interleave_up_gate", "decode_num_splits", "decode_use_balanced", "decode_algorithmic_bytes", "reserve_graph_workspace", "l2_prefetch", "table_advance", "B200Error", "LIB_PATH",
]


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _req_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise AssertionError("b200ops: all tensors must be CUDA tensors (no CPU fallback exists)")


def _wdt(w: torch.Tensor) -> int:
    if w.dtype == torch.bfloat16:
        return _lib.DT_BF16
    if w.dtype == torch.float32:
        return _lib.DT_F32
    raise AssertionError(f"b200ops: norm weight dtype {w.dtype} unsupported (bf16 / fp32)")


# ------------------------------------------------------------------ elementwise
def swiglu(gate_up: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[r, c] = silu(gate_up[r, d+c]) * gate_up[r, c]   (kernel/swiglu.py:18-46)."""
    _req_cuda(gate_up, out)
    assert gate_up.dtype == torch.bfloat16 and out.dtype == torch.bfloat16, "swiglu: bf16 only"
    d2 = gate_up.shape[-1]
    assert d2 % 2 == 0
    d = d2 // 2
    gu = gate_up.reshape(-1, d2)
    o = out.view(-1, d)
    assert gu.stride(1) == 1 and o.stride(1) == 1 and gu.shape[0] == o.shape[0]
    check(lib.b200_swiglu(gu.data_ptr(), o.data_ptr(), gu.shape[0], d, gu.stride(0), o.stride(0), _stream()),
          "b200_swiglu")
    return out


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, out: torch.Tensor, eps: float) -> torch.Tensor:
    _req_cuda(x, weight, out)
    assert x.dtype == torch.bfloat16 and out.dtype == torch.bfloat16, "rmsnorm: bf16 activations only"
    n = x.shape[-1]
    x2, y2 = x.view(-1, n), out.view(-1, n)
    assert x2.stride(1) == 1 and y2.stride(1) == 1 and weight.is_contiguous()
    check(lib.b200_rmsnorm(x2.data_ptr(), weight.data_ptr(), y2.data_ptr(), x2.shape[0], n,
                           x2.stride(0), y2.stride(0), float(eps), _wdt(weight), _stream()), "b200_rmsnorm")
    return out


def rmsnorm_residual(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                     out: torch.Tensor, res_out: torch.Tensor, eps: float):
    _req_cuda(x, residual, weight, out, res_out)
    for t in (x, residual, out, res_out):
        assert t.dtype == torch.bfloat16, "rmsnorm_residual: bf16 activations only"
    n = x.shape[-1]
    x2, r2, y2, o2 = (t.view(-1, n) for t in (x, residual, out, res_out))
    assert all(t.stride(1) == 1 for t in (x2, r2, y2, o2)) and weight.is_contiguous()
    check(lib.b200_rmsnorm_residual(x2.data_ptr(), r2.data_ptr(), weight.data_ptr(), y2.data_ptr(), o2.data_ptr(),
                                    x2.shape[0], n, x2.stride(0), r2.stride(0), y2.stride(0), o2.stride(0),
                                    float(eps), _wdt(weight), _stream()), "b200_rmsnorm_residual")
    return out, res_out


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor, block_size: int) -> None:
    _req_cuda(key, value, k_cache, v_cache, slot_mapping)
    T, H, D = key.shape
    assert key.dtype == value.dtype == k_cache.dtype == v_cache.dtype and key.element_size() == 2, \
        "store_kvcache: 16-bit key/value/cache of one dtype"
    assert slot_mapping.dtype == torch.int32 and slot_mapping.is_contiguous()
    assert key.stride(2) == 1 and value.stride(2) == 1 and k_cache.stride(3) == 1 and v_cache.stride(3) == 1
    assert k_cache.stride() == v_cache.stride() and k_cache.shape[1] == block_size
    check(lib.b200_store_kvcache(key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                 slot_mapping.data_ptr(), T, H, D,
                                 key.stride(0), key.stride(1), value.stride(0), value.stride(1),
                                 k_cache.shape[0], block_size,
                                 k_cache.stride(0), k_cache.stride(1), k_cache.stride(2), _stream()),
          "b200_store_kvcache")


def table_advance(cache_seqlens: torch.Tensor, batch: int, delta: int = 1, block_table: Optional[torch.Tensor] = None,
                  updates: Optional[torch.Tensor] = None) -> None:
    """cache_seqlens[:batch] += delta and apply the pending block-table entries of the device-side list
    ``updates`` = int32 ``[count, (row, col, block) * count]`` (count is reset to 0).  Current stream; capturable."""
    _req_cuda(cache_seqlens, block_table, updates)
    assert cache_seqlens.dtype == torch.int32 and cache_seqlens.is_contiguous() and cache_seqlens.numel() >= batch
    bt_ptr, stride, rows, cols, up_ptr, max_up = None, 0, 0, 0, None, 0
    if updates is not None:
        assert block_table is not None and block_table.dtype == torch.int32 and block_table.stride(1) == 1
        assert updates.dtype == torch.int32 and updates.is_contiguous() and updates.numel() >= 4
        bt_ptr, stride, (rows, cols) = block_table.data_ptr(), block_table.stride(0), block_table.shape
        up_ptr, max_up = updates.data_ptr(), (updates.numel() - 1) // 3
    check(lib.b200_table_advance(cache_seqlens.data_ptr(), int(batch), int(delta), bt_ptr, stride, rows, cols,
                                 up_ptr, max_up, _stream()), "b200_table_advance")


def l2_prefetch(t: torch.Tensor, nbytes: Optional[int] = None, per_row: bool = False) -> None:
    """Ask the memory system to bring ``nbytes`` of the 2-D row-major tensor ``t`` (all of it by default) into L2
    (current stream; capturable; touches no data): the first ``nbytes`` of its storage, or — ``per_row`` — the
    first ``nbytes / rows`` bytes of EVERY row (what a GEMM that walks the rows' leading columns first reads)."""
    _req_cuda(t)
    assert t.dim() == 2 and t.stride(1) == 1, "l2_prefetch: 2-D, last dim contiguous"
    es = t.element_size()
    rows, row_bytes, stride = t.shape[0], t.shape[1] * es, t.stride(0) * es
    total = rows * row_bytes
    n = total if nbytes is None else max(0, min(int(nbytes), total))
    if n == 0:
        return
    if stride == row_bytes and (not per_row or n == total):
        check(lib.b200_l2_prefetch(t.data_ptr(), n & ~15, 1, 0, _stream()), "b200_l2_prefetch")
    elif per_row:
        per = (n // rows) & ~127
        if per:
            check(lib.b200_l2_prefetch(t.data_ptr(), per, rows, stride, _stream()), "b200_l2_prefetch")
    else:
        full = n // row_bytes
        if full:
            check(lib.b200_l2_prefetch(t.data_ptr(), row_bytes & ~15, full, stride, _stream()), "b200_l2_prefetch")


# ------------------------------------------------------------------ decode attention
_SM_COUNT: dict[int, int] = {}
_WORKSPACES: dict[tuple, torch.Tensor] = {}


def sm_count(device: torch.device) -> int:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _SM_COUNT:
        with torch.cuda.device(idx):
            _SM_COUNT[idx] = max(1, lib.b200_sm_count())
    return _SM_COUNT[idx]


def decode_num_splits(batch: int, kv_heads: int, capacity_tokens: int, n_sm: int = 148) -> int:
    """Split count from launch-time constants only (CUDA-graph capture fixes the grid).

    Measured on B200 (profiles/r01/tune_*.json): one CTA per (sequence, kv head) saturates HBM
    as soon as there are at least as many CTAs as SMs, every extra split costs a prologue and a
    partial merge, and a grid that spills into a second partial wave loses more than it gains.
    So: no split when batch * kv_heads >= #SMs; otherwise the largest split count that keeps the
    grid within ~1.75 CTAs per SM (one wave), and never less than 256 tokens per split of what
    the block table can hold."""
    base = max(1, batch * kv_heads)
    if base >= n_sm:
        return 1
    want = max(1, int(1.75 * n_sm) // base)
    cap = max(1, capacity_tokens // 256)
    return int(max(1, min(want, cap, 128)))


def decode_algorithmic_bytes(seqlens_after, Hq: int, Hkv: int, D: int, block_size: int = 256, elt: int = 2) -> int:
    """Algorithmic HBM bytes of one decode attention call: K and V read once per KV
    head over the ``len_b + 1`` keys of every row, the new K / V rows written, q read and out written,
    the sequence lengths and the USED block-table entries read.  Split-KV scratch is not counted."""
    B = len(seqlens_after)
    total = 0
    for n in seqlens_after:
        total += int(n) * 2 * Hkv * D * elt + 4 * (-(-int(n) // block_size))
    return total + B * 2 * Hkv * D * elt + 2 * B * Hq * D * elt + 4 * B


_BALANCED_ENV = None
_TUNE: dict = {}


def _tune(name: str) -> int:
    """Integer tuning override from the environment (read once); 0 = unset."""
    v = _TUNE.get(name)
    if v is None:
        import os
        try:
            v = int(os.environ.get(name, "0"))
        except ValueError:
            v = 0
        _TUNE[name] = v
    return v


def decode_use_balanced(batch: int, kv_heads: int, n_sm: int = 148, q_per_kv: int = 1) -> bool:
    """Whether the persistent stream-K form is the default for this launch shape (launch-time
    constants only).  ``B200_DECODE_BALANCED`` = 0 / 1 forces it off / on (tuning)."""
    global _BALANCED_ENV
    if _BALANCED_ENV is None:
        import os
        _BALANCED_ENV = os.environ.get("B200_DECODE_BALANCED", "")
    if batch > _lib.DECODE_BALANCED_MAX_B:
        return False
    if _BALANCED_ENV in ("0", "1"):
        return _BALANCED_ENV == "1"
    # Measured on B200 (profiles/r02/bench_ragged.jsonl): a split-KV grid that needs more than one wave of the
    # 3 x #SMs resident CTAs and leaves its last wave mostly empty loses to the persistent form (Llama-3-70B heads,
    # B = 64, 8k: 512 CTAs on 444 slots, 411 us vs 343 us = 0.80 -> 0.95 of the HBM peak); a grid that fits one wave
    # does not (Qwen2.5-7B heads, B = 64, 2k: 256 CTAs, 45.5 us vs 59 us).
    groups = max(1, -(-max(1, q_per_kv) // 16))
    ctas = batch * kv_heads * groups
    slots = 3 * n_sm
    if ctas <= slots:
        return False
    waves = ctas / slots
    return waves / math.ceil(waves) < 0.85


_RETIRED: list = []


def _workspace(device: torch.device, nbytes: int, tag: str) -> torch.Tensor:
    """Per (device, stream, operator) scratch.  Zero-filled once: the decode arrival
    counters must start at 0 and the kernel resets them itself.  Outgrown buffers are
    kept alive because captured CUDA graphs may still hold their addresses.

    Launches recorded into a CUDA graph use ONE scratch per (device, operator) that every
    graph shares.  It is allocated and zeroed by the eager calls that precede a capture (the
    warm-up every capture needs anyway), never inside a capture: an allocation made while
    capturing lands in that graph's private pool and its zero fill becomes a memset node of
    that graph alone, so other graphs would find whatever the pool held."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    gkey = (idx, "graph", tag)
    if torch.cuda.is_current_stream_capturing():
        ws = _WORKSPACES.get(gkey)
        if ws is None or ws.numel() < nbytes:
            raise RuntimeError(
                f"b200ops: the '{tag}' scratch for captured launches ({nbytes} bytes) does not exist yet; run the "
                "same call once eagerly (warm-up) or call b200ops.reserve_graph_workspace() before torch.cuda.graph")
        return ws
    reserve_graph_workspace(device, nbytes, tag)
    key = (idx, torch.cuda.current_stream(device).cuda_stream, tag)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        if ws is not None:
            _RETIRED.append(ws)
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws


def reserve_graph_workspace(device: torch.device, nbytes: int, tag: str) -> None:
    """Make sure the scratch that CAPTURED launches of operator ``tag`` use holds ``nbytes``
    zeroed bytes.  Must run outside a capture; every eager launch does it implicitly."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    gkey = (idx, "graph", tag)
    ws = _WORKSPACES.get(gkey)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("b200ops.reserve_graph_workspace: cannot allocate during a CUDA graph capture")
        if ws is not None:
            _RETIRED.append(ws)
        _WORKSPACES[gkey] = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=torch.device("cuda", idx))


def paged_attn_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                      k_new: Optional[torch.Tensor], v_new: Optional[torch.Tensor],
                      rotary_cos: Optional[torch.Tensor], rotary_sin: Optional[torch.Tensor],
                      cache_seqlens: torch.Tensor, block_table: torch.Tensor,
                      softmax_scale: Optional[float] = None, rotary_interleaved: bool = False,
                      num_splits: int = 0, out: Optional[torch.Tensor] = None,
                      softmax_lse: Optional[torch.Tensor] = None, variant: int = 0,
                      pdl: int = 0, balanced: Optional[bool] = None, cluster_merge: bool = True) -> torch.Tensor:
    """q [B, Hq, D]; k_new/v_new [B, Hkv, D]; caches [n_blocks, block_size, Hkv, D];
    returns out [B, Hq, D] (allocated unless given).  ``pdl`` = 1 / 2 launches with programmatic
    stream serialization, releasing the next kernel at start / after the streaming loop (see
    B200_DECODE_FLAG_PDL in b200_abi.h for the precondition; measured slower than a plain launch on
    back-to-back decode kernels, so it is off by default).  ``balanced`` selects the persistent
    stream-K form (fixed grid, equal shares of all sequences' chunks per CTA, schedule derived on
    the device from ``cache_seqlens``: right for ragged and small batches); ``None`` = the
    :func:`decode_use_balanced` heuristic; an explicit ``num_splits`` > 0 keeps the split-KV grid.
    ``cluster_merge``: split counts 2..8 run the splits of a segment as one thread-block cluster and merge through
    distributed shared memory; False selects the workspace + last-arriver form (B200_DECODE_FLAG_NO_CLUSTER)."""
    _req_cuda(q, k_cache, v_cache, k_new, v_new, rotary_cos, rotary_sin, cache_seqlens, block_table)
    B, Hq, D = q.shape
    n_blocks, block_size, Hkv, Dc = k_cache.shape
    assert Dc == D and v_cache.shape == k_cache.shape
    assert q.dtype == torch.bfloat16 and k_cache.dtype == torch.bfloat16 and v_cache.dtype == torch.bfloat16, \
        "paged_attn_decode: bf16 only"
    assert cache_seqlens.dtype == torch.int32 and block_table.dtype == torch.int32
    assert cache_seqlens.is_contiguous() and cache_seqlens.shape[0] >= B
    assert block_table.shape[0] >= B and block_table.stride(1) == 1
    assert q.stride(2) == 1 and k_cache.stride(3) == 1 and v_cache.stride(3) == 1
    if out is None:
        out = torch.empty((B, Hq, D), dtype=q.dtype, device=q.device)
    assert out.stride(2) == 1
    a = DecodeArgs()
    a.q, a.k_cache, a.v_cache, a.out = q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), out.data_ptr()
    if k_new is not None:
        assert v_new is not None and k_new.shape == (B, Hkv, D) and v_new.shape == (B, Hkv, D)
        assert k_new.dtype == torch.bfloat16 and v_new.dtype == torch.bfloat16
        assert k_new.stride(2) == 1 and v_new.stride(2) == 1
        a.k_new, a.v_new = k_new.data_ptr(), v_new.data_ptr()
        a.k_stride_b, a.k_stride_h = k_new.stride(0), k_new.stride(1)
        a.v_stride_b, a.v_stride_h = v_new.stride(0), v_new.stride(1)
    if rotary_cos is not None:
        assert rotary_sin is not None and rotary_cos.shape == rotary_sin.shape
        assert rotary_cos.dtype == torch.bfloat16 and rotary_sin.dtype == torch.bfloat16
        assert rotary_cos.is_contiguous() and rotary_sin.is_contiguous()
        a.rotary_cos, a.rotary_sin = rotary_cos.data_ptr(), rotary_sin.data_ptr()
        a.rot_rows, a.rotary_dim = rotary_cos.shape[0], 2 * rotary_cos.shape[1]
    a.rotary_interleaved = 1 if rotary_interleaved else 0
    a.cache_seqlens, a.block_table = cache_seqlens.data_ptr(), block_table.data_ptr()
    if softmax_lse is not None:
        assert softmax_lse.dtype == torch.float32 and softmax_lse.is_contiguous() and softmax_lse.numel() >= B * Hq
        a.softmax_lse = softmax_lse.data_ptr()
    a.B, a.Hq, a.Hkv, a.D = B, Hq, Hkv, D
    a.n_blocks, a.block_size, a.max_blocks_per_seq = n_blocks, block_size, block_table.shape[1]
    a.q_stride_b, a.q_stride_h = q.stride(0), q.stride(1)
    a.o_stride_b, a.o_stride_h = out.stride(0), out.stride(1)
    a.kc_stride_blk, a.kc_stride_tok, a.kc_stride_h = k_cache.stride(0), k_cache.stride(1), k_cache.stride(2)
    a.vc_stride_blk, a.vc_stride_tok, a.vc_stride_h = v_cache.stride(0), v_cache.stride(1), v_cache.stride(2)
    a.bt_stride_b = block_table.stride(0)
    if balanced is None:
        balanced = num_splits <= 0 and not pdl and decode_use_balanced(B, Hkv, sm_count(q.device), Hq // Hkv)
    if num_splits <= 0:
        num_splits = 1 if balanced else decode_num_splits(B, Hkv, block_table.shape[1] * block_size,
                                                          sm_count(q.device))
        if not balanced and _tune("B200_DECODE_SPLITS") > 0:          # tuning override (A/B runs), launch-time constant
            num_splits = _tune("B200_DECODE_SPLITS")
    if variant == 0 and _tune("B200_DECODE_VARIANT") > 0:
        variant = _tune("B200_DECODE_VARIANT")
    a.num_splits, a.variant = int(num_splits), int(variant)
    a.flags = 0 if not pdl else (_lib.DECODE_FLAG_PDL | (_lib.DECODE_FLAG_PDL_LATE if int(pdl) == 2 else 0))
    if not cluster_merge or _tune("B200_DECODE_NO_CLUSTER") > 0:
        a.flags |= _lib.DECODE_FLAG_NO_CLUSTER
    if balanced:
        assert not pdl, "paged_attn_decode: the balanced form has no PDL variant"
        a.flags |= _lib.DECODE_FLAG_BALANCED
        need = lib.b200_decode_balanced_workspace_bytes(D)
        ws = _workspace(q.device, need, "decode")
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
    elif num_splits > 1:
        need = lib.b200_decode_workspace_bytes(B, Hq, Hkv, D, num_splits)
        ws = _workspace(q.device, need, "decode")
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 1.0 / math.sqrt(D)
    check(lib.b200_paged_attn_decode(C.byref(a), _stream()), "b200_paged_attn_decode")
    return out


def lse_combine(o_part: torch.Tensor, lse_part: torch.Tensor, out: Optional[torch.Tensor] = None,
                lse_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """o_part [B, Hq, S, D] fp32 (normalised per split), lse_part [B, Hq, S] fp32 -> out [B, Hq, D] bf16."""
    _req_cuda(o_part, lse_part)
    B, Hq, S, D = o_part.shape
    assert o_part.dtype == torch.float32 and lse_part.dtype == torch.float32
    assert o_part.is_contiguous() and lse_part.is_contiguous() and lse_part.shape == (B, Hq, S)
    if out is None:
        out = torch.empty((B, Hq, D), dtype=torch.bfloat16, device=o_part.device)
    check(lib.b200_lse_combine(o_part.data_ptr(), lse_part.data_ptr(), out.data_ptr(),
                               lse_out.data_ptr() if lse_out is not None else None,
                               B, Hq, S, D, out.stride(0), out.stride(1), _stream()), "b200_lse_combine")
    return out


# ------------------------------------------------------------------ prefill attention
def paged_attn_prefill(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                       k_cache: torch.Tensor, v_cache: torch.Tensor,
                       rotary_cos: Optional[torch.Tensor], rotary_sin: Optional[torch.Tensor],
                       cache_seqlens: torch.Tensor, block_table: torch.Tensor,
                       softmax_scale: Optional[float] = None, causal: bool = True,
                       rotary_interleaved: bool = False, max_prefix: Optional[int] = None,
                       out: Optional[torch.Tensor] = None, variant: int = 0) -> torch.Tensor:
    """q [B, S, Hq, D]; k, v [B, S, Hkv, D]; appends k/v at cache_seqlens[b] + i and attends
    causally over [0, cache_seqlens[b] + S).  Returns out [B, S, Hq, D]."""
    _req_cuda(q, k, v, k_cache, v_cache, rotary_cos, rotary_sin, cache_seqlens, block_table)
    B, S, Hq, D = q.shape
    n_blocks, block_size, Hkv, Dc = k_cache.shape
    assert Dc == D and v_cache.shape == k_cache.shape and k.shape == (B, S, Hkv, D) and v.shape == k.shape
    for t in (q, k, v, k_cache, v_cache):
        assert t.dtype == torch.bfloat16, "paged_attn_prefill: bf16 only"
        assert t.stride(-1) == 1
    assert cache_seqlens.dtype == torch.int32 and block_table.dtype == torch.int32
    assert cache_seqlens.is_contiguous() and block_table.stride(1) == 1
    if out is None:
        out = torch.empty((B, S, Hq, D), dtype=q.dtype, device=q.device)
    a = PrefillArgs()
    a.q, a.k, a.v, a.k_cache, a.v_cache, a.out = (t.data_ptr() for t in (q, k, v, k_cache, v_cache, out))
    if rotary_cos is not None:
        assert rotary_sin is not None and rotary_cos.shape == rotary_sin.shape
        assert rotary_cos.dtype == torch.bfloat16 and rotary_cos.is_contiguous() and rotary_sin.is_contiguous()
        a.rotary_cos, a.rotary_sin = rotary_cos.data_ptr(), rotary_sin.data_ptr()
        a.rot_rows, a.rotary_dim = rotary_cos.shape[0], 2 * rotary_cos.shape[1]
    a.rotary_interleaved = 1 if rotary_interleaved else 0
    a.cache_seqlens, a.block_table = cache_seqlens.data_ptr(), block_table.data_ptr()
    a.B, a.S, a.Hq, a.Hkv, a.D = B, S, Hq, Hkv, D
    a.n_blocks, a.block_size, a.max_blocks_per_seq = n_blocks, block_size, block_table.shape[1]
    cap = block_table.shape[1] * block_size
    a.max_prefix = int(min(cap, max_prefix if max_prefix is not None else cap))
    a.causal = 1 if causal else 0
    a.q_stride_b, a.q_stride_s, a.q_stride_h = q.stride(0), q.stride(1), q.stride(2)
    a.k_stride_b, a.k_stride_s, a.k_stride_h = k.stride(0), k.stride(1), k.stride(2)
    a.v_stride_b, a.v_stride_s, a.v_stride_h = v.stride(0), v.stride(1), v.stride(2)
    a.o_stride_b, a.o_stride_s, a.o_stride_h = out.stride(0), out.stride(1), out.stride(2)
    a.kc_stride_blk, a.kc_stride_tok, a.kc_stride_h = k_cache.stride(0), k_cache.stride(1), k_cache.stride(2)
    a.vc_stride_blk, a.vc_stride_tok, a.vc_stride_h = v_cache.stride(0), v_cache.stride(1), v_cache.stride(2)
    a.bt_stride_b = block_table.stride(0)
    a.variant = int(variant)
    a.softmax_scale = float(softmax_scale) if softmax_scale is not None else 1.0 / math.sqrt(D)
    need = lib.b200_prefill_workspace_bytes(B, S, Hq, D)
    if need > 0:
        ws = _workspace(q.device, need, "prefill")
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
    check(lib.b200_paged_attn_prefill(C.byref(a), _stream()), "b200_paged_attn_prefill")
    return out


# ------------------------------------------------------------------ greedy sampling
def greedy_argmax(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[r] = first index of max(logits[r, :]) as int64 (core/layer/sampler.py:21-43 at top_k = 1).
    logits [rows, vocab] bf16 or fp32, last dim contiguous; capturable (scratch is per device/stream,
    zeroed once, left zeroed by the kernel)."""
    _req_cuda(logits, out)
    assert logits.dim() == 2 and logits.stride(1) == 1, "greedy_argmax: [rows, vocab], last dim contiguous"
    assert logits.dtype in (torch.bfloat16, torch.float32), "greedy_argmax: bf16 / fp32 logits"
    rows, vocab = logits.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() >= rows
    ws = _workspace(logits.device, int(lib.b200_greedy_argmax_workspace_bytes(max(rows, 1))), "argmax")
    check(lib.b200_greedy_argmax(logits.data_ptr(), _wdt(logits), rows, vocab, logits.stride(0) if rows > 1 else max(logits.stride(0), vocab),
                                 out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "b200_greedy_argmax")
    return out[:rows]


def topk_topp_sample(logits: torch.Tensor, temperatures: Optional[torch.Tensor], top_ps: Optional[torch.Tensor],
                     top_k: int, uniforms: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                     debug: bool = False):
    """The reference sampler body (core/layer/sampler.py:21-43) in one launch: top-k of ``logits / temp``, softmax,
    top-p prefix mask (first candidate always kept), renormalise, inverse-CDF draw with ``uniforms [rows]`` in [0, 1)
    (``None`` -> 0: the arg-max).  ``debug`` also returns (candidate ids [rows, k] int32 sorted by probability,
    renormalised probabilities [rows, k] fp32) — what the parity tests compare with the eager restatement."""
    _req_cuda(logits, temperatures, top_ps, uniforms, out)
    assert logits.dim() == 2 and logits.stride(1) == 1, "topk_topp_sample: [rows, vocab], last dim contiguous"
    assert logits.dtype in (torch.bfloat16, torch.float32), "topk_topp_sample: bf16 / fp32 logits"
    rows, vocab = logits.shape
    k = min(int(top_k), vocab)
    for name, t in (("temperatures", temperatures), ("top_ps", top_ps), ("uniforms", uniforms)):
        assert t is None or (t.dtype == torch.float32 and t.is_contiguous() and t.numel() == rows), \
            f"topk_topp_sample: {name} must be fp32 [rows]"
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == rows
    di = torch.empty((rows, k), dtype=torch.int32, device=logits.device) if debug else None
    dp = torch.empty((rows, k), dtype=torch.float32, device=logits.device) if debug else None
    ptr = lambda t: t.data_ptr() if t is not None else None
    check(lib.b200_topk_topp_sample(logits.data_ptr(), _wdt(logits), rows, vocab,
                                    logits.stride(0) if rows > 1 else max(logits.stride(0), vocab), k,
                                    ptr(temperatures), ptr(top_ps), ptr(uniforms), out.data_ptr(), ptr(di), ptr(dp),
                                    _stream()), "b200_topk_topp_sample")
    return (out, di, dp) if debug else out


# ------------------------------------------------------------------ decode-sized GEMM (tcgen05)
STREAM_MAX_M = 128


def stream_gemm_ok(m: int, k: int, n: int) -> bool:
    """Shapes ``b200_stream_gemm`` accepts (1..128 rows, K % 64 == 0, N % 8 == 0)."""
    return 1 <= m <= STREAM_MAX_M and k >= 64 and k % 64 == 0 and n >= 8 and n % 8 == 0


def interleave_up_gate(up_w: torch.Tensor, gate_w: torch.Tensor) -> torch.Tensor:
    """Weight layout of the fused gate_up + SwiGLU product: rows ``2j`` / ``2j + 1`` = ``up_w[j]`` / ``gate_w[j]``
    (both ``[d, K]``, nn.Linear layout); the pair of an output column then sits in adjacent TMEM lanes."""
    assert up_w.shape == gate_w.shape and up_w.dim() == 2
    return torch.stack([up_w, gate_w], dim=1).reshape(2 * up_w.shape[0], up_w.shape[1]).contiguous()


def stream_gemm(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None,
                bias: Optional[torch.Tensor] = None, swiglu: bool = False, pdl: bool = False,
                grid: int = 0, l2_ahead_kb: int = 0) -> torch.Tensor:
    """out = x @ w.T (+ bias | SwiGLU): x [M, K] bf16 with M <= 128, w [N, K] bf16 (the nn.Linear layout).
    Streams ``w`` once at HBM speed on tcgen05 tensor cores (TMA ring, TMEM accumulators, stream-K over one
    persistent CTA per SM, deterministic); capturable.  ``bias`` is added as a second bf16 rounding like the
    reference's separate ``qkv + bias`` (core/layer/model_graph.py:147).  ``swiglu``: ``w`` is the interleaved
    up/gate matrix of :func:`interleave_up_gate` and ``out [M, N/2] = silu(gate) * up`` with the rounding points
    of ``matmul`` -> ``swiglu_gemm`` (:183-184).  ``pdl``: programmatic dependent launch (the kernel before this
    one in the stream must not write ``w`` / ``bias``); ``l2_ahead_kb`` then also prefetches that much of every CTA's
    weight share into L2 before the wait."""
    _req_cuda(x, w, out, bias)
    assert x.dim() == 2 and w.dim() == 2 and x.shape[1] == w.shape[1], "stream_gemm: x [M, K] @ w [N, K]^T"
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16, "stream_gemm: bf16 only"
    M, K = x.shape
    N = w.shape[0]
    assert stream_gemm_ok(M, K, N), f"stream_gemm: unsupported shape M={M} K={K} N={N}"
    assert x.stride(1) == 1 and w.stride(1) == 1
    assert not (swiglu and bias is not None), "stream_gemm: one epilogue at a time"
    n_out = N // 2 if swiglu else N
    if swiglu:
        assert N % 16 == 0
    if out is None:
        out = torch.empty((M, n_out), dtype=torch.bfloat16, device=x.device)
    assert out.dtype == torch.bfloat16 and out.shape == (M, n_out) and out.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.is_contiguous() and bias.numel() == N
    with torch.cuda.device(x.device):
        need = int(lib.b200_stream_gemm_workspace_bytes(M, N, int(grid)))
    ws = _workspace(x.device, need, "sgemm")
    epi = _lib.GEMM_EPI_SWIGLU if swiglu else (_lib.GEMM_EPI_BIAS if bias is not None else _lib.GEMM_EPI_NONE)
    # with PDL: 16 KB weight boxes per CTA pulled into L2 before the wait on the previous kernel (l2_ahead_kb per CTA)
    flags = (_lib.GEMM_FLAG_PDL | ((min(int(l2_ahead_kb) // 16, 0xffff)) << 8)) if pdl else 0
    check(lib.b200_stream_gemm(x.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None,
                               out.data_ptr(), ws.data_ptr(), ws.numel(), M, K, N,
                               x.stride(0) if M > 1 else max(x.stride(0), K), w.stride(0),
                               out.stride(0) if M > 1 else max(out.stride(0), n_out),
                               epi, flags, int(grid), _stream()), "b200_stream_gemm")
    return out


# ------------------------------------------------------------------ decode-sized GEMM (legacy mma.sync form)
SKINNY_MAX_M = 64


def skinny_gemm_ok(m: int, k: int, n: int) -> bool:
    """Shapes ``b200_skinny_gemm`` accepts (1..64 rows, K % 64 == 0, N % 64 == 0)."""
    return 1 <= m <= SKINNY_MAX_M and k > 0 and n > 0 and k % 64 == 0 and n % 64 == 0


def skinny_gemm(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None,
                bias: Optional[torch.Tensor] = None, ksplit: int = 0, variant: int = 0) -> torch.Tensor:
    """out = x @ w (+ bias): x [M, K] bf16 with M <= 64, w [K, N] bf16 row-major (the ``[in, out]`` layout of
    ``prepare_weights``, core/layer/model_graph.py:95-116).  Streams ``w`` once at HBM speed (split-K over
    all SMs, deterministic merge); capturable.  The bias is added as a second bf16 rounding, like the
    reference's separate ``qkv + bias`` (:147)."""
    _req_cuda(x, w, out, bias)
    assert x.dim() == 2 and w.dim() == 2 and x.shape[1] == w.shape[0], "skinny_gemm: x [M, K] @ w [K, N]"
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16, "skinny_gemm: bf16 only"
    M, K = x.shape
    N = w.shape[1]
    assert skinny_gemm_ok(M, K, N), f"skinny_gemm: unsupported shape M={M} K={K} N={N}"
    assert x.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    assert out.dtype == torch.bfloat16 and out.shape == (M, N) and out.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.is_contiguous() and bias.numel() == N
    if ksplit <= 0:
        with torch.cuda.device(x.device):
            ksplit = int(lib.b200_skinny_gemm_splits(M, K, N, int(variant)))
    ksplit = max(1, min(int(ksplit), K // 64))
    ws_ptr, ws_bytes = None, 0
    if ksplit > 1:
        ws = _workspace(x.device, int(lib.b200_skinny_gemm_workspace_bytes(M, N, ksplit)), "gemm")
        ws_ptr, ws_bytes = ws.data_ptr(), ws.numel()
    check(lib.b200_skinny_gemm(x.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None,
                               out.data_ptr(), ws_ptr, ws_bytes, M, K, N,
                               x.stride(0) if M > 1 else max(x.stride(0), K), w.stride(0),
                               out.stride(0) if M > 1 else max(out.stride(0), N),
                               int(ksplit), int(variant), _stream()), "b200_skinny_gemm")
    return out
