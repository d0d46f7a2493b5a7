"""Paged-KV attention operator on hand-written sm_100a kernels.

Drop-in for the reference's ``core/paged_attention.py`` and for the
``flash_attn_with_kvcache`` call sites it wraps:

  * ``flash_attn_with_kvcache(...)`` — same keyword surface the reference uses at
    ``core/layer/model_graph.py:161-168`` (decode), ``core/paged_attention.py:281-293``
    (class API) and ``core/layer/model_prefill.py:49-52`` (prefill).  seqlen_q == 1 goes
    to the fused RoPE + append + split-KV decode kernel, seqlen_q > 1 to the tcgen05
    prefill kernel.  Features the reference never uses raise instead of falling back.
  * ``PagedAttention`` — the class API (``paged_attention.py:140-295``): bf16 rotary
    pools ``_cos_pool/_sin_pool``, static q/k/v staging buffers, ``forward``.
  * ``PrecomputedRotaryEmbedding`` — the bf16 table, bit-identical to the reference's
    (``paged_attention.py:42-62``), quirk included: positions are ``arange(dtype=bf16)``,
    so only 897 distinct rows exist among the first 8192.

There is no flash-attn import, no Triton and no CPU path: without libb200ops.so
importing this module fails.
"""
from __future__ import annotations

import math
from typing import List, Optional, Union

import torch
import torch.nn as nn

import b200ops
from core.cache_manager import KVCacheManager, store_kvcache, store_kvcache_batch  # noqa: F401 (reference re-exports)

__all__ = ["flash_attn_with_kvcache", "PagedAttention", "PrecomputedRotaryEmbedding", "build_rotary_tables",
           "SPLIT_POLICY"]

# "auto": ignore the caller's ``num_splits`` and size the split-KV grid for 148 SMs from
#         (batch, kv heads, table capacity) — the reference's formula max(1, 32 // (4*bs))
#         (model_graph.py:167) leaves 256 CTAs for 148 SMs at bs=32.
# "exact": honour a caller-supplied ``num_splits > 0`` like flash-attn does.
SPLIT_POLICY = "auto"


def build_rotary_tables(dim: int, max_position: int, base: float = 10000.0):
    """cos/sin ``[max_position, dim]`` in bf16 with the reference's exact op sequence
    (paged_attention.py:50-62).  Always evaluated on the CPU so that every rank and the
    test oracle see bit-identical tables."""
    exponent = torch.arange(0, dim, 2).to(torch.bfloat16) / dim
    inv_freq = 1.0 / (base ** exponent)                                   # bf16
    t = torch.arange(max_position, dtype=inv_freq.dtype)                  # bf16 positions (the quirk)
    freqs = torch.einsum("i,j->ij", t, inv_freq)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().contiguous(), emb.sin().contiguous()


class PrecomputedRotaryEmbedding(nn.Module):
    def __init__(self, dim, max_position=8192, base=10000, device=None):
        super().__init__()
        self.dim = dim
        self.device = device or torch.device("cpu")
        self.max_position = max_position
        cos, sin = build_rotary_tables(dim, max_position, base)
        self.register_buffer("cos_cache", cos[None, None].to(self.device).contiguous())   # [1, 1, P, dim]
        self.register_buffer("sin_cache", sin[None, None].to(self.device).contiguous())

    def forward(self, x, positions):
        """Eager NeoX rotation of ``x [B, H, S, D]`` (reference :65-76); not on the hot path."""
        B, H, S, D = x.shape
        positions = positions.to(self.device)
        cos = self.cos_cache[:, :, positions].view(B, 1, S, D)
        sin = self.sin_cache[:, :, positions].view(B, 1, S, D)
        x1, x2 = x[..., : self.dim // 2], x[..., self.dim // 2:]
        return x * cos + torch.cat((-x2, x1), dim=-1) * sin


def flash_attn_with_kvcache(
        q: torch.Tensor,
        k_cache: torch.Tensor,
        v_cache: torch.Tensor,
        k: Optional[torch.Tensor] = None,
        v: Optional[torch.Tensor] = None,
        rotary_cos: Optional[torch.Tensor] = None,
        rotary_sin: Optional[torch.Tensor] = None,
        cache_seqlens: Optional[Union[int, torch.Tensor]] = None,
        cache_batch_idx: Optional[torch.Tensor] = None,
        cache_leftpad: Optional[torch.Tensor] = None,
        block_table: Optional[torch.Tensor] = None,
        softmax_scale: Optional[float] = None,
        causal: bool = False,
        window_size=(-1, -1),
        softcap: float = 0.0,
        rotary_interleaved: bool = True,
        alibi_slopes=None,
        num_splits: int = 0,
        return_softmax_lse: bool = False,
):
    """``q [B, Sq, Hq, D]``, paged caches ``[n_blocks, page, Hkv, D]``, ``k/v [B, Sq, Hkv, D]``.
    Updates the caches in place at ``cache_seqlens`` through ``block_table`` and returns
    ``[B, Sq, Hq, D]`` — the contract of flash-attn's function of the same name."""
    if alibi_slopes is not None:
        raise NotImplementedError("alibi_slopes is not supported by the B200 paged attention path")
    if tuple(window_size) != (-1, -1):
        raise NotImplementedError("sliding-window attention is not supported by the B200 paged attention path")
    if softcap != 0.0:
        raise NotImplementedError("softcap is not supported by the B200 paged attention path")
    if cache_batch_idx is not None or cache_leftpad is not None:
        raise NotImplementedError("cache_batch_idx / cache_leftpad are not supported by the B200 paged attention path")
    if block_table is None:
        raise NotImplementedError("only the paged KV cache (block_table) is supported")
    if cache_seqlens is None:
        raise ValueError("cache_seqlens is required")
    assert q.dim() == 4, "q must be [B, Sq, Hq, D]"
    B, Sq, Hq, D = q.shape
    if isinstance(cache_seqlens, int):
        cache_seqlens = torch.full((B,), cache_seqlens, dtype=torch.int32, device=q.device)
    if (k is None) != (v is None):
        raise ValueError("k and v must be passed together")
    if softmax_scale is None:
        softmax_scale = 1.0 / math.sqrt(D)
    cos = rotary_cos if k is not None else None          # rotary only applies when k/v are appended
    sin = rotary_sin if k is not None else None

    if Sq == 1:
        if return_softmax_lse:
            lse = torch.empty((B, Hq), dtype=torch.float32, device=q.device)
        else:
            lse = None
        splits = num_splits if (SPLIT_POLICY == "exact" and num_splits > 0) else 0
        out = b200ops.paged_attn_decode(
            q[:, 0], k_cache, v_cache,
            k[:, 0] if k is not None else None, v[:, 0] if v is not None else None,
            cos, sin, cache_seqlens, block_table,
            softmax_scale=softmax_scale, rotary_interleaved=rotary_interleaved,
            num_splits=splits, softmax_lse=lse)
        out = out.unsqueeze(1)
        return (out, lse.unsqueeze(-1)) if return_softmax_lse else out

    if return_softmax_lse:
        raise NotImplementedError("return_softmax_lse is only available for seqlen_q == 1")
    if k is None:
        raise NotImplementedError("seqlen_q > 1 without new k/v (pure cached multi-query attention) is not supported")
    return b200ops.paged_attn_prefill(q, k, v, k_cache, v_cache, cos, sin, cache_seqlens, block_table,
                                      softmax_scale=softmax_scale, causal=causal,
                                      rotary_interleaved=rotary_interleaved)


class PagedAttention(nn.Module):
    """Class API of the reference (paged_attention.py:140-295).

    ``rotary_interleaved`` makes explicit what the reference leaves to flash-attn's default:
    its ``_forward_`` (:281-293) does not pass the flag, so the class path rotates GPT-J pairs
    (default ``True``) while the engine's decode call site passes ``False`` (model_graph.py:166).
    """

    def __init__(self, num_heads: int, head_size: int, kv_num_heads: int, device: str = "auto",
                 max_batch_size=16, max_blocks=32, max_position=4096, max_tokens=8192, block_size=256,
                 rotary_interleaved: bool = True):
        super().__init__()
        self.block_size = block_size
        self.max_tokens = max_tokens
        self.max_seq_blocks = (max_tokens + block_size - 1) // block_size
        self.num_heads = num_heads
        self.kv_num_heads = kv_num_heads
        self.head_size = head_size
        self.scale = head_size ** -0.5
        self.rotary_interleaved = rotary_interleaved
        if device == "auto":
            if not torch.cuda.is_available():
                raise RuntimeError("PagedAttention needs a CUDA device (the B200 path has no CPU fallback)")
            device = "cuda"
        self.device = torch.device(device)
        self.max_batch_size = max_batch_size
        self.max_blocks = max_blocks

        max_kv_capacity = max_blocks * 256                        # reference :172 (hard-coded 256)
        self.rotary_emb = PrecomputedRotaryEmbedding(head_size, max_position=max_kv_capacity, device=self.device)
        half = head_size // 2
        self._cos_pool = self.rotary_emb.cos_cache[0, 0, :max_kv_capacity, :half].contiguous()
        self._sin_pool = self.rotary_emb.sin_cache[0, 0, :max_kv_capacity, :half].contiguous()
        self._block_table_pool = torch.full((max_batch_size, max_blocks), -1, dtype=torch.int32, device=self.device)

        # static staging buffers (fixed addresses for CUDA-graph capture); K/V use the KV head
        # count — the reference sizes them with num_heads because it only runs MHA (:198-210)
        bf16 = torch.bfloat16
        self._query_buffer = torch.empty((max_batch_size, num_heads, head_size), dtype=bf16, device=self.device)
        self._key_buffer = torch.empty((max_batch_size, kv_num_heads, head_size), dtype=bf16, device=self.device)
        self._value_buffer = torch.empty((max_batch_size, kv_num_heads, head_size), dtype=bf16, device=self.device)
        self._output_buffer = torch.empty_like(self._query_buffer)

    def forward(self, query: torch.Tensor, cache_manager: KVCacheManager, seq_ids: List[int],
                context_lens: List[int], layer_idx: int,
                key: Optional[torch.Tensor] = None, value: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``query [B, H, D]``, new ``key/value [B, Hkv, D]`` -> ``[B, H, D]``.  As in the reference,
        ``seq_ids`` / ``context_lens`` are not consulted: the block table and sequence lengths
        must already be staged with ``cache_manager.cache_batch_data`` (:264-265)."""
        B = query.shape[0]
        self._query_buffer[:B].copy_(query)
        self._key_buffer[:B].copy_(key)
        self._value_buffer[:B].copy_(value)
        return self._forward_(self._query_buffer[:B], self._key_buffer[:B], self._value_buffer[:B],
                              cache_manager, layer_idx,
                              cache_manager._block_table_buffer[:B], cache_manager._cache_seqlens_buffer[:B])

    def _forward_(self, query, key, value, cache_manager, layer_idx, block_table, cache_seqlens):
        k_cache, v_cache = cache_manager.get(layer_idx)
        B = query.shape[0]
        return b200ops.paged_attn_decode(
            query, k_cache, v_cache, key, value, self._cos_pool, self._sin_pool,
            cache_seqlens, block_table, softmax_scale=self.scale,
            rotary_interleaved=self.rotary_interleaved, out=self._output_buffer[:B])
