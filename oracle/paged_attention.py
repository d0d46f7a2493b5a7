"""fp32 gather + softmax restatement of paged attention with in-place KV append
(test infrastructure only).

Follows the reference's call sites — decode core/layer/model_graph.py:150-172 (NeoX rotary,
causal no-op at Sq = 1) and core/paged_attention.py:281-293 (interleaved default), prefill
core/layer/model_prefill.py:45-52 — and flash-attn's published contract for
``flash_attn_with_kvcache`` (flash-attn 2.8.3, flash_attn_interface.py docstring):
  * k_new is rotated at positions cache_seqlens[b] + i, q at the same positions (causal),
  * rotated k_new and raw v_new are written in place at those positions through
    block_table[b, pos // page] / pos % page,
  * q head h attends kv head h // (Hq / Hkv),
  * causal mask is bottom-right aligned; softmax_scale defaults to D ** -0.5,
  * scores run over positions [0, cache_seqlens[b] + S) INCLUDING slots that were never
    written (whatever bytes the pages hold).
All arithmetic is fp32; the only roundings are the bf16 stores the operator itself performs
(rotated q / k into bf16, output left in fp32 for tolerance checks).
"""
import math
from typing import Optional

import torch

from .rotary import rotate_fp32


def store_kvcache_oracle(key, value, k_cache, v_cache, slot_mapping, block_size):
    """core/cache_manager.py:171-178 — per-token scatter, slot -1 skipped."""
    for i, slot in enumerate(slot_mapping.tolist()):
        if slot != -1:
            blk, off = divmod(slot, block_size)
            k_cache[blk, off] = key[i]
            v_cache[blk, off] = value[i]


def _gather(cache: torch.Tensor, table_row: torch.Tensor, n: int) -> torch.Tensor:
    """Rows [0, n) of one sequence, [n, Hkv, D], through its block-table row."""
    page = cache.shape[1]
    n_pages = -(-n // page)
    ids = table_row[:n_pages].long()
    return cache[ids].reshape(n_pages * page, *cache.shape[2:])[:n]


def paged_attention_oracle(q, k_cache, v_cache, k_new, v_new, cos, sin, cache_seqlens, block_table,
                           softmax_scale: Optional[float] = None, causal: bool = True,
                           interleaved: bool = False, return_lse: bool = False):
    """q [B, S, Hq, D] (bf16), k_new/v_new [B, S, Hkv, D] or None, caches
    [n_blocks, page, Hkv, D] (bf16, UPDATED IN PLACE), cos/sin [rows, rd/2] or None.
    Returns out fp32 [B, S, Hq, D] (and lse [B, Hq, S] natural log)."""
    B, S, Hq, D = q.shape
    Hkv = k_cache.shape[2]
    G = Hq // Hkv
    page = k_cache.shape[1]
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(D)
    out = torch.zeros((B, S, Hq, D), dtype=torch.float32)
    lse = torch.full((B, Hq, S), float("-inf"), dtype=torch.float32)
    lens = cache_seqlens.tolist()
    for b in range(B):
        P = int(lens[b])
        qb = q[b]
        if k_new is not None:
            pos = torch.arange(P, P + S)
            if cos is not None:
                pidx = pos.clamp(max=cos.shape[0] - 1)
                c, s = cos[pidx][:, None, :], sin[pidx][:, None, :]
                k_rot = rotate_fp32(k_new[b], c, s, interleaved).to(k_cache.dtype)
                qb = rotate_fp32(q[b], c, s, interleaved).to(q.dtype)
            else:
                k_rot = k_new[b]
            for i in range(S):
                p_i = P + i
                blk = int(block_table[b, p_i // page])
                k_cache[blk, p_i % page] = k_rot[i]
                v_cache[blk, p_i % page] = v_new[b, i]
            n = P + S
        else:
            n = P
        if n == 0:
            continue
        K = _gather(k_cache, block_table[b], n).float()           # [n, Hkv, D]
        V = _gather(v_cache, block_table[b], n).float()
        K = K.repeat_interleave(G, dim=1)                          # head h -> kv head h // G
        V = V.repeat_interleave(G, dim=1)
        scores = torch.einsum("shd,nhd->hsn", qb.float(), K) * scale     # [Hq, S, n]
        if causal:
            # bottom-right aligned: query i (absolute position n - S + i) sees keys j <= n - S + i
            qpos = torch.arange(n - S, n)[:, None]
            kpos = torch.arange(n)[None, :]
            scores = scores.masked_fill(kpos > qpos, float("-inf"))
        lse_b = torch.logsumexp(scores, dim=-1)                    # [Hq, S]
        probs = torch.exp(scores - lse_b[..., None])
        probs = torch.nan_to_num(probs, nan=0.0)                   # fully masked rows -> 0 output
        out[b] = torch.einsum("hsn,nhd->shd", probs, V)
        lse[b] = lse_b
    return (out, lse) if return_lse else out


def decode_algorithmic_bytes(seqlens_after, Hq, Hkv, D, block_size=256, elt=2):
    """SURVEY.md §8 M4 / BASELINE.md §3: K and V read once per KV head over len_b + 1 tokens,
    new K,V written, q read + out written, seqlens + used block-table entries."""
    total = 0
    B = len(seqlens_after)
    for n in seqlens_after:
        total += n * 2 * Hkv * D * elt
        total += 4 * (-(-n // block_size))
    total += B * 2 * Hkv * D * elt
    total += 2 * B * Hq * D * elt
    total += 4 * B
    return total
