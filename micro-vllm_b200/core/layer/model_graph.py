"""Decode runner: static buffers, per-layer op sequence, CUDA-graph capture per batch size.

Drop-in for the reference's ``core/layer/model_graph.py`` (``ModelGraphRunner`` :25-260) with
the same constructor, ``prepare_weights`` / ``capture`` / ``forward`` / ``decode`` surface and
the same Qwen-1-style module tree (``transformer.h[i].attn.c_attn`` ...).  Differences, all on
the hot path this repo owns:

  * attention is ``b200ops.paged_attn_decode`` (fused RoPE + KV append + split-KV + merge,
    one launch) writing a STATIC output buffer, instead of ``flash_attn_with_kvcache``
    (:161-168) which allocates its output inside the captured region;
  * q/k/v are split GQA-aware (``Hq`` + 2 x ``Hkv`` heads); the reference's
    ``reshape(bs, 3, H, D)`` (:152) only works for MHA;
  * the MLP always runs matmul -> ``swiglu_gemm`` -> matmul (the reference's
    ``fast_mode=False`` branch, :183-186); the Inductor-compiled ``_mlp`` (:51, :199) is gone —
    no torch.compile on the path;
  * an un-captured batch size runs eagerly against the REAL block table (the reference
    substitutes an all-zero table there, :158/:257-258, which is only right during warm-up).

  * the two all-reduces of a layer and the elementwise kernels behind them are ONE launch each
    (``b200_allreduce_residual_rmsnorm``): ``all_reduce(attn_out)`` + ``rmsnorm_residual_gemm``
    (:210-216) and ``all_reduce(mlp_out).add_(residual)`` + the NEXT layer's ``rmsnorm_`` (:219,
    :143; the final ``ln_f`` :222 for the last layer).  With TP the o_proj / down_proj GEMMs write
    straight into NVLink peer-memory slots and the kernel sums all ranks' slots itself
    (``b200ops.comm.PeerComm``); ``B200_FUSED_ALLREDUCE=0`` keeps NCCL for the reduction.  The
    arithmetic (rounding points included) is the reference sequence's.

  * the four GEMMs of a decode layer (and ``lm_head``) run, for batches of up to 128 rows, on the
    tcgen05 weight-streaming kernel ``b200_stream_gemm`` over nn.Linear-layout
    (``[out, in]``) per-rank weight slices, with the QKV bias add (:145-147) and the SwiGLU launch
    (:183-184) folded into the epilogues; ``B200_STREAM_GEMM=0`` keeps ``torch.matmul`` (cuBLAS) on the
    reference's ``[in, out]`` weights plus the separate elementwise launches.
"""
import logging
import os
from typing import Dict, List

import torch

import b200ops
from core.paged_attention import PagedAttention
from b200ops.comm import allreduce_residual_rmsnorm_local
from core.parallel_config import all_reduce, get_rank, get_world_size, peer_comm
from kernel.rmsnorm_add import rmsnorm_
from kernel.swiglu import swiglu_gemm
from .rope import RoPE

logger = logging.getLogger(__name__)


def select_ar_mode(want: str, world_size: int, has_multicast: bool, ln_f_kernel_norm: bool, max_rows: int) -> str:
    """Form of the fused all-reduce + residual + RMSNorm the decode step uses (``B200_AR_MODE``, default "auto").

    ``ln_f_kernel_norm``: the model's final norm is a kernel norm, so EVERY reduction of the step is followed by a norm
    this library runs.  The last layer's reduction of a model whose ``ln_f`` is a plain module carries no norm
    (weight = None), and only the one-shot peer-load form accepts that: "sharded", "ll" and all NVLS forms (one-hop
    included — ``b200_allreduce_residual_rmsnorm`` refuses them without a weight) need the norm, so such a model always
    gets "oneshot".  The owner-rank forms ("sharded", "ll", "nvls", "nvlsp": each row reduced and normalised on ONE
    rank) are further limited to ``max_rows`` <= 1024; the one-hop form "nvls1" keeps the one-shot contract.

    Measured (profiles/r02/allreduce_tp{2,8}.json, 64 x 8192 bf16): 2 ranks — peer-load one-shot 7.0 us, one-hop NVLS
    9.2, two-hop NVLS 13.4, NCCL 19.1; 8 ranks — two-hop NVLS 13.5 us, one-hop NVLS 18.0 (every rank pulls the whole
    tensor through the switch: 8 MB leave each GPU), one-shot 21.7, LL 33.5, NCCL 34.1.  Whole L70-shaped step at
    TP = 8: 2.56 / 2.72 / 2.84 / 3.20 ms (nvls / nvls1 / oneshot / ll)."""
    if not ln_f_kernel_norm:
        return "oneshot"
    owner_ok = max_rows <= 1024
    if want == "auto":
        if world_size >= 4:
            want = ("nvls" if owner_ok else "nvls1") if has_multicast else "ll"
        else:
            want = "oneshot"
    if want in ("nvls", "nvls1", "nvlsp") and not has_multicast:
        want = "ll" if world_size >= 4 else "oneshot"
    if want == "nvls1" or (want in ("ll", "sharded", "nvls", "nvlsp") and owner_ok):
        return want
    return "oneshot"


class ModelGraphRunner:
    def __init__(self, model, num_layers: int, num_heads: int, head_size: int,
                 kv_num_heads: int, hidden_dim: int, intermediate_size: int,
                 device: str, max_batch_size: int = 32, dtype: torch.dtype = torch.bfloat16,
                 top_k: int = 1000, max_blocks: int = 32):
        """``num_heads`` / ``kv_num_heads`` are the LOCAL (per-rank) head counts and
        ``intermediate_size`` is the local width of the packed gate|up activation divided the
        way the engine does it (engine.py:108-114): i.e. this rank's ``2 * I / world_size``
        when the caller follows the reference and passes ``config.intermediate_size // ws``
        of a Qwen-1 config, whose ``intermediate_size`` already counts gate + up."""
        self.model = model
        self.num_layers = num_layers
        self.rank, self.world_size = get_rank(), get_world_size()
        self.num_heads = num_heads
        self.kv_num_heads = kv_num_heads
        self.intermediate_size = intermediate_size
        self.head_size = head_size
        self.hidden_dim = hidden_dim
        self.vocab_size = model.config.vocab_size
        self.device = device
        self.max_bs = max_batch_size
        self.dtype = dtype
        self.top_k = top_k

        # The kernel launched right before a layer's attention is its QKV projection, which never
        # writes the lengths, the block table or cached pages, so programmatic dependent launch
        # would be legal here; measured on B200 it does not pay (profiles/r01), so it stays off.
        self.use_pdl = int(os.environ.get("B200_DECODE_PDL", "0"))      # 1 / 2: see B200_DECODE_FLAG_PDL(_LATE); A/B switch
        # Decode GEMMs: B200_SKINNY_GEMM = "auto" (default) sends a product through the
        # weight-streaming split-K kernel b200_skinny_gemm only where it was measured faster than cuBLAS on B200
        # (profiles/r01/skinny_gemm_*.json): deep-K, narrow-N products at <= 32 rows — down_proj, where cuBLAS
        # launches N/64 CTAs without split-K and reaches 0.5 of the HBM roofline (36.7 us vs 25.7 us for
        # [32, 14336] x [14336, 4096]); "1" forces it for every shape it supports, "0" keeps torch.matmul everywhere.
        self._skinny = os.environ.get("B200_SKINNY_GEMM", "auto")
        # B200_STREAM_GEMM: "auto" (default) sends a decode GEMM of <= 128 rows through the tcgen05 weight-streaming
        # kernel where it was measured faster than cuBLAS on B200 (profiles/r02/stream_gemm_*.json): deep-K, narrow-N
        # products (K >= 4 N: the QKV projection of a TP shard, where cuBLAS has too few tiles and no split-K and sits at
        # 0.26 of the HBM roofline; K >= 2 N at <= 32 rows: down_proj of the 8B shape); "1" uses it for every shape it
        # accepts, with the bias / SwiGLU epilogues fused;
        # "0": cuBLAS + separate launches everywhere.  B200_STREAM_GEMM_PDL = 1 launches them with programmatic
        # dependent launch (weight boxes requested before the previous kernel has drained).
        self._stream_mode = os.environ.get("B200_STREAM_GEMM", "auto")
        self._stream = self._stream_mode != "0" and torch.device(device).type == "cuda"
        self._stream_pdl = os.environ.get("B200_STREAM_GEMM_PDL", "1") == "1"
        # With PDL a stream GEMM that follows a cross-GPU reduction (QKV, gate_up) also pulls this many KB of every
        # CTA's weight share into L2 before it waits for the reduction (B200_GEMM_FLAG_L2_AHEAD); the reductions release
        # their successor at their first instruction.  B200_AR_PDL = 1 launches the reductions themselves with PDL.
        self._l2_ahead_kb = int(os.environ.get("B200_GEMM_L2_AHEAD_KB", "256")) if self.world_size > 1 else 0
        self._ar_pdl = os.environ.get("B200_AR_PDL", "0") == "1"
        # TP only: while the NVLink reduction after o_proj / down_proj runs (a latency-bound kernel on a handful of SMs
        # that leaves HBM idle), a side stream asks L2 for the weights of the GEMMs behind it (b200_l2_prefetch), so the
        # reduction's latency is paid in parallel with weight bytes that have to come from HBM anyway.
        # B200_AR_PREFETCH_MB = how many MB per reduction (0 = off, the default: measured on 8 B200s the side-stream
        # launches cost more than the prefetch saves, 2.79 vs 2.72 ms per step; the PDL form of the stream GEMM does
        # the same job from inside the kernel that needs the weights).
        self._prefetch_bytes = int(float(os.environ.get("B200_AR_PREFETCH_MB", "0")) * 1e6) if self.world_size > 1 else 0
        self._side_stream = None
        self.attention = PagedAttention(num_heads, head_size, kv_num_heads, device, max_batch_size,
                                        max_blocks=max_blocks, rotary_interleaved=False)
        self.rope = RoPE()
        self.prepare_weights()
        self._alloc_bufs()
        self._graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self._graphs_balanced: Dict[int, torch.cuda.CUDAGraph] = {}
        self._graphs_multi: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._tok_hists = {}                # (bs, steps) -> id history buffer owned by that multi-step graph
        self._is_graph_ready = False

    # ------------------------------------------------------------------ init
    @staticmethod
    def _mlp(x, gu_weight, d_weight):
        """Eager statement of the MLP (reference :75-81): columns [0, d) of gate_up are ``up``."""
        gate_up = x @ gu_weight
        up, gate = gate_up.chunk(2, dim=-1)
        return (torch.nn.functional.silu(gate) * up) @ d_weight

    def prepare_weights(self):
        """Per-rank slices, transposed to [in, out] contiguous (reference :87-118)."""
        cfg = self.model.config
        g_heads = cfg.num_attention_heads
        g_kv = getattr(cfg, "num_key_value_heads", g_heads)
        q_dim, kv_dim = g_heads * self.head_size, g_kv * self.head_size
        ws, r = self.world_size, self.rank
        for block in self.model.transformer.h:
            w1, w2 = block.mlp.w1.weight, block.mlp.w2.weight
            block.mlp._gu = torch.cat([w1.chunk(ws, dim=0)[r], w2.chunk(ws, dim=0)[r]], dim=0).t().contiguous()
            block.mlp._d = block.mlp.c_proj.weight.chunk(ws, dim=1)[r].t().contiguous()
            block.attn._o = block.attn.c_proj.weight.chunk(ws, dim=1)[r].t().contiguous()
            w_q, w_k, w_v = block.attn.c_attn.weight.split([q_dim, kv_dim, kv_dim], dim=0)
            block.attn._qkv_w = torch.cat([w.chunk(ws, dim=0)[r] for w in (w_q, w_k, w_v)], dim=0).t().contiguous()
            bias = block.attn.c_attn.bias
            block.attn._qkv_b = None
            if bias is not None:
                b_q, b_k, b_v = bias.split([q_dim, kv_dim, kv_dim], dim=0)
                block.attn._qkv_b = torch.cat([b.chunk(ws, dim=0)[r] for b in (b_q, b_k, b_v)], dim=0)
            if self._stream:
                # nn.Linear layout ([out, in], K contiguous) per-rank slices for b200_stream_gemm, built only for the
                # products the policy sends there; at TP = 1 the q/k/v, o and down matrices are the model's own tensors
                # (no copy); gate_up is interleaved row-wise (up_j, gate_j).
                # (`.data`: plain tensors that keep the storage alive even if a caller later rebinds the Parameter)
                def want(n, k):
                    return self._stream_mode == "1" or k >= 2 * n       # _use_stream decides per batch size
                H = block.attn.c_attn.weight.shape[1]
                if want((q_dim + 2 * kv_dim) // ws, H):
                    block.attn._qkv_nk = (block.attn.c_attn.weight.data if ws == 1 else
                                          torch.cat([w.chunk(ws, dim=0)[r] for w in (w_q, w_k, w_v)], dim=0).contiguous())
                if want(H, q_dim // ws):
                    block.attn._o_nk = block.attn.c_proj.weight.data.chunk(ws, dim=1)[r].contiguous()
                if want(H, w1.shape[0] // ws):
                    block.mlp._d_nk = block.mlp.c_proj.weight.data.chunk(ws, dim=1)[r].contiguous()
                if want(2 * w1.shape[0] // ws, H):
                    block.mlp._gu_nk = b200ops.interleave_up_gate(w1.chunk(ws, dim=0)[r], w2.chunk(ws, dim=0)[r])

    def _alloc_bufs(self):
        mb, dev, dt = self.max_bs, self.device, self.dtype
        blk = self.model.transformer.h[0]
        qkv_dim, o_in = blk.attn._qkv_w.shape[1], blk.attn._o.shape[0]
        gu_dim = blk.mlp._gu.shape[1]
        self._input_ids = torch.empty(mb, dtype=torch.long, device=dev)
        self._next_ids = torch.zeros(mb, dtype=torch.long, device=dev)       # greedy ids (capture(greedy=True))
        self._logits = torch.empty(mb, self.vocab_size, dtype=dt, device=dev)
        self._normed = torch.empty((mb, self.hidden_dim), dtype=dt, device=dev)
        self._qkv = torch.empty((mb, qkv_dim), dtype=dt, device=dev)
        self._attn = torch.empty((mb, o_in), dtype=dt, device=dev)
        self._attn_out = torch.empty((mb, self.hidden_dim), dtype=dt, device=dev)
        self._residual = torch.empty((mb, self.hidden_dim), dtype=dt, device=dev)
        self._gate_up = torch.empty((mb, gu_dim), dtype=dt, device=dev)
        self._swiglu_out = torch.empty((mb, gu_dim // 2), dtype=dt, device=dev)
        self._mlp_out = torch.empty((mb, self.hidden_dim), dtype=dt, device=dev)
        self._h = torch.empty((mb, self.hidden_dim), dtype=dt, device=dev)
        # TP: the two GEMMs whose outputs are summed over ranks write into peer-visible slots
        # Vocab-parallel lm_head (B200_HEAD_SHARD, default on under TP): the reference keeps lm_head replicated
        # (models/adapter.py:25) — at TP = 8 on the 70B shape that is 2.1 GB streamed per step on every rank, 13 % of
        # the step.  Here rank r multiplies by rows [r V/ws, (r+1) V/ws) of the SAME weight and the logits are
        # assembled on every rank by b200_allgather_cols (peer stores / multimem.st): same [B, vocab] result.
        head = getattr(self.model, "lm_head", None)
        hw = getattr(head, "weight", None)
        shard_head = (self.world_size > 1 and os.environ.get("B200_HEAD_SHARD", "1") != "0" and hw is not None
                      and getattr(head, "bias", None) is None and hw.dim() == 2 and hw.is_contiguous()
                      and hw.shape[0] == self.vocab_size and self.vocab_size % (8 * self.world_size) == 0
                      and hw.dtype == torch.bfloat16)
        gather_bytes = mb * self.vocab_size * 2 if shard_head else 0
        self._comm = peer_comm(mb * self.hidden_dim * 2, gather_bytes) if self.world_size > 1 else None
        if self._comm is not None and not self._comm.fits(mb, self.hidden_dim):
            self._comm = None
        self._head_w = None
        if shard_head and self._comm is not None and self._comm.gather_bytes >= gather_bytes:
            vs = self.vocab_size // self.world_size
            self._head_w = hw.data[self.rank * vs:(self.rank + 1) * vs]          # a view: rows of the replicated matrix
            self._logits_local = torch.empty((mb, vs), dtype=dt, device=dev)
            self._logits = self._comm.gather_buf(mb, self.vocab_size)
        # Reduction form (B200_AR_MODE = oneshot | sharded | ll | nvls | nvlsp | nvls1; nvls1 = one-hop in-switch form,
        # nvlsp = two-hop in-switch form whose second hop carries flag-in-data packets instead of ending in a barrier).  Default: "nvls" (in-switch reduction,
        # multimem.ld_reduce / multimem.st on the multicast mapping of the symmetric slabs) when the communicator has
        # a multicast range and the final norm can be fused; else "ll" from 4 ranks up; else "oneshot".  "nvls", "ll"
        # and "sharded" reduce + normalise each row on ONE owner rank and keep the residual stream row-sharded, so
        # they need ln_f fused (a kernel norm); "ll" moves data as flag-in-data packets with remote stores only.
        self._ar_mode = "oneshot"
        if self._comm is not None:
            self._attn_out = self._comm.slot(0, mb, self.hidden_dim)
            self._mlp_out = self._comm.slot(1, mb, self.hidden_dim)
            want = os.environ.get("B200_AR_MODE", "auto")
            kernel_ln_f = bool(getattr(self.model.transformer.ln_f, "b200_kernel_norm", False))
            self._ar_mode = select_ar_mode(want, self.world_size, self._comm.has_multicast, kernel_ln_f, mb)
            if self._ar_mode in ("sharded", "nvls"):
                self._normed = self._comm.ybuf(mb, self.hidden_dim)
        self._zero_table = torch.zeros((mb, self.attention.max_blocks), dtype=torch.int32, device=dev)

    # ------------------------------------------------------------------ per-layer pieces
    def _mm(self, x, w, out, bias=None):
        """out = x @ w (+ bias as a separate bf16 add, reference :145-147).  Decode-sized products take the
        weight-streaming kernel when enabled; everything else is cuBLAS through torch.matmul."""
        m, (k, n) = x.shape[0], w.shape
        if self._skinny != "0" and b200ops.skinny_gemm_ok(m, k, n):
            if self._skinny == "1":
                return b200ops.skinny_gemm(x, w, out=out, bias=bias)
            if m <= 32 and k >= 2 * n:
                return b200ops.skinny_gemm(x, w, out=out, bias=bias, variant=1)      # 64-column slices
        torch.matmul(x, w, out=out)
        if bias is not None:
            out.add_(bias)
        return out

    def _use_stream(self, m, w_nk):
        if not (self._stream and w_nk is not None and b200ops.stream_gemm_ok(m, w_nk.shape[1], w_nk.shape[0])):
            return False
        # measured (profiles/r02/stream_gemm_*.json): faster than cuBLAS for deep-K narrow-N products — K >= 4 N at any
        # decode batch (QKV of a TP shard), K >= 2 N at <= 32 rows (Llama-3-8B down_proj: 24.6 us vs cuBLAS 28.6-32.5 us
        # and 25.6 us for the round-1 mma.sync kernel, which this retires from the default path)
        k, n = w_nk.shape[1], w_nk.shape[0]
        return self._stream_mode == "1" or k >= 4 * n or (m <= 32 and k >= 2 * n)

    def _qkv_proj(self, block, bs):
        """QKV projection of ``self._normed[:bs]`` (the ln_1 norm was fused into the previous launch)."""
        w_nk = getattr(block.attn, "_qkv_nk", None)
        if self._use_stream(bs, w_nk):
            return b200ops.stream_gemm(self._normed[:bs], w_nk, out=self._qkv[:bs], bias=block.attn._qkv_b,
                                       pdl=self._stream_pdl, l2_ahead_kb=self._l2_ahead_kb)
        return self._mm(self._normed[:bs], block.attn._qkv_w, self._qkv[:bs], block.attn._qkv_b)

    def _flash_attn(self, qkv, block, layer_idx, bs, cache_manager, use_graph_cache, balanced=False):
        Hq, Hkv, D = self.num_heads, self.kv_num_heads, self.head_size
        q = qkv[:, : Hq * D].view(bs, Hq, D)
        k = qkv[:, Hq * D: (Hq + Hkv) * D].view(bs, Hkv, D)
        v = qkv[:, (Hq + Hkv) * D:].view(bs, Hkv, D)
        k_cache, v_cache = cache_manager.get(layer_idx)
        cache_lens = cache_manager._cache_seqlens_buffer[:bs]
        # warm-up before capture runs against an all-zero table, like the reference (:158)
        block_table = cache_manager._block_table_buffer[:bs] if use_graph_cache else self._zero_table[:bs]
        attn = self._attn[:bs].view(bs, Hq, D)
        b200ops.paged_attn_decode(q, k_cache, v_cache, k, v,
                                  self.attention._cos_pool, self.attention._sin_pool,
                                  cache_lens, block_table, softmax_scale=self.attention.scale,
                                  rotary_interleaved=False, out=attn, pdl=0 if balanced else self.use_pdl,
                                  balanced=True if balanced else None)    # None: shape heuristic (tail waves)
        w_nk = getattr(block.attn, "_o_nk", None)
        if self._use_stream(bs, w_nk):
            return b200ops.stream_gemm(self._attn[:bs], w_nk, out=self._attn_out[:bs], pdl=self._stream_pdl)
        return self._mm(self._attn[:bs], block.attn._o, self._attn_out[:bs])

    def _ffn(self, x_normed, block, bs, fast_mode=False):
        gu_nk, d_nk = getattr(block.mlp, "_gu_nk", None), getattr(block.mlp, "_d_nk", None)
        if self._use_stream(bs, gu_nk) and gu_nk.shape[0] % 16 == 0:
            # gate_up projection + SwiGLU in one launch (reference :183-184)
            b200ops.stream_gemm(x_normed, gu_nk, out=self._swiglu_out[:bs], swiglu=True, pdl=self._stream_pdl,
                                l2_ahead_kb=self._l2_ahead_kb)
        else:
            self._mm(x_normed, block.mlp._gu, self._gate_up[:bs])
            swiglu_gemm(self._gate_up[:bs], self._swiglu_out[:bs])
        if self._use_stream(bs, d_nk):
            return b200ops.stream_gemm(self._swiglu_out[:bs], d_nk, out=self._mlp_out[:bs], pdl=self._stream_pdl)
        return self._mm(self._swiglu_out[:bs], block.mlp._d, self._mlp_out[:bs])

    def head_policy(self) -> str:
        if self._head_w is not None:
            return (f"vocab-parallel: {self._head_w.shape[0]} of {self.vocab_size} rows per rank + b200_allgather_cols "
                    f"({'multimem.st' if self._comm.has_multicast else 'peer stores'})")
        return "replicated (reference: models/adapter.py:25)"

    def _lm_head(self, x):
        """``lm_head`` (reference :223): the weight is already in the nn.Linear layout the stream GEMM takes."""
        if self._head_w is not None:
            bs = x.shape[0]
            torch.matmul(x, self._head_w.t(), out=self._logits_local[:bs])
            return self._comm.allgather_cols(self._logits_local[:bs], self._logits[:bs])
        head = self.model.lm_head
        w = getattr(head, "weight", None)
        if (isinstance(head, torch.nn.Linear) and head.bias is None and w is not None and w.dtype == torch.bfloat16
                and w.is_contiguous() and self._use_stream(x.shape[0], w) and os.environ.get("B200_STREAM_LM_HEAD", "1") != "0"):
            return b200ops.stream_gemm(x, w, out=self._logits[:x.shape[0]], pdl=self._stream_pdl)
        return head(x)

    # ------------------------------------------------------------------ weight prefetch beside the reductions
    def _prefetch_fork(self, weights):
        """Start pulling ``weights`` (list of (tensor, per_row)) into L2 on the side stream, ordered after everything
        already queued on the current stream (the GEMM whose output is about to be reduced).  Returns True if a join
        is owed."""
        budget = self._prefetch_bytes
        if budget <= 0 or not weights:
            return False
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(device=self.device)
        cur = torch.cuda.current_stream()
        self._side_stream.wait_stream(cur)
        with torch.cuda.stream(self._side_stream):
            for w, per_row in weights:
                if budget <= 0 or w is None:
                    continue
                n = min(budget, w.numel() * w.element_size())
                b200ops.l2_prefetch(w, n, per_row=per_row)
                budget -= n
        return True

    def _prefetch_join(self, owed):
        if owed:
            torch.cuda.current_stream().wait_stream(self._side_stream)

    def _gemm_weight(self, mod, nk_name, kn_name, bs):
        """(tensor, per_row) the GEMM of this module will stream for a batch of ``bs`` rows: the nn.Linear-layout slice
        when the stream GEMM takes the product (stream-K shares are contiguous per CTA: prefetch a k-slab of every
        row), else the reference-layout [in, out] matrix for cuBLAS (tiles walk k from row 0: prefetch a prefix)."""
        w_nk = getattr(mod, nk_name, None)
        if self._use_stream(bs, w_nk):
            full = w_nk.numel() * w_nk.element_size() <= self._prefetch_bytes
            return (w_nk, not full)
        return (getattr(mod, kn_name), False)

    # ------------------------------------------------------------------ decode step
    def _reduce_add_norm(self, slot, x, residual, norm, y, res_out, ss_rounded):
        """res_out = all_reduce(x) + residual;  y = rmsnorm(res_out) * norm.weight — one launch.
        ``norm`` None skips the norm.  ``x`` is peer slot ``slot`` when the communicator is up."""
        w = norm.weight if norm is not None else None
        eps = norm.eps if norm is not None else 0.0
        if self._comm is not None:
            self._comm.allreduce_residual_rmsnorm(slot, x.shape[0], self.hidden_dim, residual, w,
                                                  y if norm is not None else None, res_out, eps, ss_rounded,
                                                  sharded=self._ar_mode == "sharded", ll=self._ar_mode == "ll",
                                                  nvls=self._ar_mode in ("nvls", "nvlsp"), pdl=self._ar_pdl,
                                                  nvls1=self._ar_mode == "nvls1",
                                                  packet_tail=self._ar_mode == "nvlsp")
        else:
            all_reduce(x)                                   # NCCL when TP without peer memory; no-op at TP=1
            allreduce_residual_rmsnorm_local(x, residual, w, y if norm is not None else None, res_out, eps,
                                             ss_rounded)

    def decode(self, input_ids, bs, cache_manager, use_graph_cache, balanced=False):
        """``balanced`` selects the persistent stream-K attention schedule (ragged batches)."""
        tr = self.model.transformer
        h = tr.wte(input_ids)
        normed, residual, h_buf = self._normed[:bs], self._residual[:bs], self._h[:bs]
        rmsnorm_(h, tr.h[0].ln_1.weight, normed, tr.h[0].ln_1.eps)
        ln_f = tr.ln_f
        fuse_ln_f = bool(getattr(ln_f, "b200_kernel_norm", False))      # any other module is called as is
        for layer_idx in range(self.num_layers):
            block = tr.h[layer_idx]
            qkv = self._qkv_proj(block, bs)
            attn_out = self._flash_attn(qkv, block, layer_idx, bs, cache_manager, use_graph_cache, balanced)
            owed = self._prefetch_fork([self._gemm_weight(block.mlp, "_gu_nk", "_gu", bs)]) if self._comm is not None else False
            # all_reduce + rmsnorm_residual_gemm (reference :210-216): sum of squares of the unrounded sum
            self._reduce_add_norm(0, attn_out, h, block.ln_2, normed, residual, ss_rounded=False)
            self._prefetch_join(owed)
            mlp_out = self._ffn(normed, block, bs)
            # all_reduce(mlp_out).add_(residual) + the next rmsnorm_ (reference :219, :143 / :222)
            last = layer_idx + 1 == self.num_layers
            nxt = (ln_f if fuse_ln_f else None) if last else tr.h[layer_idx + 1].ln_1
            owed = False
            if self._comm is not None:
                if last:
                    lw = getattr(self.model.lm_head, "weight", None)
                    owed = self._prefetch_fork([(lw.data, False)] if lw is not None and lw.dim() == 2 and lw.is_contiguous() else [])
                else:
                    nb = tr.h[layer_idx + 1]
                    owed = self._prefetch_fork([self._gemm_weight(nb.attn, "_qkv_nk", "_qkv_w", bs),
                                                self._gemm_weight(nb.attn, "_o_nk", "_o", bs)])
            self._reduce_add_norm(1, mlp_out, residual, nxt, normed, h_buf, ss_rounded=True)
            self._prefetch_join(owed)
            h = h_buf
        return self._lm_head(normed if fuse_ln_f else ln_f(h))

    # A batch whose longest sequence exceeds this multiple of the mean length is "ragged": the split-KV grid
    # (one launch-constant split count for every row) leaves most SMs idle behind the long rows, the balanced
    # persistent schedule does not (profiles/r01: 64 seqs log-uniform 128-32k, 413 us -> 147 us per layer call;
    # on uniform batches the split grid stays ~20 % faster, so it remains the default).
    RAGGED_RATIO = 1.5

    def capture(self, cache_manager, batch_sizes: List[int] = [1, 2, 4, 8, 16, 32], ragged: bool = False,
                greedy: bool = False):
        """3 eager warm-ups then ``torch.cuda.graph`` per batch size (reference :225-253).  ``ragged`` also
        captures, for every batch size > 1, a second graph whose attention uses the balanced schedule;
        ``forward`` replays it when the cache manager reports a ragged batch.  ``greedy`` appends the row
        arg-max of the logits (``b200_greedy_argmax``: the reference sampler at top_k = 1, sampler.py:21-43)
        as the last launch of every graph, so ``forward_greedy`` returns token ids without the logits
        leaving the device."""
        if self._is_graph_ready:
            return
        self._greedy_in_graph = bool(greedy)
        for bs in batch_sizes:
            for balanced in ((False, True) if ragged and bs > 1 and bs <= 512 else (False,)):
                dummy = torch.randint(0, self.vocab_size, (bs,), device=self.device)
                with torch.no_grad():
                    for _ in range(3):
                        lg = self.decode(dummy, bs, cache_manager, use_graph_cache=False, balanced=balanced)
                        if greedy:                                   # the warm-up also reserves the arg-max scratch
                            b200ops.greedy_argmax(lg, self._next_ids[:bs])
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.no_grad(), torch.cuda.graph(g):
                    lg = self.decode(self._input_ids[:bs], bs, cache_manager, use_graph_cache=True, balanced=balanced)
                    if lg.data_ptr() != self._logits.data_ptr():     # lm_head through cuBLAS allocates its own output
                        self._logits[:bs] = lg
                    if greedy:
                        b200ops.greedy_argmax(self._logits[:bs], self._next_ids[:bs])
                if balanced:
                    self._graphs_balanced[bs] = g
                else:
                    self._graphs[bs] = g
                logger.info("captured decode graph for batch size %d%s", bs, " (balanced)" if balanced else "")
        self._is_graph_ready = True

    def _is_ragged(self, cache_manager, batch_size: int) -> bool:
        ratio = getattr(cache_manager, "len_max_over_mean", None)     # set by KVCacheManager.cache_batch_data
        return batch_size > 1 and ratio is not None and ratio >= self.RAGGED_RATIO

    def gemm_policy(self) -> str:
        """Which kernels run the four decode GEMMs (reported by the bench tools)."""
        if self._stream:
            which = "every decode GEMM" if self._stream_mode == "1" else "K >= 4N products (QKV of a TP shard) and K >= 2N at <= 32 rows (down_proj), cuBLAS elsewhere"
            return (f"tcgen05 stream_gemm for {which}" +
                    (f" + PDL (L2 ahead {self._l2_ahead_kb} KB/CTA behind reductions)" if self._stream_pdl else "") +
                    (" + PDL reductions" if self._ar_pdl and self.world_size > 1 else ""))
        return f"cuBLAS (torch.matmul), B200_SKINNY_GEMM={self._skinny}"

    def forward(self, input_ids: torch.Tensor, cache_manager, batch_size: int) -> torch.Tensor:
        if self._comm is not None:
            # a peer that missed a reduction (flag wait timed out) leaves this rank's residual stream wrong:
            # surface it here, without waiting for the device (reads the word an earlier async copy brought back)
            self._comm.poll_error()
        self._input_ids[:batch_size] = input_ids
        ragged = self._is_ragged(cache_manager, batch_size)
        g = (self._graphs_balanced.get(batch_size) if ragged else None) or self._graphs.get(batch_size)
        self._replayed = g is not None
        if g is None:
            with torch.no_grad():
                return self.decode(input_ids, batch_size, cache_manager, use_graph_cache=True,
                                   balanced=ragged and batch_size <= 512)
        g.replay()
        return self._logits[:batch_size]

    # ------------------------------------------------------------------ several greedy steps in ONE graph
    def capture_multistep(self, cache_manager, batch_size: int, steps: int):
        """Capture ``steps`` greedy decode steps of ``batch_size`` rows as ONE CUDA graph.  Between the steps nothing
        returns to the host: the greedy ids (``b200_greedy_argmax``) feed the next step's input on the device, the
        sequence lengths advance on the device and the pages the steps will need are written into the device block
        table by the graph's first node from the update list ``KVCacheManager.append_on_device`` uploaded
        (``b200_table_advance``).  Replaces ``steps`` round trips of engine.py:248-286 (append + cache_batch_data +
        replay + sampler + broadcast)."""
        bs = batch_size
        key = (bs, steps)
        if key in self._graphs_multi:
            return
        # one history buffer per captured graph, kept for the graph's lifetime: a graph holds raw addresses, so a
        # buffer it writes must never be replaced (and freed) when a later capture needs more rows
        hist = torch.zeros((steps, bs), dtype=torch.long, device=self.device)
        seqlens, table = cache_manager._cache_seqlens_buffer, cache_manager._block_table_buffer
        dummy = torch.randint(0, self.vocab_size, (bs,), device=self.device)
        with torch.no_grad():
            for _ in range(3):                                   # eager warm-up against the all-zero table, as capture()
                lg = self.decode(dummy, bs, cache_manager, use_graph_cache=False)
                b200ops.greedy_argmax(lg, self._next_ids[:bs])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(g):
            b200ops.table_advance(seqlens, bs, 0, table, cache_manager._updates_dev)    # pages handed out by the host
            for s in range(steps):
                lg = self.decode(self._input_ids[:bs], bs, cache_manager, use_graph_cache=True)
                if lg.data_ptr() != self._logits.data_ptr():
                    self._logits[:bs] = lg
                b200ops.greedy_argmax(self._logits[:bs], self._next_ids[:bs])
                hist[s].copy_(self._next_ids[:bs])
                self._input_ids[:bs].copy_(self._next_ids[:bs])
                b200ops.table_advance(seqlens, bs, 1)              # the token just cached is part of the context now
        self._graphs_multi[key] = g
        self._tok_hists[key] = hist

    def forward_greedy_multistep(self, input_ids: torch.Tensor, cache_manager, seq_ids: list, steps: int) -> torch.Tensor:
        """``steps`` greedy decode steps for the sequences ``seq_ids`` (rows of the batch in this order, lengths and
        table already on the device from ``cache_batch_data``): host allocator bookkeeping for all steps, ONE small
        upload of the new pages (if any), ONE graph replay.  Returns the ids ``[batch, steps]`` (a view of a static
        buffer); the device lengths have advanced by ``steps``."""
        bs = len(seq_ids)
        key = (bs, steps)
        if key not in self._graphs_multi:
            self.capture_multistep(cache_manager, bs, steps)
        if self._comm is not None:
            self._comm.poll_error()
        if cache_manager.append_on_device(seq_ids, steps) < 0:
            raise RuntimeError("OOM: KV cache pool exhausted")           # engine.py:220 convention
        self._input_ids[:bs] = input_ids
        self._graphs_multi[key].replay()
        cache_manager.mark_updates_applied()                 # the graph's first node consumed the update list
        return self._tok_hists[key].t()

    def forward_greedy(self, input_ids: torch.Tensor, cache_manager, batch_size: int) -> torch.Tensor:
        """One decode step returning the greedy next-token ids ``[batch_size]`` int64 (a view of a static
        buffer).  Inside the graph when captured with ``greedy=True``; otherwise one extra launch."""
        logits = self.forward(input_ids, cache_manager, batch_size)
        if not (self._replayed and getattr(self, "_greedy_in_graph", False)):
            b200ops.greedy_argmax(logits, self._next_ids[:batch_size])
        return self._next_ids[:batch_size]
