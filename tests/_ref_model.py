"""Reference-semantics forward of the synthetic model tree (test infrastructure).

Same GEMM path as the product (torch.matmul in the model dtype on the same device) with every
op on the hot path replaced by its oracle restatement — i.e. what the reference's runners
compute (core/layer/model_prefill.py:23-62, core/layer/model_graph.py:192-223) with
flash-attn / Triton swapped for oracle/*.  Driven exactly like the engine (engine.py:218-222,
:259-262): alloc(P), prefill with cache_seqlens = 0, then per decode step
append + cache_batch_data(current_position) with current_position = P + 1 + step."""
import torch

from oracle.elementwise import rmsnorm_oracle, rmsnorm_residual_oracle, swiglu_oracle
from oracle.paged_attention import paged_attention_oracle
from oracle.rotary import rope_eager_oracle


class RefPathModel:
    def __init__(self, runner, cos_pool, sin_pool):
        self.r = runner                      # a constructed ModelGraphRunner (weights prepared, TP sliced)
        self.cos, self.sin = cos_pool, sin_pool

    def _split(self, qkv, lead):
        r = self.r
        Hq, Hkv, D = r.num_heads, r.kv_num_heads, r.head_size
        q = qkv[..., : Hq * D].reshape(*lead, Hq, D)
        k = qkv[..., Hq * D: (Hq + Hkv) * D].reshape(*lead, Hkv, D)
        v = qkv[..., (Hq + Hkv) * D:].reshape(*lead, Hkv, D)
        return q, k, v

    def _mlp(self, normed, block):
        return torch.matmul(swiglu_oracle(torch.matmul(normed, block.mlp._gu)), block.mlp._d)

    @torch.no_grad()
    def prefill(self, input_ids, k_caches, v_caches, block_table):
        """input_ids [B, S]; caches: per-layer CPU/GPU tensors updated in place."""
        m = self.r.model
        B, S = input_ids.shape
        h = m.transformer.wte(input_ids)
        zeros = torch.zeros(B, dtype=torch.int32)
        for li, block in enumerate(m.transformer.h):
            normed = rmsnorm_oracle(h, block.ln_1.weight, block.ln_1.eps)
            qkv = torch.matmul(normed, block.attn._qkv_w)
            if block.attn._qkv_b is not None:
                qkv = qkv + block.attn._qkv_b
            q, k, v = self._split(qkv, (B, S))
            q, k = rope_eager_oracle(q, k, self.cos, self.sin)              # model_prefill.py:46 (bf16 eager)
            attn = paged_attention_oracle(q.cpu(), k_caches[li], v_caches[li], k.cpu(), v.cpu(), None, None,
                                          zeros, block_table, None, True, False).to(h.dtype).to(h.device)
            out = torch.matmul(attn.reshape(B, S, -1), block.attn._o)
            normed, residual = rmsnorm_residual_oracle(out, h, block.ln_2.weight, block.ln_2.eps)
            h = self._mlp(normed, block) + residual
        h = rmsnorm_oracle(h, m.transformer.ln_f.weight, m.transformer.ln_f.eps)
        return m.lm_head(h)

    @torch.no_grad()
    def decode(self, input_ids, k_caches, v_caches, block_table, cache_seqlens):
        """input_ids [B]; one decode step (model_graph.py:192-223)."""
        m = self.r.model
        B = input_ids.shape[0]
        h = m.transformer.wte(input_ids)
        for li, block in enumerate(m.transformer.h):
            normed = rmsnorm_oracle(h, block.ln_1.weight, block.ln_1.eps)
            qkv = torch.matmul(normed, block.attn._qkv_w)
            if block.attn._qkv_b is not None:
                qkv = qkv + block.attn._qkv_b
            q, k, v = self._split(qkv, (B,))
            attn = paged_attention_oracle(q[:, None].cpu(), k_caches[li], v_caches[li], k[:, None].cpu(),
                                          v[:, None].cpu(), self.cos.cpu(), self.sin.cpu(), cache_seqlens,
                                          block_table, None, True, False)[:, 0].to(h.dtype).to(h.device)
            out = torch.matmul(attn.reshape(B, -1), block.attn._o)
            normed, residual = rmsnorm_residual_oracle(out, h, block.ln_2.weight, block.ln_2.eps)
            h = self._mlp(normed, block) + residual
        h = rmsnorm_oracle(h, m.transformer.ln_f.weight, m.transformer.ln_f.eps)
        return m.lm_head(h)
