"""Fixed-length batched prefill runner.

Drop-in for the reference's ``core/layer/model_prefill.py`` (``ModelPrefillRunner`` :14-62):
``forward(input_ids [B, S], cache_manager, batch_size) -> logits [B, S, vocab]``.  The
reference rotates q/k with eager bf16 ``RoPE.forward`` (:46) and calls
``flash_attn_with_kvcache`` with ``cache_seqlens`` zeroed (:32, :49-52); here rotation, the
page scatter and causal attention are one operator (``b200ops.paged_attn_prefill``:
fused RoPE + append kernel, then the tcgen05/TMEM attention kernel), and the MLP is
matmul -> SwiGLU kernel -> matmul instead of the Inductor-compiled ``_mlp`` (:58).
"""
import torch

import b200ops
from core.parallel_config import all_reduce
from kernel.rmsnorm_add import rmsnorm, rmsnorm_residual_fused
from kernel.swiglu import swiglu_fused
from .model_graph import ModelGraphRunner


class ModelPrefillRunner(ModelGraphRunner):
    def forward(self, input_ids: torch.Tensor, cache_manager, batch_size: int) -> torch.Tensor:
        B, S = input_ids.shape
        Hq, Hkv, D = self.num_heads, self.kv_num_heads, self.head_size
        h = self.model.transformer.wte(input_ids)
        cache_lens = cache_manager._cache_seqlens_buffer[:batch_size].zero_()     # reference :32
        block_table = cache_manager._block_table_buffer[:batch_size]
        for layer_idx in range(self.num_layers):
            block = self.model.transformer.h[layer_idx]
            normed = rmsnorm(h, block.ln_1.weight, block.ln_1.eps)
            qkv = torch.matmul(normed, block.attn._qkv_w)
            if block.attn._qkv_b is not None:
                qkv += block.attn._qkv_b
            q = qkv[..., : Hq * D].view(B, S, Hq, D)
            k = qkv[..., Hq * D: (Hq + Hkv) * D].view(B, S, Hkv, D)
            v = qkv[..., (Hq + Hkv) * D:].view(B, S, Hkv, D)
            k_cache, v_cache = cache_manager.get(layer_idx)
            attn = b200ops.paged_attn_prefill(q, k, v, k_cache, v_cache,
                                              self.attention._cos_pool, self.attention._sin_pool,
                                              cache_lens, block_table, softmax_scale=self.attention.scale,
                                              causal=True, rotary_interleaved=False, max_prefix=0)
            out = all_reduce(torch.matmul(attn.view(B, S, -1), block.attn._o))
            normed, residual = rmsnorm_residual_fused(out, h, block.ln_2.weight, block.ln_2.eps)
            mlp_out = torch.matmul(swiglu_fused(torch.matmul(normed, block.mlp._gu)), block.mlp._d)
            h = all_reduce(mlp_out).add_(residual)
        return self.model.lm_head(self.model.transformer.ln_f(h))
