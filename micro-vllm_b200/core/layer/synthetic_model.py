"""Random-init module trees in the layout the runners consume.

The reference's runners walk a Qwen-1 style tree (``transformer.wte``, ``transformer.h[i]``
with ``ln_1 / attn.c_attn / attn.c_proj / ln_2 / mlp.w1 / mlp.w2 / mlp.c_proj``,
``transformer.ln_f``, ``lm_head``; core/layer/model_graph.py:95-116, :196-223).  BASELINE.json's
configs are GQA Llama-3 / Qwen2.5 *shapes* with random-init weights, so this builds that tree
for an arbitrary shape (no checkpoints, no network).  ``mlp.w1`` is the ``up`` projection and
``mlp.w2`` the ``gate`` projection (model_graph.py:78-81, :98-101).
"""
from __future__ import annotations

from dataclasses import dataclass
from types import SimpleNamespace

import torch
import torch.nn as nn


@dataclass(frozen=True)
class Shape:
    name: str
    hidden: int
    layers: int
    heads: int
    kv_heads: int
    head_dim: int
    ffn: int
    vocab: int
    qkv_bias: bool
    eps: float = 1e-6


SHAPES = {
    "llama3-8b": Shape("llama3-8b", 4096, 32, 32, 8, 128, 14336, 128256, False, 1e-5),
    "llama3-70b": Shape("llama3-70b", 8192, 80, 64, 8, 128, 28672, 128256, False, 1e-5),
    "qwen2.5-7b": Shape("qwen2.5-7b", 3584, 28, 28, 4, 128, 18944, 152064, True, 1e-6),
    "qwen2.5-0.5b": Shape("qwen2.5-0.5b", 896, 24, 14, 2, 64, 4864, 151936, True, 1e-6),
}


class RMSNormModule(nn.Module):
    """``weight`` + ``eps`` holder; calling it runs the sm_100a RMSNorm kernel (used as ``ln_f``)."""
    b200_kernel_norm = True        # same arithmetic as kernel.rmsnorm_add.rmsnorm: the decode runner may fuse it

    def __init__(self, hidden: int, eps: float, dtype, device):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device), requires_grad=False)
        self.eps = eps

    def forward(self, x):
        from kernel.rmsnorm_add import rmsnorm
        return rmsnorm(x, self.weight, self.eps)


def _randn(shape, gen, device, fast):
    if fast:                                   # device-side generator: seconds instead of minutes for 70B-shaped layers
        return torch.randn(*shape, generator=gen, device=device, dtype=torch.float32)
    return torch.randn(*shape, generator=gen, dtype=torch.float32)


def _linear(i, o, bias, dtype, device, gen, std, fast=False):
    lin = nn.Linear(i, o, bias=bias, dtype=dtype, device=device)
    with torch.no_grad():
        lin.weight.copy_((_randn((o, i), gen, device, fast) * std).to(dtype))
        if bias:
            lin.bias.copy_((_randn((o,), gen, device, fast) * std).to(dtype))
    lin.requires_grad_(False)
    return lin


def build_synthetic_model(shape: Shape, n_layers: int | None = None, device="cuda", dtype=torch.bfloat16,
                          seed: int = 0, std: float = 0.02, vocab: int | None = None,
                          norm_weight_jitter: float = 0.1, fast_init: bool = False, peaky: bool = False):
    """Deterministic random-init tree; the same seed gives bit-identical weights on every rank
    (CPU generator by default; ``fast_init`` draws on the device for the big bench shapes)."""
    gen = (torch.Generator(device=device) if fast_init else torch.Generator()).manual_seed(seed)
    L = n_layers if n_layers is not None else shape.layers
    V = vocab if vocab is not None else shape.vocab
    H, D = shape.hidden, shape.head_dim
    qkv_out = (shape.heads + 2 * shape.kv_heads) * D
    model = nn.Module()
    model.config = SimpleNamespace(vocab_size=V, num_attention_heads=shape.heads,
                                   num_key_value_heads=shape.kv_heads, hidden_size=H,
                                   num_hidden_layers=L, intermediate_size=2 * shape.ffn, head_dim=D)
    tr = nn.Module()
    tr.wte = nn.Embedding(V, H, dtype=dtype, device=device)
    with torch.no_grad():
        tr.wte.weight.copy_(_randn((V, H), gen, device, fast_init).to(dtype))
    tr.wte.requires_grad_(False)
    blocks = []
    for _ in range(L):
        blk = nn.Module()
        blk.ln_1 = RMSNormModule(H, shape.eps, dtype, device)
        blk.ln_2 = RMSNormModule(H, shape.eps, dtype, device)
        with torch.no_grad():
            for ln in (blk.ln_1, blk.ln_2):
                ln.weight.copy_((1.0 + norm_weight_jitter * _randn((H,), gen, device, fast_init)).to(dtype))
        blk.attn = nn.Module()
        blk.attn.c_attn = _linear(H, qkv_out, shape.qkv_bias, dtype, device, gen, std, fast_init)
        blk.attn.c_proj = _linear(shape.heads * D, H, False, dtype, device, gen, std, fast_init)
        blk.mlp = nn.Module()
        blk.mlp.w1 = _linear(H, shape.ffn, False, dtype, device, gen, std, fast_init)       # up
        blk.mlp.w2 = _linear(H, shape.ffn, False, dtype, device, gen, std, fast_init)       # gate
        blk.mlp.c_proj = _linear(shape.ffn, H, False, dtype, device, gen, std, fast_init)
        blocks.append(blk)
    tr.h = nn.ModuleList(blocks)
    tr.ln_f = RMSNormModule(H, shape.eps, dtype, device)
    model.transformer = tr
    model.lm_head = _linear(H, V, False, dtype, device, gen, std, fast_init)
    if peaky:
        # Separated logits for greedy-parity tests: tie the head to a PERMUTED embedding table,
        # lm_head[pi(t)] = wte[t].  With small residual-branch weights the hidden state stays
        # close to the input embedding, so the logit of pi(t) stands hundreds of units above the
        # rest (like a trained model's peaked distribution) while the layers' context-dependent
        # contribution still decides the winner whenever it is large enough.
        pi = torch.randperm(V, generator=torch.Generator().manual_seed(seed + 1)).to(device)
        with torch.no_grad():
            model.lm_head.weight.index_copy_(0, pi, tr.wte.weight)
    return model
