"""fp32 restatements of the reference's Triton elementwise kernels (test infrastructure only)."""
import torch


def _bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16)


def swiglu_oracle(gate_up: torch.Tensor) -> torch.Tensor:
    """kernel/swiglu.py:18-46 — up = cols [0, d), gate = cols [d, 2d);
    out = bf16(gate * sigmoid(gate) * up) evaluated in fp32."""
    d = gate_up.shape[-1] // 2
    up = gate_up[..., :d].float()
    gate = gate_up[..., d:].float()
    silu = gate * torch.sigmoid(gate)          # :40-41
    return _bf16(silu * up)                    # :42, :45


def rmsnorm_oracle(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """kernel/rmsnorm_add.py:13-34 / :144-167 — y = bf16(x * rsqrt(mean(x^2) + eps) * w) in fp32."""
    xf = x.float()
    n = x.shape[-1]
    rrms = torch.rsqrt(xf.pow(2).sum(-1, keepdim=True) / n + eps)
    return _bf16(xf * rrms * w.float())


def rmsnorm_residual_oracle(x: torch.Tensor, r: torch.Tensor, w: torch.Tensor, eps: float):
    """kernel/rmsnorm_residual.py:92-124 — s = x + r (fp32); res_out = bf16(s) (:106);
    mean of squares from the UNROUNDED s (:107); second pass re-loads the ROUNDED residual
    (:119): y = bf16(f32(res_out) * rrms * w)."""
    s = x.float() + r.float()
    res = _bf16(s)
    n = x.shape[-1]
    rrms = torch.rsqrt(s.pow(2).sum(-1, keepdim=True) / n + eps)
    y = _bf16(res.float() * rrms * w.float())
    return y, res


def mlp_oracle(x: torch.Tensor, gu_weight: torch.Tensor, d_weight: torch.Tensor) -> torch.Tensor:
    """core/layer/model_graph.py:75-81 — gate_up = x @ gu; up, gate = chunk(2); silu(gate) * up @ d."""
    gate_up = x @ gu_weight
    up, gate = gate_up.chunk(2, dim=-1)
    return (torch.nn.functional.silu(gate) * up) @ d_weight


def bf16_ulp_diff(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Distance in units of bf16 representable values (sign-magnitude aware)."""
    def key(t):
        i = t.contiguous().view(torch.int16).to(torch.int32) & 0xFFFF
        return torch.where(i >= 0x8000, 0x8000 - i, i)
    return (key(a.to(torch.bfloat16)) - key(b.to(torch.bfloat16))).abs()


def allreduce_residual_rmsnorm_oracle(parts, residual, w, eps: float, ss_rounded: bool):
    """The two all-reduce call sites of the decode loop with what follows them:
    core/layer/model_graph.py:210 + :213-216 (``all_reduce(attn_out)`` then
    ``rmsnorm_residual_gemm``: sum of squares from the UNROUNDED x + r, ``ss_rounded=False``) and
    :219 + :143 / :222 (``all_reduce(mlp_out).add_(residual)`` then ``rmsnorm_``: the norm reads the
    bf16-ROUNDED sum, ``ss_rounded=True``).  ``parts`` = the per-rank partial sums (bf16); the
    all-reduce result is a bf16 tensor (core/parallel_config.py:65-69), summed here in fp32 in rank
    order and rounded once (NCCL's own association order is unspecified).  Returns (y, res_out);
    ``w`` None skips the norm (y None)."""
    s = parts[0].float()
    for p in parts[1:]:
        s = s + p.float()
    if len(parts) > 1:
        s = _bf16(s).float()
    if residual is not None:
        s = s + residual.float()
    res = _bf16(s)
    if w is None:
        return None, res
    n = res.shape[-1]
    t = res.float() if ss_rounded else s
    rrms = torch.rsqrt(t.pow(2).sum(-1, keepdim=True) / n + eps)
    return _bf16(res.float() * rrms * w.float()), res
