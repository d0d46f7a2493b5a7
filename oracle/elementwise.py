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
