"""RMSNorm and fused residual-add + RMSNorm.

Drop-in for the reference's ``kernel/rmsnorm_add.py``: ``rmsnorm`` (:37-53), ``rmsnorm_``
(:170-206), ``rmsnorm_residual_fused`` (:89-113), ``ProRMSNorm`` (:116-127).  The Triton row
kernels (:9-34, :56-86, :133-167) are replaced by the sm_100a ``b200_rmsnorm`` /
``b200_rmsnorm_residual`` kernels, which keep a row in registers (one HBM read) and
reproduce the reference's rounding order.
"""
from typing import Tuple

import torch

import b200ops

__all__ = ["rmsnorm", "rmsnorm_", "rmsnorm_residual_fused", "ProRMSNorm"]


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    assert x.is_cuda, "Input must be on CUDA device"
    assert weight.is_cuda, "Weight must be on CUDA device"
    assert x.shape[-1] == weight.shape[0], f"Hidden dim mismatch: {x.shape[-1]} vs {weight.shape[0]}"
    y = torch.empty_like(x)
    b200ops.rmsnorm(x, weight, y, eps)
    return y


def rmsnorm_(x: torch.Tensor, weight: torch.Tensor, out_buffer: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    assert x.is_cuda and weight.is_cuda and out_buffer.is_cuda, "All must be on CUDA"
    assert x.shape[-1] == weight.shape[0], "Hidden dim mismatch"
    assert x.shape == out_buffer.shape, "Output buffer shape mismatch"
    b200ops.rmsnorm(x, weight, out_buffer, eps)
    return out_buffer


def rmsnorm_residual_fused(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                           eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    assert x.is_cuda, "Input must be on CUDA device"
    assert residual.is_cuda, "Residual must be on CUDA device"
    assert weight.is_cuda, "Weight must be on CUDA device"
    assert x.shape == residual.shape, f"Shape mismatch: x={x.shape}, residual={residual.shape}"
    assert x.shape[-1] == weight.shape[0], f"Hidden dim mismatch: {x.shape[-1]} vs {weight.shape[0]}"
    y = torch.empty_like(x)
    res_out = torch.empty_like(x)
    b200ops.rmsnorm_residual(x, residual, weight, y, res_out, eps)
    return y, res_out


class ProRMSNorm(torch.nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6):
        super().__init__()
        self.hidden_size = hidden_size
        self.eps = eps
        self.weight = torch.nn.Parameter(torch.ones(hidden_size))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return rmsnorm(x, self.weight, self.eps)

    def extra_repr(self) -> str:
        return f"{self.hidden_size}, eps={self.eps}"
