"""Fused residual-add + RMSNorm (alloc / two-buffer / hybrid variants).

Drop-in for the reference's ``kernel/rmsnorm_residual.py``: ``rmsnorm_residual_fused``
(:44-68), ``rmsnorm_residual_gemm`` (:127-166), ``rmsnorm_residual_hybrid`` (:169-204), all on
the sm_100a ``b200_rmsnorm_residual`` kernel:

    s = f32(x) + f32(r); res_out = bf16(s); rrms = rsqrt(mean(s*s) + eps)
    y = bf16(f32(res_out) * rrms * f32(w))        # second pass reads the ROUNDED residual (:119)
"""
from typing import Tuple

import torch

import b200ops
from kernel.rmsnorm_add import rmsnorm_residual_fused  # same op, same name (the reference defines it twice)

__all__ = ["rmsnorm_residual_fused", "rmsnorm_residual_gemm", "rmsnorm_residual_hybrid"]


def rmsnorm_residual_gemm(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                          out_normed_buffer: torch.Tensor, out_residual_buffer: torch.Tensor,
                          eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    assert x.is_cuda and residual.is_cuda and weight.is_cuda, "All must be on CUDA"
    assert out_normed_buffer.is_cuda and out_residual_buffer.is_cuda, "Buffers must be on CUDA"
    assert x.shape == residual.shape, f"Shape mismatch: x={x.shape}, residual={residual.shape}"
    assert x.shape == out_normed_buffer.shape, "Output buffer shape mismatch"
    assert x.shape == out_residual_buffer.shape, "Output buffer shape mismatch"
    assert x.shape[-1] == weight.shape[0], "Hidden dim mismatch"
    assert out_normed_buffer.dtype == x.dtype, f"Dtype mismatch: {out_normed_buffer.dtype} vs {x.dtype}"
    assert out_residual_buffer.dtype == x.dtype, f"Dtype mismatch: {out_residual_buffer.dtype} vs {x.dtype}"
    b200ops.rmsnorm_residual(x, residual, weight, out_normed_buffer, out_residual_buffer, eps)
    return out_normed_buffer, out_residual_buffer


def rmsnorm_residual_hybrid(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                            out_normed_buffer: torch.Tensor, eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    assert x.is_cuda and residual.is_cuda and weight.is_cuda, "All must be on CUDA"
    assert out_normed_buffer.is_cuda, "Buffer must be on CUDA"
    assert x.shape == residual.shape, f"Shape mismatch: x={x.shape}, residual={residual.shape}"
    assert x.shape == out_normed_buffer.shape, "Output buffer shape mismatch"
    assert x.shape[-1] == weight.shape[0], "Hidden dim mismatch"
    assert out_normed_buffer.dtype == x.dtype, "Dtype mismatch"
    res_out = torch.empty_like(x)
    b200ops.rmsnorm_residual(x, residual, weight, out_normed_buffer, res_out, eps)
    return out_normed_buffer, res_out
