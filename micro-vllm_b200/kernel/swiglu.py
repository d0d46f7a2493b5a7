"""SwiGLU on a packed ``[..., 2d]`` tensor (up = first half, gate = second half).

Drop-in for the reference's ``kernel/swiglu.py`` (``swiglu_fused`` :49-93, ``swiglu_gemm``
:100-148); the Triton ``_swiglu_kernel`` (:6-46) is replaced by the sm_100a
``b200_swiglu`` kernel: ``out = bf16(silu(f32(gate)) * f32(up))``.
"""
import torch

import b200ops

__all__ = ["swiglu_fused", "swiglu_gemm"]


def swiglu_fused(gate_up: torch.Tensor) -> torch.Tensor:
    """Allocating variant (reference :49-93).  Output dtype follows the input, which must be
    bf16 — the reference kernel hard-casts its result to bf16 (:45)."""
    assert gate_up.is_cuda, "Input must be on CUDA"
    last_dim = gate_up.shape[-1]
    out = torch.empty(gate_up.shape[:-1] + (last_dim // 2,), dtype=gate_up.dtype, device=gate_up.device)
    b200ops.swiglu(gate_up.contiguous(), out)
    return out


def swiglu_gemm(gate_up: torch.Tensor, out_buffer: torch.Tensor) -> torch.Tensor:
    """Writes the caller's buffer — the next matmul's input (reference :100-148)."""
    assert gate_up.is_cuda and out_buffer.is_cuda, "All must be on CUDA"
    assert gate_up.dtype == out_buffer.dtype, f"Dtype mismatch: {gate_up.dtype} vs {out_buffer.dtype}"
    expected = gate_up.shape[:-1] + (gate_up.shape[-1] // 2,)
    assert out_buffer.shape == expected, f"Shape mismatch: {out_buffer.shape} vs {expected}"
    assert out_buffer.is_contiguous(), "out_buffer must be contiguous"
    b200ops.swiglu(gate_up.contiguous(), out_buffer)
    return out_buffer
