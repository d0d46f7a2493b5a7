"""Eager NeoX half-rotation RoPE on ``[B, S, H, D]`` at positions ``0..S-1``.

Same contract as the reference's ``core/layer/rope.py`` (:10-39).  The B200 prefill path does
not call it — rotation is fused into the prefill kernels (fp32, one rounding) — it is kept so
code written against the reference's ``RoPE().forward(q, k, cos, sin)`` still works.
"""
import torch


class RoPE:
    def forward(self, q, k, cos, sin):
        S = q.shape[1]
        cos = cos[:S][None, :, None, :]
        sin = sin[:S][None, :, None, :]
        if cos.shape[-1] == q.shape[-1] // 2:
            q1, q2 = q.chunk(2, dim=-1)
            k1, k2 = k.chunk(2, dim=-1)
            q_rot = torch.cat([q1 * cos - q2 * sin, q2 * cos + q1 * sin], dim=-1)
            k_rot = torch.cat([k1 * cos - k2 * sin, k2 * cos + k1 * sin], dim=-1)
        else:
            q_rot = q * cos + self._rotate_half(q) * sin
            k_rot = k * cos + self._rotate_half(k) * sin
        return q_rot, k_rot

    @staticmethod
    def _rotate_half(x):
        half = x.shape[-1] // 2
        return torch.cat((-x[..., half:], x[..., :half]), dim=-1)
