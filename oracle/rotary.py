"""Restatement of the reference's rotary tables and rotations (test infrastructure only)."""
import torch


def rotary_table_oracle(dim: int, max_position: int, base: float = 10000.0):
    """core/paged_attention.py:50-62, :187-190 — the table is built ENTIRELY in bf16:
    inv_freq from a bf16 arange, positions arange(dtype=bf16) (so integers above 256 collapse),
    bf16 outer product, bf16 cos/sin.  Returns the pools [max_position, dim/2]."""
    inv_freq = 1.0 / (base ** (torch.arange(0, dim, 2).to(torch.bfloat16) / dim))     # :50
    t = torch.arange(max_position, dtype=inv_freq.dtype)                               # :51
    freqs = torch.einsum("i,j->ij", t, inv_freq)                                       # :52
    emb = torch.cat((freqs, freqs), dim=-1)                                            # :53
    cos, sin = emb.cos(), emb.sin()                                                    # :59-60
    return cos[:, : dim // 2].contiguous(), sin[:, : dim // 2].contiguous()            # :187-190


def rotate_fp32(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, interleaved: bool) -> torch.Tensor:
    """flash-attn rotary contract (flash_attn_interface.py, ``rotary_interleaved``): fp32 math on
    the first rotary_dim = 2 * cos.shape[-1] features, pass-through above.  NeoX pairs
    (i, i + rd/2); interleaved pairs (2i, 2i + 1).  x [..., D]; cos/sin broadcastable [..., rd/2].
    Returns fp32 (callers round to bf16 once, as the kernel stores bf16)."""
    rd = 2 * cos.shape[-1]
    xf = x.float()
    xr, xp = xf[..., :rd], xf[..., rd:]
    c, s = cos.float(), sin.float()
    if interleaved:
        x0, x1 = xr[..., 0::2], xr[..., 1::2]
        o0, o1 = x0 * c - x1 * s, x0 * s + x1 * c
        rot = torch.stack((o0, o1), dim=-1).flatten(-2)
    else:
        x0, x1 = xr[..., : rd // 2], xr[..., rd // 2:]
        rot = torch.cat((x0 * c - x1 * s, x0 * s + x1 * c), dim=-1)
    return torch.cat((rot, xp), dim=-1)


def rope_eager_oracle(q, k, cos, sin):
    """core/layer/rope.py:21-33 — eager rotation in the INPUT dtype (bf16: every op rounds),
    positions 0..S-1, tables [S, D/2]."""
    S = q.shape[1]
    cos = cos[:S][None, :, None, :]
    sin = sin[:S][None, :, None, :]
    q1, q2 = q.chunk(2, dim=-1)
    k1, k2 = k.chunk(2, dim=-1)
    q_rot = torch.cat([q1 * cos - q2 * sin, q2 * cos + q1 * sin], dim=-1)
    k_rot = torch.cat([k1 * cos - k2 * sin, k2 * cos + k1 * sin], dim=-1)
    return q_rot, k_rot
