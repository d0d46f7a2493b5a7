"""tcgen05 weight-streaming GEMM against an fp32 matmul of the same bf16 inputs (floating
point: the reference is cuBLAS behind torch.matmul, core/layer/model_graph.py:145-185, whose summation order is
unspecified; tolerance = 1 bf16 ulp of the exact result plus fp32 accumulation noise) and, for the fused
epilogues, against the reference's two-step sequences: bf16 matmul -> bf16 bias add (:145-147) and bf16 matmul ->
swiglu (:183-184, kernel/swiglu.py restated in oracle/elementwise.py)."""
import pytest
import torch

import b200ops
from oracle.elementwise import bf16_ulp_diff, swiglu_oracle

pytestmark = pytest.mark.gpu


def check(out, x, w_nk, bias=None):
    mm = x.float() @ w_nk.float().t()                             # fp32 on the GPU (TF32 is off by default)
    scale = x.float().abs() @ w_nk.float().abs().t()
    want = mm.to(torch.bfloat16)
    exact, mag = mm, mm.abs()
    if bias is not None:
        want = (want.float() + bias.float()).to(torch.bfloat16)   # reference: bf16 matmul result, then a bf16 add
        exact = mm + bias.float()
        mag = mm.abs() + exact.abs()
    err = (out.float() - exact).abs()
    tol = 2.0 ** -8 * mag + 1e-5 * scale
    assert int((err > tol).sum()) == 0, f"{int((err > tol).sum())} elements out of tolerance, max err/tol {float((err / tol).max()):.2f}"
    d = bf16_ulp_diff(out.cpu(), want.cpu())
    assert float((d == 0).float().mean()) > 0.97, f"only {float((d == 0).float().mean()):.4f} bit-identical to the rounded fp32 product"


SHAPES = [(32, 4096, 6144), (32, 14336, 4096), (64, 3584, 8192), (64, 1024, 8192), (64, 8192, 1280),
          (1, 4096, 4096), (8, 896, 1152), (17, 4864, 896), (33, 8192, 1280), (16, 64, 64), (64, 128, 448),
          (128, 1024, 2048), (100, 2048, 1000), (5, 64, 8), (48, 8192, 7168)]


@pytest.mark.parametrize("M,K,N", SHAPES)
def test_stream_gemm_matches_fp32(M, K, N):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + K + N)
    x = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    out = b200ops.stream_gemm(x, w)
    check(out, x, w)
    again = b200ops.stream_gemm(x, w)
    assert torch.equal(out, again), "the stream-K reduction order is fixed: identical bits every launch"


@pytest.mark.parametrize("grid", [1, 2, 3, 7, 37, 100, 148])
def test_stream_gemm_grid_sizes_bias_strides(grid):
    """Any grid (= any split of the units between CTAs, incl. tiles shared by many CTAs and CTAs whose whole
    share lies inside one tile), strided x / out views, bias epilogue, untouched bytes around the output."""
    M, K, N = 24, 2048, 1536
    g = torch.Generator(device="cuda").manual_seed(grid)
    xb = torch.randn(40, K + 64, device="cuda", generator=g).to(torch.bfloat16)
    x = xb[:M, 8:8 + K]                                            # strided rows, 16-byte aligned start
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g).to(torch.bfloat16)
    ob = torch.full((M, N + 128), 7.0, device="cuda", dtype=torch.bfloat16)
    out = ob[:, 64:64 + N]
    ret = b200ops.stream_gemm(x, w, out=out, bias=bias, grid=grid)
    assert ret.data_ptr() == out.data_ptr()
    check(out, x, w, bias)
    assert float((ob[:, :64] - 7).abs().max()) == 0 and float((ob[:, 64 + N:] - 7).abs().max()) == 0
    assert torch.equal(b200ops.stream_gemm(x, w, bias=bias, grid=grid), out)


@pytest.mark.parametrize("M,K,d", [(32, 4096, 14336), (64, 8192, 3584), (1, 896, 4864), (20, 512, 64), (64, 256, 8)])
@pytest.mark.parametrize("grid", [0, 5])
def test_stream_gemm_swiglu_epilogue(M, K, d, grid):
    """gate_up GEMM + SwiGLU in one launch == bf16 matmul followed by the swiglu kernel (model_graph.py:183-184):
    the epilogue rounds gate and up to bf16 first, then applies kernel/swiglu.py's arithmetic."""
    g = torch.Generator(device="cuda").manual_seed(K + d)
    x = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    up = (torch.randn(d, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    gate = (torch.randn(d, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    w = b200ops.interleave_up_gate(up, gate)
    assert torch.equal(w[0::2], up) and torch.equal(w[1::2], gate)
    out = b200ops.stream_gemm(x, w, swiglu=True, grid=grid)
    assert out.shape == (M, d)
    # Reference: the reference's two-step sequence on the exact fp32 products — gate and up rounded to bf16, then
    # kernel/swiglu.py's arithmetic (oracle).  The kernel's fp32 accumulation order differs from the exact sum, so the
    # bf16 rounding of gate / up may flip by one ulp (2^-8 relative each): tolerance 2^-6 relative + fp32 noise.
    gu32 = x.float() @ torch.cat([up, gate], dim=0).float().t()
    ref = swiglu_oracle(gu32.to(torch.bfloat16).cpu()).float()
    err = (out.float().cpu() - ref).abs()
    mag = (x.float().abs() @ torch.cat([up, gate], dim=0).float().abs().t()).cpu()
    tol = 2.0 ** -6 * ref.abs() + 2e-5 * (mag[:, :d] + mag[:, d:]) * (1 + gu32[:, :d].abs().cpu())
    assert int((err > tol).sum()) == 0, f"{int((err > tol).sum())} out of tolerance, worst {float((err / tol).max()):.2f}"
    d_ulp = bf16_ulp_diff(out.cpu(), ref.to(torch.bfloat16))
    assert float((d_ulp == 0).float().mean()) > 0.9, "most outputs must be bit-identical to the two-step reference"
    assert torch.equal(out, b200ops.stream_gemm(x, w, swiglu=True, grid=grid)), "deterministic"


def test_stream_gemm_swiglu_bitexact_vs_unfused_when_tiles_are_not_shared():
    """One CTA (grid = 1) accumulates every tile over the whole K range in k order, fused or not: the fused epilogue
    must then equal matmul -> swiglu kernel bit for bit (same rounding points as model_graph.py:183-184)."""
    M, K, d = 24, 512, 192
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    up = (torch.randn(d, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    gate = (torch.randn(d, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    out = b200ops.stream_gemm(x, b200ops.interleave_up_gate(up, gate), swiglu=True, grid=1)
    gu = b200ops.stream_gemm(x, torch.cat([up, gate], dim=0), grid=1)
    two_step = torch.empty(M, d, device="cuda", dtype=torch.bfloat16)
    b200ops.swiglu(gu, two_step)
    assert torch.equal(out, two_step)


def test_stream_gemm_graph_replay_pdl_and_errors():
    M, K, N = 32, 4096, 4096
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    w2 = (torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16)
    mid = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    for pdl, ahead in ((False, 0), (True, 0), (True, 64), (True, 4096)):     # 4096 KB/CTA: beyond every CTA's share
        b200ops.stream_gemm(x, w, out=mid, pdl=pdl, l2_ahead_kb=ahead)
        b200ops.stream_gemm(mid, w2, out=out, pdl=pdl, l2_ahead_kb=ahead)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            b200ops.stream_gemm(x, w, out=mid, pdl=pdl, l2_ahead_kb=ahead)   # a chain: the second product consumes the first
            b200ops.stream_gemm(mid, w2, out=out, pdl=pdl, l2_ahead_kb=ahead)
        for _ in range(3):
            x.copy_(torch.randn(M, K, device="cuda").to(torch.bfloat16))
            mid.zero_()
            out.zero_()
            g.replay()
            torch.cuda.synchronize()
            check(mid, x, w)
            check(out, mid, w2)
    assert not b200ops.stream_gemm_ok(129, 4096, 4096) and not b200ops.stream_gemm_ok(8, 100, 4096)
    with pytest.raises(AssertionError):
        b200ops.stream_gemm(torch.zeros(129, 64, device="cuda", dtype=torch.bfloat16),
                            torch.zeros(64, 64, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(AssertionError):
        b200ops.stream_gemm(torch.zeros(8, 64), torch.zeros(64, 64))          # CPU tensors: no fallback


def test_stream_gemm_capture_uses_preallocated_workspace():
    """scratch for captured launches is allocated and zeroed OUTSIDE the capture and shared by every
    graph, so graphs may replay in any order (a later-captured graph first)."""
    xs = {m: torch.randn(m, 1024, device="cuda").to(torch.bfloat16) for m in (8, 1)}
    w = (torch.randn(512, 1024, device="cuda") * 0.05).to(torch.bfloat16)
    outs = {m: torch.empty(m, 512, device="cuda", dtype=torch.bfloat16) for m in xs}
    graphs = {}
    for m in (1, 8):
        b200ops.stream_gemm(xs[m], w, out=outs[m], grid=16)           # eager warm-up creates the shared scratch
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            b200ops.stream_gemm(xs[m], w, out=outs[m], grid=16)
        graphs[m] = g
    for m in (8, 1, 8):
        outs[m].zero_()
        graphs[m].replay()
        torch.cuda.synchronize()
        check(outs[m], xs[m], w)
