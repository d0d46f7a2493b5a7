"""Decode-sized weight-streaming GEMM against an fp32 matmul of the same bf16 inputs
(floating point: the reference is cuBLAS behind torch.matmul, core/layer/model_graph.py:145-185, whose
summation order is unspecified; tolerance = 1 bf16 ulp of the exact result plus fp32 accumulation noise)."""
import pytest
import torch

import b200ops
from oracle.elementwise import bf16_ulp_diff

pytestmark = pytest.mark.gpu


def check(out, x, w, bias=None):
    """|out - exact| <= the bf16 rounding error of the exact result (half an ulp = 2^-8 relative; with a bias the
    two roundings of the reference's bf16 matmul + bf16 add) plus fp32 accumulation-order noise (1e-5 of
    sum |x||w|); and almost every element bit-identical to the rounded fp32 product (differences only where
    the exact value sits on a rounding boundary)."""
    mm = x.float() @ w.float()                                    # fp32 on the GPU (TF32 is off by default)
    scale = x.float().abs() @ w.float().abs()
    want = mm.to(torch.bfloat16)
    exact, mag = mm, mm.abs()
    if bias is not None:
        want = (want.float() + bias.float()).to(torch.bfloat16)   # reference: bf16 matmul result, then a bf16 add
        exact = mm + bias.float()
        mag = mm.abs() + exact.abs()                               # one rounding of the product, one of the sum
    err = (out.float() - exact).abs()
    tol = 2.0 ** -8 * mag + 1e-5 * scale
    assert int((err > tol).sum()) == 0, f"{int((err > tol).sum())} elements out of tolerance, max err/tol {float((err / tol).max()):.2f}"
    d = bf16_ulp_diff(out.cpu(), want.cpu())
    assert float((d == 0).float().mean()) > 0.97, f"only {float((d == 0).float().mean()):.4f} bit-identical to the rounded fp32 product"


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("M,K,N", [(32, 4096, 6144), (32, 14336, 4096), (64, 3584, 8192), (64, 1024, 8192),
                                    (1, 4096, 4096), (8, 896, 1152), (17, 4864, 896), (33, 8192, 1280),
                                    (16, 64, 64), (64, 128, 448)])
def test_skinny_gemm_matches_fp32(M, K, N, variant):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + K + N)
    x = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(K, N, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    out = b200ops.skinny_gemm(x, w, variant=variant)
    check(out, x, w)


@pytest.mark.parametrize("ksplit", [1, 2, 3, 7, 16])
def test_skinny_gemm_split_counts_bias_strides_and_determinism(ksplit):
    M, K, N = 24, 2048, 1536
    g = torch.Generator(device="cuda").manual_seed(ksplit)
    xb = torch.randn(40, K + 64, device="cuda", generator=g).to(torch.bfloat16)
    x = xb[:M, 8:8 + K]                                            # strided rows, 16-byte aligned start
    w = (torch.randn(K, N, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g).to(torch.bfloat16)
    ob = torch.full((M, N + 128), 7.0, device="cuda", dtype=torch.bfloat16)
    out = ob[:, 64:64 + N]
    ret = b200ops.skinny_gemm(x, w, out=out, bias=bias, ksplit=ksplit)
    assert ret.data_ptr() == out.data_ptr()
    check(out, x, w, bias)
    assert float((ob[:, :64] - 7).abs().max()) == 0 and float((ob[:, 64 + N:] - 7).abs().max()) == 0
    again = b200ops.skinny_gemm(x, w, bias=bias, ksplit=ksplit)
    assert torch.equal(again, out), "same split count must give identical bits"


def test_skinny_gemm_graph_replay_and_errors():
    M, K, N = 32, 4096, 4096
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    b200ops.skinny_gemm(x, w, out=out)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        b200ops.skinny_gemm(x, w, out=out)
    for _ in range(3):
        x.copy_(torch.randn(M, K, device="cuda").to(torch.bfloat16))
        out.zero_()
        g.replay()
        check(out, x, w)
    assert not b200ops.skinny_gemm_ok(65, 4096, 4096) and not b200ops.skinny_gemm_ok(8, 100, 4096)
    with pytest.raises(AssertionError):
        b200ops.skinny_gemm(torch.zeros(65, 64, device="cuda", dtype=torch.bfloat16),
                            torch.zeros(64, 64, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(AssertionError):
        b200ops.skinny_gemm(torch.zeros(4, 64, dtype=torch.bfloat16), torch.zeros(64, 64, dtype=torch.bfloat16))
