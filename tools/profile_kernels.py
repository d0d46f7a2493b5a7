"""Fixed small workload for ncu: the bandwidth-bound elementwise kernels at HBM-bound sizes and
the tcgen05 prefill pair at one Llama-3-8B prompt, each launched twice (the second launch of
each kernel is the one to read).  GPU box only.

    python tools/profile_kernels.py                 # all
    python tools/profile_kernels.py prefill         # only the prefill pair
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = torch.device("cuda")
torch.manual_seed(0)
bf = torch.bfloat16
acc = 0.0

if what in ("all", "elementwise"):
    rows, d, n = 8192, 14336, 4096
    gu = torch.randn(rows, 2 * d, device=dev, dtype=bf)
    so = torch.empty(rows, d, device=dev, dtype=bf)
    x = torch.randn(rows, n, device=dev, dtype=bf)
    r = torch.randn(rows, n, device=dev, dtype=bf)
    w = torch.ones(n, device=dev, dtype=bf)
    y, ro = torch.empty_like(x), torch.empty_like(x)
    T, H, D = 8192, 8, 128
    key = torch.randn(T, H, D, device=dev, dtype=bf)
    val = torch.randn(T, H, D, device=dev, dtype=bf)
    kc = torch.zeros(T // 256 + 1, 256, H, D, device=dev, dtype=bf)
    vc = torch.zeros_like(kc)
    slots = torch.randperm(T, device=dev).to(torch.int32)
    for _ in range(2):
        b200ops.swiglu(gu, so)
        b200ops.rmsnorm(x, w, y, 1e-6)
        b200ops.rmsnorm_residual(x, r, w, y, ro, 1e-6)
        b200ops.store_kvcache(key, val, kc, vc, slots, 256)
    torch.cuda.synchronize()
    acc += float(so.float().abs().mean()) + float(y.float().abs().mean())

if what in ("all", "prefill"):
    B, S, P, Hq, Hkv, D, page = 1, 8192, 0, 32, 8, 128, 256
    per_seq = -(-(P + S) // page)
    n_blocks = B * per_seq + 2
    kc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
    vc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
    table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
    lens = torch.full((B,), P, dtype=torch.int32, device=dev)
    q = torch.randn(B, S, Hq, D, device=dev, dtype=bf)
    k = torch.randn(B, S, Hkv, D, device=dev, dtype=bf)
    v = torch.randn(B, S, Hkv, D, device=dev, dtype=bf)
    cos, sin = build_rotary_tables(D, P + S + 1)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    out = torch.empty_like(q)
    for _ in range(2):
        b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out)
    torch.cuda.synchronize()
    acc += float(out.float().abs().mean())

print("done", what, acc)
