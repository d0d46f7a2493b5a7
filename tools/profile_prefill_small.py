"""Fixed small workload for ncu: the prefill launch pair (rope/append + attention) at the sizes the
reference really runs (engine.py:45 caps prefill at 2048 tokens), three calls per case.  GPU box only.

    python tools/profile_prefill_small.py [variant]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
CASES = [("l8_b1_s2048", 1, 2048, 32, 8, 128), ("l8_b4_s512", 4, 512, 32, 8, 128),
         ("q05_b4_s2048", 4, 2048, 14, 2, 64), ("l8_b1_s512", 1, 512, 32, 8, 128)]
dev = torch.device("cuda")
bf = torch.bfloat16
torch.manual_seed(0)
acc = 0.0
for name, B, S, Hq, Hkv, D in CASES:
    page = 256
    per_seq = -(-S // page)
    n_blocks = B * per_seq + 2
    kc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
    vc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
    table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
    lens = torch.zeros(B, dtype=torch.int32, device=dev)
    q, k, v = (torch.randn(B, S, h, D, device=dev, dtype=bf) for h in (Hq, Hkv, Hkv))
    cos, sin = build_rotary_tables(D, S + 1)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    o = torch.empty_like(q)
    for _ in range(3):
        b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=o, variant=variant)
    torch.cuda.synchronize()
    acc += float(o.float().abs().mean())
print("done", acc)
