"""Small fixed workload for ncu: a handful of decode-attention launches at BASELINE.json
configs[1] shape (one layer, rotating over 3 cache sets).  GPU box only."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

B, Hq, Hkv, D, ctx, page = 32, 32, 8, 128, 4096, 256
n_launch = int(sys.argv[1]) if len(sys.argv) > 1 else 6
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 0
splits = int(sys.argv[3]) if len(sys.argv) > 3 else 0
balanced = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
dev = torch.device("cuda")
per_seq = ctx // page
n_blocks = B * per_seq + 4
torch.manual_seed(0)
caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16),
           torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16)) for _ in range(3)]
table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
lens = torch.full((B,), ctx - 1, dtype=torch.int32, device=dev)
q = torch.randn(B, Hq, D, device=dev, dtype=torch.bfloat16)
k = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
v = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
cos, sin = build_rotary_tables(D, 8192)
cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
out = torch.empty(B, Hq, D, device=dev, dtype=torch.bfloat16)
for i in range(n_launch):
    kc, vc = caches[i % 3]
    b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, out=out, variant=variant, num_splits=splits,
                              balanced=balanced)
torch.cuda.synchronize()
print("done", n_launch, float(out.float().abs().mean()))
