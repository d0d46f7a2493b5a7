"""Small fixed workload for ncu: a few b200_skinny_gemm launches (Llama-3-8B down_proj with 64-column slices and
4 k-ranges — the shape the runner policy uses — and gate_up without split-K, M = 32, rotating weight copies so L2 does not help).  GPU box only."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

dev = torch.device("cuda")
M = 32
for K, N, ks, variant in [(14336, 4096, 4, 1), (4096, 28672, 1, 0)]:
    ws = [(torch.randn(K, N, device=dev) * 0.05).to(torch.bfloat16) for _ in range(3)]
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for i in range(3):
        b200ops.skinny_gemm(x, ws[i], out=y, ksplit=ks, variant=variant)
    torch.cuda.synchronize()
    del ws
print("done")
