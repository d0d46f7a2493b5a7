"""cuBLAS (torch.matmul) time of the four decode GEMMs of a layer at the per-rank shapes of a TP run:
weight-streaming, M = batch rows.  Weights rotate over several copies so L2 (126 MB) never helps.
Reports us and weight GB/s per GEMM against the HBM roofline — the part of a TP=8 layer that is NOT
this repo's kernels.  GPU box only; writes gpurun_out/gemm_shapes.json.

    python tools/bench_gemm_shapes.py [--shape llama3-70b] [--tp 8] [--batch 64]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="llama3-70b")
    ap.add_argument("--tp", default="8,4,2,1")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--lt", action="store_true", help="prefer cuBLASLt")
    a = ap.parse_args()
    if a.lt:
        torch.backends.cuda.preferred_blas_library("cublaslt")
    from core.layer.synthetic_model import SHAPES
    sh = SHAPES[a.shape]
    dev = torch.device("cuda")
    out = []
    for tp in [int(x) for x in a.tp.split(",")]:
        Hq, Hkv, ffn, D, H = sh.heads // tp, sh.kv_heads // tp, sh.ffn // tp, sh.head_dim, sh.hidden
        gemms = [("qkv", H, (Hq + 2 * Hkv) * D), ("o_proj", Hq * D, H), ("gate_up", H, 2 * ffn), ("down", ffn, H)]
        for name, K, N in [(n + lay, k, nn) for n, k, nn in gemms for lay in ("", "_TN")]:
            tn = name.endswith("_TN")                       # weights kept [out, in] (nn.Linear layout) instead of [in, out]
            wbytes = K * N * 2
            copies = max(2, min(16, int(400e6 // wbytes) + 1))
            ws = [torch.randn((N, K) if tn else (K, N), device=dev, dtype=torch.bfloat16) for _ in range(copies)]
            x = torch.randn(a.batch, K, device=dev, dtype=torch.bfloat16)
            y = torch.empty(a.batch, N, device=dev, dtype=torch.bfloat16)

            def mm(i):
                torch.matmul(x, ws[i].t() if tn else ws[i], out=y)
            for i in range(copies):
                mm(i)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for i in range(copies):
                    mm(i)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
            row = {"tp": tp, "gemm": name, "M": a.batch, "K": K, "N": N, "weight_MB": round(wbytes / 1e6, 1),
                   "us": round(us, 2), "GBps": round(wbytes / us / 1e3, 1)}
            out.append(row)
            print(row, flush=True)
            del ws
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"gemm_shapes_{a.shape}_b{a.batch}{'_lt' if a.lt else ''}.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
