"""SwiGLU / RMSNorm / residual+RMSNorm / store_kvcache: latency at decode sizes (us) and GB/s at
prefill sizes against the measured HBM peak.  Algorithmic bytes per SURVEY.md §8 M4.  An L2 flush
(256 MB memset) runs between timed launches of the small shapes.  GPU box only."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

dev = torch.device("cuda")
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=20, flush=True):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        if flush:
            flush_buf.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        total += e0.elapsed_time(e1)
    return total / iters * 1e3     # us


def main():
    peak = 6524.6
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    res = []
    for rows in (1, 32, 64, 2048, 8192, 16384):
        for d in (14336, 3584, 4864):
            gu = torch.randn(rows, 2 * d, device=dev, dtype=torch.bfloat16)
            out = torch.empty(rows, d, device=dev, dtype=torch.bfloat16)
            us = timeit(lambda: b200ops.swiglu(gu, out))
            nbytes = rows * d * 2 * 3
            res.append(dict(op="swiglu", rows=rows, n=d, us=round(us, 2), gbs=round(nbytes / us / 1e3, 1),
                            frac=round(nbytes / us / 1e3 / peak, 3)))
            print(res[-1], flush=True)
        for n in (4096, 8192, 3584, 896):
            x = torch.randn(rows, n, device=dev, dtype=torch.bfloat16)
            r = torch.randn(rows, n, device=dev, dtype=torch.bfloat16)
            w = torch.ones(n, device=dev, dtype=torch.bfloat16)
            y = torch.empty_like(x)
            ro = torch.empty_like(x)
            us = timeit(lambda: b200ops.rmsnorm(x, w, y, 1e-6))
            nbytes = rows * n * 2 * 2 + n * 2
            res.append(dict(op="rmsnorm", rows=rows, n=n, us=round(us, 2), gbs=round(nbytes / us / 1e3, 1),
                            frac=round(nbytes / us / 1e3 / peak, 3)))
            print(res[-1], flush=True)
            us = timeit(lambda: b200ops.rmsnorm_residual(x, r, w, y, ro, 1e-6))
            nbytes = rows * n * 2 * 4 + n * 2
            res.append(dict(op="rmsnorm_residual", rows=rows, n=n, us=round(us, 2), gbs=round(nbytes / us / 1e3, 1),
                            frac=round(nbytes / us / 1e3 / peak, 3)))
            print(res[-1], flush=True)
    for T in (64, 2048, 16384):
        H, D = 8, 128
        key = torch.randn(T, H, D, device=dev, dtype=torch.bfloat16)
        val = torch.randn(T, H, D, device=dev, dtype=torch.bfloat16)
        nb = T // 256 + 2
        kc = torch.zeros(nb, 256, H, D, device=dev, dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(nb * 256, device=dev)[:T].to(torch.int32)
        us = timeit(lambda: b200ops.store_kvcache(key, val, kc, vc, slots, 256))
        nbytes = T * H * D * 2 * 4 + 4 * T
        res.append(dict(op="store_kvcache", rows=T, n=H * D, us=round(us, 2), gbs=round(nbytes / us / 1e3, 1),
                        frac=round(nbytes / us / 1e3 / peak, 3)))
        print(res[-1], flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "elementwise_bench.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
