"""b200_skinny_gemm vs cuBLAS (torch.matmul) on the four decode GEMMs of a layer at per-rank TP shapes:
weight-streaming, M = batch rows, weights rotate over several copies so the 126 MB L2 never helps, CUDA-graph
replay, CUDA events.  Sweeps slice width (variant) x split count.  GPU box only; writes
gpurun_out/skinny_gemm_<shape>_tp<tp>_b<batch>.json.

    python tools/bench_skinny_gemm.py --shape llama3-8b --tp 1 --batch 32 [--quick]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

PEAK = 6524.6
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:  # noqa: BLE001
    pass


def timed_graph(fn, copies, reps=5):
    for i in range(copies):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(copies):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * copies)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="llama3-8b")
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    from core.layer.synthetic_model import SHAPES
    sh = SHAPES[a.shape]
    tp = a.tp
    Hq, Hkv, ffn, D, H = sh.heads // tp, sh.kv_heads // tp, sh.ffn // tp, sh.head_dim, sh.hidden
    gemms = [("qkv", H, (Hq + 2 * Hkv) * D), ("o_proj", Hq * D, H), ("gate_up", H, 2 * ffn), ("down", ffn, H)]
    dev = torch.device("cuda")
    M = a.batch
    out_rows = []
    for name, K, N in gemms:
        wbytes = K * N * 2
        copies = max(3, min(12, int(500e6 // wbytes) + 1))
        ws = [(torch.randn(K, N, device=dev) * 0.05).to(torch.bfloat16) for _ in range(copies)]
        x = torch.randn(M, K, device=dev).to(torch.bfloat16)
        y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        t_cublas = timed_graph(lambda i: torch.matmul(x, ws[i], out=y), copies)
        row = {"gemm": name, "M": M, "K": K, "N": N, "weight_MB": round(wbytes / 1e6, 1),
               "cublas_us": round(t_cublas, 2), "cublas_frac": round(wbytes / t_cublas / 1e3 / PEAK, 3), "skinny": []}
        if b200ops.skinny_gemm_ok(M, K, N):
            units = K // 64
            for variant in (0, 1, 2):
                auto = int(b200ops.lib.b200_skinny_gemm_splits(M, K, N, variant))
                cands = sorted({auto} | ({1, 2, 3, 4, 6, 8, 12, 16, 24} if not a.quick else {max(1, auto // 2), auto * 2}))
                for ks in cands:
                    if ks < 1 or ks > max(1, units // 2):
                        continue
                    try:
                        t = timed_graph(lambda i: b200ops.skinny_gemm(x, ws[i], out=y, ksplit=ks, variant=variant), copies)
                    except Exception as e:  # noqa: BLE001
                        row["skinny"].append({"variant": variant, "ksplit": ks, "error": str(e)[:200]})
                        continue
                    row["skinny"].append({"variant": variant, "ksplit": ks, "auto": ks == auto, "us": round(t, 2),
                                          "frac": round(wbytes / t / 1e3 / PEAK, 3)})
            ok = [r for r in row["skinny"] if "us" in r]
            if ok:
                best = min(ok, key=lambda r: r["us"])
                autos = [r for r in ok if r.get("auto") and r["variant"] == 0]
                row["best"] = best
                row["auto_v0"] = autos[0] if autos else None
                # sanity: same numbers as cuBLAS up to rounding
                torch.matmul(x, ws[0], out=y)
                ref = y.clone()
                b200ops.skinny_gemm(x, ws[0], out=y, ksplit=best["ksplit"], variant=best["variant"])
                row["max_abs_diff_vs_cublas"] = float((y.float() - ref.float()).abs().max())
        out_rows.append(row)
        print(json.dumps({k: v for k, v in row.items() if k != "skinny"}), flush=True)
        del ws
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", f"skinny_gemm_{a.shape}_tp{tp}_b{M}.json")
    json.dump({"shape": a.shape, "tp": tp, "batch": M, "hbm_peak_gbs": PEAK, "gemms": out_rows}, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
