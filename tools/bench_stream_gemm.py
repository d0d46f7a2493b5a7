"""b200_stream_gemm (tcgen05, swap-AB, stream-K) vs cuBLAS (torch.matmul) vs the legacy mma.sync kernel on the four
decode GEMMs of a layer (+ lm_head) at per-rank TP shapes: weight-streaming, M = batch rows, weights rotate over
several copies so the 126 MB L2 never helps, CUDA-graph replay, CUDA events.  GPU box only; writes
gpurun_out/stream_gemm_<shape>_tp<tp>_b<batch>.json.

    python tools/bench_stream_gemm.py --shape llama3-70b --tp 8 --batch 64
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

PEAK = 6454.6
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
    ap.add_argument("--shape", default="llama3-70b")
    ap.add_argument("--tp", type=int, default=8)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--grids", default="0")
    ap.add_argument("--lm-head", action="store_true")
    a = ap.parse_args()
    from core.layer.synthetic_model import SHAPES
    sh = SHAPES[a.shape]
    tp = a.tp
    Hq, Hkv, ffn, D, H = sh.heads // tp, sh.kv_heads // tp, sh.ffn // tp, sh.head_dim, sh.hidden
    gemms = [("qkv", H, (Hq + 2 * Hkv) * D, "bias" if sh.qkv_bias else "none"), ("o_proj", Hq * D, H, "none"),
             ("gate_up", H, 2 * ffn, "swiglu"), ("down", ffn, H, "none")]
    if a.lm_head:
        gemms.append(("lm_head", H, sh.vocab, "none"))
    dev = torch.device("cuda")
    M = a.batch
    rows = []
    for name, K, N, epi in gemms:
        wbytes = K * N * 2
        copies = max(3, min(12, int(600e6 // wbytes) + 1))
        ws_nk = [(torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16) for _ in range(copies)]
        x = torch.randn(M, K, device=dev).to(torch.bfloat16)
        y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        bias = torch.randn(N, device=dev).to(torch.bfloat16) if epi == "bias" else None
        # cuBLAS: the reference's [in, out] layout (x @ w) and the nn.Linear layout (x @ w_nk.T), + the separate launches
        ws_kn = [w.t().contiguous() for w in ws_nk[:3]]
        ys = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)

        def cublas_kn(i):
            torch.matmul(x, ws_kn[i % 3], out=y)
            if bias is not None:
                y.add_(bias)
            if epi == "swiglu":
                b200ops.swiglu(y, ys)

        def cublas_nk(i):
            torch.matmul(x, ws_nk[i].t(), out=y)
            if bias is not None:
                y.add_(bias)
            if epi == "swiglu":
                b200ops.swiglu(y, ys)
        t_kn = timed_graph(cublas_kn, 3) if wbytes * 3 > 300e6 else None
        t_nk = timed_graph(cublas_nk, copies)
        del ws_kn
        row = {"gemm": name, "M": M, "K": K, "N": N, "epilogue": epi, "weight_MB": round(wbytes / 1e6, 1),
               "roofline_us": round(wbytes / PEAK / 1e3, 2),
               "cublas_nk_us": round(t_nk, 2), "cublas_nk_frac": round(wbytes / t_nk / 1e3 / PEAK, 3),
               "cublas_kn_us": round(t_kn, 2) if t_kn else None, "stream": []}
        out = ys if epi == "swiglu" else y
        for grid in [int(g) for g in a.grids.split(",")]:
            for pdl in (False, True):
                t = timed_graph(lambda i: b200ops.stream_gemm(x, ws_nk[i], out=out, bias=bias, swiglu=epi == "swiglu",
                                                              pdl=pdl, grid=grid), copies)
                row["stream"].append({"grid": grid, "pdl": pdl, "us": round(t, 2), "frac": round(wbytes / t / 1e3 / PEAK, 3)})
        best = min(row["stream"], key=lambda r: r["us"])
        row["best"] = best
        row["speedup_vs_cublas"] = round(min(t_nk, t_kn or 1e9) / best["us"], 3)
        # sanity: same numbers as cuBLAS up to rounding
        if epi != "swiglu":
            torch.matmul(x, ws_nk[0].t(), out=y)
            if bias is not None:
                y.add_(bias)
            ref = y.clone()
            b200ops.stream_gemm(x, ws_nk[0], out=y, bias=bias)
            row["max_abs_diff_vs_cublas"] = float((y.float() - ref.float()).abs().max())
        rows.append(row)
        print(json.dumps(row), flush=True)
        del ws_nk
        torch.cuda.empty_cache()
    tot = {"roofline_us": round(sum(r["roofline_us"] for r in rows if r["gemm"] != "lm_head"), 2),
           "cublas_us": round(sum(min(r["cublas_nk_us"], r["cublas_kn_us"] or 1e9) for r in rows if r["gemm"] != "lm_head"), 2),
           "stream_us": round(sum(r["best"]["us"] for r in rows if r["gemm"] != "lm_head"), 2)}
    print(json.dumps({"layer_total": tot}), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", f"stream_gemm_{a.shape}_tp{tp}_b{M}.json")
    json.dump({"shape": a.shape, "tp": tp, "batch": M, "hbm_peak_gbs": PEAK, "gemms": rows, "layer_total": tot},
              open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
