"""Per-CTA timeline of b200_stream_gemm (debug hook b200_debug_stream_gemm_trace): %globaltimer stamps of every
CTA at  0 entry | 1 setup done | 2 all loads issued | 3 first stage landed | 4 last MMA committed | 5 last accumulator
drained | 6 peers' partials complete | 7 exit, reported relative to the earliest entry.  GPU box only.

    python tools/trace_stream_gemm.py M K N [swiglu|bias|none]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

NAMES = ["entry", "setup_done", "loads_issued", "first_stage", "last_mma", "acc_drained", "peers_done", "exit"]


def main():
    M, K, N = (int(a) for a in sys.argv[1:4])
    epi = sys.argv[4] if len(sys.argv) > 4 else "none"
    dev = torch.device("cuda")
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    ws = [(torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16) for _ in range(4)]
    bias = torch.randn(N, device=dev).to(torch.bfloat16) if epi == "bias" else None
    out = torch.empty(M, N // 2 if epi == "swiglu" else N, device=dev, dtype=torch.bfloat16)
    for w in ws:
        b200ops.stream_gemm(x, w, out=out, bias=bias, swiglu=epi == "swiglu")
    torch.cuda.synchronize()
    grid = b200ops.sm_count(dev)
    trace = torch.zeros(grid * 8, dtype=torch.int64, device=dev)
    b200ops.lib.b200_debug_stream_gemm_trace(trace.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    b200ops.stream_gemm(x, ws[0], out=out, bias=bias, swiglu=epi == "swiglu")      # cold weights (4 x rotated)
    e1.record()
    torch.cuda.synchronize()
    b200ops.lib.b200_debug_stream_gemm_trace(None)
    t = trace.view(grid, 8).double().cpu()
    t0 = t[:, 0][t[:, 0] > 0].min()
    res = {"M": M, "K": K, "N": N, "epilogue": epi, "weight_MB": round(K * N * 2 / 1e6, 1),
           "event_us": round(e0.elapsed_time(e1) * 1e3, 2)}
    for i, name in enumerate(NAMES):
        col = t[:, i]
        ok = col > 0
        if ok.any():
            v = (col[ok] - t0) / 1e3
            res[name] = {"mean_us": round(float(v.mean()), 2), "min_us": round(float(v.min()), 2),
                         "max_us": round(float(v.max()), 2), "ctas": int(ok.sum())}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
