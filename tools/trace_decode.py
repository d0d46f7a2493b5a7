"""Per-CTA timeline of the split-KV decode kernel (debug hook b200_debug_decode_trace): %globaltimer stamps of every
CTA at  0 entry | 1 first loads issued | 2 prologue done | 3 first stage landed | 4 streaming loop done | 5 partial
written | 6 exit (not the merging CTA) | 7 merge done (the last-arriving CTA of a segment), relative to the earliest
entry.  GPU box only.

    python tools/trace_decode.py [--shape l70tp8] [--splits 4] [--variant 0]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

SHAPES = {"l8": dict(B=32, Hq=32, Hkv=8, D=128, ctx=4096), "l70tp8": dict(B=64, Hq=8, Hkv=1, D=128, ctx=8192),
          "l70tp2": dict(B=64, Hq=32, Hkv=4, D=128, ctx=8192), "l8tp8": dict(B=32, Hq=4, Hkv=1, D=128, ctx=4096)}
NAMES = ["entry", "loads_issued", "prologue_done", "first_stage", "loop_done", "partial_written", "exit", "merge_done"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="l70tp8")
    ap.add_argument("--splits", type=int, default=0)
    ap.add_argument("--variant", type=int, default=0)
    a = ap.parse_args()
    sh = SHAPES[a.shape]
    B, Hq, Hkv, D, ctx = sh["B"], sh["Hq"], sh["Hkv"], sh["D"], sh["ctx"]
    dev = torch.device("cuda")
    page = 256
    per_seq = -(-(ctx + 8) // page)
    n_blocks = B * per_seq + 1
    g = torch.Generator(device="cpu").manual_seed(0)
    perm = torch.randperm(n_blocks - 1, generator=g)[: B * per_seq].view(B, per_seq).to(torch.int32) + 1
    table = perm.to(dev)
    caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev).to(torch.bfloat16),
               torch.randn(n_blocks, page, Hkv, D, device=dev).to(torch.bfloat16)) for _ in range(3)]
    lens = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    q = torch.randn(B, Hq, D, device=dev).to(torch.bfloat16)
    k = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16)
    v = torch.randn(B, Hkv, D, device=dev).to(torch.bfloat16)
    cos, sin = build_rotary_tables(D, ctx + 16)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    S = a.splits or b200ops.decode_num_splits(B, Hkv, per_seq * page, b200ops.sm_count(dev))
    out = torch.empty_like(q)

    def call(i):
        kc, vc = caches[i % 3]
        b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, out=out, num_splits=S, variant=a.variant,
                                  balanced=False)

    for i in range(6):
        call(i)
    torch.cuda.synchronize()
    ngroups = -(-(Hq // Hkv) // 16)
    ctas = S * Hkv * ngroups * B
    trace = torch.zeros(ctas * 8, dtype=torch.int64, device=dev)
    b200ops.lib.b200_debug_decode_trace(trace.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    call(0)
    e1.record()
    torch.cuda.synchronize()
    b200ops.lib.b200_debug_decode_trace(None)
    t = trace.view(ctas, 8).double().cpu()
    t0 = t[:, 0][t[:, 0] > 0].min()
    nbytes = b200ops.decode_algorithmic_bytes([ctx + 1] * B, Hq, Hkv, D)
    res = {"shape": a.shape, "splits": S, "variant": a.variant, "ctas": ctas, "event_us": round(e0.elapsed_time(e1) * 1e3, 2),
           "algorithmic_MB": round(nbytes / 1e6, 1), "ideal_us_at_6454GBps": round(nbytes / 6454.6e3, 2)}
    for i, name in enumerate(NAMES):
        col = t[:, i]
        ok = col > 0
        if ok.any():
            vv = (col[ok] - t0) / 1e3
            res[name] = {"mean_us": round(float(vv.mean()), 2), "min_us": round(float(vv.min()), 2),
                         "max_us": round(float(vv.max()), 2), "p10_us": round(float(vv.quantile(0.1)), 2),
                         "p90_us": round(float(vv.quantile(0.9)), 2), "ctas": int(ok.sum())}
    ok = (t[:, 3] > 0) & (t[:, 4] > 0)
    res["stream_phase_us_per_cta"] = {"mean": round(float(((t[ok, 4] - t[ok, 3]) / 1e3).mean()), 2),
                                      "min": round(float(((t[ok, 4] - t[ok, 3]) / 1e3).min()), 2),
                                      "max": round(float(((t[ok, 4] - t[ok, 3]) / 1e3).max()), 2)}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
