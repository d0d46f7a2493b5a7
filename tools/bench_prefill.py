"""Prefill attention throughput (tensor-pipe bound): TFLOP/s of the tcgen05 kernel vs the
flash-attn wheel on the same inputs.  flops = 4 * Hq * D * sum_b S_b * (P_b + (S_b + 1) / 2)
(causal, prefix P_b).  GPU box only; writes gpurun_out/prefill_bench.json."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

CASES = [  # name, B, S, prefix, Hq, Hkv, D
    ("l8_b1_s2048", 1, 2048, 0, 32, 8, 128),
    ("l8_b1_s8192", 1, 8192, 0, 32, 8, 128),
    ("l8_b4_s2048", 4, 2048, 0, 32, 8, 128),
    ("l8_b1_s2048_p6144", 1, 2048, 6144, 32, 8, 128),
    ("q7_b2_s4096", 2, 4096, 0, 28, 4, 128),
    ("q05_b4_s2048", 4, 2048, 0, 14, 2, 64),
    ("l8_b4_s512", 4, 512, 0, 32, 8, 128),
    ("l8_b8_s1024", 8, 1024, 0, 32, 8, 128),
    ("q7_b1_s2048", 1, 2048, 0, 28, 4, 128),
]


def main():
    only = sys.argv[1].split(",") if len(sys.argv) > 1 else None
    dev = torch.device("cuda")
    page = 256
    res = []
    try:
        from flash_attn import flash_attn_with_kvcache as fa
    except Exception:  # noqa: BLE001
        fa = None
    for name, B, S, P, Hq, Hkv, D in CASES:
        if only and name not in only:
            continue
        per_seq = -(-(P + S) // page)
        n_blocks = B * per_seq + 2
        kc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16)
        vc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16)
        table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
        lens = torch.full((B,), P, dtype=torch.int32, device=dev)
        q = torch.randn(B, S, Hq, D, device=dev, dtype=torch.bfloat16)
        k = torch.randn(B, S, Hkv, D, device=dev, dtype=torch.bfloat16)
        v = torch.randn(B, S, Hkv, D, device=dev, dtype=torch.bfloat16)
        cos, sin = build_rotary_tables(D, P + S + 1)
        cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
        out = torch.empty_like(q)
        flops = 4 * Hq * D * B * S * (P + (S + 1) / 2)

        def timeit(fn, iters=10):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters

        ms = timeit(lambda: b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out))
        # the same two launches replayed from a CUDA graph: no Python / ctypes / tensor-map encode time
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out)
        ms_graph = timeit(g.replay)
        rec = dict(case=name, ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1), graph_ms=round(ms_graph, 4),
                   graph_tflops=round(flops / ms_graph / 1e9, 1))
        for var, key in ((4, "sm_poly25"), (6, "pair_poly25"), (9, "persist_poly25"), (10, "persist_packed"), (11, "persist_poly50")):
            msv = timeit(lambda: b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out, variant=var))
            rec[key + "_tflops"] = round(flops / msv / 1e9, 1)
        if fa is not None:
            ms_fa = timeit(lambda: fa(q, kc, vc, k=k, v=v, rotary_cos=cos, rotary_sin=sin, cache_seqlens=lens,
                                      block_table=table, causal=True, rotary_interleaved=False))
            rec.update(fa_ms=round(ms_fa, 4), fa_tflops=round(flops / ms_fa / 1e9, 1))
        print(rec, flush=True)
        res.append(rec)
        del kc, vc, q, k, v, out
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "prefill_bench.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
