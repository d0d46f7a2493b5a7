"""Balanced (stream-K) decode form on ragged batches: BASELINE.json configs[2] (56 decode rows, context log-uniform
128..32768, Qwen2.5-7B heads), a Llama-3-8B-headed ragged batch and a uniform batch, graph replay over two cache sets.
B200_DECODE_SK_OV selects the per-segment weight (read once per process).  GPU box only.

    B200_DECODE_SK_OV=2 python tools/bench_ragged.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

dev = torch.device("cuda")
bf = torch.bfloat16


def case(name, lens_after, Hq, Hkv, D, balanced, calls=8, n_sets=2):
    page = 256
    need = [-(-n // page) for n in lens_after]
    n_blocks = sum(need) + 4
    g = torch.Generator().manual_seed(0)
    perm = iter(torch.randperm(n_blocks, generator=g).tolist())
    table = torch.full((len(lens_after), max(need)), -1, dtype=torch.int32)
    for i, nb in enumerate(need):
        for j in range(nb):
            table[i, j] = next(perm)
    table = table.to(dev)
    caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf),
               torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)) for _ in range(n_sets)]
    cos, sin = build_rotary_tables(D, max(lens_after) + 8)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    B = len(lens_after)
    lens = torch.tensor([n - 1 for n in lens_after], dtype=torch.int32, device=dev)
    q = torch.randn(B, Hq, D, device=dev, dtype=bf)
    k = torch.randn(B, Hkv, D, device=dev, dtype=bf)
    v = torch.randn(B, Hkv, D, device=dev, dtype=bf)
    o = torch.empty(B, Hq, D, device=dev, dtype=bf)

    def fn():
        for i in range(calls):
            kc, vc = caches[i % n_sets]
            b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, out=o, balanced=balanced)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        fn()
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 10 / calls * 1e3
    nbytes = b200ops.decode_algorithmic_bytes(lens_after, Hq, Hkv, D)
    return {"case": name, "balanced": balanced, "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1)}


def main():
    g = torch.Generator().manual_seed(0)
    u = torch.rand(64, generator=g)
    ctx = [int(128 * (32768 / 128) ** float(x)) for x in u]
    order = sorted(range(64), key=lambda i: -ctx[i])
    pre_ids = set(order[1:9])
    dec = [ctx[i] for i in range(64) if i not in pre_ids]
    res = {"sk_ov": os.environ.get("B200_DECODE_SK_OV", "default")}
    rows = [case("cfg3_q7_56_ragged", dec, 28, 4, 128, True),
            case("cfg3_q7_56_ragged", dec, 28, 4, 128, None),
            case("l8_64_ragged_128_32k", ctx, 32, 8, 128, True),
            case("q7_b64_2k_uniform", [2048] * 64, 28, 4, 128, True),
            case("q7_b64_2k_uniform", [2048] * 64, 28, 4, 128, False),
            case("l70tp1_b64_8k_uniform", [8192] * 64, 64, 8, 128, True, calls=4),
            case("l70tp1_b64_8k_uniform", [8192] * 64, 64, 8, 128, False, calls=4)]
    res["rows"] = rows
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
