"""BASELINE.json configs[2]: Qwen2.5-7B-shaped continuous batching, 64 ragged sequences
(context lengths log-uniform on [128, 32768], seed 0), one MIXED step = 56 decode rows + 8 prefill
chunks (512 - 2048 new tokens each on top of a cached prefix), attention path of all 28 layers.

Per layer: ONE balanced (stream-K) decode launch over the 56 ragged rows + one tcgen05 prefill
call per distinct chunk length (the prefill entry point takes a uniform S per call; prefixes are
ragged).  Reports step time (CUDA events around a graph replay of all layers), decode KV GB/s
against the algorithmic bytes, prefill TFLOP/s against 4*Hq*D*sum S*(P+(S+1)/2), and the same
step through flash-attn 2.8.3 (the library the reference calls) on the same inputs.
GPU box only; writes gpurun_out/continuous_q7.json.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402
decode_algorithmic_bytes = b200ops.decode_algorithmic_bytes

Hq, Hkv, D, LAYERS, PAGE = 28, 4, 128, 28, 256
N_SEQ, N_PREFILL = 64, 8
CHUNKS = [512, 512, 512, 512, 1024, 1024, 2048, 2048]


def run(layers=LAYERS, with_flash_attn=True, verbose=True):
    dev = torch.device("cuda")
    bf = torch.bfloat16
    g = torch.Generator().manual_seed(0)
    u = torch.rand(N_SEQ, generator=g)
    ctx = [int(128 * (32768 / 128) ** float(x)) for x in u]            # log-uniform on [128, 32768]
    # the 8 longest-but-one sequences are mid-prefill: their context = cached prefix + this step's chunk
    order = sorted(range(N_SEQ), key=lambda i: -ctx[i])
    pre_ids = order[1:1 + N_PREFILL]
    dec_ids = [i for i in range(N_SEQ) if i not in pre_ids]
    chunk = dict(zip(pre_ids, CHUNKS))
    prefix = {i: max(0, ctx[i] - chunk[i]) for i in pre_ids}
    need = [-(-(ctx[i] + 1) // PAGE) for i in range(N_SEQ)]
    n_blocks = sum(need) + 4
    width = max(need)
    perm = iter(torch.randperm(n_blocks, generator=g).tolist())
    table = torch.full((N_SEQ, width), -1, dtype=torch.int32)
    for i in range(N_SEQ):
        for j in range(need[i]):
            table[i, j] = next(perm)
    n_sets = 2                                                         # 2 cache sets alternate across layers: > L2 each
    caches = [(torch.randn(n_blocks, PAGE, Hkv, D, device=dev, dtype=bf),
               torch.randn(n_blocks, PAGE, Hkv, D, device=dev, dtype=bf)) for _ in range(n_sets)]
    cos, sin = build_rotary_tables(D, 32768 + 8)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)

    # decode rows
    Bd = len(dec_ids)
    d_table = table[dec_ids].contiguous().to(dev)
    d_lens = torch.tensor([ctx[i] - 1 for i in dec_ids], dtype=torch.int32, device=dev)
    dq = torch.randn(Bd, Hq, D, device=dev, dtype=bf)
    dk = torch.randn(Bd, Hkv, D, device=dev, dtype=bf)
    dv = torch.randn(Bd, Hkv, D, device=dev, dtype=bf)
    d_out = torch.empty_like(dq)
    # prefill groups by chunk length
    groups = []
    for S in sorted(set(CHUNKS)):
        ids = [i for i in pre_ids if chunk[i] == S]
        groups.append(dict(S=S, ids=ids, table=table[ids].contiguous().to(dev),
                           lens=torch.tensor([prefix[i] for i in ids], dtype=torch.int32, device=dev),
                           q=torch.randn(len(ids), S, Hq, D, device=dev, dtype=bf),
                           k=torch.randn(len(ids), S, Hkv, D, device=dev, dtype=bf),
                           v=torch.randn(len(ids), S, Hkv, D, device=dev, dtype=bf),
                           max_prefix=max(prefix[i] for i in ids)))
        groups[-1]["out"] = torch.empty_like(groups[-1]["q"])

    dec_bytes = decode_algorithmic_bytes([ctx[i] for i in dec_ids], Hq, Hkv, D, PAGE)
    pre_flops = sum(4 * Hq * D * chunk[i] * (prefix[i] + (chunk[i] + 1) / 2) for i in pre_ids)

    def ours_decode(l, balanced=True):
        kc, vc = caches[l % n_sets]
        b200ops.paged_attn_decode(dq, kc, vc, dk, dv, cos, sin, d_lens, d_table, out=d_out, balanced=balanced)

    def ours_prefill(l):
        kc, vc = caches[l % n_sets]
        for gr in groups:
            b200ops.paged_attn_prefill(gr["q"], gr["k"], gr["v"], kc, vc, cos, sin, gr["lens"], gr["table"],
                                       max_prefix=gr["max_prefix"], out=gr["out"])

    def time_graph(fn, reps=5):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    res = {"config": "qwen2.5-7b shape, 64 ragged seqs (128-32k ctx, log-uniform, seed 0): 56 decode rows + 8 prefill chunks",
           "layers": layers, "decode_rows": Bd, "decode_ctx_tokens": sum(ctx[i] for i in dec_ids),
           "prefill_chunks": [(chunk[i], prefix[i]) for i in pre_ids],
           "decode_bytes_per_layer": dec_bytes, "prefill_flops_per_layer": pre_flops}
    t_dec_bal = time_graph(lambda: [ours_decode(l, True) for l in range(layers)])
    t_dec_split = time_graph(lambda: [ours_decode(l, False) for l in range(layers)])
    t_pre = time_graph(lambda: [ours_prefill(l) for l in range(layers)])
    t_mixed = time_graph(lambda: [(ours_decode(l, True), ours_prefill(l)) for l in range(layers)])
    new_tokens = Bd + sum(CHUNKS)
    res["b200"] = {
        "decode_ms_balanced": round(t_dec_bal, 4), "decode_ms_split_grid": round(t_dec_split, 4),
        "decode_kv_GBps_balanced": round(dec_bytes * layers / t_dec_bal / 1e6, 1),
        "decode_kv_GBps_split_grid": round(dec_bytes * layers / t_dec_split / 1e6, 1),
        "prefill_ms": round(t_pre, 4), "prefill_TFLOPs": round(pre_flops * layers / t_pre / 1e9, 1),
        "mixed_step_ms": round(t_mixed, 4),
        "attention_tokens_per_s": round(new_tokens / (t_mixed * LAYERS / layers) * 1e3, 1)}
    if verbose:
        print(json.dumps(res["b200"]), flush=True)
    try:
        if not with_flash_attn:
            raise RuntimeError("skipped")
        from flash_attn import flash_attn_with_kvcache as fa

        def fa_decode(l):
            kc, vc = caches[l % n_sets]
            fa(dq.unsqueeze(1), kc, vc, k=dk.unsqueeze(1), v=dv.unsqueeze(1), rotary_cos=cos, rotary_sin=sin,
               cache_seqlens=d_lens, block_table=d_table, causal=True, rotary_interleaved=False, num_splits=0)

        def fa_prefill(l):
            kc, vc = caches[l % n_sets]
            for gr in groups:
                fa(gr["q"], kc, vc, k=gr["k"], v=gr["v"], rotary_cos=cos, rotary_sin=sin, cache_seqlens=gr["lens"],
                   block_table=gr["table"], causal=True, rotary_interleaved=False)
        f_dec = time_graph(lambda: [fa_decode(l) for l in range(layers)])
        f_pre = time_graph(lambda: [fa_prefill(l) for l in range(layers)])
        f_mixed = time_graph(lambda: [(fa_decode(l), fa_prefill(l)) for l in range(layers)])
        res["flash_attn_2.8.3"] = {
            "decode_ms": round(f_dec, 4), "decode_kv_GBps": round(dec_bytes * layers / f_dec / 1e6, 1),
            "prefill_ms": round(f_pre, 4), "prefill_TFLOPs": round(pre_flops * layers / f_pre / 1e9, 1),
            "mixed_step_ms": round(f_mixed, 4),
            "attention_tokens_per_s": round(new_tokens / (f_mixed * LAYERS / layers) * 1e3, 1)}
        if verbose:
            print(json.dumps(res["flash_attn_2.8.3"]), flush=True)
    except Exception as e:  # noqa: BLE001
        res["flash_attn_2.8.3"] = {"unavailable": str(e)[:200]}
    return res


def main():
    layers = int(sys.argv[1]) if len(sys.argv) > 1 else LAYERS
    res = run(layers)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "continuous_q7.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
