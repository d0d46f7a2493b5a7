"""Sweep decode-kernel variants x split counts at a BASELINE.json shape and time the library the
reference calls (flash-attn wheel) beside them.  GPU box only; writes gpurun_out/tune_<tag>.json.

    python tools/tune_decode.py [--shape l8|l70tp8|q7|long] [--variants 1,2,..] [--splits 1,2,4,..]
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
decode_algorithmic_bytes = b200ops.decode_algorithmic_bytes

SHAPES = {
    "l8": dict(B=32, Hq=32, Hkv=8, D=128, ctx=4096),
    "l70tp8": dict(B=64, Hq=8, Hkv=1, D=128, ctx=8192),
    "l70tp2": dict(B=64, Hq=32, Hkv=4, D=128, ctx=8192),
    "l8tp8": dict(B=32, Hq=4, Hkv=1, D=128, ctx=4096),        # per-GPU shard of bench.py --gpus 8
    "l8tp4": dict(B=32, Hq=8, Hkv=2, D=128, ctx=4096),
    "q7": dict(B=64, Hq=28, Hkv=4, D=128, ctx=2048),
    "long": dict(B=8, Hq=32, Hkv=8, D=128, ctx=131072),
    "q05": dict(B=1, Hq=14, Hkv=2, D=64, ctx=192),
    "l8b1": dict(B=1, Hq=32, Hkv=8, D=128, ctx=4096),
    "l8b8": dict(B=8, Hq=32, Hkv=8, D=128, ctx=4096),
    "l8b16": dict(B=16, Hq=32, Hkv=8, D=128, ctx=4096),
    # configs[2]: 64 ragged sequences, context lengths log-uniform on [128, 32768]
    "q7ragged": dict(B=64, Hq=28, Hkv=4, D=128, ctx=32768, ragged=True),
    "l8ragged": dict(B=32, Hq=32, Hkv=8, D=128, ctx=16384, ragged=True),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="l8")
    ap.add_argument("--variants", default="1,2,3,4,5,6,7,8")
    ap.add_argument("--splits", default="1,2,3,4,5,6,8,12,16")
    ap.add_argument("--iters", type=int, default=40)
    ap.add_argument("--no-fa", action="store_true")
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    sh = SHAPES[args.shape]
    B, Hq, Hkv, D, ctx = (sh[k] for k in ("B", "Hq", "Hkv", "D", "ctx"))
    page = 256
    per_seq = ctx // page
    dev = torch.device("cuda")
    if sh.get("ragged"):
        gl = torch.Generator().manual_seed(0)
        u = torch.rand(B, generator=gl)
        lens_list = [int(128 * (ctx / 128) ** float(x)) for x in u]          # log-uniform on [128, ctx]
    else:
        lens_list = [ctx] * B
    kv_bytes = sum(lens_list) * Hkv * D * 2 * 2
    n_sets = max(2, min(8, int(1.5e9 // kv_bytes) + 1))            # rotate caches so L2 never helps
    need = [-(-n // page) for n in lens_list]
    n_blocks = sum(need) + 4
    g = torch.Generator().manual_seed(0)
    caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16),
               torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=torch.bfloat16)) for _ in range(n_sets)]
    perm = torch.randperm(n_blocks, generator=g).tolist()
    table = torch.full((B, per_seq), -1, dtype=torch.int32)
    it = iter(perm)
    for b in range(B):
        for j in range(need[b]):
            table[b, j] = next(it)
    table = table.to(dev)
    lens = torch.tensor([n - 1 for n in lens_list], dtype=torch.int32, device=dev)
    q = torch.randn(B, Hq, D, device=dev, dtype=torch.bfloat16)
    k = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
    v = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
    cos, sin = build_rotary_tables(D, max(8192, ctx + 1))
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    out = torch.empty(B, Hq, D, device=dev, dtype=torch.bfloat16)
    nbytes = decode_algorithmic_bytes(lens_list, Hq, Hkv, D, page)

    def timeit(fn, iters=args.iters):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / iters            # us per call

    def timeit_graph(fn, iters=args.iters):
        """Same launches replayed from a CUDA graph: removes the Python/ctypes launch gap."""
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for i in range(n_sets):
                fn(i)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(n_sets):
                fn(i)
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(1, iters // n_sets)
        e0.record()
        for _ in range(reps):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * n_sets)

    results = []
    variants = [int(x) for x in args.variants.split(",")] if D == 128 else [1, 2, 3]
    for var in variants:
        for S in [int(x) for x in args.splits.split(",")]:
            def fn(i, var=var, S=S):
                kc, vc = caches[i % n_sets]
                if S == 0:          # 0 = balanced persistent (stream-K) schedule
                    b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, variant=var, out=out, balanced=True)
                else:
                    b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, num_splits=S, variant=var, out=out,
                                              pdl=args.pdl, balanced=False)
            try:
                us = timeit_graph(fn)
            except Exception as e:  # noqa: BLE001
                print(f"variant {var} splits {S}: {e}")
                continue
            gbs = nbytes / us / 1e3
            results.append(dict(variant=var, splits=S, us=round(us, 2), gbs=round(gbs, 1)))
            print(f"variant {var} splits {S:3d}: {us:8.2f} us  {gbs:7.1f} GB/s", flush=True)
    fa = {}
    if not args.no_fa:
        try:
            from flash_attn import flash_attn_with_kvcache
            for name, ns in (("reference_formula", max(1, 32 // max(1, B * 4))), ("heuristic", 0)):
                def fn(i, ns=ns):
                    kc, vc = caches[i % n_sets]
                    flash_attn_with_kvcache(q.unsqueeze(1), kc, vc, k=k.unsqueeze(1), v=v.unsqueeze(1), rotary_cos=cos,
                                            rotary_sin=sin, cache_seqlens=lens, block_table=table, causal=True,
                                            rotary_interleaved=False, num_splits=ns)
                us = timeit_graph(fn)
                fa[name] = dict(num_splits=ns, us=round(us, 2), gbs=round(nbytes / us / 1e3, 1))
                print(f"flash-attn 2.8.3 num_splits={ns} ({name}): {us:8.2f} us  {nbytes / us / 1e3:7.1f} GB/s", flush=True)
        except Exception as e:  # noqa: BLE001
            print("flash-attn timing unavailable:", e)
    best = min(results, key=lambda r: r["us"]) if results else None
    print("best:", best)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"tune_{args.shape}{'_pdl%d' % args.pdl if args.pdl else ''}{args.tag}.json"), "w") as f:
        json.dump(dict(shape=sh, algorithmic_bytes=nbytes, results=results, flash_attn=fa, best=best,
                       gpu=torch.cuda.get_device_name(0)), f, indent=1)


if __name__ == "__main__":
    main()
