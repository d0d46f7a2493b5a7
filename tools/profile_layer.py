"""Per-kernel time table of one graph-replayed decode step at the PER-RANK shapes of a
tensor-parallel run, on ONE GPU (no collectives: world size 1, the all-reduces are no-ops).
It answers "where does a TP=k layer spend its time apart from NVLink": the KV-head-sharded
attention, the sharded GEMMs and the launch-bound elementwise kernels.

    python tools/profile_layer.py --shape llama3-70b --tp 8 --layers 4 --batch 64 --ctx 8192

Uses torch.profiler (CUPTI) around graph replays; writes gpurun_out/layer_<shape>_tp<k>.json.
"""
import argparse
import json
import os
import sys
from dataclasses import replace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="llama3-70b")
    ap.add_argument("--tp", type=int, default=8)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--no-profile", action="store_true", help="event timing only (A/B runs of env switches)")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    from torch.profiler import ProfilerActivity, profile
    from core.cache_manager import KVCacheManager
    from core.layer.model_graph import ModelGraphRunner
    from core.layer.synthetic_model import SHAPES, build_synthetic_model

    full = SHAPES[a.shape]
    assert full.heads % a.tp == 0 and full.kv_heads % a.tp == 0 and full.ffn % a.tp == 0
    # the rank-local slice of a TP=k run has exactly these dimensions
    shape = replace(full, heads=full.heads // a.tp, kv_heads=full.kv_heads // a.tp, ffn=full.ffn // a.tp,
                    vocab=32000)
    dev = "cuda:0"
    torch.cuda.set_device(0)
    model = build_synthetic_model(shape, n_layers=a.layers, device=dev, seed=0, fast_init=True)
    page = 256
    per_seq = -(-(a.ctx + 64) // page)
    runner = ModelGraphRunner(model, a.layers, shape.heads, shape.head_dim, shape.kv_heads, shape.hidden,
                              2 * shape.ffn, dev, max_batch_size=a.batch, max_blocks=max(32, per_seq + 1))
    mgr = KVCacheManager(a.batch * per_seq + 4, page, a.layers, shape.kv_heads, shape.head_dim,
                         dtype=torch.bfloat16, device=dev, max_tokens=per_seq * page, max_batch_size=a.batch)
    order = torch.randperm(mgr.n_blocks, generator=torch.Generator().manual_seed(7)).tolist()
    mgr._free.clear()
    mgr._free.extend(order)
    ids = list(range(a.batch))
    for s in ids:
        ok, _ = mgr.alloc(s, a.ctx - 1)
        assert ok
    for l in range(a.layers):
        mgr.k_caches[l].normal_()
        mgr.v_caches[l].normal_()
    mgr.cache_batch_data(ids, [a.ctx - 1] * a.batch)
    runner.capture(mgr, batch_sizes=[a.batch])
    tok = torch.randint(0, shape.vocab, (a.batch,), device=dev)
    for _ in range(3):
        runner.forward(tok, mgr, a.batch)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        runner.forward(tok, mgr, a.batch)
    e1.record()
    torch.cuda.synchronize()
    step_ms = e0.elapsed_time(e1) / a.steps
    if a.no_profile:
        print(json.dumps({"tag": a.tag, "shape": a.shape, "tp_local_of": a.tp, "layers": a.layers, "batch": a.batch,
                          "ctx": a.ctx, "step_ms_events": round(step_ms, 4),
                          "us_per_layer_events": round(step_ms * 1e3 / a.layers, 2), "gemm": runner.gemm_policy()}))
        return
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(a.steps):
            runner.forward(tok, mgr, a.batch)
        torch.cuda.synchronize()
    rows = []
    for ev in prof.key_averages():
        t = getattr(ev, "device_time_total", None)
        if t is None:
            t = getattr(ev, "cuda_time_total", 0)
        if t and ev.device_type is not None and "cuda" in str(ev.device_type).lower():
            rows.append({"kernel": ev.key[:110], "calls_per_step": ev.count / a.steps,
                         "us_per_call": round(t / ev.count, 2),
                         "us_per_layer": round(t / a.steps / a.layers, 2)})
    rows.sort(key=lambda r: -r["us_per_layer"])
    busy = sum(r["us_per_call"] * r["calls_per_step"] for r in rows)
    res = {"shape": a.shape, "tp_local_of": a.tp, "layers": a.layers, "batch": a.batch, "ctx": a.ctx,
           "step_ms_events": round(step_ms, 4), "us_per_layer_events": round(step_ms * 1e3 / a.layers, 2),
           "kernel_busy_us_per_step": round(busy, 1), "kernels": rows}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", f"layer_{a.shape}_tp{a.tp}_b{a.batch}.json")
    with open(path, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps({k: v for k, v in res.items() if k != "kernels"}))
    for r in rows[:16]:
        print(f"{r['us_per_layer']:9.2f} us/layer  {r['us_per_call']:8.2f} us x {r['calls_per_step']:6.1f}  {r['kernel']}")


if __name__ == "__main__":
    main()
