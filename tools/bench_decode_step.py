"""Decode tokens/s of a reference-shaped layer stack through ModelGraphRunner (CUDA graph,
TP = WORLD_SIZE, NCCL all-reduce after o_proj / down_proj) — BASELINE.json configs[3]/[1].

    python tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \\
        tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402


def run(shape_name="llama3-70b", layers=16, batch=64, ctx=8192, steps=20, warmup=3, verbose=False, profile=False):
    import torch.distributed as dist
    from core import parallel_config as pc
    from core.cache_manager import KVCacheManager
    from core.layer.model_graph import ModelGraphRunner
    from core.layer.synthetic_model import SHAPES, build_synthetic_model

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    pc.setup()
    shape = SHAPES[shape_name]
    assert shape.heads % world == 0 and shape.kv_heads % world == 0, "TP degree must divide the head counts"
    t0 = time.time()
    model = build_synthetic_model(shape, n_layers=layers, device=dev, seed=0, fast_init=True)
    Hq, Hkv, D = shape.heads // world, shape.kv_heads // world, shape.head_dim
    page = 256
    per_seq = -(-(ctx + 64) // page)
    runner = ModelGraphRunner(model, layers, Hq, D, Hkv, shape.hidden, 2 * shape.ffn // world, dev,
                              max_batch_size=batch, max_blocks=max(32, per_seq + 1))
    # the full-size originals are no longer needed once the per-rank slices exist
    for blk in model.transformer.h:
        for mod in (blk.attn.c_attn, blk.attn.c_proj, blk.mlp.w1, blk.mlp.w2, blk.mlp.c_proj):
            mod.weight.data = torch.empty(0, device=dev, dtype=torch.bfloat16)
    torch.cuda.empty_cache()
    mgr = KVCacheManager(batch * per_seq + 4, page, layers, Hkv, D, dtype=torch.bfloat16, device=dev,
                         max_tokens=per_seq * page, max_batch_size=batch)
    g = torch.Generator().manual_seed(7)                    # same permutation on every rank (replicated tables)
    order = torch.randperm(mgr.n_blocks, generator=g).tolist()
    mgr._free.clear()
    mgr._free.extend(order)
    ids = list(range(batch))
    for s in ids:
        ok, _ = mgr.alloc(s, ctx - 1)
        assert ok
    for l in range(layers):
        mgr.k_caches[l].normal_()
        mgr.v_caches[l].normal_()
    mgr.cache_batch_data(ids, [ctx - 1] * batch)
    runner.capture(mgr, batch_sizes=[batch])
    setup_s = time.time() - t0
    tok = torch.randint(0, shape.vocab, (batch,), device=dev)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        runner.forward(tok, mgr, batch)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        runner.forward(tok, mgr, batch)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # lm_head + final norm alone, through the runner's own head path (replicated, or vocab-parallel + all-gather —
    # every rank runs this loop, so the collective inside matches up): needed to extrapolate 16 -> 80 layers (§8d M2)
    h = torch.randn(batch, shape.hidden, device=dev, dtype=torch.bfloat16)
    with torch.no_grad():
        for _ in range(3):
            runner._lm_head(model.transformer.ln_f(h))
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            runner._lm_head(model.transformer.ln_f(h))
        e1.record()
    torch.cuda.synchronize()
    head_ms = e0.elapsed_time(e1) / 20
    layers_ms = max(ms - head_ms, 0.0)
    full_ms = layers_ms * shape.layers / layers + head_ms
    res = {"shape": shape_name, "tp": world, "layers_run": layers, "layers_model": shape.layers, "batch": batch,
           "ctx": ctx, "ms_per_step_run": round(ms, 4), "head_ms": round(head_ms, 4),
           "tokens_per_s_run": round(batch / (ms * 1e-3), 1),
           "tokens_per_s_full_depth_extrapolated": round(batch / (full_ms * 1e-3), 1),
           "setup_s": round(setup_s, 1),
           "allreduce": ("none (TP=1)" if world == 1 else
                         ("fused-" + runner._ar_mode) if runner._comm is not None else "nccl"),
           "gemm": runner.gemm_policy(), "ar_prefetch_MB": round(runner._prefetch_bytes / 1e6, 1),
           "lm_head": runner.head_policy()}
    if profile:
        # per-kernel table of the REAL TP run on rank 0 (torch.profiler / CUPTI around graph replays); under the
        # profiler kernels are serialised less tightly, so compare SHARES
        from torch.profiler import ProfilerActivity, profile as tprofile
        with tprofile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for _ in range(steps):
                runner.forward(tok, mgr, batch)
            torch.cuda.synchronize()
        rows = []
        for ev in prof.key_averages():
            t = getattr(ev, "device_time_total", None)
            if t is None:
                t = getattr(ev, "cuda_time_total", 0)
            if t and ev.device_type is not None and "cuda" in str(ev.device_type).lower():
                rows.append({"kernel": ev.key[:110], "calls_per_step": ev.count / steps,
                             "us_per_call": round(t / ev.count, 2), "us_per_layer": round(t / steps / layers, 2)})
        rows.sort(key=lambda r_: -r_["us_per_layer"])
        res["kernels"] = rows
        res["kernel_busy_us_per_layer"] = round(sum(r_["us_per_layer"] for r_ in rows), 1)
        res["us_per_layer_events"] = round(ms * 1e3 / layers, 2)
        if rank == 0:
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", f"layer_real_{shape_name}_tp{world}_b{batch}.json"), "w") as f:
                json.dump(res, f, indent=1)
    comm = getattr(runner, "_comm", None)
    if comm is not None:
        res["allreduce_timeout_flag"] = comm.error_flag()
    if rank == 0 and verbose:
        print(json.dumps(res), flush=True)
    # graphs that captured NCCL kernels must be gone before the communicator is destroyed
    del runner, mgr, model
    import gc
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="llama3-70b")
    ap.add_argument("--layers", type=int, default=16)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--profile", action="store_true", help="also write the per-kernel table of rank 0 (torch.profiler)")
    a = ap.parse_args()
    r = run(a.shape, a.layers, a.batch, a.ctx, a.steps, verbose=False, profile=a.profile)
    if int(os.environ.get("RANK", 0)) == 0:
        print(json.dumps({k: v for k, v in r.items() if k != "kernels"}), flush=True)
        for row in r.get("kernels", [])[:14]:
            print(f"{row['us_per_layer']:9.2f} us/layer  {row['us_per_call']:8.2f} us x {row['calls_per_step']:6.1f}  {row['kernel']}")
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
