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


def run(shape_name="llama3-70b", layers=16, batch=64, ctx=8192, steps=20, warmup=3, verbose=False):
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
    # lm_head + final norm alone (replicated): needed to extrapolate 16 -> 80 layers
    h = torch.randn(batch, shape.hidden, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        model.lm_head(model.transformer.ln_f(h))
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        model.lm_head(model.transformer.ln_f(h))
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
           "gemm": runner.gemm_policy()}
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
    a = ap.parse_args()
    run(a.shape, a.layers, a.batch, a.ctx, a.steps, verbose=True)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
