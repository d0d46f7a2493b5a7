#!/usr/bin/env python
"""bench.py — paged decode attention KV GB/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload: BASELINE.json configs[1] — Llama-3-8B shape (32 q heads, 8 kv heads, d = 128), bf16,
batch 32, 4k context (cache_seqlens = 4095, one token appended -> 4096), block = 256, pages drawn
from a seeded random permutation of the pool.  One STEP = the attention of one decode step: one
fused RoPE + KV-append + split-KV + merge launch per layer for all 32 layers, each layer with its
own KV cache (16 GiB of KV per step >> 126 MB L2, nothing is served from cache between steps).

  value   : algorithmic bytes of the WHOLE (unsharded) workload / device time,
            inputs resident in HBM, the 32 launches replayed from one CUDA graph (the way
            ModelGraphRunner drives the op); MAX time over ranks.
  e2e     : same metric through the public operator API with HOST inputs: per step the q/k/v of
            every layer come from pinned host memory (H2D), block tables and sequence lengths are
            rebuilt from Python lists by KVCacheManager.cache_batch_data, the outputs of every
            layer are read back (D2H); wall clock between device synchronisations.  The KV cache
            is device-resident state (as in the reference, core/cache_manager.py:311-323).
  N > 1   : the reference's tensor parallelism (core/engine.py:108-114): the SAME workload is
            sharded by KV head — rank r owns Hq/N query heads, Hkv/N kv heads and a cache
            [n_blocks, 256, Hkv/N, D]; block tables and lengths are replicated (every rank runs the
            same deterministic allocator); attention needs no collective.  Strong scaling:
            value = unsharded bytes / MAX-over-ranks time.
  decode_step (every N): decode tokens/s of a Llama-3-70B-shaped 16-layer stack (BASELINE.json
            configs[3]: batch 64, 8k context) through ModelGraphRunner at TP = N — KV-head sharded
            attention, o_proj / down_proj partial sums reduced over NVLink inside the CUDA graph
            (core/layer/model_graph.py:210, :219) — outside the headline's timed region.
  tp_check (N > 1): fused NVLink reduction vs NCCL all_reduce on the same weights: ranks
            bit-identical, relative rms difference, no flag-wait timeout.
  configs (N = 1): BASELINE.json configs[2] (64 ragged sequences: the 56 decode rows alone and the whole mixed
            prefill + decode step) and configs[4] (batch 8, 128k context) decode attention, prefill attention
            TFLOP/s, elementwise GB/s.  (N > 1): the per-GPU shard of configs[4] at TP = N.
  --impl reference : the reference semantics on the host CPU cores (oracle port, torch eager,
            all threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]

import torch  # noqa: E402

METRIC = "paged_decode_attn_kv_GBps"
UNIT = "GB/s"
CFG = dict(B=32, Hq=32, Hkv=8, D=128, ctx=4096, layers=32, page=256)


def workload_name():
    return (f"llama3-8b-shape decode attention, batch {CFG['B']}, ctx {CFG['ctx']}, "
            f"{CFG['layers']} layers/step, block {CFG['page']}, bf16")


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            j = json.load(f)
        return (float(j["hbm_gbs"]), float(j.get("bf16_tflops_sustained", 1400.0)),
                "measured (MEASURED_PEAKS.json)")
    return 6650.0, 1400.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:  # noqa: BLE001
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for n, bit in names.items():
                    if mask & bit:
                        self.reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---------------------------------------------------------------------------------- CPU arm
class CpuArm:
    """Reference semantics on the host cores: oracle port of the decode call site (rotate q/k at
    cache_seqlens, append into the page, gather pages by block table, softmax . V), torch eager.
    One layer call over `rows` of the 32 batch rows.  The ONLY user of oracle/ in this file."""

    def __init__(self, rows: int, threads: int):
        from oracle.paged_attention import decode_algorithmic_bytes
        from oracle.rotary import rotary_table_oracle
        torch.set_num_threads(threads)
        g = torch.Generator().manual_seed(0)
        B, Hq, Hkv, D, page, ctx = rows, CFG["Hq"], CFG["Hkv"], CFG["D"], CFG["page"], CFG["ctx"]
        per_seq = ctx // page
        n_blocks = B * per_seq
        self.table = torch.randperm(n_blocks, generator=g).view(B, per_seq).to(torch.int32)
        self.k_cache = torch.randn(n_blocks, page, Hkv, D, generator=g).to(torch.bfloat16)
        self.v_cache = torch.randn(n_blocks, page, Hkv, D, generator=g).to(torch.bfloat16)
        self.q = torch.randn(B, 1, Hq, D, generator=g).to(torch.bfloat16)
        self.k = torch.randn(B, 1, Hkv, D, generator=g).to(torch.bfloat16)
        self.v = torch.randn(B, 1, Hkv, D, generator=g).to(torch.bfloat16)
        self.cos, self.sin = rotary_table_oracle(D, 8192)
        self.lens = torch.full((B,), ctx - 1, dtype=torch.int32)
        self.nbytes = decode_algorithmic_bytes([ctx] * B, Hq, Hkv, D, page)

    def call(self):
        from oracle.paged_attention import paged_attention_oracle
        return paged_attention_oracle(self.q, self.k_cache, self.v_cache, self.k, self.v, self.cos, self.sin,
                                      self.lens, self.table, None, True, False)

    def time(self, reps: int) -> float:
        t0 = time.perf_counter()
        for _ in range(reps):
            self.call()
        return (time.perf_counter() - t0) / reps


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    rows, reps = 2, 1
    # one "step" of this arm = `reps` layer calls over `rows` batch rows (bounded sample)
    arm = CpuArm(rows, threads)
    for _ in range(args.warmup):
        arm.call()
    times = [arm.time(reps) for _ in range(args.steps)]
    dt = sum(times) / len(times)
    val = arm.nbytes / dt / 1e9
    sample = f"{rows} of {CFG['B']} batch rows, 1 of {CFG['layers']} layers per call, {reps} calls per step"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 4), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * reps * 1e3, 3),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(),
                   "note": "reference semantics in torch eager on host CPU cores (oracle port)"},
        "cpu_baseline": {"value": round(val, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(val, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------- helpers
def _graph_of(fn):
    """Warm `fn` up on a side stream, then capture it on torch's capture stream."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g


def _time_graph(g, reps=10, warm=3):
    for _ in range(warm):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def _release():
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------- configs[4] shard (N > 1)
def long_context_shard(dev, world, peak_gbs, emulate_rank_count=None):
    """BASELINE.json configs[4] at TP = world: the per-GPU shard (Hq/world query heads, Hkv/world kv heads) of the batch-8,
    128k-context decode attention call; every rank times its own shard (graph replay over two cache sets, CUDA events),
    the figure is this rank's.  `whole_job_kv_GBps` = unsharded bytes / this rank's time."""
    import b200ops
    from core.paged_attention import build_rotary_tables
    bf = torch.bfloat16
    n = emulate_rank_count or world
    Hq, Hkv, D, page, B, ctx = 32 // n, 8 // n, 128, 256, 8, 131072
    per_seq = ctx // page
    n_blocks = B * per_seq + 4
    g = torch.Generator().manual_seed(0)
    table = torch.randperm(n_blocks, generator=g)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
    caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf),
               torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)) for _ in range(2)]
    cos, sin = build_rotary_tables(D, ctx + 8)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    lens = torch.full((B,), ctx - 1, dtype=torch.int32, device=dev)
    q = torch.randn(B, Hq, D, device=dev, dtype=bf)
    k = torch.randn(B, Hkv, D, device=dev, dtype=bf)
    v = torch.randn(B, Hkv, D, device=dev, dtype=bf)
    o = torch.empty(B, Hq, D, device=dev, dtype=bf)
    calls = 4

    def fn():
        for i in range(calls):
            kc, vc = caches[i % 2]
            b200ops.paged_attn_decode(q, kc, vc, k, v, cos, sin, lens, table, out=o)
    ms = _time_graph(_graph_of(fn)) / calls
    local = b200ops.decode_algorithmic_bytes([ctx] * B, Hq, Hkv, D)
    total = b200ops.decode_algorithmic_bytes([ctx] * B, 32, 8, D)
    rec = {"tp": n, "Hq_local": Hq, "Hkv_local": Hkv, "us_per_call": round(ms * 1e3, 2),
           "kv_GBps_per_gpu": round(local / ms / 1e6, 1), "frac_of_measured_hbm": round(local / ms / 1e6 / peak_gbs, 4),
           "whole_job_kv_GBps": round(total / ms / 1e6, 1), "algorithmic_bytes_per_call_per_gpu": local}
    del caches
    _release()
    return rec


# ---------------------------------------------------------------------------------- secondary configs (N = 1)
def secondary_configs(dev, peak_gbs, peak_tf):
    """BASELINE.json configs[2] / configs[4], prefill and elementwise figures; each timed on its own with
    CUDA events around graph replays, inputs larger than L2 or rotated so that L2 never helps."""
    import b200ops
    from core.paged_attention import build_rotary_tables
    bf = torch.bfloat16
    out = {}

    def paged_case(lens_after, Hq, Hkv, D, n_sets, rot_rows):
        page = 256
        need = [-(-n // page) for n in lens_after]
        n_blocks = sum(need) + 4
        g = torch.Generator().manual_seed(0)
        perm = iter(torch.randperm(n_blocks, generator=g).tolist())
        table = torch.full((len(lens_after), max(need)), -1, dtype=torch.int32)
        for i, nb in enumerate(need):
            for j in range(nb):
                table[i, j] = next(perm)
        caches = [(torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf),
                   torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)) for _ in range(n_sets)]
        cos, sin = build_rotary_tables(D, rot_rows)
        B = len(lens_after)
        return dict(table=table.to(dev), caches=caches, cos=cos[:, : D // 2].contiguous().to(dev),
                    sin=sin[:, : D // 2].contiguous().to(dev),
                    lens=torch.tensor([n - 1 for n in lens_after], dtype=torch.int32, device=dev),
                    q=torch.randn(B, Hq, D, device=dev, dtype=bf), k=torch.randn(B, Hkv, D, device=dev, dtype=bf),
                    v=torch.randn(B, Hkv, D, device=dev, dtype=bf), o=torch.empty(B, Hq, D, device=dev, dtype=bf))

    def decode_line(name, lens_after, Hq, Hkv, D, n_sets, balanced, rot_rows, calls):
        c = paged_case(lens_after, Hq, Hkv, D, n_sets, rot_rows)

        def fn():
            for i in range(calls):
                kc, vc = c["caches"][i % n_sets]
                b200ops.paged_attn_decode(c["q"], kc, vc, c["k"], c["v"], c["cos"], c["sin"], c["lens"], c["table"],
                                          out=c["o"], balanced=balanced)
        ms = _time_graph(_graph_of(fn)) / calls
        nbytes = b200ops.decode_algorithmic_bytes(lens_after, Hq, Hkv, D)
        gbs = nbytes / ms / 1e6
        rec = {"kv_GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / peak_gbs, 4), "us_per_call": round(ms * 1e3, 2),
               "algorithmic_bytes_per_call": nbytes, "schedule": "balanced stream-K" if balanced else "split-KV grid"}
        try:
            from flash_attn import flash_attn_with_kvcache as fa

            def ffn():
                for i in range(calls):
                    kc, vc = c["caches"][i % n_sets]
                    fa(c["q"].unsqueeze(1), kc, vc, k=c["k"].unsqueeze(1), v=c["v"].unsqueeze(1), rotary_cos=c["cos"],
                       rotary_sin=c["sin"], cache_seqlens=c["lens"], block_table=c["table"], causal=True,
                       rotary_interleaved=False, num_splits=0)
            rec["flash_attn_2.8.3_us_per_call"] = round(_time_graph(_graph_of(ffn)) / calls * 1e3, 2)
        except Exception as e:  # noqa: BLE001
            rec["flash_attn_2.8.3_us_per_call"] = None
            rec["flash_attn_note"] = f"{type(e).__name__}"[:80]
        out[name] = rec
        del c
        _release()

    # configs[2]: 64 ragged sequences, context log-uniform on [128, 32768] (seed 0); the 56 decode rows
    g = torch.Generator().manual_seed(0)
    u = torch.rand(64, generator=g)
    ctx = [int(128 * (32768 / 128) ** float(x)) for x in u]
    order = sorted(range(64), key=lambda i: -ctx[i])
    pre_ids = set(order[1:9])
    dec = [ctx[i] for i in range(64) if i not in pre_ids]
    decode_line("cfg3_qwen2.5-7b_56_ragged_decode_rows", dec, 28, 4, 128, 2, True, 32768 + 8, 8)
    # configs[4]: batch 8, 128k context (4 GiB of KV per call, two cache sets)
    decode_line("cfg5_llama3-8b_b8_128k", [131072] * 8, 32, 8, 128, 2, None, 131072 + 8, 4)
    # L70 TP=8 shard shape (one kv head per GPU) — the per-GPU attention call of configs[3]
    decode_line("cfg4_llama3-70b_tp8_shard_b64_8k", [8192] * 64, 8, 1, 128, 4, None, 8200, 8)

    # prefill attention at the reference's size (engine.py:45 caps prefill at 2048 tokens)
    def prefill_line(name, B, S, Hq, Hkv, D):
        page = 256
        per_seq = -(-S // page)
        n_blocks = B * per_seq + 2
        kc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
        vc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
        table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
        lens = torch.zeros(B, dtype=torch.int32, device=dev)
        q, k, v = (torch.randn(B, S, h, D, device=dev, dtype=bf) for h in (Hq, Hkv, Hkv))
        cos, sin = build_rotary_tables(D, S + 1)
        cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
        o = torch.empty_like(q)
        ms = _time_graph(_graph_of(lambda: b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=o)))
        flops = 4 * Hq * D * B * S * ((S + 1) / 2)
        tf = flops / ms / 1e9
        out[name] = {"tflops": round(tf, 1), "frac_of_sustained_bf16": round(tf / peak_tf, 4), "ms": round(ms, 4)}
        _release()

    prefill_line("prefill_llama3-8b_b1_s2048", 1, 2048, 32, 8, 128)
    prefill_line("prefill_llama3-8b_b4_s512", 4, 512, 32, 8, 128)
    prefill_line("prefill_qwen2.5-0.5b_b4_s2048", 4, 2048, 14, 2, 64)

    # configs[2] as ONE mixed step: 56 ragged decode rows (balanced launch) + 8 prefill chunks (512-2048 new tokens on
    # cached prefixes) per layer, 4 of the 28 layers in one CUDA graph (tools/bench_continuous.py), flash-attn beside it
    try:
        from bench_continuous import run as run_mixed
        mr = run_mixed(layers=4, with_flash_attn=True, verbose=False)
        rec = dict(mr["b200"])
        rec["layers_timed"] = 4
        fa_ = mr.get("flash_attn_2.8.3", {})
        rec["flash_attn_2.8.3_mixed_step_ms"] = fa_.get("mixed_step_ms")
        rec["prefill_frac_of_sustained_bf16"] = round(rec["prefill_TFLOPs"] / peak_tf, 4)
        rec["decode_frac_of_measured_hbm"] = round(rec["decode_kv_GBps_balanced"] / peak_gbs, 4)
        out["cfg3_mixed_step_qwen2.5-7b_56_decode_rows_plus_8_prefill_chunks"] = rec
    except Exception as e:  # noqa: BLE001
        out["cfg3_mixed_step_qwen2.5-7b_56_decode_rows_plus_8_prefill_chunks"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    _release()

    # elementwise at 8192 rows (HBM-bound sizes), Llama-3-8B widths
    rows = 8192
    gu = torch.randn(rows, 2 * 14336, device=dev, dtype=bf)
    so = torch.empty(rows, 14336, device=dev, dtype=bf)
    ms = _time_graph(_graph_of(lambda: b200ops.swiglu(gu, so)))
    out["swiglu_8192x14336"] = {"GBps": round(rows * 14336 * 6 / ms / 1e6, 1),
                                "frac_of_measured_hbm": round(rows * 14336 * 6 / ms / 1e6 / peak_gbs, 4)}
    del gu, so
    xs = [torch.randn(rows, 4096, device=dev, dtype=bf) for _ in range(4)]
    rs = [torch.randn(rows, 4096, device=dev, dtype=bf) for _ in range(4)]
    w = torch.ones(4096, device=dev, dtype=bf)
    y, ro = torch.empty_like(xs[0]), torch.empty_like(xs[0])

    def norm4():
        for i in range(4):
            b200ops.rmsnorm(xs[i], w, y, 1e-6)

    def resnorm4():
        for i in range(4):
            b200ops.rmsnorm_residual(xs[i], rs[i], w, y, ro, 1e-6)
    ms = _time_graph(_graph_of(norm4)) / 4
    nb = rows * 4096 * 4 + 8192
    out["rmsnorm_8192x4096"] = {"GBps": round(nb / ms / 1e6, 1), "frac_of_measured_hbm": round(nb / ms / 1e6 / peak_gbs, 4)}
    ms = _time_graph(_graph_of(resnorm4)) / 4
    nb = rows * 4096 * 8 + 8192
    out["rmsnorm_residual_8192x4096"] = {"GBps": round(nb / ms / 1e6, 1),
                                         "frac_of_measured_hbm": round(nb / ms / 1e6 / peak_gbs, 4)}
    del xs, rs
    _release()
    return out


# ---------------------------------------------------------------------------------- B200 arm
def run_b200_arm(args):
    import torch.distributed as dist
    import b200ops
    from core.cache_manager import KVCacheManager
    from core.paged_attention import build_rotary_tables
    from core import parallel_config

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        parallel_config.setup()

    B, Hq, Hkv, D, ctx, L, page = (CFG[k] for k in ("B", "Hq", "Hkv", "D", "ctx", "layers", "page"))
    if Hq % world or Hkv % world:
        raise SystemExit(f"bench.py: TP degree {world} must divide the head counts ({Hq} q / {Hkv} kv)")
    Hq_l, Hkv_l = Hq // world, Hkv // world          # core/engine.py:108-114
    per_seq = ctx // page
    n_blocks = B * per_seq + 8
    mgr = KVCacheManager(n_blocks, page, L, Hkv_l, D, dtype=torch.bfloat16, device=dev, max_tokens=ctx,
                         max_batch_size=B)
    # non-contiguous pages: hand the allocator a seeded permutation of the pool — the SAME on every rank
    # (block tables are replicated: every rank runs the same allocator on the same sequence list)
    g = torch.Generator().manual_seed(1234)
    order = torch.randperm(n_blocks, generator=g).tolist()
    mgr._free.clear()
    mgr._free.extend(order)
    seq_ids = list(range(B))
    for sid in seq_ids:
        ok, _ = mgr.alloc(sid, ctx - 1)
        assert ok
    lens = [ctx - 1] * B            # the launch appends token 4095 -> 4096 keys per sequence
    torch.manual_seed(rank)
    for l in range(L):
        mgr.k_caches[l].normal_()
        mgr.v_caches[l].normal_()
    cos_full, sin_full = build_rotary_tables(D, 8192)
    cos = cos_full[:, : D // 2].contiguous().to(dev)
    sin = sin_full[:, : D // 2].contiguous().to(dev)

    qkv_host = torch.randn(L, B, (Hq_l + 2 * Hkv_l) * D, dtype=torch.float32).to(torch.bfloat16).pin_memory()
    out_host = torch.empty(L, B, Hq_l * D, dtype=torch.bfloat16).pin_memory()
    # two device-side sets of step inputs / outputs: the e2e leg copies step t+1 in and step t-1 out while step t runs
    qkv_devs = [torch.empty_like(qkv_host, device=dev) for _ in range(2)]
    out_devs = [torch.empty(L, B, Hq_l, D, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    qkv_dev, out_dev = qkv_devs[0], out_devs[0]
    for t_ in qkv_devs:
        t_.copy_(qkv_host)
    block_table, seqlens = mgr.cache_batch_data(seq_ids, lens)
    num_splits = args.splits if args.splits > 0 else 0

    def layer_call(l, s_=0):
        x = qkv_devs[s_][l]
        q = x[:, : Hq_l * D].view(B, Hq_l, D)
        k = x[:, Hq_l * D: (Hq_l + Hkv_l) * D].view(B, Hkv_l, D)
        v = x[:, (Hq_l + Hkv_l) * D:].view(B, Hkv_l, D)
        b200ops.paged_attn_decode(q, mgr.k_caches[l], mgr.v_caches[l], k, v, cos, sin, seqlens, block_table,
                                  rotary_interleaved=False, out=out_devs[s_][l], num_splits=num_splits,
                                  variant=args.variant, pdl=args.pdl)

    def layers(lo, hi, s_=0):
        for l in range(lo, hi):
            layer_call(l, s_)

    graph = _graph_of(lambda: layers(0, L))      # the 32 launches of one decode step in one graph
    n_chunks = args.e2e_chunks                   # e2e: host<->device copies pipelined against compute
    bounds = [(c * L // n_chunks, (c + 1) * L // n_chunks) for c in range(n_chunks)]
    chunk_graphs = [[_graph_of(lambda lo=lo, hi=hi, s_=s_: layers(lo, hi, s_)) for lo, hi in bounds] for s_ in (0, 1)]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    bytes_layer_total = b200ops.decode_algorithmic_bytes([ctx] * B, Hq, Hkv, D, page)       # unsharded workload
    bytes_layer_local = b200ops.decode_algorithmic_bytes([ctx] * B, Hq_l, Hkv_l, D, page)   # what one launch moves
    bytes_step = bytes_layer_total * L

    # ---- value: device-resident, graph replay, CUDA events on the launching stream
    for _ in range(args.warmup):
        graph.replay()
    sync_all()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        ev0.record()
        for _ in range(args.steps):
            graph.replay()
        ev1.record()
        torch.cuda.synchronize()
    ms_total = ev0.elapsed_time(ev1)
    sync_all()
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = bytes_step / (ms_step * 1e-3) / 1e9
    kernel_us = ms_step * 1e3 / L
    achieved = bytes_layer_local / (kernel_us * 1e-6) / 1e9

    # ---- e2e: host buffers in, host buffers out, metadata rebuilt from Python lists every step.
    # The layers are driven in chunks (--e2e-chunks, default 2: measured 0.968 of `value` at 2, 0.945 at 8,
    # profiles/r02/README.md) so that the H2D copy of chunk c+1 and the D2H copy of chunk c-1 overlap the
    # attention of chunk c (two copy streams + events), and the device-side input / output sets alternate between
    # steps so that the copies of neighbouring steps overlap too; every byte still crosses PCIe inside the timed region.
    main = torch.cuda.current_stream()
    h2d_s, d2h_s = torch.cuda.Stream(), torch.cuda.Stream()
    ev_in = [[torch.cuda.Event() for _ in range(n_chunks)] for _ in (0, 1)]
    ev_out = [[torch.cuda.Event() for _ in range(n_chunks)] for _ in (0, 1)]
    ev_consumed = [torch.cuda.Event() for _ in (0, 1)]      # the step that used input set s has read all of it
    ev_drained = [torch.cuda.Event() for _ in (0, 1)]       # output set s has been copied back to the host
    out_views = [o.view(L, B, Hq_l * D) for o in out_devs]
    step_no = [0]

    def e2e_step():
        s_ = step_no[0] & 1
        step_no[0] += 1
        mgr.cache_batch_data(seq_ids, lens)
        h2d_s.wait_event(ev_consumed[s_])
        with torch.cuda.stream(h2d_s):
            for c, (lo, hi) in enumerate(bounds):
                qkv_devs[s_][lo:hi].copy_(qkv_host[lo:hi], non_blocking=True)
                ev_in[s_][c].record(h2d_s)
        main.wait_event(ev_drained[s_])                 # output set s_ (written two steps ago) is back on the host
        for c in range(n_chunks):
            main.wait_event(ev_in[s_][c])
            chunk_graphs[s_][c].replay()
            ev_out[s_][c].record(main)
        ev_consumed[s_].record(main)
        with torch.cuda.stream(d2h_s):
            for c, (lo, hi) in enumerate(bounds):
                d2h_s.wait_event(ev_out[s_][c])
                out_host[lo:hi].copy_(out_views[s_][lo:hi], non_blocking=True)
            ev_drained[s_].record(d2h_s)

    for e_ in ev_consumed + ev_drained:
        e_.record(main)
    for _ in range(max(3, args.warmup)):
        e2e_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()                     # includes the last step's device->host copies
    e2e_s = time.perf_counter() - t0
    sync_all()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = bytes_step / (float(t.item()) / args.steps) / 1e9
    # per-step host<->device bytes of the WHOLE job (all ranks)
    h2d = world * (qkv_host.numel() * 2 + block_table.numel() * 4 + seqlens.numel() * 4)
    d2h = world * out_host.numel() * 2
    splits_used = num_splits or b200ops.decode_num_splits(B, Hkv_l, per_seq * page, b200ops.sm_count(dev))

    # ---- the reference's own GPU operator on the SAME workload, same caches, same graph-replay timing:
    # flash_attn_with_kvcache called as /root/reference core/layer/model_graph.py:161-168 calls it, including its
    # num_splits = max(1, 32 // (4 * bs)).  Context for the headline, never part of `value`; library code, not ours.
    ref_op = None
    if rank == 0 and not args.no_configs:
        try:
            from flash_attn import flash_attn_with_kvcache as fa
            fa_splits = max(1, 32 // (4 * B))

            def fa_layers():
                x_all = qkv_devs[0]
                for l in range(L):
                    x = x_all[l]
                    fa(x[:, : Hq_l * D].view(B, 1, Hq_l, D), mgr.k_caches[l], mgr.v_caches[l],
                       k=x[:, Hq_l * D: (Hq_l + Hkv_l) * D].view(B, 1, Hkv_l, D),
                       v=x[:, (Hq_l + Hkv_l) * D:].view(B, 1, Hkv_l, D), rotary_cos=cos, rotary_sin=sin,
                       cache_seqlens=seqlens, block_table=block_table, causal=True, rotary_interleaved=False,
                       num_splits=fa_splits)
            fa_ms = _time_graph(_graph_of(fa_layers), reps=max(3, min(args.steps, 10)))
            ref_op = {"operator": "flash_attn 2.8.3 flash_attn_with_kvcache (the reference's decode call, model_graph.py:161-168)",
                      "num_splits": fa_splits, "ms_per_step": round(fa_ms, 4), "us_per_call": round(fa_ms * 1e3 / L, 2),
                      "kv_GBps": round(bytes_step / world / (fa_ms * 1e-3) / 1e9, 1),
                      "b200_kernel_us_per_call": round(kernel_us, 2), "ratio_us": round(fa_ms * 1e3 / L / kernel_us, 3),
                      "timing": "CUDA graph of the step's launches, CUDA events, this rank's shard"}
            del fa_layers
        except Exception as e:  # noqa: BLE001 - context figure only
            ref_op = {"error": f"{type(e).__name__}: {e}"[:200]}

    # ---- free the headline's memory before the auxiliary figures
    del graph, chunk_graphs, mgr, qkv_dev, out_dev, qkv_devs, out_devs, out_views, qkv_host, out_host, block_table, seqlens
    _release()

    peak, peak_tf, peak_src = measured_peaks()
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    # decode tokens/s of the Llama-3-70B-shaped stack at TP = world (BASELINE.json configs[3]) — on EVERY line
    step = None
    if not args.no_decode_step:
        try:
            from bench_decode_step import run as run_step
            r = run_step("llama3-70b", layers=16, batch=64, ctx=8192, steps=20, warmup=3)
            step = {"tp": world, "shape": "llama3-70b", "layers": 16, "layers_model": 80, "batch": 64, "ctx": 8192,
                    "tokens_per_s": r["tokens_per_s_run"], "ms_per_step": r["ms_per_step_run"],
                    "tokens_per_s_full_depth_extrapolated": r["tokens_per_s_full_depth_extrapolated"],
                    "allreduce": r.get("allreduce"), "gemm": r.get("gemm"),
                    "parallelism": f"TP={world}: Hq_local={64 // world}, Hkv_local={8 // world}, ffn_local={28672 // world}; "
                                   "sum over ranks after o_proj and down_proj inside the CUDA graph",
                    "workload": "llama3-70b-shape 16-layer decode step through ModelGraphRunner (CUDA graph), "
                                "random-init weights, lm_head included"}
            ref = os.path.join(ROOT, "profiles", "r02", "step_l70_tp1.json")
            if os.path.exists(ref):
                with open(ref) as f:
                    t1 = json.load(f)["tokens_per_s_run"]
                step["tp_efficiency_vs_n1"] = round(r["tokens_per_s_run"] / (world * t1), 4)
                step["n1_reference"] = {"tokens_per_s": t1, "source": "profiles/r02/step_l70_tp1.json (committed N=1 run)"}
        except Exception as e:  # noqa: BLE001 - auxiliary figure must never take the headline line down
            step = {"error": f"{type(e).__name__}: {e}"[:300]}
        _release()
    tp_check = None
    if world > 1 and not args.no_tp_check:
        try:
            from check_tp import run_check
            tp_check = run_check()
        except Exception as e:  # noqa: BLE001
            tp_check = {"error": f"{type(e).__name__}: {e}"[:300]}
        _release()
    long_ctx = None
    if world > 1 and not args.no_configs:
        try:
            long_ctx = long_context_shard(dev, world, peak)
        except Exception as e:  # noqa: BLE001
            long_ctx = {"error": f"{type(e).__name__}: {e}"[:300]}
        _release()

    if rank != 0:
        return
    cpu = None
    step_l8 = None
    configs = None
    if world == 1:
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rows, reps = 4, 3
            arm = CpuArm(rows, threads)
            arm.call()
            dt, nb = arm.time(reps), arm.nbytes
            cpu = {"value": round(nb / dt / 1e9, 4), "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"{rows} of {B} batch rows, 1 layer call, mean of {reps} calls (oracle port, torch eager)"}
        if not args.no_decode_step:
            try:
                from bench_decode_step import run as run_step
                r = run_step("llama3-8b", layers=L, batch=B, ctx=per_seq * page, steps=20, warmup=3)
                step_l8 = {"tokens_per_s": r["tokens_per_s_run"], "ms_per_step": r["ms_per_step_run"], "tp": 1,
                           "gemm": r.get("gemm"),
                           "workload": f"llama3-8b-shape {L}-layer decode step through ModelGraphRunner (CUDA graph), "
                                       f"batch {B}, ctx {per_seq * page}, random-init weights, lm_head included"}
            except Exception as e:  # noqa: BLE001
                step_l8 = {"error": f"{type(e).__name__}: {e}"[:300]}
            _release()
        if not args.no_configs:
            try:
                configs = secondary_configs(dev, peak, peak_tf)
            except Exception as e:  # noqa: BLE001
                configs = {"error": f"{type(e).__name__}: {e}"[:300]}
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "decode_traffic.json")
    if os.path.exists(tpath) and world == 1:
        with open(tpath) as f:
            tj = json.load(f)
        traffic = tj.get("dram_bytes_per_launch")
        traffic_src = tj.get("source", "profiles/decode_traffic.json (committed ncu --set full capture, not this run)")
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(), "kv_bytes_per_step": bytes_step,
                   "l2": "inputs (16 GiB KV per step over all ranks) larger than L2; no flush needed",
                   "parallelism": (f"tp{world}: KV-head shard of ONE workload, Hq_local={Hq_l}, Hkv_local={Hkv_l}, "
                                   f"cache [n_blocks,256,{Hkv_l},{D}] per rank, replicated block tables, "
                                   "no collective inside attention (core/engine.py:108-114)"),
                   "num_splits": splits_used, "variant": args.variant, "pdl": args.pdl,
                   "e2e_pipeline": f"{n_chunks} chunks of {L // n_chunks} layers, H2D/D2H on side streams, two alternating device-side step buffers"},
        "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": UNIT,
                     "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "paged_decode_kernel", "kernel_us": round(kernel_us, 3),
                     "algorithmic_bytes_per_launch": bytes_layer_local, "peak_source": peak_src,
                     "frac_of_nominal_8TBps": round(achieved / 8000.0, 4),
                     "note": "per-GPU: this rank's shard bytes per launch / its launch time"},
        "cpu_baseline": cpu,
        "e2e": {"value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": args.steps * L,
        "clocks": clk.summary(),
        "tokens_per_s_attention_only": round(B / (ms_step * 1e-3), 1),
        "reference_gpu_operator": ref_op,
        "decode_step": step,
        "tp_check": tp_check,
        "decode_step_l8": step_l8,
        "configs": configs if world == 1 else {"cfg5_llama3-8b_b8_128k_shard": long_ctx},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--splits", type=int, default=0, help="override the split-KV count (0 = heuristic)")
    ap.add_argument("--variant", type=int, default=0, help="decode kernel tuning variant")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-decode-step", action="store_true", help="skip the whole-model decode tokens/s figures")
    ap.add_argument("--no-tp-check", action="store_true", help="skip the fused-vs-NCCL TP check (N > 1)")
    ap.add_argument("--no-configs", action="store_true", help="skip the secondary BASELINE.json configs (N = 1)")
    ap.add_argument("--e2e-chunks", type=int, default=2, choices=[1, 2, 4, 8, 16, 32],
                    help="e2e leg: the 32 layer calls run as this many graphs with H2D / D2H overlapped chunk by chunk")
    ap.add_argument("--pdl", type=int, default=0, help="programmatic dependent launch: 0 off, 1 early trigger, 2 late trigger")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            gc.collect()
            torch.cuda.synchronize()
            dist.barrier()
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
