#!/usr/bin/env python
"""bench.py — paged decode attention KV GB/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (N = 1): BASELINE.json configs[1] — Llama-3-8B shape (32 q heads, 8 kv heads, d = 128),
bf16, batch 32, 4k context (cache_seqlens = 4095, one token appended -> 4096), block = 256,
pages drawn from a seeded random permutation of the pool.  One STEP = the attention of one
decode step: one fused RoPE + KV-append + split-KV + merge launch per layer for all 32
layers, each layer with its own 512 MiB KV cache (16 GiB of KV per step >> 126 MB L2, so
nothing is served from cache between steps or layers).

  value   : algorithmic bytes / device time, inputs resident in HBM, the 32
            launches replayed from one CUDA graph (the way ModelGraphRunner drives the op).
  e2e     : same metric through the public operator API with HOST inputs: per step the q/k/v
            of every layer come from pinned host memory (H2D), block tables and sequence
            lengths are rebuilt from Python lists by KVCacheManager.cache_batch_data, the
            outputs of every layer are read back (D2H); timed by wall clock between device
            synchronisations.  The KV cache itself is device-resident state (as in the
            reference, core/cache_manager.py:311-323).
  N > 1   : attention shards by KV head with no collective (core/parallel_config TP):
            every rank runs the same per-GPU workload on its own 8-kv-head shard (weak
            scaling), barrier + device sync on both sides, MAX time over ranks.
  --impl reference : the reference semantics on the host CPU cores (oracle port, torch eager,
            all threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
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


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes_per_layer():
    from oracle.paged_attention import decode_algorithmic_bytes
    return decode_algorithmic_bytes([CFG["ctx"]] * CFG["B"], CFG["Hq"], CFG["Hkv"], CFG["D"], CFG["page"])


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
    One layer call over `rows` of the 32 batch rows."""

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
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(), "note": "reference semantics in torch eager on host CPU cores (oracle port)"},
        "cpu_baseline": {"value": round(val, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(val, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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
    per_seq = ctx // page
    n_blocks = B * per_seq + 8
    mgr = KVCacheManager(n_blocks, page, L, Hkv, D, dtype=torch.bfloat16, device=dev, max_tokens=ctx,
                         max_batch_size=B)
    # non-contiguous pages: hand the allocator a seeded permutation of the pool
    g = torch.Generator().manual_seed(1234 + rank)
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

    qkv_host = torch.randn(L, B, (Hq + 2 * Hkv) * D, dtype=torch.float32).to(torch.bfloat16).pin_memory()
    out_host = torch.empty(L, B, Hq * D, dtype=torch.bfloat16).pin_memory()
    qkv_dev = torch.empty_like(qkv_host, device=dev)
    out_dev = torch.empty(L, B, Hq, D, dtype=torch.bfloat16, device=dev)
    qkv_dev.copy_(qkv_host)
    block_table, seqlens = mgr.cache_batch_data(seq_ids, lens)
    num_splits = args.splits if args.splits > 0 else 0

    def layer_call(l):
        x = qkv_dev[l]
        q = x[:, : Hq * D].view(B, Hq, D)
        k = x[:, Hq * D: (Hq + Hkv) * D].view(B, Hkv, D)
        v = x[:, (Hq + Hkv) * D:].view(B, Hkv, D)
        # pdl: inside a decode step the kernel launched before a layer's attention never writes
        # the lengths, the block table or that layer's cached pages (B200_DECODE_FLAG_PDL)
        b200ops.paged_attn_decode(q, mgr.k_caches[l], mgr.v_caches[l], k, v, cos, sin, seqlens, block_table,
                                  rotary_interleaved=False, out=out_dev[l], num_splits=num_splits,
                                  variant=args.variant, pdl=args.pdl)

    def layers(lo, hi):
        for l in range(lo, hi):
            layer_call(l)

    def capture(lo, hi):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            layers(lo, hi)                       # warm-up on a side stream
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            layers(lo, hi)
        return g

    graph = capture(0, L)                        # the 32 launches of one decode step in one graph
    n_chunks = 8                                 # e2e: host<->device copies pipelined against compute
    bounds = [(c * L // n_chunks, (c + 1) * L // n_chunks) for c in range(n_chunks)]
    chunk_graphs = [capture(lo, hi) for lo, hi in bounds]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    bytes_layer = algorithmic_bytes_per_layer()
    bytes_step = bytes_layer * L

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
    value = world * bytes_step / (ms_step * 1e-3) / 1e9
    kernel_us = ms_step * 1e3 / L
    achieved = bytes_layer / (kernel_us * 1e-6) / 1e9

    # ---- e2e: host buffers in, host buffers out, metadata rebuilt from Python lists every step.
    # The layers are driven in 8 chunks so that the H2D copy of chunk c+1 and the D2H copy of
    # chunk c-1 overlap the attention of chunk c (two copy streams + events); every byte still
    # crosses PCIe inside the timed region.
    main = torch.cuda.current_stream()
    h2d_s, d2h_s = torch.cuda.Stream(), torch.cuda.Stream()
    ev_in = [torch.cuda.Event() for _ in range(n_chunks)]
    ev_out = [torch.cuda.Event() for _ in range(n_chunks)]
    ev_step = torch.cuda.Event()
    out_view = out_dev.view(L, B, Hq * D)

    def e2e_step():
        mgr.cache_batch_data(seq_ids, lens)
        h2d_s.wait_event(ev_step)                       # previous step's compute has consumed qkv_dev
        with torch.cuda.stream(h2d_s):
            for c, (lo, hi) in enumerate(bounds):
                qkv_dev[lo:hi].copy_(qkv_host[lo:hi], non_blocking=True)
                ev_in[c].record(h2d_s)
        for c in range(n_chunks):
            main.wait_event(ev_in[c])
            chunk_graphs[c].replay()
            ev_out[c].record(main)
        ev_step.record(main)
        with torch.cuda.stream(d2h_s):
            for c, (lo, hi) in enumerate(bounds):
                d2h_s.wait_event(ev_out[c])
                out_host[lo:hi].copy_(out_view[lo:hi], non_blocking=True)
        main.wait_stream(d2h_s)

    ev_step.record(main)
    for _ in range(max(3, args.warmup)):
        e2e_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sync_all()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = world * bytes_step / (float(t.item()) / args.steps) / 1e9
    h2d = qkv_host.numel() * 2 + block_table.numel() * 4 + seqlens.numel() * 4
    d2h = out_host.numel() * 2

    if rank != 0:
        return
    peak, peak_src = measured_peak()
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        rows, reps = 4, 3
        arm = CpuArm(rows, threads)
        arm.call()
        dt, nb = arm.time(reps), arm.nbytes
        cpu = {"value": round(nb / dt / 1e9, 4), "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{rows} of {B} batch rows, 1 layer call, mean of {reps} calls (oracle port, torch eager)"}
    # auxiliary (N = 1 only, outside the timed region of the headline metric): decode tokens/s of the whole
    # Llama-3-8B-shaped layer stack through the reference-shaped runner, i.e. the second half of BASELINE.json's metric
    step = None
    if world == 1 and not args.no_decode_step:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            from bench_decode_step import run as run_step
            r = run_step("llama3-8b", layers=L, batch=B, ctx=per_seq * page, steps=20, warmup=3)
            step = {"tokens_per_s": r["tokens_per_s_run"], "ms_per_step": r["ms_per_step_run"], "tp": 1,
                    "workload": f"llama3-8b-shape {L}-layer decode step through ModelGraphRunner (CUDA graph), batch {B}, "
                                f"ctx {per_seq * page}, random-init weights, lm_head included",
                    "skinny_gemm": os.environ.get("B200_SKINNY_GEMM", "auto")}
        except Exception as e:  # noqa: BLE001 - auxiliary figure must never take the headline line down
            step = {"error": f"{type(e).__name__}: {e}"[:300]}
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "decode_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(), "kv_bytes_per_step": bytes_step,
                   "l2": "inputs (16 GiB KV per step) larger than L2; no flush needed",
                   "parallelism": f"kv-head shard per rank x{world}, no collective in attention",
                   "num_splits": num_splits or b200ops.decode_num_splits(B, Hkv, per_seq * page, b200ops.sm_count(dev)),
                   "variant": args.variant, "pdl": args.pdl,
                   "e2e_pipeline": f"{n_chunks} chunks of {L // n_chunks} layers, H2D/D2H on side streams"},
        "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": UNIT,
                     "frac": round(achieved / peak, 4), "traffic": traffic, "kernel": "paged_decode_kernel",
                     "kernel_us": round(kernel_us, 3), "algorithmic_bytes_per_launch": bytes_layer,
                     "peak_source": peak_src, "frac_of_nominal_8TBps": round(achieved / 8000.0, 4)},
        "cpu_baseline": cpu,
        "e2e": {"value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": args.steps * L,
        "clocks": clk.summary(),
        "tokens_per_s_attention_only": round(world * B / (ms_step * 1e-3), 1),
        "decode_step": step,
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
    ap.add_argument("--no-decode-step", action="store_true", help="skip the auxiliary whole-model decode tokens/s figure")
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
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
