"""Latency of the decode-path reduction alone on WORLD_SIZE GPUs (no GEMMs, no skew source):
[rows, n] bf16 partial sums -> all-reduce + residual + RMSNorm, back to back in a CUDA graph.

  one-shot : b200_allreduce_residual_rmsnorm (every rank loads all ranks' slots)
  sharded  : row-sharded form (owner reduces its rows, pushes y to all ranks)
  ll       : LL push form (flag-in-data packets, remote stores only)
  nvls     : in-switch reduction (multimem.ld_reduce / multimem.st on the multicast mapping; needs the symm back end)
  nvlsp    : nvls with a packet tail (y as flag-in-data packets; no release fence, no done counter)
  nvls1    : one-hop in-switch form (every rank pulls the sum of every row; no second hop, no trailing barrier)
  nccl     : dist.all_reduce + the local fused residual+norm kernel (B200_FUSED_ALLREDUCE=0 path)

Also prints the in-kernel %globaltimer phases of the fused forms (mean over CTAs of rank 0, last
call): start -> barrier passed -> rows reduced -> end.  Time = CUDA events around graph replays,
max over ranks.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402


def main():
    import torch.distributed as dist
    from b200ops.comm import allreduce_residual_rmsnorm_local
    from core import parallel_config as pc
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    pc.setup()
    n = 8192
    max_rows = 64
    comm = pc.peer_comm(max_rows * n * 2)
    calls = 40                                              # per graph (even: slots alternate 0, 1, 0, 1 ...)
    out = []
    w = torch.ones(n, dtype=torch.bfloat16, device=dev)
    for rows in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else '64,32,8,1').split(',')]:
        res_a = torch.zeros(rows, n, dtype=torch.bfloat16, device=dev)
        res_b = torch.zeros_like(res_a)
        y_loc = torch.empty_like(res_a)
        yb = comm.ybuf(max_rows, n)[:rows]
        trace = torch.zeros(4 * rows, dtype=torch.int64, device=dev)
        for s in (0, 1):
            comm.slot(s, max_rows, n)[:rows].normal_()
        xs = [comm.slot(0, max_rows, n)[:rows], comm.slot(1, max_rows, n)[:rows]]
        nccl_x = [torch.randn(rows, n, device=dev, dtype=torch.bfloat16) for _ in range(2)]

        def fused(mode, tr=None):
            for c in range(calls):
                comm.allreduce_residual_rmsnorm(c & 1, rows, n, res_a if c & 1 == 0 else res_b, w,
                                                yb if mode in ("sharded", "nvls") else y_loc,
                                                res_b if c & 1 == 0 else res_a, 1e-5,
                                                bool(c & 1), sharded=mode == "sharded", ll=mode == "ll",
                                                nvls=mode in ("nvls", "nvlsp"), nvls1=mode == "nvls1",
                                                packet_tail=mode == "nvlsp",
                                                trace=tr if c == calls - 1 else None)

        def nccl():
            for c in range(calls):
                x = nccl_x[c & 1]
                dist.all_reduce(x)
                allreduce_residual_rmsnorm_local(x, res_a if c & 1 == 0 else res_b, w, y_loc,
                                                 res_b if c & 1 == 0 else res_a, 1e-5, bool(c & 1))

        def time_graph(fn, reps=10):
            comm._last_mode = None        # phases are separated by barriers: nobody still reads either slot
            fn()
            torch.cuda.synchronize()
            dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            g.replay()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (reps * calls)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.barrier()
            del g
            return float(t.item())

        row = {"world": world, "rows": rows, "n": n, "bytes_per_rank": rows * n * 2, "backend": comm.backend,
               "multicast": comm.has_multicast}
        modes = ("oneshot", "sharded", "ll") + (("nvls", "nvlsp", "nvls1") if comm.has_multicast else ())
        # correctness of every form against an fp32 NCCL sum of the same slot (one call, slot 0, no residual stream reuse)
        R = -(-rows // world)
        own = slice(rank * R, min(rows, (rank + 1) * R))
        ref = xs[0].float().clone()
        dist.all_reduce(ref)
        resid = torch.randn(rows, n, device=dev, dtype=torch.bfloat16)
        dist.broadcast(resid, 0)                             # the residual stream is replicated (owners read their rows)
        want_res = (ref.bfloat16().float() + resid.float()).bfloat16()
        want_y = (want_res.float() * torch.rsqrt(want_res.float().pow(2).mean(-1, keepdim=True) + 1e-5)).bfloat16()
        for name in modes:
            ro = torch.zeros(rows, n, dtype=torch.bfloat16, device=dev)
            yo = yb if name in ("sharded", "nvls") else y_loc
            yo.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            comm._last_mode = None        # every rank is past the barrier above: nobody still reads either slot
            comm.allreduce_residual_rmsnorm(1, rows, n, resid, w, yo, torch.empty_like(ro), 1e-5, True,
                                            sharded=name == "sharded", ll=name == "ll", nvls=name in ("nvls", "nvlsp"),
                                            nvls1=name == "nvls1", packet_tail=name == "nvlsp")  # slot alternation
            comm.allreduce_residual_rmsnorm(0, rows, n, resid, w, yo, ro, 1e-5, True, sharded=name == "sharded",
                                            ll=name == "ll", nvls=name in ("nvls", "nvlsp"), nvls1=name == "nvls1",
                                            packet_tail=name == "nvlsp")
            torch.cuda.synchronize()
            dist.barrier()
            sl = slice(0, rows) if name in ("oneshot", "nvls1") else own
            # bound in units of the operands: the switch's sum may sit one bf16 ulp OF THE SUM from the rank-order one
            tol = (want_res[sl].float().abs() + ref[sl].abs() + resid[sl].float().abs()) * 2 ** -7 + 1e-3
            bad_res = int((ro[sl].float() - want_res[sl].float()).abs().gt(tol).sum())
            rr = torch.rsqrt(want_res.float().pow(2).mean(-1, keepdim=True) + 1e-5)
            tol_y = (want_y.float().abs() * 2 ** -6 + 2e-3) + (ref.abs() + resid.float().abs()) * rr * 2 ** -7
            bad_y = int((yo.float() - want_y.float()).abs().gt(tol_y).sum())
            ys = [torch.empty_like(yo) for _ in range(world)]
            dist.all_gather(ys, yo.contiguous())
            same = all(torch.equal(ys[0], t_) for t_ in ys)
            row[name + "_check"] = {"res_out_bad": bad_res, "y_bad": bad_y, "ranks_bit_identical_y": bool(same)}
        if comm.has_multicast:
            # which rounding does the switch apply to the fp32 sum?  zero residual: res_out IS the switch's bf16 sum
            zero = torch.zeros(rows, n, dtype=torch.bfloat16, device=dev)
            ro = torch.zeros_like(zero)
            torch.cuda.synchronize()
            dist.barrier()
            comm._last_mode = None
            comm.allreduce_residual_rmsnorm(0, rows, n, zero, w, yb, ro, 1e-5, True, nvls=True)
            torch.cuda.synchronize()
            dist.barrier()
            rne = ref.bfloat16()
            rtz = (ref.view(torch.int32) & -65536).view(torch.float32).bfloat16()
            g_ = ro[own]
            row["nvls_sum_rounding"] = {"equals_rne_of_fp32_sum": round(float((g_ == rne[own]).float().mean()), 5),
                                        "equals_rtz_of_fp32_sum": round(float((g_ == rtz[own]).float().mean()), 5),
                                        "max_ulp_vs_rne": int((g_.view(torch.int16).int() - rne[own].view(torch.int16).int()).abs().max()) if g_.numel() else 0}
        for name in modes:
            row[name + "_us"] = round(time_graph(lambda: fused(name)), 2)
            trace.zero_()
            comm._last_mode = None
            fused(name, trace)
            torch.cuda.synchronize()
            dist.barrier()
            t = trace.view(rows, 4).double()
            own = t[:, 2] > 0                                # CTAs that stamped the "reduced" phase
            if name == "nvls1":
                own = t[:, 1] > 0
            if own.any():
                tt = t[own]
                row[name + "_phases_us"] = {"barrier": round(float((tt[:, 1] - tt[:, 0]).mean()) / 1e3, 2),
                                            "reduce": round(float((tt[:, 2] - tt[:, 1]).mean()) / 1e3, 2),
                                            "tail": round(float((tt[:, 3] - tt[:, 2]).mean()) / 1e3, 2),
                                            "kernel_span": round(float(tt[:, 3].max() - tt[:, 0].min()) / 1e3, 2)}
        row["nccl_us"] = round(time_graph(nccl), 2)
        out.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    assert comm.error_flag() == 0
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"allreduce_tp{world}.json"), "w") as f:
            json.dump(out, f, indent=1)
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
