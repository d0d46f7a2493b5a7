"""Greedy arg-max kernel vs torch.argmax on decode-sized logits (L2 flushed between launches).  GPU box only."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402

PEAK = 6524.6
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:  # noqa: BLE001
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=30):
    a = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    b = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    for i in range(reps + 3):
        flush.zero_()
        if i >= 3:
            a[i - 3].record()
        fn()
        if i >= 3:
            b[i - 3].record()
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) * 1e3 for x, y in zip(a, b))
    return ts[len(ts) // 2]


rows_out = []
for rows, vocab in [(1, 152064), (32, 152064), (32, 128256), (64, 128256), (512, 152064)]:
    x = torch.randn(rows, vocab, device="cuda").to(torch.bfloat16)
    out = torch.empty(rows, dtype=torch.int64, device="cuda")
    t_our = timed(lambda: b200ops.greedy_argmax(x, out=out))
    t_torch = timed(lambda: torch.argmax(x, dim=-1))
    gbs = rows * vocab * 2 / t_our / 1e3
    rows_out.append(dict(rows=rows, vocab=vocab, b200_us=round(t_our, 2), torch_argmax_us=round(t_torch, 2),
                         GBps=round(gbs, 1), frac_of_measured_hbm=round(gbs / PEAK, 3)))
    print(json.dumps(rows_out[-1]))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rows_out, open(os.path.join(ROOT, "gpurun_out", "sampling_bench.json"), "w"), indent=1)
