"""Pipeline trace of the tcgen05 prefill kernel: clock64 stamps of every mbarrier hand-off of the
heaviest CTA (blockIdx 0,0,0) -> per-tile durations of each softmax phase and the MMA / TMA waits.
GPU box only.   python tools/trace_prefill.py [variant] [S]   -> gpurun_out/prefill_trace_v<variant>.json"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
import torch  # noqa: E402
import b200ops  # noqa: E402
from b200ops import _lib  # noqa: E402
from core.paged_attention import build_rotary_tables  # noqa: E402

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
S = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
B_arg = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dev = torch.device("cuda")
torch.manual_seed(0)
bf = torch.bfloat16
B, P, Hq, Hkv, D, page = B_arg, 0, 32, 8, 128, 256
per_seq = -(-(P + S) // page)
n_blocks = B * per_seq + 2
kc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
vc = torch.randn(n_blocks, page, Hkv, D, device=dev, dtype=bf)
table = torch.randperm(n_blocks)[: B * per_seq].view(B, per_seq).to(torch.int32).to(dev)
lens = torch.full((B,), P, dtype=torch.int32, device=dev)
q = torch.randn(B, S, Hq, D, device=dev, dtype=bf)
k = torch.randn(B, S, Hkv, D, device=dev, dtype=bf)
v = torch.randn(B, S, Hkv, D, device=dev, dtype=bf)
cos, sin = build_rotary_tables(D, P + S + 1)
cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
out = torch.empty_like(q)
for _ in range(2):
    b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out, variant=variant)
torch.cuda.synchronize()
buf = torch.zeros(16 * 64, dtype=torch.int64, device=dev)
_lib.lib.b200_debug_prefill_trace(buf.data_ptr())
b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, out=out, variant=variant)
torch.cuda.synchronize()
_lib.lib.b200_debug_prefill_trace(None)
t = buf.cpu().view(16, 64)
nt = min(64, -(-S // 128), int((t[11] != 0).sum()))    # CTA 0's first work item may not be the heaviest one
t0 = int(t[9, 0])
names = {0: "sm_S_ready", 1: "sm_S_loaded", 2: "sm_max_done", 3: "sm_exp_done", 4: "sm_PE_passed", 5: "sm_P_published",
         6: "mma_QK_issue", 7: "mma_PV_enter", 8: "mma_PV_issue", 9: "tma_K_issue", 10: "tma_V_issue", 11: "sm_tile_top"}
rel = {names[e]: [int(t[e, j]) - t0 for j in range(nt)] for e in names}
# steady-state per-tile durations (tiles 8..nt-2)
lo, hi = min(8, nt - 3), nt - 2
def avg(f):
    xs = [f(j) for j in range(lo, hi)]
    return round(sum(xs) / max(1, len(xs)), 1)
r = rel
summ = {
    "tile_period": avg(lambda j: r["sm_tile_top"][j + 1] - r["sm_tile_top"][j]),
    "wait_S": avg(lambda j: r["sm_S_ready"][j] - r["sm_tile_top"][j]),
    "ld_S": avg(lambda j: r["sm_S_loaded"][j] - r["sm_S_ready"][j]),
    "mask_max_xchg": avg(lambda j: r["sm_max_done"][j] - r["sm_S_loaded"][j]),
    "exp_sum_pack": avg(lambda j: r["sm_exp_done"][j] - r["sm_max_done"][j]),
    "wait_PE": avg(lambda j: r["sm_PE_passed"][j] - r["sm_exp_done"][j]),
    "st_P_publish": avg(lambda j: r["sm_P_published"][j] - r["sm_PE_passed"][j]),
    "P_published_to_PV_issue": avg(lambda j: r["mma_PV_issue"][j] - r["sm_P_published"][j]),
    "PV_issue_to_next_QK_issue": avg(lambda j: r["mma_QK_issue"][j + 2] - r["mma_PV_issue"][j]) if nt > lo + 3 else None,
    "QK_issue_to_S_ready": avg(lambda j: r["sm_S_ready"][j] - r["mma_QK_issue"][j]),
    "QK_issue_lead_over_softmax_start": avg(lambda j: r["sm_tile_top"][j] - r["mma_QK_issue"][j]),
    "K_issue_lead_over_QK_issue": avg(lambda j: r["mma_QK_issue"][j] - r["tma_K_issue"][j]),
}
item_ends = [int(x) - t0 for x in t[12].tolist() if int(x) != 0]
print(json.dumps({"variant": variant, "S": S, "tiles": nt, "steady_state_clk": summ,
                  "cta0_item_epilogue_done_clk": item_ends,
                  "cta0_item_epilogue_enter_clk": [int(x) - t0 for x in t[13].tolist() if int(x) != 0],
                  "cta0_item_last_PV_retired_clk": [int(x) - t0 for x in t[14].tolist() if int(x) != 0],
                  "cta0_item_first_S_ready_clk": [int(x) - t0 for x in t[15].tolist() if int(x) != 0]}))
for j in list(range(0, min(6, nt))) + list(range(max(6, nt - 3), nt)):
    print(j, {kk: vv[j] for kk, vv in rel.items()})
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", f"prefill_trace_v{variant}.json"), "w") as f:
    json.dump({"variant": variant, "S": S, "summary": summ, "events": rel}, f)
