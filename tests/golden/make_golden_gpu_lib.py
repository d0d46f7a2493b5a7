"""Helpers shared by make_golden_gpu.py (generator, B200 box) and the test that reads its output."""
import hashlib

import torch


def inputs_digest(case):
    h = hashlib.sha256()
    for k in ("q", "k", "v", "k_cache", "v_cache", "cos", "sin", "seqlens", "table"):
        t = case[k].contiguous()
        h.update(t.view(torch.uint8).numpy().tobytes() if t.dtype != torch.bfloat16 else t.view(torch.int16).numpy().tobytes())
    return h.hexdigest()


def appended_rows(cache, case):
    """[B, S, Hkv, D] rows the operator wrote: positions seqlens[b] + i through the block table."""
    B, S = case["q"].shape[:2]
    page = case["page"]
    rows = []
    for b in range(B):
        P = int(case["seqlens"][b])
        rows.append(torch.stack([cache[int(case["table"][b, (P + i) // page]), (P + i) % page] for i in range(S)]))
    return torch.stack(rows)
