"""The oracle restatements are pinned against vectors recorded from the reference itself
(tests/golden/make_golden.py).  CPU only."""
import gzip
import hashlib
import json
import os

import pytest
import torch

from oracle.cache_manager import AllocatorOracle
from oracle.elementwise import mlp_oracle
from oracle.paged_attention import store_kvcache_oracle
from oracle.rotary import rope_eager_oracle, rotary_table_oracle


@pytest.fixture(scope="module")
def traces(golden_dir):
    with gzip.open(os.path.join(golden_dir, "allocator_traces.json.gz"), "rt") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def ref(golden_dir):
    return torch.load(os.path.join(golden_dir, "reference_cpu.pt"), weights_only=False)


def _state(o):
    s = o.snapshot()
    return {"free": s["free"], "blocks": {str(k): v for k, v in s["blocks"].items()},
            "pos": {str(k): v for k, v in s["pos"].items()}}


def replay(m, ops, *, alloc, append, free, get_slots, batch, stats, state):
    for i, op in enumerate(ops):
        kind = op["op"]
        if kind == "alloc":
            ok, slots = alloc(m, op["seq"], op["n"])
            assert ok == op["ok"], f"op {i}"
            assert slots == op["slots"], f"op {i}: slot mapping differs"
        elif kind == "append":
            assert append(m, op["seq"]) == op["ret"], f"op {i}"
        elif kind == "free":
            free(m, op["seq"])
        elif kind == "get_slots":
            assert get_slots(m, op["seq"], op["positions"]) == op["ret"], f"op {i}"
        elif kind == "cache_batch_data":
            bt, sl = batch(m, op["seqs"], op["lens"])
            assert bt == op["block_table"], f"op {i}: block table differs"
            assert sl == op["seqlens"], f"op {i}: seqlens differ"
        elif kind == "stats":
            assert stats(m) == op["ret"], f"op {i}"
        assert state(m) == op["state"], f"op {i} ({kind}): allocator state differs"


def test_allocator_oracle_matches_reference_traces(traces):
    for tr in traces["traces"]:
        o = AllocatorOracle(tr["n_blocks"], tr["block_size"], tr["max_tokens"], tr["max_batch_size"])
        replay(o, tr["ops"],
               alloc=lambda m, s, n: m.alloc(s, n), append=lambda m, s: m.append(s), free=lambda m, s: m.free(s),
               get_slots=lambda m, s, p: m.get_slots(s, p), batch=lambda m, s, l: m.cache_batch_data(s, l),
               stats=lambda m: m.stats(), state=_state)


def test_allocator_oracle_engine_trace(traces):
    """alloc(128) then 200 x (append, cache_batch_data): slot 255 is skipped, the table grows one
    step early."""
    eng = traces["engine"]
    o = AllocatorOracle(8, 256, 1024, 4)
    ok, slots = o.alloc(0, eng["P"])
    assert ok and slots == eng["alloc_slots"]
    seen = []
    for step, rec in enumerate(eng["steps"]):
        slot = o.append(0)
        bt, sl = o.cache_batch_data([0], [eng["P"] + 1 + step])
        assert slot == rec["slot"] and bt == rec["block_table"] and sl == rec["seqlens"]
        seen.append(slot)
    assert 255 not in seen and 254 in seen and 256 in seen
    assert _state(o) == eng["state"]


@pytest.mark.parametrize("dim", [64, 128])
def test_rotary_table_bit_exact(ref, dim):
    cos, sin = rotary_table_oracle(dim, 8192)
    assert hashlib.sha256(cos.view(torch.int16).numpy().tobytes()).hexdigest() == ref[f"rot{dim}_cos_sha256"]
    assert hashlib.sha256(sin.view(torch.int16).numpy().tobytes()).hexdigest() == ref[f"rot{dim}_sin_sha256"]
    rows = ref[f"rot{dim}_rows"]
    assert torch.equal(cos[rows], ref[f"rot{dim}_cos"]) and torch.equal(sin[rows], ref[f"rot{dim}_sin"])
    distinct = torch.unique(torch.cat((cos, sin), dim=1).view(torch.int16), dim=0).shape[0]
    assert distinct == ref[f"rot{dim}_distinct_rows"]
    if dim == 128:
        assert distinct == 897            # SURVEY.md up-front finding 5: bf16 positions collapse


@pytest.mark.parametrize("dim", [64, 128])
def test_rope_eager_bit_exact(ref, dim):
    cos, sin = rotary_table_oracle(dim, 8192)
    q_rot, k_rot = rope_eager_oracle(ref[f"rope{dim}_q"], ref[f"rope{dim}_k"], cos, sin)
    assert torch.equal(q_rot, ref[f"rope{dim}_q_rot"]) and torch.equal(k_rot, ref[f"rope{dim}_k_rot"])


def test_store_kvcache_bit_exact(ref):
    kc = torch.zeros_like(ref["store_kc"])
    vc = torch.zeros_like(ref["store_vc"])
    store_kvcache_oracle(ref["store_key"], ref["store_val"], kc, vc, ref["store_slots"], 4)
    assert torch.equal(kc, ref["store_kc"]) and torch.equal(vc, ref["store_vc"])


def test_mlp_packing(ref):
    out = mlp_oracle(ref["mlp_x"], ref["mlp_gu"], ref["mlp_d"])
    assert torch.equal(out, ref["mlp_out"])
