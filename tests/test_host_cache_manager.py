"""Product KVCacheManager (host logic, device='cpu') against the traces recorded from the
reference's own class: block tables, slot maps, seqlens and allocator state must be bit-exact."""
import gzip
import json
import os

import pytest
import torch

from core.cache_manager import KVCacheManager, store_kvcache
from core.paged_attention import build_rotary_tables, PrecomputedRotaryEmbedding
from core.layer.rope import RoPE
from test_oracle_golden import replay


@pytest.fixture(scope="module")
def traces(golden_dir):
    with gzip.open(os.path.join(golden_dir, "allocator_traces.json.gz"), "rt") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def ref(golden_dir):
    return torch.load(os.path.join(golden_dir, "reference_cpu.pt"), weights_only=False)


def _mk(tr):
    return KVCacheManager(n_blocks=tr["n_blocks"], block_size=tr["block_size"], n_layers=1, n_heads=1, head_size=8,
                          dtype=torch.bfloat16, device="cpu", max_tokens=tr["max_tokens"],
                          max_batch_size=tr["max_batch_size"])


def _state(m):
    return {"free": list(m._free), "blocks": {str(k): list(v) for k, v in m._blocks.items()},
            "pos": {str(k): int(v) for k, v in m._pos.items()}}


def _alloc(m, s, n):
    ok, slots = m.alloc(s, n)
    if ok:
        assert slots.dtype == torch.int32
    return ok, (slots.tolist() if ok else None)


def _batch(m, s, l):
    bt, sl = m.cache_batch_data(s, l)
    assert bt.dtype == torch.int32 and sl.dtype == torch.int32
    assert bt.data_ptr() == m._block_table_buffer.data_ptr()          # static buffers: same memory every step
    assert sl.data_ptr() == m._cache_seqlens_buffer.data_ptr()
    return bt.tolist(), sl.tolist()


def test_traces_bit_exact(traces):
    for tr in traces["traces"]:
        m = _mk(tr)
        replay(m, tr["ops"], alloc=_alloc, append=lambda m, s: m.append(s), free=lambda m, s: m.free(s),
               get_slots=lambda m, s, p: m.get_slots(s, p), batch=_batch, stats=lambda m: m.stats, state=_state)


def test_engine_trace_bit_exact(traces):
    eng = traces["engine"]
    m = KVCacheManager(8, 256, 1, 1, 8, dtype=torch.bfloat16, device="cpu", max_tokens=1024, max_batch_size=4)
    ok, slots = m.alloc(0, eng["P"])
    assert ok and slots.tolist() == eng["alloc_slots"]
    for step, rec in enumerate(eng["steps"]):
        assert m.append(0) == rec["slot"]
        bt, sl = m.cache_batch_data([0], [eng["P"] + 1 + step])
        assert bt.tolist() == rec["block_table"] and sl.tolist() == rec["seqlens"]
    assert _state(m) == eng["state"]


def test_reference_conventions():
    m = KVCacheManager(2, 256, 2, 2, 16, dtype=torch.bfloat16, device="cpu", max_tokens=512, max_batch_size=2)
    assert m.alloc(0, 3 * 256) == (False, None)                       # exhaustion -> (False, None)
    assert m.append(99) == -1                                         # unknown sequence
    m.free(99)                                                        # no-op
    ok, _ = m.alloc(0, 256)
    assert ok and m._pos[m._blocks[0][0]] == 0                        # n % block_size == 0 quirk
    assert m.get_blocks(0) is m._blocks[0]                            # live reference
    assert m.get_blocks(5) == []
    k, v = m.get(1)
    assert k.shape == (2, 256, 2, 16) and k.data_ptr() == m.k_caches[1].data_ptr() and v is m.v_caches[1]
    assert m.get(0, 1)[0].shape == (256, 2, 16)
    assert m.max_seq_blocks == 2 and m._block_table_buffer.shape == (2, 2)
    assert int(m._block_table_buffer.min()) == -1 and float(m.k_caches[0].abs().sum()) == 0.0
    m.reset()
    assert list(m._free) == [0, 1] and not m._blocks and not m._pos
    with pytest.raises(AssertionError):
        store_kvcache(torch.zeros(1, 2, 16), torch.zeros(1, 2, 16), k, v, torch.zeros(1, dtype=torch.int32), 256)


@pytest.mark.parametrize("dim", [64, 128])
def test_rotary_pools_bit_exact(ref, dim):
    cos, sin = build_rotary_tables(dim, 8192)
    rows = ref[f"rot{dim}_rows"]
    assert torch.equal(cos[rows, : dim // 2], ref[f"rot{dim}_cos"])
    assert torch.equal(sin[rows, : dim // 2], ref[f"rot{dim}_sin"])
    emb = PrecomputedRotaryEmbedding(dim, max_position=8192)
    assert emb.cos_cache.shape == (1, 1, 8192, dim) and torch.equal(emb.cos_cache[0, 0], cos)


@pytest.mark.parametrize("dim", [64, 128])
def test_rope_module_bit_exact(ref, dim):
    cos, sin = build_rotary_tables(dim, 8192)
    q_rot, k_rot = RoPE().forward(ref[f"rope{dim}_q"], ref[f"rope{dim}_k"], cos[:, : dim // 2], sin[:, : dim // 2])
    assert torch.equal(q_rot, ref[f"rope{dim}_q_rot"]) and torch.equal(k_rot, ref[f"rope{dim}_k_rot"])
