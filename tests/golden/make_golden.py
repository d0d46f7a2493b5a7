"""Generate golden vectors by running THE REFERENCE ITSELF (CPU-runnable parts).

Run in the build container only (the reference tree does not exist on the GPU box):

    python tests/golden/make_golden.py [/root/reference]

Writes tests/golden/allocator_traces.json.gz and tests/golden/reference_cpu.pt.  The process
puts ONLY the reference on sys.path, so ``core.*`` / ``kernel.*`` below are the reference's
modules, not this repo's same-named ones.

Covered (everything on the path that the reference can execute without a GPU):
  * KVCacheManager alloc / append / free / get_slots / cache_batch_data / stats traces
    (core/cache_manager.py:262-691), block_size 256 and small block sizes;
  * PrecomputedRotaryEmbedding bf16 tables -> PagedAttention pools (core/paged_attention.py:42-62,
    :187-190): sha256 of the full table + sampled rows;
  * RoPE.forward (core/layer/rope.py:10-39);
  * store_kvcache CPU path (core/cache_manager.py:171-178);
  * ModelGraphRunner._mlp packing (core/layer/model_graph.py:75-81).
"""
import gzip
import hashlib
import json
import os
import random
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != os.path.dirname(os.path.dirname(HERE))]
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

import torch  # noqa: E402

from core.cache_manager import KVCacheManager, store_kvcache  # noqa: E402  (reference)
from core.paged_attention import PrecomputedRotaryEmbedding  # noqa: E402  (reference)
from core.layer.rope import RoPE  # noqa: E402  (reference)
from core.layer.model_graph import ModelGraphRunner  # noqa: E402  (reference)

assert os.path.abspath(sys.modules["core.cache_manager"].__file__).startswith(os.path.abspath(REF))


def snapshot(m):
    return {"free": list(m._free), "blocks": {str(k): list(v) for k, v in m._blocks.items()},
            "pos": {str(k): int(v) for k, v in m._pos.items()}}


def allocator_trace(seed, n_blocks, block_size, max_tokens, max_bs, n_ops):
    rng = random.Random(seed)
    m = KVCacheManager(n_blocks=n_blocks, block_size=block_size, n_layers=1, n_heads=1, head_size=8,
                       dtype=torch.bfloat16, device="cpu", max_tokens=max_tokens, max_batch_size=max_bs)
    live, next_id, ops = [], 0, []
    for _ in range(n_ops):
        r = rng.random()
        if r < 0.25 or not live:
            choices = [1, block_size - 1, block_size, block_size + 1, 2 * block_size, rng.randint(1, max_tokens)]
            n = min(rng.choice(choices), max_tokens)
            ok, slots = m.alloc(next_id, n)
            ops.append({"op": "alloc", "seq": next_id, "n": n, "ok": bool(ok),
                        "slots": slots.tolist() if ok else None})
            if ok:
                live.append(next_id)
            next_id += 1
        elif r < 0.70:
            sid = rng.choice(live + [10 ** 6])            # sometimes an unknown id
            room = sid in m._blocks and len(m._blocks[sid]) < m.max_seq_blocks
            if sid in m._blocks and not room and m._pos[m._blocks[sid][-1]] >= block_size - 1:
                continue                                   # would outgrow the table; the engine never does this
            ops.append({"op": "append", "seq": sid, "ret": int(m.append(sid))})
        elif r < 0.80:
            sid = rng.choice(live + [10 ** 6])
            if sid in live:
                live.remove(sid)
            m.free(sid)
            ops.append({"op": "free", "seq": sid})
        elif r < 0.88:
            sid = rng.choice(live)
            pos = [rng.randint(0, max_tokens + 5) for _ in range(6)] + [0, block_size - 1, block_size]
            ops.append({"op": "get_slots", "seq": sid, "positions": pos, "ret": m.get_slots(sid, pos)})
        elif r < 0.97:
            k = min(len(live), max_bs, rng.randint(1, 4))
            ids = rng.sample(live, k)
            if rng.random() < 0.3:
                ids = ids + [ids[0]]                       # scheduler-style duplicate padding
                ids = ids[:max_bs]
            lens = [sum(m._pos[b] for b in m._blocks[s]) for s in ids]
            bt, sl = m.cache_batch_data(ids, lens)
            ops.append({"op": "cache_batch_data", "seqs": ids, "lens": lens,
                        "block_table": bt.tolist(), "seqlens": sl.tolist()})
        else:
            ops.append({"op": "stats", "ret": m.stats})
        ops[-1]["state"] = snapshot(m)
    return {"seed": seed, "n_blocks": n_blocks, "block_size": block_size, "max_tokens": max_tokens,
            "max_batch_size": max_bs, "ops": ops}


def engine_style_trace():
    """The sequence the engine actually drives (engine.py:218-222, :259-262): alloc(P), then per
    decode step append + cache_batch_data(current_position) with current_position = P + 1 + step."""
    m = KVCacheManager(n_blocks=8, block_size=256, n_layers=1, n_heads=1, head_size=8, dtype=torch.bfloat16,
                       device="cpu", max_tokens=1024, max_batch_size=4)
    P = 128
    ok, slots = m.alloc(0, P)
    steps = []
    for step in range(200):
        slot = m.append(0)
        bt, sl = m.cache_batch_data([0], [P + 1 + step])
        steps.append({"slot": int(slot), "block_table": bt.tolist(), "seqlens": sl.tolist()})
    return {"P": P, "alloc_slots": slots.tolist(), "steps": steps, "state": snapshot(m)}


def main():
    traces = [
        allocator_trace(0, 12, 256, 1024, 4, 400),
        allocator_trace(1, 40, 256, 2048, 8, 600),
        allocator_trace(2, 6, 16, 64, 4, 500),
        allocator_trace(3, 3, 256, 1024, 4, 300),        # exhaustion-heavy
    ]
    with gzip.open(os.path.join(HERE, "allocator_traces.json.gz"), "wt") as f:
        json.dump({"traces": traces, "engine": engine_style_trace()}, f)

    out = {}
    for dim in (64, 128):
        emb = PrecomputedRotaryEmbedding(dim, max_position=8192, device=torch.device("cpu"))
        cos = emb.cos_cache[0, 0, :8192, : dim // 2].contiguous()
        sin = emb.sin_cache[0, 0, :8192, : dim // 2].contiguous()
        rows = list(range(0, 320)) + [511, 512, 513, 1023, 1024, 2047, 4095, 4096, 8191]
        out[f"rot{dim}_rows"] = torch.tensor(rows)
        out[f"rot{dim}_cos"] = cos[rows].clone()
        out[f"rot{dim}_sin"] = sin[rows].clone()
        out[f"rot{dim}_cos_sha256"] = hashlib.sha256(cos.view(torch.int16).numpy().tobytes()).hexdigest()
        out[f"rot{dim}_sin_sha256"] = hashlib.sha256(sin.view(torch.int16).numpy().tobytes()).hexdigest()
        out[f"rot{dim}_distinct_rows"] = int(torch.unique(torch.cat((cos, sin), dim=1).view(torch.int16), dim=0).shape[0])

        g = torch.Generator().manual_seed(100 + dim)
        q = torch.randn(2, 7, 3, dim, generator=g).to(torch.bfloat16)
        k = torch.randn(2, 7, 3, dim, generator=g).to(torch.bfloat16)
        q_rot, k_rot = RoPE().forward(q, k, cos, sin)
        out[f"rope{dim}_q"], out[f"rope{dim}_k"] = q, k
        out[f"rope{dim}_q_rot"], out[f"rope{dim}_k_rot"] = q_rot, k_rot

    g = torch.Generator().manual_seed(7)
    key = torch.randn(9, 2, 16, generator=g).to(torch.bfloat16)
    val = torch.randn(9, 2, 16, generator=g).to(torch.bfloat16)
    kc = torch.zeros(3, 4, 2, 16, dtype=torch.bfloat16)
    vc = torch.zeros(3, 4, 2, 16, dtype=torch.bfloat16)
    slots = torch.tensor([5, -1, 0, 11, 3, -1, 7, 8, 2], dtype=torch.int32)
    store_kvcache(key, val, kc, vc, slots, 4)
    out.update(store_key=key, store_val=val, store_slots=slots, store_kc=kc, store_vc=vc)

    g = torch.Generator().manual_seed(8)
    x = torch.randn(5, 32, generator=g)
    gu = torch.randn(32, 48, generator=g) * 0.2
    dw = torch.randn(24, 32, generator=g) * 0.2
    out.update(mlp_x=x, mlp_gu=gu, mlp_d=dw, mlp_out=ModelGraphRunner._mlp(x, gu, dw))

    torch.save(out, os.path.join(HERE, "reference_cpu.pt"))
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
