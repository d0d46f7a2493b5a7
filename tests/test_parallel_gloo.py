"""Tensor-parallel host logic on CPU with gloo, world_size = 2: core.parallel_config helpers,
ModelGraphRunner.prepare_weights slicing (model_graph.py:87-118), and KV-head sharded attention
+ row-parallel o_proj / down_proj with a SUM all-reduce == the unsharded computation."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, result_q):
    try:
        os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                          LOCAL_RANK=str(rank))
        for p in (os.path.join(ROOT, "micro-vllm_b200"), ROOT, os.path.join(ROOT, "tests")):
            if p not in sys.path:
                sys.path.insert(0, p)
        torch.set_num_threads(2)
        from core import parallel_config as pc
        from core.layer.model_graph import ModelGraphRunner
        from core.layer.synthetic_model import Shape, build_synthetic_model
        from oracle.elementwise import mlp_oracle
        from oracle.paged_attention import paged_attention_oracle
        from _cases import make_paged_case

        pc.setup()
        assert pc.get_world_size() == world and pc.get_rank() == rank and pc.rank0() == (rank == 0)
        assert pc.get_group() is not None
        t = torch.full((4,), float(rank + 1))
        assert pc.all_reduce(t) is t and t.tolist() == [3.0] * 4          # in place, returns its argument
        pc.barrier()

        shape = Shape("tiny", 128, 1, 8, 4, 16, 96, 64, True)
        model = build_synthetic_model(shape, device="cpu", dtype=torch.float32, seed=5)
        Hq, Hkv, D = shape.heads // world, shape.kv_heads // world, shape.head_dim
        runner = ModelGraphRunner(model, 1, Hq, D, Hkv, shape.hidden, 2 * shape.ffn // world, "cpu",
                                  max_batch_size=4, dtype=torch.float32)
        blk = model.transformer.h[0]
        assert blk.attn._qkv_w.shape == (128, (Hq + 2 * Hkv) * D) and blk.attn._o.shape == (Hq * D, 128)
        assert blk.mlp._gu.shape == (128, 2 * shape.ffn // world) and blk.mlp._d.shape == (shape.ffn // world, 128)
        # slices really are this rank's rows / columns of the full weights
        wq = blk.attn.c_attn.weight[: shape.heads * D]
        assert torch.equal(blk.attn._qkv_w[:, : Hq * D], wq[rank * Hq * D:(rank + 1) * Hq * D].t())
        assert torch.equal(blk.mlp._d, blk.mlp.c_proj.weight[:, rank * shape.ffn // world:(rank + 1) * shape.ffn // world].t())

        # one layer, decode step, sharded vs unsharded (fp32, oracle ops): attention needs no collective
        B = 3
        case = make_paged_case(B, shape.heads, shape.kv_heads, D, [5, 300, 255], seed=9)
        g = torch.Generator().manual_seed(11)
        x = torch.randn(B, shape.hidden, generator=g)
        full_qkv = x @ blk.attn.c_attn.weight.t() + blk.attn.c_attn.bias
        fq, fk, fv = full_qkv.split([shape.heads * D, shape.kv_heads * D, shape.kv_heads * D], dim=-1)

        def attn(q, k, v, kc, vc):
            return paged_attention_oracle(q.reshape(B, 1, -1, D).bfloat16(), kc, vc, k.reshape(B, 1, -1, D).bfloat16(),
                                          v.reshape(B, 1, -1, D).bfloat16(), case["cos"], case["sin"], case["seqlens"],
                                          case["table"], None, True, False)[:, 0]
        full = attn(fq, fk, fv, case["k_cache"].clone(), case["v_cache"].clone()).reshape(B, -1) @ blk.attn.c_proj.weight.t()
        qkv = x @ blk.attn._qkv_w + blk.attn._qkv_b
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        kc = case["k_cache"][:, :, rank * Hkv:(rank + 1) * Hkv].clone()     # this rank's KV-head shard of the cache
        vc = case["v_cache"][:, :, rank * Hkv:(rank + 1) * Hkv].clone()
        part = attn(q, k, v, kc, vc).reshape(B, -1) @ blk.attn._o
        pc.all_reduce(part)
        assert torch.allclose(part, full, atol=1e-4, rtol=1e-4), float((part - full).abs().max())
        # MLP: column-parallel gate|up, row-parallel down, SUM all-reduce
        full_gu = torch.cat([blk.mlp.w1.weight, blk.mlp.w2.weight], dim=0).t()
        full_mlp = mlp_oracle(x, full_gu, blk.mlp.c_proj.weight.t())
        part = mlp_oracle(x, blk.mlp._gu, blk.mlp._d)
        pc.all_reduce(part)
        assert torch.allclose(part, full_mlp, atol=1e-4, rtol=1e-4)
        pc.barrier()
        result_q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        result_q.put((rank, "FAIL: " + traceback.format_exc()))


def test_tp2_gloo_sharding_equals_unsharded():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_single_process_defaults():
    sys.path.insert(0, os.path.join(ROOT, "micro-vllm_b200"))
    from core import parallel_config as pc
    env = {k: os.environ.pop(k, None) for k in ("RANK", "WORLD_SIZE")}
    try:
        pc.setup()
        assert pc.get_world_size() == 1 and pc.get_rank() == 0 and pc.rank0() and pc.get_group() is None
        t = torch.ones(2)
        assert pc.all_reduce(t) is t
        pc.barrier()
    finally:
        for k, v in env.items():
            if v is not None:
                os.environ[k] = v


def test_reduction_form_selection_respects_a_plain_final_norm():
    """A model whose ln_f is a plain module sends weight=None with the last layer's reduction; only the one-shot form
    accepts that (b200_allreduce_residual_rmsnorm refuses every NVLS / LL / sharded call without a norm), so such a
    model must never be given another form, whatever the node offers or B200_AR_MODE asks for."""
    from core.layer.model_graph import select_ar_mode
    for want in ("auto", "nvls", "nvls1", "nvlsp", "ll", "sharded", "oneshot"):
        for world in (2, 4, 8):
            for mc in (False, True):
                assert select_ar_mode(want, world, mc, False, 64) == "oneshot"
    # kernel ln_f: the measured defaults
    assert select_ar_mode("auto", 8, True, True, 64) == "nvls"
    assert select_ar_mode("auto", 4, False, True, 64) == "ll"
    assert select_ar_mode("auto", 2, True, True, 64) == "oneshot"
    assert select_ar_mode("auto", 8, True, True, 2048) == "nvls1"        # owner-rank forms stop at 1024 rows
    assert select_ar_mode("nvls1", 8, True, True, 2048) == "nvls1"
    assert select_ar_mode("nvls", 8, True, True, 2048) == "oneshot"
    assert select_ar_mode("nvls", 8, False, True, 64) == "ll"            # no multicast range: packet form
    assert select_ar_mode("nvls1", 2, False, True, 64) == "oneshot"
    assert select_ar_mode("bogus", 8, True, True, 64) == "oneshot"
