"""The C-ABI shared library loads on a CPU-only machine, exports every symbol that
include/b200_abi.h declares, and rejects bad arguments before touching the device."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_abi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from b200ops import _lib
    names = declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in b200_abi.h but not exported by libb200ops.so"
    assert sorted(_lib.SIGNATURES) == names, "ctypes binding and header disagree"
    assert _lib.lib.b200_version() == _lib.ABI_VERSION


def test_struct_layout_matches_header():
    """ctypes struct sizes must match the C structs (checked by compiling a tiny C probe)."""
    import subprocess, tempfile
    from b200ops import _lib
    src = '#include <stdio.h>\n#include "b200_abi.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(b200_decode_args), sizeof(b200_prefill_args), sizeof(b200_ar_args));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "p.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "p")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        dec, pre, ar = map(int, subprocess.check_output([exe]).split())
    assert ctypes.sizeof(_lib.DecodeArgs) == dec
    assert ctypes.sizeof(_lib.PrefillArgs) == pre
    assert ctypes.sizeof(_lib.ArArgs) == ar


def test_argument_errors_without_device():
    from b200ops import _lib
    lib = _lib.lib
    assert lib.b200_swiglu(None, None, 1, 8, 16, 8, None) == _lib.E_NULL
    assert b"null" in lib.b200_last_error()
    assert lib.b200_rmsnorm(None, None, None, 1, 8, 8, 8, 1e-6, 0, None) == _lib.E_NULL
    assert lib.b200_rmsnorm_residual(None, None, None, None, None, 1, 8, 8, 8, 8, 8, 1e-6, 0, None) == _lib.E_NULL
    assert lib.b200_store_kvcache(None, None, None, None, None, 1, 1, 8, 8, 8, 8, 8, 1, 256, 8, 8, 8, None) == _lib.E_NULL
    assert lib.b200_paged_attn_decode(None, None) == _lib.E_NULL
    assert lib.b200_paged_attn_prefill(None, None) == _lib.E_NULL
    assert lib.b200_greedy_argmax(None, 0, 1, 8, 8, None, None, 0, None) == _lib.E_NULL
    assert lib.b200_greedy_argmax(0x1000, 0, 1, 8, 4, 0x1000, None, 0, None) == _lib.E_RANGE          # stride < vocab
    assert lib.b200_greedy_argmax(0x1000, 7, 1, 8, 8, 0x1000, None, 0, None) == _lib.E_UNSUPPORTED    # dtype
    assert lib.b200_greedy_argmax(0x1000, 0, 4, 100000, 100000, 0x1000, None, 0, None) == _lib.E_NULL  # scratch needed
    assert lib.b200_greedy_argmax(0x1000, 0, 4, 100000, 100000, 0x1000, 0x1000, 8, None) == _lib.E_RANGE
    assert lib.b200_greedy_argmax(0x1000, 0, 0, 100000, 100000, 0x1000, None, 0, None) == 0            # no rows: no-op
    assert lib.b200_greedy_argmax_workspace_bytes(32) == 512
    gemm = lambda M, K, N, sx, sw, so, ws=None, wsb=0, ks=1: lib.b200_skinny_gemm(  # noqa: E731
        0x1000, 0x1000, None, 0x1000, ws, wsb, M, K, N, sx, sw, so, ks, 0, None)
    assert lib.b200_skinny_gemm(None, None, None, None, None, 0, 1, 64, 64, 64, 64, 64, 1, 0, None) == _lib.E_NULL
    assert gemm(65, 64, 64, 64, 64, 64) == _lib.E_SHAPE                 # more than 64 rows
    assert gemm(8, 100, 64, 104, 64, 64) == _lib.E_SHAPE                # K % 64
    assert gemm(8, 64, 72, 64, 72, 72) == _lib.E_SHAPE                  # N % 64
    assert gemm(8, 64, 64, 60, 64, 64) == _lib.E_ALIGN                  # stride_x < K
    assert gemm(8, 64, 64, 64, 64, 66) == _lib.E_ALIGN                  # stride_out % 4
    assert gemm(8, 256, 64, 256, 64, 64, None, 0, 2) == _lib.E_RANGE    # split-K without a workspace
    assert gemm(0, 64, 64, 64, 64, 64) == 0                             # no rows: no-op
    assert lib.b200_skinny_gemm_workspace_bytes(32, 4096, 1) == 16384
    assert lib.b200_skinny_gemm_workspace_bytes(32, 4096, 9) == 16384 + 9 * 32 * 4096 * 4
    assert lib.b200_skinny_gemm_splits(32, 100, 4096, 0) == -1
    assert 1 <= lib.b200_skinny_gemm_splits(32, 14336, 4096, 0) <= 64
    a = _lib.DecodeArgs()
    for f in ("q", "k_cache", "v_cache", "cache_seqlens", "block_table", "out"):
        setattr(a, f, 0x1000)
    a.B, a.Hq, a.Hkv, a.D = 1, 6, 4, 128
    assert lib.b200_paged_attn_decode(ctypes.byref(a), None) == _lib.E_SHAPE       # Hq % Hkv
    a.Hq, a.D = 8, 96
    assert lib.b200_paged_attn_decode(ctypes.byref(a), None) == _lib.E_SHAPE       # head_dim
    a.D, a.block_size, a.n_blocks, a.max_blocks_per_seq, a.num_splits = 128, 256, 4, 4, 0
    assert lib.b200_paged_attn_decode(ctypes.byref(a), None) == _lib.E_RANGE       # num_splits
    assert lib.b200_decode_workspace_bytes(32, 32, 8, 128, 4) == 65536 * 4 + 32 * 32 * 4 * 4 + 32 * 32 * 4 * 128 * 4
    assert lib.b200_decode_workspace_bytes(32, 30, 8, 128, 4) == -1
    assert lib.b200_prefill_workspace_bytes(2, 128, 32, 128) == 2 * 128 * 32 * 128 * 2 + 256   # rotated Q + work counter


def test_allreduce_norm_argument_errors_without_device():
    from b200ops import _lib
    lib = _lib.lib
    assert lib.b200_allreduce_residual_rmsnorm(None, None) == _lib.E_NULL
    a = _lib.ArArgs()
    a.world, a.rank, a.res_out, a.epoch = 3, 0, 0x1000, 0x1000
    for k in range(3):
        a.x_peers[k] = a.flag_peers[k] = 0x1000
    a.rows, a.n, a.stride_x, a.stride_res = 1, 64, 64, 64
    assert lib.b200_allreduce_residual_rmsnorm(ctypes.byref(a), None) == _lib.E_UNSUPPORTED     # world 3
    a.world, a.n = 2, 60
    assert lib.b200_allreduce_residual_rmsnorm(ctypes.byref(a), None) == _lib.E_SHAPE           # n % 8
    a.n, a.rank = 64, 2
    assert lib.b200_allreduce_residual_rmsnorm(ctypes.byref(a), None) == _lib.E_RANGE           # rank >= world
    a.rank, a.weight = 0, 0x1000
    assert lib.b200_allreduce_residual_rmsnorm(ctypes.byref(a), None) == _lib.E_NULL            # weight without y
    assert lib.b200_comm_alloc(0, None) == _lib.E_NULL
    assert lib.b200_ipc_export(None, None) == _lib.E_NULL and lib.b200_ipc_import(None, None) == _lib.E_NULL
    assert lib.b200_comm_free(None) == 0 and lib.b200_ipc_close(None) == 0


def test_peer_comm_requires_process_group():
    import torch
    from b200ops.comm import PeerComm, allreduce_residual_rmsnorm_local
    with pytest.raises(AssertionError):
        PeerComm(1 << 20)                                   # no torch.distributed: nothing to rendezvous with
    x = torch.zeros(2, 16, dtype=torch.bfloat16)
    with pytest.raises(AssertionError):                     # CPU tensors are refused, no fallback
        allreduce_residual_rmsnorm_local(x, x, torch.ones(16, dtype=torch.bfloat16), x.clone(), x.clone(), 1e-6, False)


def test_ops_refuse_cpu_tensors():
    import torch
    from kernel.swiglu import swiglu_fused, swiglu_gemm
    from kernel.rmsnorm_add import rmsnorm, rmsnorm_
    from kernel.rmsnorm_residual import rmsnorm_residual_gemm
    x = torch.zeros(2, 16, dtype=torch.bfloat16)
    w = torch.ones(16, dtype=torch.bfloat16)
    with pytest.raises(AssertionError):
        swiglu_fused(x)
    with pytest.raises(AssertionError):
        swiglu_gemm(x, torch.zeros(2, 8, dtype=torch.bfloat16))
    with pytest.raises(AssertionError):
        rmsnorm(x, w)
    with pytest.raises(AssertionError):
        rmsnorm_(x, w, torch.zeros_like(x))
    with pytest.raises(AssertionError):
        rmsnorm_residual_gemm(x, x, w, torch.zeros_like(x), torch.zeros_like(x))


def test_split_heuristic_is_launch_constant():
    from b200ops import decode_num_splits
    assert decode_num_splits(32, 8, 16 * 256) == 1           # 256 CTAs >= 148 SMs: no split
    assert decode_num_splits(1, 2, 4 * 256) == 4             # capped by what the table can hold
    assert decode_num_splits(8, 8, 512 * 256) == 4           # 256 CTAs, one wave
    assert decode_num_splits(1, 8, 16 * 256) == 16
    assert decode_num_splits(64, 8, 32 * 256) >= 1


def test_sampler_and_new_ops_refuse_cpu_and_unsupported_modes():
    import torch
    import b200ops
    from core.layer.sampler import Sampler
    logits = torch.zeros(2, 64, dtype=torch.bfloat16)
    ones = torch.ones(2)
    with pytest.raises(AssertionError):                     # top-k / top-p: CUDA tensors only, no fallback
        Sampler()(logits, ones, ones, 50)
    with pytest.raises(AssertionError):
        b200ops.topk_topp_sample(logits, ones, ones, 50)
    lib = b200ops._lib.lib
    E = b200ops._lib
    assert lib.b200_topk_topp_sample(None, 0, 1, 8, 8, 4, None, None, None, None, None, None, None) == E.E_NULL
    assert lib.b200_topk_topp_sample(0x1000, 0, 1, 8, 4, 4, None, None, None, 0x1000, None, None, None) == E.E_RANGE
    assert lib.b200_topk_topp_sample(0x1000, 0, 1, 8, 8, 0, None, None, None, 0x1000, None, None, None) == E.E_RANGE
    assert lib.b200_topk_topp_sample(0x1000, 0, 1, 8, 8, 1025, None, None, None, 0x1000, None, None, None) == E.E_RANGE
    assert lib.b200_topk_topp_sample(0x1000, 7, 1, 8, 8, 4, None, None, None, 0x1000, None, None, None) == E.E_UNSUPPORTED
    assert lib.b200_topk_topp_sample(0x1000, 0, 1, 8, 8, 4, None, None, None, 0x1000, 0x1000, None, None) == E.E_NULL
    assert lib.b200_topk_topp_sample(0x1000, 0, 0, 8, 8, 4, None, None, None, 0x1000, None, None, None) == 0   # no rows
    with pytest.raises(AssertionError):                     # greedy: CUDA tensors only, no fallback
        Sampler()(logits, ones, ones, 1)
    with pytest.raises(AssertionError):
        b200ops.greedy_argmax(logits)
    with pytest.raises(AssertionError):
        b200ops.skinny_gemm(torch.zeros(4, 64, dtype=torch.bfloat16), torch.zeros(64, 64, dtype=torch.bfloat16))
    assert b200ops.skinny_gemm_ok(32, 14336, 4096) and b200ops.skinny_gemm_ok(1, 64, 64)
    assert not b200ops.skinny_gemm_ok(65, 4096, 4096) and not b200ops.skinny_gemm_ok(0, 64, 64)
    assert not b200ops.skinny_gemm_ok(8, 4096 + 32, 4096) and not b200ops.skinny_gemm_ok(8, 4096, 4096 + 8)


def test_balanced_form_heuristic_is_a_launch_time_function_of_shape():
    """b200ops.decode_use_balanced (pure host logic): the persistent stream-K form for split-KV grids that need more than
    one wave of the 3 x #SMs resident CTAs and leave the last wave mostly empty (measured: profiles/r02/bench_ragged.jsonl),
    the plain grid when everything fits one wave; never above the kernel's batch limit."""
    import b200ops
    f = b200ops.decode_use_balanced
    assert f(64, 8, 148, 8) is True            # Llama-3-70B heads, B = 64: 512 CTAs on 444 slots
    assert f(32, 8, 148, 4) is False           # config[1]: 256 CTAs, one wave
    assert f(64, 4, 148, 7) is False           # Qwen2.5-7B heads, B = 64: 256 CTAs
    assert f(111, 8, 148, 4) is False          # 888 CTAs = exactly two full waves
    assert f(128, 8, 148, 4) is True           # 1024 CTAs: 2.3 waves
    assert f(4096, 8, 148, 4) is False         # above DECODE_BALANCED_MAX_B
    assert f(64, 1, 148, 40) is False          # 3 groups of 16 q heads: 192 CTAs
