"""b200_greedy_argmax_noalloc: the arg-max for hosts that cannot allocate device memory.

Some users of libb200ops.so have no scratch buffer to lend a kernel: an engine whose KV pool was sized to fill the
GPU (a later cudaMalloc fails), a decode loop that cannot take cudaMalloc's device-wide synchronisation, or a
cgo / JNI / ctypes binding that owns nothing but the logits and the id buffer it passes in.  For them the library
carries the arg-max merge scratch itself (64 KiB per GPU, B200_ARGMAX_NOALLOC_MAX_ROWS = 4096 rows) and this entry
point uses it.  The checks below go straight through the C-ABI: the ids must be IDENTICAL (integer output, so the
tolerance is zero, as in tests/test_gpu_sampling.py) to what b200_greedy_argmax returns with a caller workspace,
on two calls in a row - the second call sees whatever the first left in the shared array.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem here must not hide the
parity tests.  It passes alone in a fresh process and after the rest of the suite."""
import pytest
import torch

from b200ops import _lib

lib = _lib.lib
MAX_ROWS = 4096          # B200_ARGMAX_NOALLOC_MAX_ROWS in include/b200_abi.h


def _dt(t):
    return _lib.DT_BF16 if t.dtype == torch.bfloat16 else _lib.DT_F32


def _with_workspace(logits, rows, vocab, stride):
    out = torch.full((rows,), -7, dtype=torch.int64, device=logits.device)
    nbytes = int(lib.b200_greedy_argmax_workspace_bytes(rows))
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=logits.device)
    rc = lib.b200_greedy_argmax(logits.data_ptr(), _dt(logits), rows, vocab, stride, out.data_ptr(),
                                ws.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.b200_last_error()
    return out


def _no_alloc(logits, rows, vocab, stride):
    out = torch.full((rows,), -7, dtype=torch.int64, device=logits.device)
    rc = lib.b200_greedy_argmax_noalloc(logits.data_ptr(), _dt(logits), rows, vocab, stride, out.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.b200_last_error()
    return out


def test_no_alloc_argument_errors_need_no_device():
    """Bad arguments are refused before the device is touched (runs on the CPU-only box too)."""
    assert lib.b200_greedy_argmax_noalloc(None, 0, 1, 8, 8, None, None) == _lib.E_NULL
    assert lib.b200_greedy_argmax_noalloc(0x1000, 0, 1, 8, 4, 0x1000, None) == _lib.E_RANGE           # stride < vocab
    assert lib.b200_greedy_argmax_noalloc(0x1000, 7, 1, 8, 8, 0x1000, None) == _lib.E_UNSUPPORTED     # dtype
    assert lib.b200_greedy_argmax_noalloc(0x1000, 0, MAX_ROWS + 1, 8, 8, 0x1000, None) == _lib.E_RANGE
    assert b"built-in scratch" in lib.b200_last_error()
    assert lib.b200_greedy_argmax_noalloc(0x1000, 0, 0, 8, 8, 0x1000, None) == 0                      # no rows: no-op


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_no_alloc_matches_workspace_entry_twice_in_a_row(dtype):
    """Two consecutive calls (different data, different row counts) equal the workspace form bit for bit."""
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(1234)
    # vocab sizes that need many column chunks per row (the scratch is only used when chunks > 1), one ragged
    for rows, vocab, pad in [(32, 152064, 0), (7, 128256 + 3, 5)]:
        stride = vocab + pad
        base = torch.randn(rows, stride, generator=g, dtype=torch.float32).to(dtype).to(dev)
        for call in range(2):
            logits = base.roll(call * 9973 + 1, dims=1).contiguous()
            logits[:, vocab:] = 1e4                                      # padding beyond vocab must be ignored
            want = _with_workspace(logits, rows, vocab, stride)
            got = _no_alloc(logits, rows, vocab, stride)
            torch.cuda.synchronize()
            assert torch.equal(got, want), f"rows={rows} vocab={vocab} call {call}"
            x = logits[:, :vocab].float().cpu()                           # and both are right: first index of the max
            first = (x == x.max(dim=1, keepdim=True).values).float().argmax(dim=1)
            assert torch.equal(got.cpu(), first)


@pytest.mark.gpu
def test_no_alloc_full_capacity_and_ties():
    """MAX_ROWS rows use the whole built-in array; equal maxima resolve to the first index like the workspace form."""
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    vocab = 8192 * 2 + 8                      # bf16: 8192 elements per chunk -> 3 chunks
    g = torch.Generator(device="cpu").manual_seed(7)
    logits = torch.randint(-3, 4, (MAX_ROWS, vocab), generator=g).to(torch.bfloat16).to(dev)   # many ties
    for _ in range(2):
        want = _with_workspace(logits, MAX_ROWS, vocab, vocab)
        got = _no_alloc(logits, MAX_ROWS, vocab, vocab)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
        logits = logits.flip(1).contiguous()
    # one row beyond the limit is refused, and the refusal names the limit
    out = torch.empty(MAX_ROWS + 1, dtype=torch.int64, device=dev)
    big = torch.zeros(MAX_ROWS + 1, 8, dtype=torch.bfloat16, device=dev)
    rc = lib.b200_greedy_argmax_noalloc(big.data_ptr(), _lib.DT_BF16, MAX_ROWS + 1, 8, 8, out.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.E_RANGE and b"4096" in lib.b200_last_error()
