"""Greedy token selection kernel: the reference sampler evaluated at top_k = 1
(core/layer/sampler.py:21-43 -> index of the row maximum) against an exact CPU restatement —
integer output, so bit-exact, ties resolved to the FIRST index."""
import pytest
import torch

import b200ops

pytestmark = pytest.mark.gpu


def first_argmax_cpu(logits: torch.Tensor) -> torch.Tensor:
    x = logits.float().cpu()
    m = x.max(dim=-1, keepdim=True).values
    return torch.tensor([int((row == mv).nonzero()[0]) for row, mv in zip(x, m)], dtype=torch.int64)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,vocab", [(1, 151936), (32, 152064), (64, 32000), (5, 2048), (3, 1001), (2, 7),
                                         (7, 8192), (4, 8193), (1, 128256)])
def test_greedy_argmax_matches_cpu(rows, vocab, dtype):
    g = torch.Generator().manual_seed(rows * 1000003 + vocab)
    logits = (torch.randn(rows, vocab, generator=g) * 4).to(dtype)      # bf16: many exact ties near the top
    want = first_argmax_cpu(logits)
    got = b200ops.greedy_argmax(logits.cuda())
    assert got.dtype == torch.int64 and got.shape == (rows,)
    assert torch.equal(got.cpu(), want)


def test_greedy_argmax_ties_edges_and_strided_rows():
    V = 40000
    x = torch.zeros(6, V, dtype=torch.bfloat16)
    x[1, V - 1] = 1.0                       # maximum is the very last element
    x[2] = float("-inf")                    # every element -inf -> index 0
    x[3, 12345] = 2.0
    x[3, 30000] = 2.0                       # tie across chunks -> first
    x[4] = -3.0
    x[4, 8191], x[4, 8192] = -1.0, -1.0     # tie across the CTA chunk boundary
    x[5, 0] = 5.0
    want = torch.tensor([0, V - 1, 0, 12345, 8191, 0])
    assert torch.equal(b200ops.greedy_argmax(x.cuda()).cpu(), want)
    # rows that are a strided view of a wider buffer, and an unaligned vocab (scalar path)
    big = torch.randn(4, 2 * 9000 + 3, generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    view = big[:, 5:9004]
    assert torch.equal(b200ops.greedy_argmax(big.cuda()[:, 5:9004]).cpu(), first_argmax_cpu(view))
    # 16-byte aligned rows whose width is not a multiple of the vector: vector body + scalar tail
    wide = torch.randn(3, 16400, generator=torch.Generator().manual_seed(2)).to(torch.bfloat16)
    wide[1, 16394] = 50.0                   # maximum inside the tail
    assert torch.equal(b200ops.greedy_argmax(wide.cuda()[:, :16395]).cpu(), first_argmax_cpu(wide[:, :16395]))
    out = torch.full((8,), -1, dtype=torch.int64, device="cuda")
    ret = b200ops.greedy_argmax(x.cuda(), out=out)
    assert ret.data_ptr() == out.data_ptr() and torch.equal(out[:6].cpu(), want) and int(out[6]) == -1


def test_greedy_argmax_graph_replay_resets_scratch():
    rows, vocab = 8, 152064
    buf = torch.zeros(rows, vocab, dtype=torch.bfloat16, device="cuda")
    ids = torch.zeros(rows, dtype=torch.int64, device="cuda")
    b200ops.greedy_argmax(buf, out=ids)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        b200ops.greedy_argmax(buf, out=ids)
    gen = torch.Generator().manual_seed(9)
    for _ in range(4):
        x = torch.randn(rows, vocab, generator=gen).to(torch.bfloat16)
        buf.copy_(x)
        g.replay()
        assert torch.equal(ids.cpu(), first_argmax_cpu(x))


def test_sampler_mirror_and_runner_greedy_graph():
    """``ModelGraphRunner.capture(greedy=True)``: the ids produced inside the graph equal the arg-max of
    the logits the same replay produced; ``Sampler`` keeps the reference call surface."""
    from core.layer.sampler import Sampler
    from test_gpu_model import make_runners
    from core.layer.synthetic_model import SHAPES
    shape = SHAPES["qwen2.5-0.5b"]
    model, dec, pre, mgr = make_runners(shape, 2, max_bs=4, seed=2)
    B, P = 4, 60
    mgr.reset()
    dec.capture(mgr, batch_sizes=[B], greedy=True)
    prompt = torch.randint(0, shape.vocab, (B, P), generator=torch.Generator().manual_seed(3))
    for s in range(B):
        assert mgr.alloc(s, P)[0]
    mgr.cache_batch_data(list(range(B)), [P] * B)
    with torch.no_grad():
        tok = pre.forward(prompt.cuda(), mgr, B)[:, -1, :].float().argmax(-1)
    sampler = Sampler()
    for i in range(5):
        for s in range(B):
            mgr.append(s)
        mgr.cache_batch_data(list(range(B)), [P + 1 + i] * B)
        ids = dec.forward_greedy(tok, mgr, B).clone()
        logits = dec._logits[:B]
        assert torch.equal(ids.cpu(), first_argmax_cpu(logits))
        temps = torch.ones(B, device="cuda")
        assert torch.equal(sampler(logits, temps, temps, 1), ids)
        tok = ids
    with pytest.raises(NotImplementedError):
        sampler(logits, temps, temps, 50)
    # a batch size without a graph takes the eager path + one extra launch
    ids2 = dec.forward_greedy(tok[:2], mgr, 2)
    assert ids2.shape == (2,) and ids2.dtype == torch.int64
