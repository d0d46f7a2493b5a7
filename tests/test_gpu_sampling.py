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
    drawn = sampler(logits, temps, temps * 0.9, 50)                      # stochastic path: one of the 50 best per row
    top50 = logits.float().topk(50, dim=-1).indices
    assert drawn.shape == (B,) and bool((top50 == drawn[:, None]).any(-1).all())
    # a batch size without a graph takes the eager path + one extra launch
    ids2 = dec.forward_greedy(tok[:2], mgr, 2)
    assert ids2.shape == (2,) and ids2.dtype == torch.int64


# ------------------------------------------------------------------ top-k / top-p sampling (sampler.py:21-43)
def _dense(ids, probs, vocab):
    out = torch.zeros(ids.shape[0], vocab, dtype=torch.float64)
    out.scatter_(1, ids.long().cpu(), probs.double().cpu())
    return out


def _check_against_oracle(logits, temps, top_ps, top_k, u, name=""):
    from oracle.sampling import decision_margin, sample_oracle
    V = logits.shape[1]
    got, gi, gp = b200ops.topk_topp_sample(logits.cuda(), temps.cuda(), top_ps.cuda(), top_k, u.cuda(), debug=True)
    ids, probs, draw = sample_oracle(logits, temps, top_ps, top_k, u)
    gi, gp, got = gi.cpu().long(), gp.cpu(), got.cpu()
    # candidate ORDER is part of the kernel's contract (value descending, token id ascending): bit-exact ids
    assert torch.equal(gi, ids), f"{name}: candidate order differs"
    safe = decision_margin(logits, temps, top_ps, top_k, u) > 2e-6          # decisions not inside fp32 summation noise
    assert float(safe.float().mean()) >= 0.8, f"{name}: test inputs too degenerate"
    assert torch.equal((gp > 0).sum(-1)[safe], (probs > 0).sum(-1)[safe]), f"{name}: top-p cut differs"
    # fp32 probabilities: expf / summation order differ -> relative 2e-5, absolute 1e-7 (written tolerance)
    assert torch.allclose(gp[safe], probs[safe], rtol=2e-5, atol=1e-7), \
        f"{name}: probabilities differ by {float((gp[safe] - probs[safe]).abs().max()):.3e}"
    assert torch.equal(got[safe], draw[safe]), f"{name}: draw differs"
    return got


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,vocab,top_k", [(8, 151936, 50), (4, 151936, 1000), (64, 32000, 50), (5, 2048, 50),
                                                (3, 1001, 7), (2, 9, 50), (1, 128256, 1024), (6, 8193, 2), (3, 40, 33)])
def test_topk_topp_sample_matches_oracle(rows, vocab, top_k, dtype):
    g = torch.Generator().manual_seed(rows * 7919 + vocab + top_k)
    logits = (torch.randn(rows, vocab, generator=g) * 3).to(dtype)          # bf16: ties inside and across the top-k cut
    temps = torch.rand(rows, generator=g) * 1.2 + 0.3
    top_ps = torch.rand(rows, generator=g) * 0.7 + 0.3
    top_ps[0] = 2.0                                                          # nothing masked
    u = torch.rand(rows, generator=g)
    _check_against_oracle(logits, temps, top_ps, top_k, u, f"{rows}x{vocab} k={top_k} {dtype}")


def test_topk_topp_sample_reference_golden():
    """The kernel against the reference's own outputs (tests/golden/sampler_reference.pt, written by running
    core/layer/sampler.py:21-43 on CPU): same probability for every token id, same draw up to exact ties."""
    import os
    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "sampler_reference.pt"), weights_only=False)
    for c in gold:
        V = c["logits"].shape[1]
        got, gi, gp = b200ops.topk_topp_sample(c["logits"].cuda(), c["temperatures"].cuda(), c["top_ps"].cuda(),
                                               c["top_k"], c["u"].cuda(), debug=True)
        assert torch.allclose(_dense(gi, gp, V), _dense(c["ref_ids"], c["ref_probs"], V), rtol=2e-5, atol=1e-7), c["name"]
        lg = c["logits"].float()
        assert torch.equal(lg.gather(1, got.cpu()[:, None]), lg.gather(1, c["ref_draw"][:, None])), c["name"]
        if c["logits"].dtype == torch.float32:
            assert torch.equal(got.cpu(), c["ref_draw"]), c["name"]


def test_topk_topp_sample_edges():
    V = 5000
    x = torch.full((6, V), -2.0, dtype=torch.bfloat16)
    x[0, 4999] = 9.0                         # one dominant candidate at the very end
    x[1] = 0.0                               # every logit equal: candidates are ids 0..k-1, uniform
    x[2] = float("-inf"); x[2, 17] = 1.0     # -inf everywhere else: probability exactly 1 on id 17
    x[3, 100], x[3, 200], x[3, 300] = 3.0, 3.0, 3.0
    x[4, :] = torch.linspace(-8, 8, V).to(torch.bfloat16)
    x[5, 0] = 4.0
    temps = torch.tensor([1.0, 0.7, 1.0, 0.5, 1.3, 1e-3])      # near-zero temperature: all mass on the maximum
    top_ps = torch.tensor([0.9, 2.0, 0.5, 0.95, 0.8, 0.9])
    for uval in (0.0, 0.37, 0.99):
        u = torch.full((6,), uval)
        got = _check_against_oracle(x, temps, top_ps, 50, u, f"edges u={uval}")
        assert int(got[0]) == 4999 and int(got[2]) == 17 and int(got[5]) == 0
        assert 0 <= int(got[1]) < 50 and int(got[3]) in (100, 200, 300)
    # defaults (no temperature / top_p / uniform) = the arg-max; strided rows; caller's output buffer
    big = torch.randn(4, 2 * 9000 + 3, generator=torch.Generator().manual_seed(3)).to(torch.bfloat16).cuda()
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    r = b200ops.topk_topp_sample(big[:, 5:9004], None, None, 50, None, out=out)
    assert r.data_ptr() == out.data_ptr()
    assert torch.equal(out.cpu(), first_argmax_cpu(big[:, 5:9004].cpu()))
    with pytest.raises(b200ops.B200Error):
        b200ops.topk_topp_sample(big, None, None, 0)
    with pytest.raises(b200ops.B200Error):
        b200ops.topk_topp_sample(big, None, None, 1025)


def test_topk_topp_sample_distribution_and_sampler_class():
    """Many draws of one row follow the renormalised probabilities (chi-square against the oracle's distribution,
    the same check one would apply to torch.multinomial), through the reference-facing ``Sampler`` call."""
    from core.layer.sampler import Sampler
    from oracle.sampling import sample_oracle
    g = torch.Generator().manual_seed(11)
    V, k, n = 32000, 50, 200000
    row = (torch.randn(1, V, generator=g) * 2.5).to(torch.bfloat16)
    temps, top_ps = torch.tensor([0.8]), torch.tensor([0.9])
    ids, probs, _ = sample_oracle(row, temps, top_ps, k)
    s = Sampler(seed=5)
    logits = row.cuda().expand(n // 100, V).contiguous()
    counts = torch.zeros(V, dtype=torch.float64)
    tt, pp = temps.cuda().expand(n // 100).contiguous(), top_ps.cuda().expand(n // 100).contiguous()
    for _ in range(100):
        d = s(logits, tt, pp, k)
        counts += torch.bincount(d.cpu(), minlength=V).double()
    expect = _dense(ids, probs, V)[0] * n
    assert float(counts[expect == 0].sum()) == 0, "a masked / non-candidate token was drawn"
    m = expect > 5
    chi2 = float(((counts[m] - expect[m]) ** 2 / expect[m]).sum())
    dof = int(m.sum()) - 1
    assert chi2 < dof + 6 * (2 * dof) ** 0.5 + 10, f"chi-square {chi2:.1f} for {dof} degrees of freedom"
    # greedy through the same class; a second sampler with the same seed repeats the stream
    assert torch.equal(s(logits[:3], tt[:3], pp[:3], 1).cpu(), first_argmax_cpu(row).expand(3))
    a, b = Sampler(seed=9), Sampler(seed=9)
    assert torch.equal(a(logits, tt, pp, k), b(logits, tt, pp, k))
