"""Golden vectors for the top-k / top-p sampler, produced by THE REFERENCE's own ``Sampler._sample_impl``
(/root/reference/core/layer/sampler.py:21-43) run eagerly on CPU.  Build container only:

    python tests/golden/make_golden_sampler.py [/root/reference]   ->  tests/golden/sampler_reference.pt

``torch.multinomial`` is the one stage whose output depends on torch's generator, so it is instrumented while the
reference runs: (a) called once per candidate position j it returns column j, which makes the reference's own
mapping code (:41-43) emit the token id it associates with sorted position j; (b) it records the renormalised
probabilities it is handed (:36); (c) for the final draw it returns the inverse-CDF sample for a stored uniform u.
Cases avoid ties across the top-k boundary and decisions closer than 1e-4 to top_p / u (the reference leaves those
to torch internals / fp32 summation order).
"""
import os
import sys
import zlib

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != os.path.dirname(os.path.dirname(HERE))]
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

import torch  # noqa: E402

torch.compile = lambda fn=None, **kw: (fn if fn is not None else (lambda f: f))      # the ctor compiles; we call eagerly
from core.layer.sampler import Sampler  # noqa: E402  (reference)

assert os.path.abspath(sys.modules["core.layer.sampler"].__file__).startswith(os.path.abspath(REF))
impl = Sampler._sample_impl
real_multinomial = torch.multinomial


def run_case(name, logits, temps, top_ps, top_k, u):
    B = logits.shape[0]
    k = min(top_k, logits.shape[1])
    seen = {}

    def column(j):
        def fake(p, n):
            seen["probs"] = p.clone()
            return torch.full((p.shape[0], 1), j, dtype=torch.int64)
        return fake

    ids = torch.empty(B, k, dtype=torch.int64)
    for j in range(k):
        torch.multinomial = column(j)
        ids[:, j] = impl(logits, temps, top_ps, top_k)
    probs = seen["probs"].float()

    def inverse_cdf(p, n):
        cn = torch.cumsum(p.float(), -1)
        nv = (p > 0).sum(-1)
        return torch.minimum((cn <= u[:, None]).sum(-1), nv - 1)[:, None]

    torch.multinomial = inverse_cdf
    draw = impl(logits, temps, top_ps, top_k)
    torch.multinomial = real_multinomial
    # margins (see the module docstring)
    x = logits.float() / temps[:, None]
    srt = torch.sort(x, -1, descending=True).values
    if k < x.shape[1]:
        assert bool((srt[:, k - 1] > srt[:, k]).all()), f"{name}: tie across the top-k boundary, pick another seed"
    cum = torch.cumsum(torch.softmax(srt[:, :k], -1), -1)
    assert float((cum - top_ps[:, None]).abs().min()) > 1e-4, f"{name}: cumulative sum too close to top_p"
    assert float((torch.cumsum(probs, -1) - u[:, None]).abs().min()) > 1e-4, f"{name}: u too close to a boundary"
    return {"name": name, "logits": logits, "temperatures": temps, "top_ps": top_ps, "top_k": top_k, "u": u,
            "ref_ids": ids, "ref_probs": probs, "ref_draw": draw}


def break_boundary_ties(logits, k):
    """bf16 rows almost always have equal values on both sides of the top-k cut; which of them torch.topk keeps is
    unspecified, so the surplus copies (those past the k-th place) are lowered by about one bf16 step."""
    x = logits.clone()
    for r in range(x.shape[0]):
        row = x[r].float()
        thr = torch.sort(row, descending=True).values[k - 1] if k <= row.numel() else None
        if thr is None:
            continue
        need = k - int((row > thr).sum())
        pos = (row == thr).nonzero().flatten()
        lower = (thr - thr.abs() * 2.0 ** -6 - 1e-30).to(x.dtype)
        assert float(lower) < float(thr)
        x[r, pos[need:]] = lower
    return x


def main():
    cases = []
    specs = [("f32_k50", torch.float32, 6, 3000, 50, 3.0), ("bf16_k50", torch.bfloat16, 6, 3000, 50, 4.0),
             ("bf16_k1000_prefill", torch.bfloat16, 3, 4096, 1000, 6.0), ("f32_k8_small_vocab", torch.float32, 4, 9, 8, 2.0),
             ("bf16_k50_peaky", torch.bfloat16, 5, 2048, 50, 12.0)]
    for name, dt, B, V, k, scale in specs:
        for seed in range(100):
            g = torch.Generator().manual_seed(zlib.crc32(name.encode()) % 100000 + seed)
            logits = break_boundary_ties((torch.randn(B, V, generator=g) * scale).to(dt), k)
            temps = torch.rand(B, generator=g) * 1.2 + 0.3
            top_ps = torch.rand(B, generator=g) * 0.6 + 0.4
            top_ps[0] = 2.0                              # nothing masked (at exactly 1.0 the last candidate hangs on fp32 rounding)
            if B > 1:
                top_ps[1] = 0.05                         # (almost) only the first candidate survives
            u = torch.rand(B, generator=g)
            try:
                cases.append(run_case(name, logits, temps, top_ps, k, u))
                break
            except AssertionError as e:
                print("retry:", e)
        else:
            raise SystemExit(f"no admissible seed for {name}")
    out = os.path.join(HERE, "sampler_reference.pt")
    torch.save(cases, out)
    print("wrote", out, os.path.getsize(out), "bytes;", [c["name"] for c in cases])


if __name__ == "__main__":
    main()
