"""CPU restatement of the reference sampler (TEST INFRASTRUCTURE — never imported by the product path).

Follows /root/reference/core/layer/sampler.py:21-43 (``Sampler._sample_impl``) stage by stage:
``logits / temp`` -> ``topk`` -> ``softmax`` -> ``sort`` descending -> ``cumsum`` -> ``cum < top_p`` with the first
candidate always kept -> renormalise -> ``multinomial(1)`` -> map back to token ids.

Two places where the reference leaves the result to torch internals are made explicit here (and in the kernel):
  * ties: ``torch.topk`` / ``torch.sort`` do not specify the order of equal values; the restatement orders equal
    values by ascending token id, and takes the lowest ids of a tie that straddles the top-k boundary;
  * the draw: ``torch.multinomial`` consumes torch's generator stream; the restatement draws by inverse CDF from a
    caller-supplied uniform ``u`` in [0, 1) — the same distribution (first candidate whose cumulative renormalised
    probability exceeds ``u``), reproducible across implementations.

PINNED against the reference itself: tests/golden/sampler_reference.pt holds the outputs of the reference's
``_sample_impl`` (imported from /root/reference by tests/golden/make_golden_sampler.py with ``torch.multinomial``
instrumented to enumerate every candidate position and to draw by the same inverse CDF); tests/test_oracle_golden.py
checks this file against it.
"""
import torch


def sample_oracle(logits: torch.Tensor, temperatures: torch.Tensor, top_ps: torch.Tensor, top_k: int,
                  u: torch.Tensor = None):
    """Returns (ids [B, k] int64 candidates in sampling order, probs [B, k] fp32 renormalised (0 = masked),
    draw [B] int64 or None).  fp32 arithmetic as the reference (bf16 logits promote on the division, :23)."""
    x = logits.float() / temperatures.float()[:, None]                     # :23
    B, V = x.shape
    k = min(int(top_k), V)
    # :26 + :30 with explicit tie order: value descending, token id ascending (a stable sort of id-ordered input)
    order = torch.sort(x, dim=-1, descending=True, stable=True).indices[:, :k]
    vals = x.gather(-1, order)
    probs = torch.softmax(vals, dim=-1)                                     # :27 (already sorted: :30 is the identity)
    cum = torch.cumsum(probs, dim=-1)                                       # :31
    valid = cum < top_ps.float()[:, None]                                   # :32
    valid[..., 0] = True                                                    # :33
    masked = probs * valid                                                  # :35
    norm = masked / masked.sum(dim=-1, keepdim=True)                        # :36
    draw = None
    if u is not None:
        cn = torch.cumsum(norm, dim=-1)
        below = ((cn <= u.float()[:, None]) & valid).sum(-1)
        pick = torch.minimum(below, valid.sum(-1) - 1)
        draw = order.gather(-1, pick[:, None]).squeeze(-1)                  # :42-43
    return order, norm, draw


def decision_margin(logits, temperatures, top_ps, top_k, u=None) -> torch.Tensor:
    """Per row: how far (absolute probability) the nearest data-dependent decision is from flipping — the distance
    of any cumulative sum to ``top_p`` and, with ``u``, of any renormalised cumulative sum to ``u``.  Rows with a
    margin below fp32 summation noise (~1e-6) may legitimately differ between two correct implementations."""
    x = logits.float() / temperatures.float()[:, None]
    k = min(int(top_k), x.shape[1])
    vals = torch.sort(x, dim=-1, descending=True, stable=True).values[:, :k]
    probs = torch.softmax(vals, dim=-1)
    cum = torch.cumsum(probs, dim=-1)
    m = (cum - top_ps.float()[:, None]).abs().min(-1).values
    if u is not None:
        _, norm, _ = sample_oracle(logits, temperatures, top_ps, top_k)
        m = torch.minimum(m, (torch.cumsum(norm, -1) - u.float()[:, None]).abs().min(-1).values)
    return m
