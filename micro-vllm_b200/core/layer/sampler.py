"""Token selection — drop-in for the reference's ``core/layer/sampler.py`` (``Sampler`` :4-43).

The reference compiles ``_sample_impl`` with Inductor (:7-13) and evaluates
``topk(logits / temp, top_k)`` -> softmax -> top-p mask -> multinomial (:21-43).  Here both settings the
engine uses are one hand-written launch each, capturable at the tail of the decode graph:

  * ``top_k == 1`` (greedy; every stage after the top-k is the identity): ``b200_greedy_argmax``
    (``ModelGraphRunner.capture(greedy=True)``);
  * ``top_k > 1`` (50 for decode, 1000 for prefill: core/engine.py:233, :278): ``b200_topk_topp_sample`` — radix
    select of the k largest, shared-memory sort, fp32 softmax, top-p prefix mask with the first candidate always
    kept, renormalisation and an inverse-CDF draw, one CTA per row.  The draw consumes one uniform number per row
    from ``torch.rand`` on this sampler's generator (plumbing): the DISTRIBUTION is that of the reference's
    ``torch.multinomial``; its random stream is torch-internal and not reproducible by another implementation.
"""
import torch

import b200ops


class Sampler:
    def __init__(self, seed: int = None):
        self._gen = None
        self._seed = seed

    def _generator(self, device):
        if self._seed is None:
            return None
        if self._gen is None or self._gen.device != device:
            self._gen = torch.Generator(device=device)
            self._gen.manual_seed(self._seed)
        return self._gen

    def __call__(self, logits: torch.Tensor, temperatures: torch.Tensor = None, top_ps: torch.Tensor = None,
                 top_k: int = 1, uniforms: torch.Tensor = None) -> torch.Tensor:
        """Same positional surface as the reference (:16): ``logits [B, vocab]``, ``temperatures`` / ``top_ps``
        ``[B]``; returns ``[B]`` int64.  ``uniforms`` (optional, fp32 ``[B]`` in [0, 1)) replaces the internal draw."""
        assert logits.is_cuda, "Sampler: logits must be a CUDA tensor (no CPU fallback exists)"
        logits = logits.view(-1, logits.shape[-1])
        if int(top_k) == 1:
            return b200ops.greedy_argmax(logits)        # a positive temperature does not move the arg-max
        rows = logits.shape[0]
        f32 = lambda t: None if t is None else t.to(device=logits.device, dtype=torch.float32).contiguous()
        if uniforms is None:
            uniforms = torch.rand(rows, device=logits.device, dtype=torch.float32, generator=self._generator(logits.device))
        return b200ops.topk_topp_sample(logits, f32(temperatures), f32(top_ps), int(top_k), f32(uniforms))

    _sample_impl = None     # the reference's Inductor-compiled body (:21-43): replaced by the kernels above
