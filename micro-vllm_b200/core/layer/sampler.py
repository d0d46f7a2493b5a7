"""Token selection — drop-in for the reference's ``core/layer/sampler.py`` (``Sampler`` :4-43) on the
configuration the hot path's parity contract names: greedy decoding (``top_k == 1``).

The reference compiles ``_sample_impl`` with Inductor (:7-13) and evaluates
``topk(logits / temp, top_k)`` -> softmax -> top-p mask -> multinomial (:21-43).  With one
candidate every stage after the top-k is the identity, so the result is the index of the row
maximum; that is one hand-written launch here (``b200_greedy_argmax``), capturable at the tail
of the decode graph (``ModelGraphRunner.capture(greedy=True)``).  Stochastic sampling
(``top_k > 1``) is outside the hot-path scope and raises instead of silently
running another code path.
"""
import torch

import b200ops


class Sampler:
    def __call__(self, logits: torch.Tensor, temperatures: torch.Tensor = None, top_ps: torch.Tensor = None,
                 top_k: int = 1) -> torch.Tensor:
        """Same positional surface as the reference (:16): ``logits [B, vocab]``; returns ``[B]`` int64.
        A positive temperature does not move the arg-max, so ``temperatures`` / ``top_ps`` are unused."""
        if int(top_k) != 1:
            raise NotImplementedError("Sampler: only greedy decoding (top_k == 1) is implemented on the B200 path")
        assert logits.is_cuda, "Sampler: logits must be a CUDA tensor (no CPU fallback exists)"
        return b200ops.greedy_argmax(logits.view(-1, logits.shape[-1]))

    _sample_impl = None     # the reference's Inductor-compiled body (:21-43) has no counterpart
