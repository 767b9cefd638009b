"""make_sampler of the reference (PIE/samplers/__init__.py:11-45) over the device-side sampler kernel.

    sampler = make_sampler(temp=0.8, top_p=0.95, seed=0)
    tokens = sampler(logits)            # bf16 logits [rows, vocab] -> int32 [rows], nothing leaves the GPU

The reference's samplers take log-probabilities; pie_sample computes them itself from the logits in the same launch
(logprobs = f32(logits) - logsumexp, PIE/engine/inference_engine.py:268-269), which is what the engine feeds them.
The random stream is Philox keyed by a seed in device memory (MLX's threefry stream cannot be reproduced without MLX):
draws are reproducible per (seed, call number), and a CUDA graph that captured the call replays with fresh numbers."""
from __future__ import annotations

from collections.abc import Callable

import torch

from . import ops


def make_sampler(temp: float = 0.0, top_p: float = 0.0, min_p: float = 0.0, min_tokens_to_keep: int = 1,
                 top_k: int = -1, seed: int = 0, device="cuda") -> Callable[[torch.Tensor], torch.Tensor]:
    if temp == 0:
        return lambda x: ops.argmax(x)
    state = ops.new_rng_state(seed, device)

    def sampler(x: torch.Tensor, **kw) -> torch.Tensor:
        return ops.sample(x, temp, top_p, min_p, min_tokens_to_keep, top_k, rng_state=state, **kw)

    sampler.rng_state = state
    sampler.params = dict(temp=temp, top_p=top_p, min_p=min_p, min_tokens_to_keep=min_tokens_to_keep, top_k=top_k)
    return sampler
