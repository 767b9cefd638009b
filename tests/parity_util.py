"""Shared helpers of the GPU parity tests and __graft_entry__.smoke(): run the CUDA path and the CPU oracle on
the same random-init quantized weights and prompts, teacher-forced, and report activation error and token agreement."""
from __future__ import annotations

import dataclasses

import numpy as np
import torch


def tiny_args(bits: int = 4, tie: bool = False, head_dim: int = 64, **kw):
    from proxy_inference_engine_b200.llama import ModelArgs

    cfg = dict(hidden_size=256, num_hidden_layers=2, intermediate_size=512, num_attention_heads=4,
               num_key_value_heads=2, head_dim=head_dim, vocab_size=640, rms_norm_eps=1e-5, rope_theta=500000.0,
               max_position_embeddings=131072, tie_word_embeddings=tie,
               rope_scaling={"factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0},
               quantization={"group_size": 64, "bits": bits})
    cfg.update(kw)
    return ModelArgs(**cfg)


def t2n(t: torch.Tensor) -> np.ndarray:
    """torch tensor -> the oracle's numpy convention (uint32 words, float32 holding bf16 values)."""
    t = t.detach().cpu()
    if t.dtype == torch.int32:
        return t.numpy().view(np.uint32)
    if t.dtype == torch.uint32:
        return t.view(torch.int32).numpy().view(np.uint32)
    return t.to(torch.float32).numpy()


def n2t_bf16(a: np.ndarray, device="cuda") -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.bfloat16).to(device)


def n2t_words(a: np.ndarray, device="cuda") -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32)).to(device)


def weights_to_oracle(weights: dict) -> dict:
    return {k: t2n(v) for k, v in weights.items()}


def rel_l2(a: np.ndarray, b: np.ndarray) -> float:
    a, b = a.astype(np.float64), b.astype(np.float64)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (den if den > 0 else 1.0))


def make_oracle(args, weights, mode="B"):
    from oracle import oracle as orc

    oargs = orc.ModelArgs(**{f.name: getattr(args, f.name) for f in dataclasses.fields(args)
                             if f.name in orc.ModelArgs.__dataclass_fields__})
    return orc.LlamaOracle(oargs, weights_to_oracle(weights), mode=mode)


def run_decode_parity(args, prompt_len=16, steps=8, batch=1, seed=0, paged=False, use_graph=True, pdl=True,
                      fused=False, kv_bits=None):
    from proxy_inference_engine_b200.checkpoint import build_model, random_weights
    from proxy_inference_engine_b200.decode import DecodeEngine

    w = random_weights(args, seed=seed, device="cuda", gen_device="cpu")
    model = build_model(args, w)
    o = make_oracle(args, w)
    prompt = torch.randint(0, args.vocab_size, (batch, prompt_len), generator=torch.Generator().manual_seed(1))
    eng = DecodeEngine(model, batch=batch, max_context=prompt_len + steps + 2, paged=paged, use_graph=use_graph,
                       pdl=pdl, fused=fused, kv_bits=kv_bits)
    eng.prefill(prompt)
    caches = o.make_cache(kv_bits=kv_bits) if kv_bits is not None else o.make_cache()
    lo = o(prompt.numpy(), caches, last_only=True)[:, -1]
    errs, match, strong_mismatch, total = [], 0, 0, 0
    for _ in range(steps):
        lg = eng.logits.float().cpu().numpy()
        errs.append(rel_l2(lg, lo))
        tok_o = lo.argmax(-1).astype(np.int32)
        tok_g = eng.tokens.cpu().numpy()
        top2 = np.sort(lo, axis=-1)[:, -2:]
        gap = top2[:, 1] - top2[:, 0]
        tol = 2e-2 * np.abs(lo).max(axis=-1)
        for b in range(batch):
            total += 1
            if tok_o[b] == tok_g[b]:
                match += 1
            elif gap[b] > tol[b]:
                strong_mismatch += 1
        eng.tokens.copy_(torch.from_numpy(tok_o))  # teacher forcing: both sides continue from the oracle's token
        eng.step(1)
        lo = o(tok_o[:, None], caches)[:, -1]
    return {"max_rel_err": max(errs), "errs": [round(e, 5) for e in errs], "token_match": match / total,
            "strong_mismatch": strong_mismatch, "launches_per_step": eng.launches_per_step}
