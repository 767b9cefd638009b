"""Checkpoint boundary of the path: MLX-layout quantized Llama checkpoints.

``load_model`` mirrors PIE/models/utils.py:51-125: read ``config.json`` + every ``model*.safetensors``, build
``Model(ModelArgs)``, ``sanitize``, and assign ``weight`` (U32 words) / ``scales`` / ``biases`` (BF16) to the
quantized layers named by ``config["quantization"]``.  ``random_weights`` is the synthetic-checkpoint generator
of the benchmarks: nn.Linear / nn.Embedding initialisers, bf16, quantized with the library's
bit-exact ``quantize``.
"""
from __future__ import annotations

import glob
import json
import math
import os

import torch

from . import ops
from .llama import Model, ModelArgs

LLAMA_PRESETS = {
    # SURVEY.md section 8 header; defaults of PIE/models/llama/language.py:13-29
    "llama-3.2-1b": dict(hidden_size=2048, num_hidden_layers=16, intermediate_size=8192, num_attention_heads=32,
                         num_key_value_heads=8, head_dim=64, vocab_size=128256, tie_word_embeddings=True,
                         rope_scaling={"factor": 32.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0}),
    "llama-3.1-8b": dict(hidden_size=4096, num_hidden_layers=32, intermediate_size=14336, num_attention_heads=32,
                         num_key_value_heads=8, head_dim=128, vocab_size=128256, tie_word_embeddings=False,
                         rope_scaling={"factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0}),
    "llama-3.1-70b": dict(hidden_size=8192, num_hidden_layers=80, intermediate_size=28672, num_attention_heads=64,
                          num_key_value_heads=8, head_dim=128, vocab_size=128256, tie_word_embeddings=False,
                          rope_scaling={"factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0}),
}


def preset_args(name: str, bits: int = 4, group_size: int = 64, **overrides) -> ModelArgs:
    cfg = dict(LLAMA_PRESETS[name], rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
               quantization={"group_size": group_size, "bits": bits})
    cfg.update(overrides)
    return ModelArgs(**cfg)


def _linear_names(args: ModelArgs):
    h, f = args.hidden_size, args.intermediate_size
    nh = args.num_attention_heads
    nkv = args.num_key_value_heads or nh
    d = args.head_dim or h // nh
    for i in range(args.num_hidden_layers):
        p = f"model.layers.{i}"
        yield p + ".self_attn.q_proj", nh * d, h
        yield p + ".self_attn.k_proj", nkv * d, h
        yield p + ".self_attn.v_proj", nkv * d, h
        yield p + ".self_attn.o_proj", h, nh * d
        yield p + ".mlp.gate_proj", f, h
        yield p + ".mlp.down_proj", h, f
        yield p + ".mlp.up_proj", f, h
    if not args.tie_word_embeddings:
        yield "lm_head", args.vocab_size, h


def random_weights(args: ModelArgs, seed: int = 0, device="cuda", gen_device=None, head_gain: float = 1.0,
                   shard: tuple | None = None) -> dict:
    """Random-init checkpoint in the MLX schema.  Linear: U(-1/sqrt(K), 1/sqrt(K)) (MLX nn/layers/linear.py:51-56);
    embedding: N(0, 1/sqrt(h)) (embedding.py:25); norm weights 1 + 0.1 N(0,1) so the weight multiply is exercised.
    ``gen_device='cpu'`` makes the values independent of the GPU (the parity fixtures use it).
    ``shard=(rank, world)``: every rank draws the same full tensors (same seed) and keeps only its tensor-parallel
    slice (tp.shard_tensor), so a 70B model never has to sit whole on one GPU."""
    from .tp import shard_tensor
    rk, wd = shard if shard else (0, 1)
    gen_device = gen_device or device
    g = torch.Generator(device=gen_device).manual_seed(seed)
    gs, bits = args.quantization["group_size"], args.quantization["bits"]
    w = {}

    def put_quant(name, t):
        wq, s, b = ops.quantize(t.to(torch.bfloat16).to(device), gs, bits)
        for suffix, v in ((".weight", wq), (".scales", s), (".biases", b)):
            w[name + suffix] = shard_tensor(name + suffix, v, rk, wd)

    h = args.hidden_size
    emb = torch.randn((args.vocab_size, h), generator=g, device=gen_device) * math.sqrt(1.0 / h)
    if args.tie_word_embeddings:
        emb = emb * head_gain
    put_quant("model.embed_tokens", emb)
    del emb
    for name, n, k in _linear_names(args):
        t = (torch.rand((n, k), generator=g, device=gen_device) * 2 - 1) * math.sqrt(1.0 / k)
        if name == "lm_head":
            t = t * head_gain
        put_quant(name, t)
        del t
    for i in range(args.num_hidden_layers):
        for nm in ("input_layernorm", "post_attention_layernorm"):
            w[f"model.layers.{i}.{nm}.weight"] = (1 + 0.1 * torch.randn((h,), generator=g, device=gen_device)).to(
                torch.bfloat16).to(device)
    w["model.norm.weight"] = (1 + 0.1 * torch.randn((h,), generator=g, device=gen_device)).to(torch.bfloat16).to(device)
    return w


def build_model(args: ModelArgs, weights: dict, device="cuda") -> Model:
    """Model shell + load_weights, skipping the random initialisers of the layer constructors."""
    model = _empty_model(args, device)
    weights = model.sanitize(dict(weights))
    model.load_weights({k: _as_param(v, device) for k, v in weights.items()}, strict=True)
    return model


def colocate_for_decode(model: Model) -> None:
    """Re-home the quantized matrices of the decode GEMVs so that one TMA copy brings what used to take several
    (csrc/pie_qmv2.h, Qmv2Args): every matrix's biases directly behind its scales, and each layer's up projection directly
    behind its gate projection.  Every array keeps the MLX layout (the modules hold views into the new storage);
    idempotent; costs one extra copy of a layer at a time while it runs."""
    def stacked(a, b):
        return b.data_ptr() - a.data_ptr() == a.numel() * a.element_size() and a._base is not None and a._base is b._base

    def scales_biases(lin):
        if stacked(lin.scales, lin.biases):
            return
        st = torch.empty((2,) + tuple(lin.scales.shape), dtype=lin.scales.dtype, device=lin.scales.device)
        st[0].copy_(lin.scales); st[1].copy_(lin.biases)
        lin.scales, lin.biases = st[0], st[1]

    for ly in model.layers:
        at, mlp = ly.self_attn, ly.mlp
        for lin in (at.q_proj, at.k_proj, at.v_proj, at.o_proj, mlp.down_proj):
            scales_biases(lin)
        g, u = mlp.gate_proj, mlp.up_proj
        if g.weight.shape != u.weight.shape:
            scales_biases(g); scales_biases(u)
            continue
        if not stacked(g.weight, u.weight):
            w = torch.empty((2,) + tuple(g.weight.shape), dtype=g.weight.dtype, device=g.weight.device)
            w[0].copy_(g.weight); w[1].copy_(u.weight)
            g.weight, u.weight = w[0], w[1]
        if not (stacked(g.scales, u.scales) and stacked(g.biases, u.biases)):
            sb = torch.empty((2, 2) + tuple(g.scales.shape), dtype=g.scales.dtype, device=g.scales.device)   # [s|b][gate|up]
            sb[0, 0].copy_(g.scales); sb[0, 1].copy_(u.scales); sb[1, 0].copy_(g.biases); sb[1, 1].copy_(u.biases)
            g.scales, u.scales, g.biases, u.biases = sb[0, 0], sb[0, 1], sb[1, 0], sb[1, 1]
    head = model.model.embed_tokens if model.args.tie_word_embeddings else model.lm_head
    scales_biases(head)


def _as_param(v: torch.Tensor, device):
    v = v.to(device)
    return v.view(torch.int32) if v.dtype == torch.uint32 else v


def _empty_model(args: ModelArgs, device) -> Model:
    """Instantiate the module tree with empty parameters (nn.quantize swaps shells in the same way,
    PIE/models/utils.py:95-111)."""
    from . import nn as pnn

    gs, bits = args.quantization["group_size"], args.quantization["bits"]
    orig_ql, orig_qe = pnn.QuantizedLinear.__init__, pnn.QuantizedEmbedding.__init__

    def ql_init(self, input_dims, output_dims, bias=True, group_size=64, bits=4, device="cuda"):
        self.group_size, self.bits = group_size, bits
        self.weight = torch.empty((output_dims, input_dims * bits // 32), dtype=torch.int32, device=device)
        self.scales = torch.empty((output_dims, input_dims // group_size), dtype=torch.bfloat16, device=device)
        self.biases = torch.empty_like(self.scales)

    def qe_init(self, num_embeddings, dims, group_size=64, bits=4, device="cuda"):
        self.group_size, self.bits, self.num_embeddings, self.dims = group_size, bits, num_embeddings, dims
        self.weight = torch.empty((num_embeddings, dims * bits // 32), dtype=torch.int32, device=device)
        self.scales = torch.empty((num_embeddings, dims // group_size), dtype=torch.bfloat16, device=device)
        self.biases = torch.empty_like(self.scales)

    pnn.QuantizedLinear.__init__, pnn.QuantizedEmbedding.__init__ = ql_init, qe_init
    try:
        return Model(args, device)
    finally:
        pnn.QuantizedLinear.__init__, pnn.QuantizedEmbedding.__init__ = orig_ql, orig_qe


def save_checkpoint(path: str, args: ModelArgs, weights: dict, max_shard_bytes: int | None = None) -> None:
    """config.json + MLX-layout safetensors.  max_shard_bytes: split into ``model-0000i-of-0000n.safetensors`` files of at
    most that size plus ``model.safetensors.index.json`` (``metadata.total_size``, ``weight_map``: tensor name -> file),
    the sharded form mlx_lm / Hugging Face write for large models and PIE's loader globs (PIE/models/utils.py:62-66)."""
    from safetensors.torch import save_file

    os.makedirs(path, exist_ok=True)
    cfg = {k: getattr(args, k) for k in args.__dataclass_fields__}
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cfg, f, indent=1)
    out = {k: (v.view(torch.uint32) if v.dtype == torch.int32 else v).cpu().contiguous() for k, v in weights.items()}
    if max_shard_bytes is None:
        save_file(out, os.path.join(path, "model.safetensors"), metadata={"format": "mlx"})
        return
    shards, cur, cur_bytes = [], {}, 0
    for k in sorted(out):                       # a layer's weight / scales / biases stay together where they fit
        nb = out[k].numel() * out[k].element_size()
        if cur and cur_bytes + nb > max_shard_bytes:
            shards.append(cur)
            cur, cur_bytes = {}, 0
        cur[k] = out[k]
        cur_bytes += nb
    shards.append(cur)
    weight_map, total = {}, 0
    for i, sh in enumerate(shards):
        name = f"model-{i + 1:05d}-of-{len(shards):05d}.safetensors"
        save_file(sh, os.path.join(path, name), metadata={"format": "mlx"})
        for k, v in sh.items():
            weight_map[k] = name
            total += v.numel() * v.element_size()
    with open(os.path.join(path, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {"total_size": total}, "weight_map": weight_map}, f, indent=1)


def read_weights(path: str) -> dict:
    """Every ``model*.safetensors`` under ``path`` (PIE/models/utils.py:62-66).  When ``model.safetensors.index.json`` is
    present it is the manifest: a shard it names that is missing, or a tensor it maps that its shard does not hold, is an
    error here (the reference would fail later, at load_weights, with a missing-parameter message)."""
    from safetensors.torch import load_file

    files = sorted(glob.glob(os.path.join(path, "model*.safetensors")))
    if not files:
        raise FileNotFoundError(f"No safetensors found in {path}")
    weights = {}
    for wf in files:
        weights.update(load_file(wf))
    index = os.path.join(path, "model.safetensors.index.json")
    if os.path.exists(index):
        with open(index) as f:
            wm = json.load(f)["weight_map"]
        have = {os.path.basename(f) for f in files}
        missing = sorted(set(wm.values()) - have)
        if missing:
            raise FileNotFoundError(f"checkpoint index names shards that are not in {path}: {missing}")
        lost = sorted(k for k in wm if k not in weights)
        if lost:
            raise ValueError(f"checkpoint index maps tensors no shard holds: {lost[:4]}{' ...' if len(lost) > 4 else ''}")
    return weights


def load_model(path: str, device="cuda") -> Model:
    """PIE/models/utils.py:51-125 for the Llama family."""
    with open(os.path.join(path, "config.json")) as f:
        config = json.load(f)
    if config.get("model_type", "llama") != "llama":
        raise ValueError(f"Model type {config.get('model_type')} not supported.")
    if "quantization" not in config:
        raise ValueError("only MLX-quantized checkpoints (config['quantization']) are on this path")
    return build_model(ModelArgs.from_dict(config), read_weights(path), device)
