"""Layer surface: the ``mlx.nn`` modules PIE's Llama model instantiates, with the same constructor arguments,
parameter names (``weight`` / ``scales`` / ``biases``) and ``__call__`` semantics, over torch CUDA tensors.

Reference: MLX/python/mlx/nn/layers/quantized.py:11-229, normalization.py:142-143 (MLX = src/pie_core/external/mlx).
"""
from __future__ import annotations

import math

import torch

from . import ops


class Module:
    """Minimal stand-in for mlx.nn.Module: attribute tree + flat parameter dict with dotted names."""

    def children(self):
        for k, v in vars(self).items():
            if isinstance(v, Module):
                yield k, v
            elif isinstance(v, list) and v and isinstance(v[0], Module):
                for i, m in enumerate(v):
                    yield f"{k}.{i}", m

    def parameters(self, prefix: str = "") -> dict:
        out = {}
        for k, v in vars(self).items():
            if isinstance(v, torch.Tensor) and not k.startswith("_"):
                out[prefix + k] = v
        for k, m in self.children():
            out.update(m.parameters(prefix + k + "."))
        return out

    def load_weights(self, weights: dict, strict: bool = True):
        """nn.Module.load_weights: assign by dotted name; strict -> every parameter must be provided."""
        mine = self.parameters()
        if strict:
            missing = sorted(set(mine) - set(weights))
            extra = sorted(set(weights) - set(mine))
            if extra:
                raise ValueError(f"Received parameters not in model: {' '.join(extra[:8])}.")
            if missing:
                raise ValueError(f"Missing parameters: {' '.join(missing[:8])}.")
        for name, value in weights.items():
            if name not in mine:
                continue
            if tuple(value.shape) != tuple(mine[name].shape):
                raise ValueError(f"Expected shape {tuple(mine[name].shape)} but received shape "
                                 f"{tuple(value.shape)} for parameter {name}")
            obj = self
            parts = name.split(".")
            for p in parts[:-1]:
                obj = getattr(obj, p) if not p.isdigit() else obj[int(p)]
            setattr(obj, parts[-1], value)
        return self


class QuantizedLinear(Module):
    """nn.QuantizedLinear(input_dims, output_dims, bias, group_size, bits) (quantized.py:135-229).
    weight: uint32 words (held as int32) [output_dims, input_dims * bits / 32]; scales, biases: bf16
    [output_dims, input_dims / group_size]."""

    def __init__(self, input_dims: int, output_dims: int, bias: bool = True, group_size: int = 64, bits: int = 4,
                 device="cuda"):
        self.group_size, self.bits = group_size, bits
        scale = math.sqrt(1.0 / input_dims)
        w = (torch.rand((output_dims, input_dims), device=device) * 2 - 1) * scale
        self.weight, self.scales, self.biases = ops.quantize(w.to(torch.bfloat16), group_size, bits)
        if bias:
            self.bias = torch.zeros((output_dims,), dtype=torch.bfloat16, device=device)

    def __call__(self, x):
        y = ops.quantized_matmul(x, self.weight, self.scales, self.biases, transpose=True,
                                 group_size=self.group_size, bits=self.bits)
        if "bias" in vars(self):
            y = y + self.bias
        return y

    @classmethod
    def from_linear(cls, weight: torch.Tensor, bias=None, group_size: int = 64, bits: int = 4):
        out_dims, in_dims = weight.shape
        ql = cls.__new__(cls)
        ql.group_size, ql.bits = group_size, bits
        ql.weight, ql.scales, ql.biases = ops.quantize(weight, group_size, bits)
        if bias is not None:
            ql.bias = bias
        return ql


class QuantizedEmbedding(Module):
    """nn.QuantizedEmbedding(num_embeddings, dims, group_size, bits) (quantized.py:60-132)."""

    def __init__(self, num_embeddings: int, dims: int, group_size: int = 64, bits: int = 4, device="cuda"):
        self.group_size, self.bits = group_size, bits
        self.num_embeddings, self.dims = num_embeddings, dims
        w = torch.randn((num_embeddings, dims), device=device) * math.sqrt(1.0 / dims)
        self.weight, self.scales, self.biases = ops.quantize(w.to(torch.bfloat16), group_size, bits)

    def __call__(self, x):
        return ops.dequantize(self.weight, self.scales, self.biases, self.group_size, self.bits, ids=x)

    def as_linear(self, x):
        """Tied output head (quantized.py:109-124)."""
        return ops.quantized_matmul(x, self.weight, self.scales, self.biases, transpose=True,
                                    group_size=self.group_size, bits=self.bits)


class RMSNorm(Module):
    """nn.RMSNorm(dims, eps) (normalization.py:105-143)."""

    def __init__(self, dims: int, eps: float = 1e-5, device="cuda"):
        self.weight = torch.ones((dims,), dtype=torch.bfloat16, device=device)
        self.eps = eps

    def __call__(self, x):
        return ops.rms_norm(x, self.weight, self.eps)
