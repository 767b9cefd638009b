"""On-disk format of the path: MLX-layout safetensors, single file or sharded with
model.safetensors.index.json, read back the way PIE's loader globs them (PIE/models/utils.py:62-66)."""
import json
import os

import pytest
import torch

from parity_util import tiny_args
from proxy_inference_engine_b200.checkpoint import read_weights, save_checkpoint


def _weights():
    g = torch.Generator().manual_seed(0)
    w = {}
    for i in range(3):
        p = f"model.layers.{i}.mlp.down_proj"
        w[p + ".weight"] = torch.randint(-2**31, 2**31 - 1, (64, 32), generator=g, dtype=torch.int64).to(torch.int32)
        w[p + ".scales"] = torch.randn((64, 4), generator=g).to(torch.bfloat16)
        w[p + ".biases"] = torch.randn((64, 4), generator=g).to(torch.bfloat16)
    w["model.norm.weight"] = torch.randn((256,), generator=g).to(torch.bfloat16)
    return w


def _same(a, b):
    assert a.keys() == b.keys()
    for k in a:
        got = b[k].view(torch.int32) if b[k].dtype == torch.uint32 else b[k]
        assert got.dtype == a[k].dtype and torch.equal(got, a[k]), k


def test_single_file_roundtrip(tmp_path):
    w = _weights()
    save_checkpoint(str(tmp_path), tiny_args(), w)
    assert sorted(os.listdir(tmp_path)) == ["config.json", "model.safetensors"]
    _same(w, read_weights(str(tmp_path)))
    from safetensors import safe_open
    with safe_open(os.path.join(tmp_path, "model.safetensors"), "pt") as f:
        assert f.metadata() == {"format": "mlx"}
        assert f.get_slice("model.layers.0.mlp.down_proj.weight").get_dtype() == "U32"   # MLX's packed words


def test_sharded_roundtrip_and_index(tmp_path):
    w = _weights()
    save_checkpoint(str(tmp_path), tiny_args(), w, max_shard_bytes=10000)
    idx = json.load(open(os.path.join(tmp_path, "model.safetensors.index.json")))
    shards = sorted(set(idx["weight_map"].values()))
    assert len(shards) > 1 and shards[0] == f"model-00001-of-{len(shards):05d}.safetensors"
    assert set(idx["weight_map"]) == set(w)
    assert idx["metadata"]["total_size"] == sum(v.numel() * v.element_size() for v in w.values())
    for s in shards:
        assert os.path.getsize(os.path.join(tmp_path, s)) < 10000 + 4096
    _same(w, read_weights(str(tmp_path)))


def test_missing_shard_and_stale_index_are_errors(tmp_path):
    w = _weights()
    save_checkpoint(str(tmp_path), tiny_args(), w, max_shard_bytes=10000)
    idx_path = os.path.join(tmp_path, "model.safetensors.index.json")
    idx = json.load(open(idx_path))
    idx["weight_map"]["model.layers.9.extra.weight"] = sorted(set(idx["weight_map"].values()))[0]
    json.dump(idx, open(idx_path, "w"))
    with pytest.raises(ValueError, match="no shard holds"):
        read_weights(str(tmp_path))
    os.remove(os.path.join(tmp_path, sorted(set(idx["weight_map"].values()))[-1]))
    with pytest.raises(FileNotFoundError, match="not in"):
        read_weights(str(tmp_path))
    with pytest.raises(FileNotFoundError):
        read_weights(str(tmp_path / "nowhere"))
