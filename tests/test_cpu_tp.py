"""Tensor-parallel host logic on CPU (gloo, world_size 2): the partition of tp.py reproduces the unsharded layer
through the oracle's arithmetic, with the all-reduce / all-gather being the only exchange steps.

Mirrors MLX's own sharded-layer test idea (QuantizedAllToShardedLinear / QuantizedShardedToAllLinear,
MLX/python/mlx/nn/layers/distributed.py:355-566): shard, apply, exchange, compare with the dense result.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_checkpoint(args, seed=0):
    """Random MLX-layout checkpoint made with the ORACLE's quantizer (CPU only; the product quantizer needs a GPU)."""
    from oracle import oracle as orc
    rng = np.random.default_rng(seed)
    gs, bits = args.quantization["group_size"], args.quantization["bits"]
    h, f = args.hidden_size, args.intermediate_size
    nh, nkv, d = args.num_attention_heads, args.num_key_value_heads, args.head_dim
    w = {}

    def put(name, n, k):
        t = ((rng.random((n, k), dtype=np.float32) * 2 - 1) / np.sqrt(k)).astype(np.float32)
        wq, s, b = orc.quantize(orc.round_bf16(t), gs, bits)
        w[name + ".weight"] = torch.from_numpy(wq.view(np.int32))
        w[name + ".scales"] = torch.from_numpy(s)
        w[name + ".biases"] = torch.from_numpy(b)

    put("model.embed_tokens", args.vocab_size, h)
    p = "model.layers.0"
    put(p + ".self_attn.q_proj", nh * d, h); put(p + ".self_attn.k_proj", nkv * d, h)
    put(p + ".self_attn.v_proj", nkv * d, h); put(p + ".self_attn.o_proj", h, nh * d)
    put(p + ".mlp.gate_proj", f, h); put(p + ".mlp.up_proj", f, h); put(p + ".mlp.down_proj", h, f)
    put("lm_head", args.vocab_size, h)
    w["model.norm.weight"] = torch.ones(h)
    return w


def _deq(w, name, gs, bits):
    from oracle import oracle as orc
    return orc.dequantize(w[name + ".weight"].numpy().view(np.uint32), w[name + ".scales"].numpy().astype(np.float32),
                          w[name + ".biases"].numpy().astype(np.float32), gs, bits)


def _tiny_args():
    from proxy_inference_engine_b200.llama import ModelArgs
    return ModelArgs(hidden_size=512, num_hidden_layers=1, intermediate_size=1024, num_attention_heads=4,
                     num_key_value_heads=2, head_dim=128, vocab_size=64, rms_norm_eps=1e-5,
                     tie_word_embeddings=False, quantization={"group_size": 64, "bits": 4})


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from proxy_inference_engine_b200 import tp as T
        args = _tiny_args()
        gs, bits = 64, 4
        full = _fake_checkpoint(args)
        mine = T.shard_weights(full, args, rank, world)
        info = T.TPInfo(rank, world)
        rng = np.random.default_rng(5)
        x = rng.standard_normal((3, args.hidden_size)).astype(np.float32)
        p = "model.layers.0"
        # column-parallel (row split): q_proj shards concatenate to the dense output
        yq = torch.from_numpy(x @ _deq(mine, p + ".self_attn.q_proj", gs, bits).T)
        yq_all = T.all_gather_last(yq, info).numpy()
        ref = x @ _deq(full, p + ".self_attn.q_proj", gs, bits).T
        ok_col = np.allclose(yq_all, ref, rtol=1e-5, atol=1e-5)
        # row-parallel (column split): MLP down_proj on this rank's ffn slice, summed by the all-reduce
        g = x @ _deq(mine, p + ".mlp.gate_proj", gs, bits).T
        u = x @ _deq(mine, p + ".mlp.up_proj", gs, bits).T
        act = (g / (1 + np.exp(-g))) * u
        part = torch.from_numpy((act @ _deq(mine, p + ".mlp.down_proj", gs, bits).T).astype(np.float32))
        T.all_reduce_sum(part, info)
        gf = x @ _deq(full, p + ".mlp.gate_proj", gs, bits).T
        uf = x @ _deq(full, p + ".mlp.up_proj", gs, bits).T
        ref = ((gf / (1 + np.exp(-gf))) * uf) @ _deq(full, p + ".mlp.down_proj", gs, bits).T
        ok_row = np.allclose(part.numpy(), ref, rtol=1e-4, atol=1e-4)
        # vocabulary split of the head + replicated tensors
        lg = T.all_gather_last(torch.from_numpy(x @ _deq(mine, "lm_head", gs, bits).T), info).numpy()
        ok_head = np.allclose(lg, x @ _deq(full, "lm_head", gs, bits).T, rtol=1e-5, atol=1e-5)
        ok_rep = torch.equal(mine["model.embed_tokens.weight"], full["model.embed_tokens.weight"])
        sa = T.shard_args(args, world)
        ok_args = (sa.num_attention_heads, sa.num_key_value_heads, sa.intermediate_size, sa.lm_head_rows,
                   sa.hidden_size, sa.vocab_size) == (2, 1, 512, 32, 512, 64)
        q.put((rank, bool(ok_col), bool(ok_row), bool(ok_head), bool(ok_rep), bool(ok_args)))
    except Exception as e:  # surface the failure instead of leaving the parent waiting on the queue
        q.put((rank, False, repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def test_tp_partition_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in res:
        assert all(r[1:]), r


def test_tp_divisibility_errors():
    from proxy_inference_engine_b200 import tp as T
    args = _tiny_args()
    with pytest.raises(ValueError):
        T.shard_args(args, 3)          # does not divide the head counts
    with pytest.raises(ValueError):
        T.shard_args(args, 4)          # 2 kv heads cannot go to 4 ranks
    assert T.shard_args(args, 1).num_attention_heads == 4
    t = torch.arange(24).reshape(4, 6)
    assert torch.equal(T.shard_tensor("model.layers.0.self_attn.q_proj.weight", t, 1, 2), t[2:])
    assert torch.equal(T.shard_tensor("model.layers.0.mlp.down_proj.scales", t, 1, 2), t[:, 3:])
    assert T.shard_tensor("model.norm.weight", t, 1, 2) is t
