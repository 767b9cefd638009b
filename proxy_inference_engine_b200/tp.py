"""Tensor parallelism of the quantized Llama path over one NVSwitch box.

The partition is MLX's own (``QuantizedAllToShardedLinear`` / ``QuantizedShardedToAllLinear``,
MLX/python/mlx/nn/layers/distributed.py:355-566): q/k/v_proj and gate/up_proj are split by output rows (whole
heads / ffn columns), o_proj and down_proj by input columns in multiples of the quantization group, so that no
(scale, bias) group is ever cut.  Each rank then owns Hq/W query heads, Hkv/W KV heads (its slice of the KV cache) and
ffn/W MLP columns; attention is fully local.  One exchange step per sub-block: an all-reduce (sum, bf16, [B, hidden])
after o_proj and after down_proj.  The lm_head is split by vocabulary rows and the logits are all-gathered.
Embedding, norms and the residual stream are replicated.

This module is pure index logic + ``torch.distributed`` calls: it runs on CPU tensors with the gloo backend (the
world_size-2 tests) exactly as on CUDA tensors with NCCL.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass

import torch
import torch.distributed as dist

ROW_SPLIT = ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "mlp.gate_proj", "mlp.up_proj", "lm_head")
COL_SPLIT = ("self_attn.o_proj", "mlp.down_proj")


@dataclass
class TPInfo:
    rank: int = 0
    world: int = 1
    group: object = None  # torch.distributed process group (None = default)

    @property
    def on(self) -> bool:
        return self.world > 1


def check_divisible(args, world: int) -> None:
    """The predicates MLX's shard_linear relies on, plus the fast-path granularity of the GEMV (one ring stage)."""
    nh = args.num_attention_heads
    nkv = args.num_key_value_heads or nh
    d = args.head_dim or args.hidden_size // nh
    gs, bits = args.quantization["group_size"], args.quantization["bits"]
    stage_k = 128 * 8 // bits  # elements of K per ring stage of the decode GEMV (csrc/pie_internal.h kRowBytes)
    if nh % world or nkv % world:
        raise ValueError(f"tensor parallel degree {world} must divide the head counts ({nh} query, {nkv} kv)")
    if args.intermediate_size % world or args.vocab_size % world:
        raise ValueError(f"tensor parallel degree {world} must divide intermediate_size and vocab_size")
    for k in (nh * d // world, args.intermediate_size // world):
        if k % gs or k % stage_k:
            raise ValueError(f"a K shard of {k} elements cuts a quantization group or a {stage_k}-element stage")


def shard_args(args, world: int):
    """ModelArgs of one rank: local head counts and ffn width; hidden size, vocabulary and head_dim unchanged."""
    check_divisible(args, world)
    nh = args.num_attention_heads
    nkv = args.num_key_value_heads or nh
    d = args.head_dim or args.hidden_size // nh
    return dataclasses.replace(args, num_attention_heads=nh // world, num_key_value_heads=nkv // world, head_dim=d,
                               intermediate_size=args.intermediate_size // world,
                               lm_head_rows=args.vocab_size // world)


def shard_tensor(name: str, t: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """The slice of checkpoint tensor ``name`` (MLX schema, PIE/models/utils.py:62-116) that rank ``rank`` owns."""
    if world == 1:
        return t
    base = name.rsplit(".", 1)[0]
    if any(base.endswith(s) for s in ROW_SPLIT):
        n = t.shape[0] // world
        return t[rank * n:(rank + 1) * n].contiguous()
    if any(base.endswith(s) for s in COL_SPLIT):
        # weight: K * bits / 32 words per row; scales / biases: K / group_size per row -- all proportional to K
        n = t.shape[1] // world
        return t[:, rank * n:(rank + 1) * n].contiguous()
    return t  # embed_tokens, norms: replicated


def shard_weights(weights: dict, args, rank: int, world: int) -> dict:
    check_divisible(args, world)
    return {k: shard_tensor(k, v, rank, world) for k, v in weights.items()}


def all_reduce_sum(x: torch.Tensor, tp: TPInfo) -> torch.Tensor:
    """The exchange step after o_proj / down_proj (in place)."""
    if tp is not None and tp.on:
        dist.all_reduce(x, op=dist.ReduceOp.SUM, group=tp.group)
    return x


def all_gather_last(x: torch.Tensor, tp: TPInfo) -> torch.Tensor:
    """Concatenate the ranks' shards along the last dimension (vocabulary shards of the logits)."""
    if tp is None or not tp.on:
        return x
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    out = torch.empty((tp.world * x2.shape[0], x2.shape[1]), dtype=x.dtype, device=x.device)  # rank-major rows
    dist.all_gather_into_tensor(out, x2, group=tp.group)
    out = out.view(tp.world, x2.shape[0], x2.shape[1]).permute(1, 0, 2)
    return out.reshape(*x.shape[:-1], tp.world * x.shape[-1])


class TPExchange:
    """One-shot exchanges over NVLink peer memory for the decode step (csrc/pie_comm.cu): every rank pushes its partial
    row into a slot of every peer's SYMMETRIC buffer and reduces the `world` slots locally -- one small kernel per
    exchange instead of an NCCL collective (8-16 KB messages are pure latency: ~11 us per ncclAllReduce at 2 GPUs,
    ~26 us at 8, measured in round 1).  NCCL / torch.distributed stay for the rendezvous and for the prefill path.
    The buffer comes from torch.distributed._symmetric_memory (CUDA IPC / fabric handles mapped into every peer)."""

    def __init__(self, tp: TPInfo, max_elems: int, gather_rows: int, gather_width: int, device):
        import ctypes as C

        import torch.distributed._symmetric_memory as symm

        from . import _lib
        self.L = _lib.load()
        self.tp = tp
        self.max_elems = int(max_elems)
        gbytes = gather_rows * gather_width * tp.world * 2
        nbytes = self.L.pie_tp_buffer_bytes(tp.world, self.max_elems, gbytes)
        group = tp.group if tp.group is not None else dist.group.WORLD
        self.buf = symm.empty(int(nbytes), dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group)
        self.epoch = torch.tensor([1, 0, 0, 0], dtype=torch.int32, device=device)  # number, departures (2 kinds)
        self.peers = _lib.TpPeers()
        for r, ptr in enumerate(self.hdl.buffer_ptrs):
            self.peers.base[r] = ptr
        self.peers.rank, self.peers.world, self.peers.max_elems = tp.rank, tp.world, self.max_elems
        self.peers.epoch = self.epoch.data_ptr()
        goff = self.L.pie_tp_gather_offset(tp.world, self.max_elems)
        self.gathered = self.buf[goff:goff + gbytes].view(torch.bfloat16).view(gather_rows, gather_width * tp.world)
        self._C = C
        torch.cuda.synchronize(device)
        dist.barrier(group=group)  # every rank's buffer is zeroed before anyone pushes into it

    def fused(self, mode: int, h_out=None):
        """pie_tp_fused descriptor for pie_qmv_fused_tp: mode 1 = producer (partial row pushed to every peer), mode 2 =
        consumer (x = residual + sum of the slots, new residual stream stored to h_out).  Keep the object alive."""
        from . import _lib
        f = _lib.TpFused()
        f.peers = self.peers
        f.mode = mode
        f.h_out = h_out.data_ptr() if h_out is not None else None
        return f

    def all_reduce(self, out, partial, residual, stream):
        """out = bf16(residual + bf16(sum over ranks of partial)); partial [rows, width] bf16."""
        from . import _lib
        rows, width = partial.shape
        _lib.check(self.L.pie_tp_allreduce(out.data_ptr(), partial.data_ptr(),
                                           residual.data_ptr() if residual is not None else None, rows, width,
                                           self._C.byref(self.peers), stream), "tp_allreduce")

    def all_gather(self, slice_, stream):
        """Every rank's [rows, width] slice -> self.gathered [rows, world * width] on every rank."""
        from . import _lib
        rows, width = slice_.shape
        _lib.check(self.L.pie_tp_allgather(slice_.data_ptr(), rows, width, self._C.byref(self.peers), stream),
                   "tp_allgather")
        return self.gathered
