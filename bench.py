"""bench.py -- decode tok/s of the B200 quantized-Llama path (BASELINE.json configs[1]: Llama-3.1-8B, MLX q4 gs64,
bs1 decode) with the HBM roofline of its dominant kernel and the reference's CPU path timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--model llama-3.1-8b] [--bits 4] [--batch 1]

A step is one decode token for every sequence of the batch (one pass of the hot path).  `value` is device-timed
(CUDA events around K graph replays, token ids resident on the device); `e2e` drives the same step from HOST
buffers: token id copied in from pinned memory, sampled token copied back and synchronised, every step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--bits", type=int, default=4)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--prompt", type=int, default=128)
    ap.add_argument("--layers", type=int, default=0, help="debug only: override the layer count (marks the line invalid)")
    ap.add_argument("--paged", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true")
    ap.add_argument("--no-70b", action="store_true", help="N > 1: skip the Llama-3.1-70B tensor-parallel sub-records")
    ap.add_argument("--no-config3", action="store_true", help="skip the 8-bit / batch-32 / paged-KV sub-record (BASELINE configs[2])")
    ap.add_argument("--no-prefill", action="store_true", help="skip the prefill GEMM figure (BASELINE config 4 shapes)")
    return ap.parse_args()


def model_numbers(args, bits):
    """Algorithmic bytes per decode step: packed W + bf16 scales + bf16 biases, read once per step."""
    h, f, L, V = args.hidden_size, args.intermediate_size, args.num_hidden_layers, args.vocab_size
    nh, nkv = args.num_attention_heads, args.num_key_value_heads
    d = args.head_dim
    per_layer = h * (nh * d) + 2 * h * (nkv * d) + (nh * d) * h + 3 * h * f
    head = V * h
    params = L * per_layer + head
    bpp = bits / 8 + 4 / 64
    kv_per_tok = L * 2 * nkv * d * 2
    return {"params": params, "weight_bytes": params * bpp, "kv_bytes_per_ctx_token": kv_per_tok,
            "embed_row_bytes": h * bpp, "bpp": bpp}


# ------------------------------------------------------------------------------------------ CPU baseline
def host_threads():
    """Threads the CPU arm uses: the cores this process may run on.  NOT omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to its workers, which made round 1's N>1 reference arms 15-20x slower than its N=1 arm."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_baseline(margs, bits, steps=1, warmup=0, budget_s=20.0, config1=True):
    """The reference's own `_qmm_t<float, bits, 64>` (MLX/mlx/backend/cpu/quantized.cpp:96-153, compiled where it lies
    into oracle/_ref) on the decode shapes of this model, M = 1; falls back to the oracle port when oracle/_ref is
    absent.  Reported figures (SURVEY 8d / BASELINE.md section 4):
      value          all host cores (rows sliced over `cores` threads, each thread runs the untouched reference loop);
      single_thread  1 thread -- MLX's CPU stream is one worker thread (MLX/mlx/scheduler.h:22-35) and `_qmm_t` has no
                     intra-op threading: this is what the reference itself would do on this host;
      bf16_exact     one matrix through `_qmm_t<bfloat16_t>` (what MLX-CPU really executes for a bf16 checkpoint on x86:
                     every product and sum rounded to bf16 -- low precision, time only);
      config1        BASELINE config 1 (Llama-3.2-1B q4, 128-token prefill, greedy decode) through the WHOLE oracle
                     (attention, norms, rope, head), prefill and decode tok/s separately.
    Sample: `n_layers_sample` transformer layers' seven matmuls, tok/s extrapolated by parameter count."""
    import numpy as np
    from oracle import oracle as orc

    R = orc.ref()
    kind = "reference" if R is not None else "port"
    nthreads = host_threads()
    h, f = margs.hidden_size, margs.intermediate_size
    nh, nkv, d = margs.num_attention_heads, margs.num_key_value_heads, margs.head_dim
    shapes = [(nh * d, h), (nkv * d, h), (nkv * d, h), (h, nh * d), (f, h), (f, h), (h, f)]
    rng = np.random.default_rng(0)
    mats = []
    for n, k in shapes:
        wq = rng.integers(0, 2**32, (n, k * bits // 32), dtype=np.uint32)
        s = (rng.random((n, k // 64), dtype=np.float32) * 0.01).astype(np.float32)
        b = (rng.random((n, k // 64), dtype=np.float32) * -0.05).astype(np.float32)
        x = rng.standard_normal((1, k)).astype(np.float32)
        mats.append((n, k, wq, s, b, x, np.empty((1, n), np.float32)))
    layer_params = sum(n * k for n, k in shapes)
    total_params = model_numbers(margs, bits)["params"]

    def one_layer(nt, only=None):
        for i, (n, k, wq, s, b, x, y) in enumerate(mats):
            if only is not None and i != only:
                continue
            if R is not None:
                R.pie_ref_qmm_t_f32_mt(y.reshape(-1), x.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1),
                                       1, n, k, bits, 64, nt)
            else:
                orc.lib().orc_set_threads(nt)
                orc.lib().orc_qmm_t(y.reshape(-1), x.reshape(-1), wq.reshape(-1), s.reshape(-1), b.reshape(-1),
                                    1, n, k, bits, 64, 0)

    t0 = time.perf_counter(); one_layer(nthreads); t_probe = time.perf_counter() - t0
    n_layers_sample = max(1, min(margs.num_hidden_layers, int(budget_s / max(t_probe, 1e-3) / max(steps + warmup, 1))))
    for _ in range(warmup):
        for _ in range(n_layers_sample):
            one_layer(nthreads)
    times = []
    for _ in range(max(steps, 1)):
        t0 = time.perf_counter()
        for _ in range(n_layers_sample):
            one_layer(nthreads)
        times.append(time.perf_counter() - t0)
    t_step = sum(times) / len(times)
    scale = total_params / (layer_params * n_layers_sample)
    out = {"value": 1.0 / (t_step * scale), "unit": "tok/s", "cores": int(nthreads), "kind": kind,
           "sample": f"{n_layers_sample} transformer layer(s) x 7 quantized matmuls at M=1 (fp32 _qmm_t, "
                     f"{layer_params * n_layers_sample / 1e6:.0f} M params of {total_params / 1e6:.0f} M), "
                     f"extrapolated by parameter count; {t_step:.3f} s per sample; threads pinned to the process's "
                     f"CPU affinity ({nthreads}), OMP_NUM_THREADS ignored",
           "ms_per_step": t_step * 1e3 * scale}
    # the reference's own threading: one worker thread (one layer's q_proj + gate_proj, scaled by parameter count)
    t0 = time.perf_counter(); one_layer(1, only=0); one_layer(1, only=4); t1 = time.perf_counter() - t0
    p1 = shapes[0][0] * shapes[0][1] + shapes[4][0] * shapes[4][1]
    out["single_thread"] = {"value": 1.0 / (t1 * total_params / p1), "unit": "tok/s", "cores": 1,
                            "sample": f"q_proj + gate_proj of one layer ({p1 / 1e6:.0f} M params), {t1:.3f} s"}
    if R is not None:  # reference-exact bf16 arithmetic of MLX-CPU on x86 (low precision; time only)
        n, k, wq = shapes[0][0], shapes[0][1], mats[0][2]
        s16 = orc.to_bf16_bits(mats[0][3]); b16 = orc.to_bf16_bits(mats[0][4]); x16 = orc.to_bf16_bits(mats[0][5])
        y16 = np.empty((1, n), np.uint16)
        t0 = time.perf_counter()
        R.pie_ref_qmm_t_bf16_mt(y16.reshape(-1), x16.reshape(-1), wq.reshape(-1), s16.reshape(-1), b16.reshape(-1),
                                1, n, k, bits, 64, 1)
        tb = time.perf_counter() - t0
        t0 = time.perf_counter(); one_layer(1, only=0); tf = time.perf_counter() - t0
        out["bf16_exact"] = {"value": 1.0 / (tb * total_params / (n * k)), "unit": "tok/s", "cores": 1,
                             "slower_than_f32": round(tb / tf, 2), "precision": "low (bf16 accumulate), time only",
                             "sample": f"q_proj {n}x{k}, _qmm_t<bfloat16_t,{bits},64>, {tb:.3f} s"}
    if config1:
        try:
            out["config1"] = cpu_config1(nthreads)
        except Exception as e:  # the whole-oracle leg must never cost the bench its line
            out["config1"] = {"error": repr(e)[:200]}
    return out


def cpu_config1(nthreads, prompt_len=128, decode_steps=3):
    """BASELINE config 1 through the whole CPU oracle (fp32 mode A = MLX-CPU on fp32 inputs): Llama-3.2-1B q4 gs64,
    bs1, 128-token prompt in ONE forward (inference_engine.py:278), then greedy decode steps.  Random packed weights
    (the path is data-independent).  Bounded: the prefill plus `decode_steps` steps."""
    import numpy as np
    from oracle import oracle as orc
    from proxy_inference_engine_b200.checkpoint import LLAMA_PRESETS
    cfg = dict(LLAMA_PRESETS["llama-3.2-1b"], rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
               quantization={"group_size": 64, "bits": 4})
    oargs = orc.ModelArgs(**{k: v for k, v in cfg.items() if k in orc.ModelArgs.__dataclass_fields__})
    rng = np.random.default_rng(1)
    h, f, V = oargs.hidden_size, oargs.intermediate_size, oargs.vocab_size
    nh, nkv, d = oargs.num_attention_heads, oargs.num_key_value_heads, oargs.head_dim

    def lin(n, k):
        return (rng.integers(0, 2**32, (n, k // 8), dtype=np.uint32),
                (rng.random((n, k // 64), dtype=np.float32) * 0.02 / np.sqrt(k)).astype(np.float32),
                (rng.random((n, k // 64), dtype=np.float32) * -0.15 / np.sqrt(k)).astype(np.float32))

    w = {}
    def put(name, n, k):
        w[name + ".weight"], w[name + ".scales"], w[name + ".biases"] = lin(n, k)
    put("model.embed_tokens", V, h)
    for i in range(oargs.num_hidden_layers):
        p = f"model.layers.{i}"
        put(p + ".self_attn.q_proj", nh * d, h); put(p + ".self_attn.k_proj", nkv * d, h)
        put(p + ".self_attn.v_proj", nkv * d, h); put(p + ".self_attn.o_proj", h, nh * d)
        put(p + ".mlp.gate_proj", f, h); put(p + ".mlp.up_proj", f, h); put(p + ".mlp.down_proj", h, f)
        w[p + ".input_layernorm.weight"] = np.ones(h, np.float32)
        w[p + ".post_attention_layernorm.weight"] = np.ones(h, np.float32)
    w["model.norm.weight"] = np.ones(h, np.float32)
    orc.lib().orc_set_threads(nthreads)
    o = orc.LlamaOracle(oargs, w, mode="A")
    cache = o.make_cache()
    ids = rng.integers(0, V, (1, prompt_len), dtype=np.int32)
    t0 = time.perf_counter(); lg = o(ids, cache, last_only=True); t_pre = time.perf_counter() - t0
    tok = int(lg[0, -1].argmax())
    t0 = time.perf_counter()
    for _ in range(decode_steps):
        lg = o(np.array([[tok]], np.int32), cache); tok = int(lg[0, -1].argmax())
    t_dec = (time.perf_counter() - t0) / decode_steps
    return {"workload": "llama-3.2-1b q4 gs64 bs1, whole oracle (fp32 inside every op), prompt 128",
            "prefill_tok_s": round(prompt_len / t_pre, 2), "decode_tok_s": round(1.0 / t_dec, 3), "cores": int(nthreads),
            "sample": f"one 128-token prefill ({t_pre:.2f} s) + {decode_steps} decode steps ({t_dec:.3f} s each)"}


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def time_qmv_launches(eng, n_steps=3):
    """Average device duration of the dominant kernel (qmv_kernel): an eager pass (no graph, no PDL overlap) with a
    CUDA-event pair around every pie_qmv_fused launch, on the launching stream."""
    import torch
    saved = (eng.tokens.clone(), eng.positions.clone(), eng.seq.clone())
    pairs = []
    orig = eng._qmv
    pdl0 = eng.pdl
    eng.pdl = 0

    def timed(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); orig(*a, **k); e1.record()
        pairs.append((e0, e1))

    eng._qmv = timed
    try:
        for _ in range(n_steps):
            eng._step_kernels()
        torch.cuda.synchronize()
    finally:
        eng._qmv, eng.pdl = orig, pdl0
        eng._restore(saved)
    per_step = len(pairs) // n_steps
    ms = sum(e0.elapsed_time(e1) for e0, e1 in pairs[per_step:]) / max(n_steps - 1, 1)  # skip the first (cold) step
    return ms * 1e-3, per_step


def time_qmv_in_graph(eng, replays=20):
    """The same launches as they run inside the step: one CUDA graph holding only the step's quantized matmuls (with
    programmatic dependent launch, back to back), replayed `replays` times between two CUDA events on the launching
    stream.  Returns (seconds per graph = per step's worth of GEMV launches, launches)."""
    import torch
    saved = (eng.tokens.clone(), eng.positions.clone(), eng.seq.clone(), eng.h.clone())
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        n = eng.enqueue_gemvs()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        eng.enqueue_gemvs()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    eng.h.copy_(saved[3])
    eng._restore(saved[:3])
    return e0.elapsed_time(e1) * 1e-3 / replays, n


def time_exchanges_in_graph(eng, replays=20):
    """Tensor parallel only: a CUDA graph holding exactly one step's exchange steps (2 all-reduces per layer + the
    gather of the logits), replayed between two CUDA events; every rank takes part.  Returns (seconds per step's
    exchanges, count)."""
    import torch
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        n = eng.enqueue_exchanges()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        eng.enqueue_exchanges()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / replays, n


def tp_subrecord(model_name, bits, batches, rank, world, prompt=128, steps=32):
    """north_star's tensor-parallel curve: Llama-3.1-70B q4 at bs1 and bs16 on this many GPUs (BASELINE config 5).
    One sharded random-init model (every rank draws the full tensors with the same seed and keeps its slice), one
    engine per batch size; device-timed like the headline value (barrier, CUDA events, max over ranks)."""
    import torch
    import torch.distributed as dist
    from proxy_inference_engine_b200 import tp as T
    from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
    from proxy_inference_engine_b200.decode import DecodeEngine
    margs = preset_args(model_name, bits=bits)
    weights = random_weights(margs, seed=0, device="cuda", shard=(rank, world))
    model = build_model(T.shard_args(margs, world), weights).set_tp(T.TPInfo(rank, world))
    del weights
    nums = model_numbers(margs, bits)
    peak, _ = measured_peaks()
    out = {}
    for B in batches:
        eng = DecodeEngine(model, batch=B, max_context=prompt + steps + 72)
        ids = torch.randint(0, margs.vocab_size, (B, prompt), generator=torch.Generator().manual_seed(1))
        eng.prefill(ids)
        eng.step(4)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng.step(steps); e1.record()
        torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0]) / steps
        xs, xn = time_exchanges_in_graph(eng)
        step_bytes = nums["weight_bytes"] + B * ((prompt + 4 + steps / 2) * nums["kv_bytes_per_ctx_token"])
        out[f"bs{B}"] = {"tok_s": round(B / (ms * 1e-3), 2), "ms_per_step": round(ms, 4), "steps": steps,
                         "hbm_frac_per_gpu": round(step_bytes / world / (ms * 1e-3) / 1e9 / peak, 4),
                         "exchange_us_per_step": round(xs * 1e6, 1), "exchanges_per_step": xn,
                         "exchange": "one-shot NVLink push" if eng.xch is not None else "NCCL"}
        del eng
        torch.cuda.empty_cache()
    return {"workload": f"{model_name} random-init MLX q{bits} gs64, tp{world}, prompt {prompt}", **out}


def config3_subrecord(model_name, steps=24, warmup=4, batch=32, prompt=2048, kv_bits=None):
    """BASELINE configs[2]: 8-bit weights, batch 32, paged KV cache, 2k context on one GPU -- the other end of the metric's
    "bs 1/32".  Random-init model of its own (the headline model is 4-bit), device-timed like the headline value."""
    import torch
    from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
    from proxy_inference_engine_b200.decode import DecodeEngine
    margs = preset_args(model_name, bits=8)
    model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
    nums = model_numbers(margs, 8)
    peak, _ = measured_peaks()
    eng = DecodeEngine(model, batch=batch, max_context=prompt + steps + warmup + 40, paged=True, kv_bits=kv_bits)
    ids = torch.randint(0, margs.vocab_size, (batch, prompt), generator=torch.Generator().manual_seed(1))
    eng.prefill(ids)
    eng.step(warmup)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.step(steps); e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    kv_tok = nums["kv_bytes_per_ctx_token"]
    if kv_bits == 8:   # quants + bf16 scale and bias per 64-group instead of bf16 values (PIE/cache/kv_cache/quantized.py)
        kv_tok = kv_tok * (64 + 4) // 128
    step_bytes = nums["weight_bytes"] + batch * ((prompt + warmup + steps / 2) * kv_tok)
    kv_name = "paged KV" if kv_bits is None else f"paged {kv_bits}-bit gs64 KV (QuantizedKVCache)"
    out = {"workload": f"{model_name} random-init MLX q8 gs64, bs{batch} greedy decode, {kv_name}, prompt {prompt}",
           "tok_s": round(batch / (ms * 1e-3), 1), "ms_per_step": round(ms, 3), "steps": steps, "warmup": warmup,
           "bytes_per_step": int(step_bytes), "hbm_frac": round(step_bytes / (ms * 1e-3) / 1e9 / peak, 4),
           "matmul_kernel": "qmi_kernel (tcgen05.mma kind::i8 over the packed 8-bit weights)"}
    del eng, model
    torch.cuda.empty_cache()
    return out


def ncu_traffic_per_launch(alg_bytes_per_launch, bits):
    """dram__bytes_read.sum + dram__bytes_write.sum of the GEMV per launch, from the committed `ncu --set full` capture
    (profiles/r01e_prof_qmv_ncu_full_summary.csv: the q4 gate+up, down, qkv and o launches of one 8B layer).  The capture
    holds four of the five shapes, so the measured traffic / algorithmic ratio of those launches scales this run's
    average launch.  None if the capture is absent or was taken at another bit width."""
    import csv
    pdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles")
    path = os.path.join(pdir, "r02c_prof_qmv_ncu_full_summary.csv")   # this round's capture (both GEMV kernels)
    if not os.path.exists(path):
        path = os.path.join(pdir, "r01e_prof_qmv_ncu_full_summary.csv")
    if bits != 4 or not os.path.exists(path):
        return None, "no ncu capture for this configuration"
    rows = {r[0]: r for r in csv.reader(open(path)) if r}
    unit = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
    try:
        rd = sum(float(x) for x in rows["dram__bytes_read.sum"][2:6]) * unit[rows["dram__bytes_read.sum"][1]]
        wr = sum(float(x) for x in rows["dram__bytes_write.sum"][2:6]) * unit[rows["dram__bytes_write.sum"][1]]
    except (KeyError, ValueError):
        return None, "ncu summary not readable"
    alg = 66.10e6 + 33.07e6 + 14.18e6 + 9.45e6  # weights + scales + biases of gate+up, down, qkv, o
    ratio = (rd + wr) / alg
    return int(alg_bytes_per_launch * ratio), (f"{os.path.basename(path)}: committed ncu --set full capture of the step's GEMV "
                                                f"launches (traffic / algorithmic = {ratio:.4f}); NOT measured in this run")


def run_b200(a):
    import torch
    import torch.distributed as dist

    from proxy_inference_engine_b200 import ops
    from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
    from proxy_inference_engine_b200.decode import DecodeEngine

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    over = {"num_hidden_layers": a.layers} if a.layers else {}
    margs = preset_args(a.model, bits=a.bits, **over)
    torch.manual_seed(0)
    if world > 1:
        # tensor parallel over the box (tp.py): same model, sharded by heads / ffn columns / vocabulary rows
        from proxy_inference_engine_b200 import tp as T
        weights = random_weights(margs, seed=0, device="cuda", shard=(rank, world))
        model = build_model(T.shard_args(margs, world), weights).set_tp(T.TPInfo(rank, world))
    else:
        weights = random_weights(margs, seed=0, device="cuda")
        model = build_model(margs, weights)
    del weights
    B = a.batch
    max_ctx = a.prompt + a.warmup + a.steps * 2 + 64
    eng = DecodeEngine(model, batch=B, max_context=max_ctx, paged=a.paged)
    prompt = torch.randint(0, margs.vocab_size, (B, a.prompt), generator=torch.Generator().manual_seed(1))
    eng.prefill(prompt)
    eng.step(max(a.warmup, 3))
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed region: K graph replays, token ids resident --------------------------------------------------
    ctx0 = int(eng.positions[0].item())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    e0.record()
    eng.step(a.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # ---- end-to-end region: host token in, host token out, every step ------------------------------------------------
    pin_in = torch.zeros((B,), dtype=torch.int32).pin_memory()
    pin_out = torch.zeros((B,), dtype=torch.int32).pin_memory()
    pin_in.copy_(eng.tokens.cpu())
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        eng.tokens.copy_(pin_in, non_blocking=True)      # H2D: this step's input token ids
        eng.step(1)
        pin_out.copy_(eng.tokens, non_blocking=True)     # D2H: the sampled token ids
        torch.cuda.current_stream().synchronize()
        pin_in.copy_(pin_out)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    # every rank takes part: the eager pass contains the tensor-parallel collectives
    qsec, nlaunch = (0.0, 0) if a.no_kernel_timing else time_qmv_launches(eng)
    gsec, gl = (0.0, 0) if a.no_kernel_timing else time_qmv_in_graph(eng)
    xsec, xn, oneshot = 0.0, 0, False
    sub70 = None
    if world > 1:
        xsec, xn = time_exchanges_in_graph(eng)
        oneshot = eng.xch is not None
        if not a.no_70b and a.model == "llama-3.1-8b":
            launches_per_step = eng.launches_per_step
            del eng, model
            torch.cuda.empty_cache()
            try:
                sub70 = tp_subrecord("llama-3.1-70b", 4, (1, 16), rank, world)
            except Exception as e:  # the sub-record must never cost the headline its line
                sub70 = {"error": repr(e)[:300]}

            class _E:  # what the line below still needs of the engine
                pass
            eng = _E(); eng.launches_per_step = launches_per_step
    if rank != 0:
        _finish(world)
        return
    nums = model_numbers(margs, a.bits)
    ctx_mid = ctx0 + a.steps / 2
    step_bytes = nums["weight_bytes"] + B * (ctx_mid * nums["kv_bytes_per_ctx_token"] + nums["embed_row_bytes"])
    peak, peak_src = measured_peaks()
    value = B * a.steps / (ms * 1e-3)   # one model instance, whatever the GPU count (tensor parallel)
    line = {
        "metric": "decode tok/s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": round(ms / a.steps, 5), "higher_is_better": True,
        "scaling": "weak" if world == 1 else "strong", "vs_baseline": None,
        "dtype": "bf16 activations / fp32 accumulate over u4 weights"
        if a.bits == 4 else "bf16 activations / fp32 accumulate over u8 weights", "data": "synthetic",
        "config": {"workload": f"{a.model} random-init MLX q{a.bits} gs64, bs{B} greedy decode, prompt {a.prompt}, "
                               f"context {ctx0}..{ctx0 + a.steps}" + (", paged KV" if a.paged else ""),
                   "parallelism": "single GPU" if world == 1 else
                   f"tp{world}: heads / ffn / vocab sharded, 2 all-reduces per layer + 1 all-gather of logits, "
                   + ("one-shot pushes over NVLink peer memory (pie_comm.cu)" if oneshot else "NCCL"),
                   "l2": "each step streams the whole weight set (>> 126 MB L2), no flush needed",
                   "layers": margs.num_hidden_layers},
        "e2e": {"value": round(B * a.steps / (e2e_ms * 1e-3), 2), "unit": "tok/s",
                "h2d_bytes_per_step": 4 * B, "d2h_bytes_per_step": 4 * B},
        "gpu_launches": eng.launches_per_step * a.steps,
        "clocks": clocks,
        "step_roofline": {"bytes_per_step_per_gpu": int(step_bytes / world),
                          "achieved_GBps_per_gpu": round(step_bytes / world / (ms / a.steps) / 1e6, 1),
                          "frac": round(step_bytes / world / (ms / a.steps) / 1e6 / peak, 4)},
    }
    if world > 1:
        h = margs.hidden_size
        sent = (world - 1) * (2 * margs.num_hidden_layers * B * h * 2 + B * margs.vocab_size // world * 2)
        line["tp_exchange"] = {"kind": "one-shot NVLink push" if oneshot else "NCCL", "exchanges_per_step": xn,
                               "allreduce_us_per_step": round(xsec * 1e6, 1), "us_per_exchange": round(xsec * 1e6 / max(xn, 1), 2),
                               "nvlink_bytes_sent_per_step_per_gpu": int(sent),
                               "how": "CUDA graph of one step's exchanges alone, 20 replays between CUDA events"}
        if sub70 is not None:
            line["llama_3_1_70b_q4"] = sub70
    if a.layers:
        line["invalid"] = "layer count overridden (debug run)"
    if not a.no_kernel_timing:
        qbytes = nums["weight_bytes"] / world  # the qmv launches of a step together read this GPU's weights once
        ach = qbytes / gsec / 1e9      # in the step's own conditions: graph replay, dependent launch, back to back
        ach_eager = qbytes / qsec / 1e9
        traffic, traffic_src = ncu_traffic_per_launch(qbytes / nlaunch, a.bits)
        line["roofline"] = {"bound": "hbm", "kernel": ("qmv2_kernel (TMA-fed GEMV)" if a.bits == 4 and a.batch <= 8 else f"qmv_kernel<{a.bits}> / qmd_kernel"), "achieved": round(ach, 1), "peak": peak,
                            "peak_source": peak_src, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": traffic,
                            "traffic_source": traffic_src,
                            "launches_per_step": gl, "avg_launch_us": round(gsec / gl * 1e6, 2),
                            "algorithmic_bytes_per_launch": int(qbytes / gl),
                            "how": "CUDA events around 20 replays of a CUDA graph holding exactly the step's qmv launches "
                                   "(programmatic dependent launch, as in the timed step)",
                            "eager": {"achieved": round(ach_eager, 1), "frac": round(ach_eager / peak, 4),
                                      "avg_launch_us": round(qsec / nlaunch * 1e6, 2),
                                      "how": "CUDA-event pair around each launch in an eager pass (no graph / PDL)"}}
    if not a.no_prefill and world == 1:
        line["prefill_qmm"] = prefill_figure(margs, a.bits)
    if not a.no_config3 and world == 1 and a.bits == 4 and a.batch == 1 and not a.layers:
        del eng
        torch.cuda.empty_cache()
        line["config3_q8_bs32_paged_ctx2k"] = config3_subrecord(a.model)
        line["config3_q8_bs32_paged_ctx2k_kv8"] = config3_subrecord(a.model, kv_bits=8)
    if not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(margs, a.bits)
    print(json.dumps(line), flush=True)
    _finish(world)


def prefill_figure(margs, bits, M=8192):
    """Tensor-pipe figure of the prefill contraction (BASELINE config 4: 8k-token prompt): one layer's seven quantized
    matmuls at M = 8192 through pie_qmm (tcgen05), useful flops 2*M*N*K, against the measured bf16 sustained peak."""
    import torch
    from proxy_inference_engine_b200 import ops
    h, f = margs.hidden_size, margs.intermediate_size
    nh, nkv, d = margs.num_attention_heads, margs.num_key_value_heads, margs.head_dim
    shapes = [(nh * d, h), (nkv * d, h), (nkv * d, h), (h, nh * d), (f, h), (f, h), (h, f)]
    mats, xs = [], {}
    for n, k in shapes:
        wq = torch.randint(-2**31, 2**31 - 1, (n, k * bits // 32), dtype=torch.int32, device="cuda")
        s = (torch.rand((n, k // 64), device="cuda") * 0.01).to(torch.bfloat16)
        b = (torch.rand((n, k // 64), device="cuda") * -0.05).to(torch.bfloat16)
        xs.setdefault(k, torch.randn((M, k), device="cuda").to(torch.bfloat16))
        mats.append((wq, s, b, k))

    def layer():
        for wq, s, b, k in mats:
            ops.quantized_matmul(xs[k], wq, s, b, True, 64, bits)

    for _ in range(3):
        layer()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        layer()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    flops = 2 * M * sum(n * k for n, k in shapes)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peak = float(json.load(fh)["bf16_tflops_sustained"])
        src = "measured sustained (MEASURED_PEAKS.json)"
    except Exception:
        peak, src = 1400.0, "fallback (B200_PROFILING.md)"
    tf = flops / ms / 1e9
    return {"workload": f"one transformer layer's 7 quantized matmuls, M = {M} (8k-token prefill), q{bits}",
            "kernel": f"qmm_tc_kernel<{bits}> (tcgen05.mma kind::f16, TMEM accumulators)", "ms_per_layer": round(ms, 3),
            "achieved": round(tf, 1), "unit": "TFLOP/s", "peak": peak, "peak_source": src, "frac": round(tf / peak, 4),
            "bound": "tensor"}


def _finish(world):
    """Multi-rank exit: NCCL teardown after CUDA-graph capture of collectives can block; leave without it."""
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        os._exit(0)


def run_reference(a):
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    from proxy_inference_engine_b200.checkpoint import preset_args
    margs = preset_args(a.model, bits=a.bits)
    steps, warm = min(a.steps, 4), min(a.warmup, 1)
    cb = cpu_baseline(margs, a.bits, steps=steps, warmup=warm, budget_s=60.0, config1=False)
    line = {"impl": "reference", "metric": "decode tok/s", "value": round(cb["value"], 4), "unit": "tok/s",
            "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": round(cb["ms_per_step"], 2),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{a.model} random-init MLX q{a.bits} gs64, bs{a.batch} greedy decode "
                                   "(reference CPU _qmm_t on the host cores; bounded sample)"},
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": round(cb["value"], 4), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
