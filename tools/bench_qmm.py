"""Prefill quantized GEMM micro-benchmark (BASELINE config 4 shapes: Llama-3.1-8B, M = 8192): achieved TFLOP/s of
pie_qmm per matrix (CUDA events; 2*M*N*K useful flops, dequant ALU work not counted)."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200 import ops

SHAPES = {"q": (4096, 4096), "kv": (1024, 4096), "gate": (14336, 4096), "down": (4096, 14336)}
ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=8192); ap.add_argument("--bits", type=int, default=4)
ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()
res = {}
x_cache = {}
for name, (N, K) in SHAPES.items():
    wq = torch.randint(-2**31, 2**31 - 1, (N, K * a.bits // 32), dtype=torch.int32, device="cuda")
    s = (torch.rand((N, K // 64), device="cuda") * 0.01).to(torch.bfloat16)
    b = (torch.rand((N, K // 64), device="cuda") * -0.05).to(torch.bfloat16)
    x = x_cache.setdefault(K, torch.randn((a.M, K), device="cuda").to(torch.bfloat16))
    for _ in range(2):
        y = ops.quantized_matmul(x, wq, s, b, True, 64, a.bits)
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); y = ops.quantized_matmul(x, wq, s, b, True, 64, a.bits); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    res[name] = {"ms": round(ms, 3), "TFLOPs": round(2 * a.M * N * K / ms / 1e9, 1)}
layer_ms = res["q"]["ms"] * 2 + res["kv"]["ms"] * 2 + res["gate"]["ms"] * 2 + res["down"]["ms"]
layer_flops = 2 * a.M * (4096 * 4096 * 2 + 1024 * 4096 * 2 + 14336 * 4096 * 3)
print(json.dumps({"M": a.M, "bits": a.bits, "res": res, "layer_ms": round(layer_ms, 3),
                  "layer_TFLOPs": round(layer_flops / layer_ms / 1e9, 1), "body_32_layers_ms": round(32 * layer_ms, 1)}))
