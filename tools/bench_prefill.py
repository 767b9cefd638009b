"""BASELINE config 4: Llama-3.1-8B q4, one 8192-token prefill (whole prompt in one forward, as
PIE/engine/inference_engine.py:278): time to first token and achieved TFLOP/s (2*M*N*K of the quantized matmuls,
lm_head on the last position only, + causal attention 4*L^2*h*layers/2)."""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
from proxy_inference_engine_b200.decode import DecodeEngine

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3.1-8b"); ap.add_argument("--bits", type=int, default=4)
ap.add_argument("--prompt", type=int, default=8192); ap.add_argument("--layers", type=int, default=0)
a = ap.parse_args()
over = {"num_hidden_layers": a.layers} if a.layers else {}
margs = preset_args(a.model, bits=a.bits, **over)
model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
eng = DecodeEngine(model, batch=1, max_context=a.prompt + 16)
prompt = torch.randint(0, margs.vocab_size, (1, a.prompt), generator=torch.Generator().manual_seed(1))
ts = []
for it in range(3):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.prefill(prompt); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = min(ts[1:])
h, f, L = margs.hidden_size, margs.intermediate_size, margs.num_hidden_layers
nh, nkv, d = margs.num_attention_heads, margs.num_key_value_heads, margs.head_dim
mm = 2 * a.prompt * L * (h * nh * d * 2 + 2 * h * nkv * d + 3 * h * f)
att = 4 * a.prompt * a.prompt * nh * d * L / 2
print(json.dumps({"config": f"{a.model} q{a.bits} prefill {a.prompt} tokens, bs1", "ttft_ms": round(ms, 1),
                  "all_ms": [round(t, 1) for t in ts], "matmul_TFLOP": round(mm / 1e12, 1), "attn_TFLOP": round(att / 1e12, 1),
                  "achieved_TFLOPs": round((mm + att) / ms / 1e9, 1), "layers": L}))
