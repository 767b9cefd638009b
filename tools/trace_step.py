"""Per-matrix timeline of the fused decode step (CTA 0's %globaltimer at phase boundaries).
   python tools/trace_step.py [--bits 4] [--batch 1] [--ctx 300] [--layers 0]"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
from proxy_inference_engine_b200.decode import DecodeEngine

ap = argparse.ArgumentParser()
ap.add_argument("--bits", type=int, default=4); ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--ctx", type=int, default=300); ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--layers", type=int, default=0)
a = ap.parse_args()
over = {"num_hidden_layers": a.layers} if a.layers else {}
margs = preset_args(a.model, bits=a.bits, **over)
model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
eng = DecodeEngine(model, batch=a.batch, max_context=a.ctx + 200, trace=True)
eng.positions.fill_(a.ctx); eng.tokens.fill_(5); eng._next_pos = a.ctx
eng.step(5); torch.cuda.synchronize()
t = eng.step_trace.cpu().numpy().astype("int64")
n = (len(t) - 4) // 2
ready, done = t[2:2 + 2 * n:2], t[3:3 + 2 * n:2]
names = ["qkv", "o", "gate_up", "down"]
agg = {}
prev = t[0]
for p in range(n):
    nm = names[p % 4] if p < n - 1 else "head"
    w, g = agg.setdefault(nm, [0, 0, 0])[:2]
    agg[nm][0] += int(ready[p] - prev); agg[nm][1] += int(done[p] - ready[p]); agg[nm][2] += 1
    prev = done[p]
out = {k: {"wait_us": round(v[0] / v[2] / 1e3, 2), "gemv_us": round(v[1] / v[2] / 1e3, 2), "n": v[2]} for k, v in agg.items()}
out["total_us"] = round(int(t[1] - t[0]) / 1e3, 1)
out["first_layers_us"] = [[round(int(ready[p] - (done[p - 1] if p else t[0])) / 1e3, 2), round(int(done[p] - ready[p]) / 1e3, 2)] for p in range(min(n, 12))]
print(json.dumps(out))
