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
eng = DecodeEngine(model, batch=a.batch, max_context=a.ctx + 200, trace=True, fused=True)
eng.positions.fill_(a.ctx); eng.tokens.fill_(5); eng._next_pos = a.ctx
eng.step(5); torch.cuda.synchronize()
t = eng.step_trace.cpu().numpy().astype("int64")
n = (len(t) - 4) // 7
# per matrix: [input ready (grid barrier passed), x staged, matrix done]; t[0] kernel start, t[1] all matrices done
ready, staged, done = t[2:2 + 3 * n:3], t[3:3 + 3 * n:3], t[4:4 + 3 * n:3]
names = ["qkv", "o", "gate_up", "down"]
agg = {}
prev = t[0]
for p in range(n):
    nm = names[p % 4] if p < n - 1 else "head"
    v = agg.setdefault(nm, [0, 0, 0, 0])
    v[0] += int(ready[p] - prev); v[1] += int(staged[p] - ready[p]); v[2] += int(done[p] - staged[p]); v[3] += 1
    prev = done[p]
out = {k: {"wait_us": round(v[0] / v[3] / 1e3, 2), "stage_us": round(v[1] / v[3] / 1e3, 2),
           "gemv_us": round(v[2] / v[3] / 1e3, 2), "n": v[3]} for k, v in agg.items()}
st4 = t[4 + 3 * n:4 + 7 * n].reshape(n, 4)
stage = {}
for p in range(n):
    nm = names[p % 4] if p < n - 1 else "head"
    if st4[p, 3] == 0:
        continue
    d = [int(st4[p, 0] - ready[p])] + [int(st4[p, i] - st4[p, i - 1]) for i in (1, 2, 3)] + [int(staged[p] - st4[p, 3])]
    v = stage.setdefault(nm, [0, 0, 0, 0, 0, 0])
    for i in range(5):
        v[i] += d[i]
    v[5] += 1
out["stage_parts_us"] = {k: [round(x / v[5] / 1e3, 2) for x in v[:5]] for k, v in stage.items()}
out["total_us"] = round(int(t[1] - t[0]) / 1e3, 1)
out["note"] = "wait = pre-phase (embedding / attention) + grid barrier(s) seen by CTA 0; o's wait includes the attention phase"
print(json.dumps(out))
