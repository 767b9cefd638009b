"""Where does a decode step go?  Times the captured step of the 8B model with parts of it left out (debug only:
the outputs are then wrong).   python tools/bench_step_parts.py [--bits 4] [--batch 1] [--ctx 300]"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
from proxy_inference_engine_b200 import decode as dec

ap = argparse.ArgumentParser()
ap.add_argument("--bits", type=int, default=4); ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--ctx", type=int, default=300); ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--layers", type=int, default=0); ap.add_argument("--unfused-attn", action="store_true"); ap.add_argument("--fused-attn", action="store_true")
a = ap.parse_args()
over = {"num_hidden_layers": a.layers} if a.layers else {}
margs = preset_args(a.model, bits=a.bits, **over)
model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
res = {}
ap2 = os.environ.get("PIE_PARTS", "")
for skip in (tuple(ap2.split(";")) if ap2 else ("", "rope", "attn", "rope,attn", "rope,attn,o,qkv", "rope,attn,o,qkv,gu", "rope,attn,o,qkv,gu,down")):
    os.environ["PIE_DEBUG_SKIP"] = skip
    eng = dec.DecodeEngine(model, batch=a.batch, max_context=a.ctx + 200, attn_fused=False if a.unfused_attn else (True if a.fused_attn else None))
    eng.positions.fill_(a.ctx); eng.tokens.fill_(5); eng._next_pos = a.ctx
    eng.step(3); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.step(50); e1.record(); torch.cuda.synchronize()
    res[skip or "full"] = round(e0.elapsed_time(e1) / 50 * 1e3, 1)
    eng_fused = eng.attn_fused
    del eng
print(json.dumps({"us_per_step": res, "layers": margs.num_hidden_layers, "batch": a.batch, "ctx": a.ctx, "bits": a.bits,
                  "attn": "fused launch" if eng_fused else "rope_kv_append + decode_attn (tensor cores)"}))
