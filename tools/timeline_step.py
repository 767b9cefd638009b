"""Timeline of the per-operator decode step from inside the kernels: every TMA GEMV / fused-attention launch of the
captured graph stamps %globaltimer per CTA (pie_debug_stamps).  Dumps the stamps (--dump) and prints, per kind of launch
(averaged over layers), the median time of each stamp relative to the moment the PREVIOUS launch's last CTA exited.
   python tools/timeline_step.py [--bits 4] [--ctx 300] [--layers 8] [--dump file.npy]
Stamps: GEMV 0 entry, 8 set-up done, 1 first loads issued, 2 dependency wait returned, 9 x loaded + reduced, 10 CTA barrier
of the staging passed, 11 scale known, 3 x staged, 4 first weights landed, 5 main loop done, 6 CTA barrier, 7 exit.
Attention: 0 entry, 1 position known + first K/V requested, 3 sines ready, 2 wait returned, 4 q rotated, 5 context loop done,
6 partial pushed, 7 exit."""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200 import _lib
from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
from proxy_inference_engine_b200.decode import DecodeEngine

COLS = [0, 8, 1, 2, 9, 10, 11, 3, 4, 5, 7]
NAMES = ["qkv", "attn", "o", "gate_up", "down"]


def summarise(t):
    """t: [launches, 256, 16] stamps of one step (launch order)."""
    acc = {}
    for j in range(5, len(t) - 1):   # skip the first layer (follows the embedding) and the head
        s = t[j]
        s = s[: int((s[:, 0] != 0).sum())]
        prev_end = t[j - 1][:, 7].max()
        row = []
        for c in COLS:
            v = s[:, c]; v = v[v != 0]
            row.append(float("nan") if len(v) == 0 else (np.median(v) - prev_end) / 1e3)
        row.append((s[:, 7].max() - prev_end) / 1e3)
        acc.setdefault(NAMES[j % 5], []).append(row)
    out, tot = {}, 0.0
    for nm in NAMES:
        with np.errstate(all="ignore"):
            m = np.nanmean(np.array(acc[nm]), axis=0)
        out[nm] = {("s%d" % c): round(float(v), 2) for c, v in zip(COLS, m[:-1])}
        out[nm]["last_exit"] = round(float(m[-1]), 2)
        tot += m[-1]
    return out, tot


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bits", type=int, default=4); ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=300); ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--layers", type=int, default=8); ap.add_argument("--dump", default="")
    a = ap.parse_args()
    margs = preset_args(a.model, bits=a.bits, num_hidden_layers=a.layers)
    model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
    eng = DecodeEngine(model, batch=a.batch, max_context=a.ctx + 200)
    per = 5 * a.layers + 1
    nslots = 4 * per
    buf = torch.zeros((nslots, 256, 16), dtype=torch.int64, device="cuda")
    _lib.check(_lib.load().pie_debug_stamps(buf.data_ptr(), buf.numel() * 8), "stamps")
    eng.positions.fill_(a.ctx); eng.tokens.fill_(5); eng._next_pos = a.ctx
    eng.step(3); torch.cuda.synchronize()          # (warm-up launches and the graph capture take slots; the graph keeps its own)
    _lib.load().pie_debug_stamps(None, 0)
    buf.zero_(); torch.cuda.synchronize()
    eng.step(1); torch.cuda.synchronize()
    t = buf.cpu().numpy().astype(np.int64)
    live = [i for i in range(nslots) if t[i, 0, 0] != 0]
    assert len(live) == per, (len(live), per)
    t = np.stack([t[i] for i in live])
    out, tot = summarise(t)
    first = t[0][:, 0]; first = first[first != 0].min()
    print(json.dumps({"median_us_after_previous_launch_last_exit": out, "per_layer_sum_of_last_exits_us": round(float(tot), 2),
                      "step_us_first_entry_to_last_exit": round(float(t[-1][:, 7].max() - first) / 1e3, 1),
                      "layers": a.layers, "ctx": a.ctx}))
    if a.dump:
        np.save(a.dump, t)


if __name__ == "__main__":
    main()
