"""Is the attention launch's first 2 us instruction-cache warm-up?  Same launch 8 times back to back in one graph."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200 import _lib
from proxy_inference_engine_b200.checkpoint import build_model, preset_args, random_weights
from proxy_inference_engine_b200.decode import DecodeEngine
from proxy_inference_engine_b200.ops import _p, _stream
margs = preset_args("llama-3.1-8b", bits=4, num_hidden_layers=2)
model = build_model(margs, random_weights(margs, seed=0, device="cuda"))
eng = DecodeEngine(model, batch=1, max_context=500)
eng.positions.fill_(300); eng.tokens.fill_(5); eng._next_pos = 300
eng.step(2); torch.cuda.synchronize()
L = eng.L
ly = model.layers[0]
N = 8
buf = torch.zeros((N, 256, 16), dtype=torch.int64, device="cuda")
def run():
    for _ in range(N):
        _lib.check(L.pie_decode_attn_fused(_p(eng.attn), _p(eng.qkv), _p(eng.kbuf[0]), _p(eng.vbuf[0]), C.byref(eng.lay[0]),
                                           _p(eng.positions), _p(ly.self_attn.rope._freqs), _p(eng.rope_tab), 1, eng.Hq, eng.Hkv,
                                           eng.D, float(ly.self_attn.scale), eng.nsplit_fused, _p(eng.attn_ws), eng.attn_ws.numel(),
                                           1, _stream()), "attn")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    run(); torch.cuda.synchronize()
    L.pie_debug_stamps(buf.data_ptr(), buf.numel() * 8)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        run()
    L.pie_debug_stamps(None, 0)
    g.replay(); torch.cuda.synchronize(); buf.zero_(); torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
t = buf.cpu().numpy().astype(np.int64)
for j in range(N):
    sj = t[j]; n = int((sj[:, 0] != 0).sum()); sj = sj[:n]
    base = sj[:, 0].min()
    med = lambda c: (np.median(sj[:, c][sj[:, c] != 0]) - base) / 1e3
    print(j, n, "s8 %.2f s9 %.2f s10 %.2f" % (med(8)-med(0), med(9)-med(0), med(10)-med(0)), "s0 %.2f s1 %.2f s3 %.2f s2 %.2f s4 %.2f s5 %.2f s6 %.2f s7 %.2f | last exit %.2f" % (
        med(0), med(1), med(3), med(2), med(4), med(5), med(6), med(7), (sj[:, 7].max() - base) / 1e3))
