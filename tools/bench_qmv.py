"""Micro-benchmark of the decode GEMV on Llama shapes: HBM GB/s per launch (CUDA events, L2 flushed by rotating
through more weight copies than fit in the 126 MB L2)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from proxy_inference_engine_b200 import _lib, ops  # noqa: E402
from proxy_inference_engine_b200.ops import _p, _stream  # noqa: E402

SHAPES = {  # name: (segments N, K, epilogue, norm)
    "qkv": ([4096, 1024, 1024], 4096, 0, True),
    "o": ([4096], 4096, 1, False),
    "gate_up": ([14336, 14336], 4096, 2, True),
    "down": ([4096], 14336, 1, False),
    "head": ([128256], 4096, 0, True),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bits", type=int, default=4)
    ap.add_argument("--M", type=int, default=1)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--shapes", default="qkv,o,gate_up,down,head")
    ap.add_argument("--ksplit", type=int, default=0)
    ap.add_argument("--graph", type=int, default=0, help="time N launches captured in one CUDA graph (no CPU launch cost)")
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--copies", type=int, default=0, help="weight copies to rotate through (0: enough to defeat L2; 1: L2-warm)")
    a = ap.parse_args()
    L = _lib.load()
    dev = "cuda"
    res = {}
    for name in a.shapes.split(","):
        Ns, K, epi, norm = SHAPES[name]
        per_copy = sum(Ns) * K * (a.bits / 8 + 4 / 64)
        copies = a.copies or max(2, int(400e6 // per_copy) + 1)
        sets = []
        for _ in range(copies):
            segs = (_lib.QmvSeg * len(Ns))()
            keep = []
            for i, n in enumerate(Ns):
                wq = torch.randint(-2**31, 2**31 - 1, (n, K * a.bits // 32), dtype=torch.int32, device=dev)
                s = (torch.rand((n, K // 64), device=dev) * 0.01).to(torch.bfloat16)
                b = (torch.rand((n, K // 64), device=dev) * -0.05).to(torch.bfloat16)
                keep += [wq, s, b]
                segs[i].wq, segs[i].scales, segs[i].biases, segs[i].N = wq.data_ptr(), s.data_ptr(), b.data_ptr(), n
            sets.append((segs, keep))
        x = torch.randn((a.M, K), device=dev).to(torch.bfloat16)
        nw = torch.ones((K,), dtype=torch.bfloat16, device=dev)
        out_n = Ns[0] if epi == 2 else sum(Ns)
        out = torch.zeros((a.M, out_n), dtype=torch.bfloat16, device=dev)
        nb = L.pie_qmv_workspace_bytes(a.M, Ns[0] if epi == 2 else sum(Ns), K, a.bits, int(epi == 2), a.ksplit)
        ws = torch.zeros(max(nb, 256), dtype=torch.uint8, device=dev)

        def run(i):
            segs = sets[i % copies][0]
            st = L.pie_qmv_fused(_p(out), out_n, _p(x), K, a.M, K, segs, len(Ns), 64, a.bits, _p(nw) if norm else None,
                                 1e-5, epi, _p(out) if epi == 1 else None, a.ksplit, _p(ws), ws.numel(), a.pdl, _stream())
            _lib.check(st, name)

        for i in range(3):
            run(i)
        torch.cuda.synchronize()
        if a.graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for i in range(a.graph):
                    run(i)
            g.replay(); torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) / a.graph)
            ts.sort()
            med = ts[len(ts) // 2]
            byts = per_copy + a.M * K * 2 + a.M * out_n * 2
            res[name] = {"us": round(med * 1e3, 2), "min_us": round(ts[0] * 1e3, 2), "GBps": round(byts / med / 1e6, 1),
                         "MB": round(byts / 1e6, 2)}
            continue
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.iters)]
        for i, (e0, e1) in enumerate(evs):
            e0.record(); run(i); e1.record()
        torch.cuda.synchronize()
        ts = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
        med = ts[len(ts) // 2]
        byts = per_copy + a.M * K * 2 + a.M * out_n * 2
        res[name] = {"us": round(med * 1e3, 2), "min_us": round(ts[0] * 1e3, 2), "GBps": round(byts / med / 1e6, 1),
                     "MB": round(byts / 1e6, 2)}
    print(json.dumps({"bits": a.bits, "M": a.M, "ksplit": a.ksplit, "env": {k: v for k, v in os.environ.items()
                                                                          if k.startswith("PIE_")}, "res": res}))


if __name__ == "__main__":
    main()
