"""Print the table of tools/timeline_step.py dumps:  python tools/tl_show.py a.npy b.npy ..."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from timeline_step import COLS, NAMES, summarise
for f in sys.argv[1:]:
    t = np.load(f).astype(np.int64)
    out, tot = summarise(t)
    first = t[0][:, 0]; first = first[first != 0].min()
    print(f, "step_us", (t[-1][:, 7].max() - first) / 1e3, "per-layer", round(float(tot), 2))
    print("        " + "".join(f"{'s' + str(c):>7s}" for c in COLS) + "  last_exit")
    for nm in NAMES:
        print(f"{nm:8s}" + "".join(f"{out[nm]['s%d' % c]:7.2f}" for c in COLS) + f"{out[nm]['last_exit']:10.2f}")
