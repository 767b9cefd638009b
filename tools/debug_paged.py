import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from parity_util import run_decode_parity, tiny_args  # noqa: E402

for cfg in [dict(bits=8, head_dim=128, batch=3, paged=False), dict(bits=4, head_dim=64, batch=1, paged=True),
            dict(bits=4, head_dim=64, batch=2, paged=True), dict(bits=4, head_dim=64, batch=1, paged=True, graph=False)]:
    r = run_decode_parity(tiny_args(bits=cfg["bits"], head_dim=cfg["head_dim"]), prompt_len=20, steps=4,
                          batch=cfg["batch"], paged=cfg["paged"], use_graph=cfg.get("graph", True))
    print(cfg, r["errs"])
