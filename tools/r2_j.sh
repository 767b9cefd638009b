#!/bin/bash
OUT=gpurun_out/r2j
mkdir -p $OUT
( PIE_QMV2=0 timeout 1200 python -m pytest tests/test_gpu_configs.py -q -k "8b_shaped" ) > $OUT/pytest_q0.log 2>&1; tail -8 $OUT/pytest_q0.log
( timeout 2400 python -m pytest tests/test_gpu_configs.py -q -s ) > $OUT/pytest_cfg.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_cfg.log; grep -E "config1 parity|passed|failed|FAILED|Error" $OUT/pytest_cfg.log | head -20
python - <<'PY' > $OUT/dbg.log 2>&1
import numpy as np, torch, sys
sys.path.insert(0,'tests')
from parity_util import *
from proxy_inference_engine_b200.checkpoint import preset_args, build_model, random_weights
args = preset_args("llama-3.1-8b", bits=4, num_hidden_layers=1, vocab_size=4096)
w = random_weights(args, seed=0, device="cuda", gen_device="cpu")
model = build_model(args, w); o = make_oracle(args, w)
for L in (1, 8, 16, 32, 33, 40, 64, 80):
    ids = torch.randint(0, args.vocab_size, (1, L), generator=torch.Generator().manual_seed(1))
    tr = {}
    lo = o(ids.numpy(), o.make_cache(), trace=tr)
    lg = model(ids.cuda().to(torch.int32), cache=model.make_cache()).float().cpu().numpy()
    print(L, "logits rel err all", rel_l2(lg, lo), "last", rel_l2(lg[:, -1], lo[:, -1]), "first", rel_l2(lg[:, 0], lo[:, 0]), flush=True)
PY
cat $OUT/dbg.log
