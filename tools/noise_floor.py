"""bf16 noise floor of the step: the engine schedule with exact fp32 arithmetic but bf16 STORAGE of every
activation (oracle/ops_ref.py RefOps(bfloat16), CPU) against the same schedule in fp32.  Usage:
python tools/noise_floor.py <golden case> [mode]   (mode overrides if_pre_sampling)"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.ops_ref import RefOps  # noqa: E402
from tests.util import case_setup, load_golden, run_engine  # noqa: E402

name = sys.argv[1]
fx = load_golden(name)
cfg, sd, batch, noise = case_setup(fx)
if len(sys.argv) > 2:
    cfg.if_pre_sampling = int(sys.argv[2])
    if cfg.if_pre_sampling != 1:
        noise = None
o32, L, g32 = run_engine(cfg, sd, batch, noise, RefOps(torch.float32), "cpu", torch.float32, fx["seeds"]["python"])
o16, L, g16 = run_engine(cfg, sd, batch, noise, RefOps(torch.bfloat16), "cpu", torch.bfloat16, fx["seeds"]["python"])
for k in ("loss_t", "loss_v", "loss_pv", "loss_itm", "loss_lpm"):
    a, b = float(o32[k]), float(o16[k])
    print(k, a, b, "rel %.2e" % (abs(a - b) / abs(a)))
rel = []
for n, (off, num, shape) in L.offsets.items():
    a, b = g32[off:off + num], g16[off:off + num]
    if float(a.norm()) < 1e-9:
        continue
    rel.append((float((a - b).norm() / a.norm()), n))
rel.sort(reverse=True)
print("per-tensor relative L2: median %.3e  p90 %.3e  worst %.3e (%s)" % (
    statistics.median([r[0] for r in rel]), rel[len(rel) // 10][0], rel[0][0], rel[0][1]))
