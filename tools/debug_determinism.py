"""Two training-mode forward+backward runs of the engine with the same RNG state: which gradient tensors differ?
Env switches bisect the cause: K3M_GEMM_GROUP=0, K3M_PARALLEL=0, K3M_WGRAD_STREAM=0."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps  # noqa: E402
from tests.util import case_setup, load_golden, run_engine  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "tiny_mean"
training = (sys.argv[2] != "eval") if len(sys.argv) > 2 else True
ops = CudaOps("cuda")
fx = load_golden(name)
cfg, sd, batch, noise = case_setup(fx)
if os.environ.get("K3M_DEBUG_WARM"):       # populate the caching allocator with the blocks of a B = 64 run first
    fxw = load_golden(os.environ["K3M_DEBUG_WARM"])
    cw, sw, bw, nw = case_setup(fxw)
    run_engine(cw, sw, bw, nw, ops, "cuda", torch.bfloat16, 1, training=False)
    torch.cuda.synchronize()
runs = []
for r in range(3):
    out, L, g = run_engine(cfg, sd, batch, noise, ops, "cuda", torch.bfloat16, 1, training=training)
    torch.cuda.synchronize()
    runs.append((out, g.clone()))
g1 = runs[0][1]
for r in (1, 2):
    g2 = runs[r][1]
    rows = []
    for n, (off, num, _) in L.offsets.items():
        a, b = g1[off:off + num], g2[off:off + num]
        d = float((a - b).abs().max())
        if d > 0:
            rows.append((d / (float(a.abs().max()) + 1e-30), d, n))
    rows.sort(reverse=True)
    print("run 0 vs run %d: %d tensors differ; loss_t %r %r; top:" % (r, len(rows), float(runs[0][0]["loss_t"]), float(runs[r][0]["loss_t"])))
    for rel, d, n in rows[:8]:
        print("   rel-to-max %.3e  abs %.3e  %s" % (rel, d, n))
