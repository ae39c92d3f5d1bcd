"""Do HBM-bound kernels run UNDER a persistent GEMM?  A grouped GEMM loop on one stream, AdamW (fat / background variant)
on another: alone and together, CUDA events around the pair.  together ~ max(alone) => overlap; ~ sum => serialised."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps  # noqa: E402

ops = CudaOps("cuda")
dev = "cuda"
probs = [(2368, 3072, 1024), (2304, 3072, 768)] * 2 + [(2304, 2304, 768)] * 2
calls = []
for M, N, K in probs:
    calls.append(dict(a=torch.randn(M, K, device=dev).bfloat16(), b=torch.randn(N, K, device=dev).bfloat16(), layout="nt",
                      out=torch.empty(M, N, device=dev, dtype=torch.bfloat16), bias=torch.randn(N, device=dev).bfloat16()))
n = 64 * (1 << 20) // 4 * 4          # 256 MB of fp32 = 64 M parameters
p32, g32, m, v = (torch.zeros(n, device=dev) for _ in range(4))
w16 = torch.zeros(n, device=dev, dtype=torch.bfloat16)
hyper = torch.tensor([1e-4, 0.9, 0.98, 1e-8, 0.01, 5.0, 1.0, 0.0], device=dev)
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()


def gemms(reps=40):
    for _ in range(reps):
        ops.gemm_group(calls)


def run(fa, fb):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sa.wait_stream(torch.cuda.current_stream())
    sb.wait_stream(torch.cuda.current_stream())
    if fa:
        with torch.cuda.stream(sa):
            fa()
    if fb:
        with torch.cuda.stream(sb):
            fb()
    torch.cuda.current_stream().wait_stream(sa)
    torch.cuda.current_stream().wait_stream(sb)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


for name, fb in (("adamw fat", lambda: ops.adamw(p32, g32, m, v, w16, hyper)),
                 ("adamw bg", lambda: ops.adamw(p32, g32, m, v, w16, hyper, background=True))):
    for _ in range(2):
        run(gemms, fb)
    ta, tb, tab = run(gemms, None), run(None, fb), run(gemms, fb)
    print("%-10s gemm alone %.3f ms  %s alone %.3f ms (%.0f GB/s)  together %.3f ms  (sum %.3f, max %.3f)" % (
        name, ta, name, tb, n * 34 / tb / 1e6, tab, ta + tb, max(ta, tb)))
