"""Microbenchmark of k3m_gemm_grouped_bf16 on the groups the K3M step launches (B = 64), against the same problems
as single launches of k3m_gemm_bf16 and as torch.matmul (cuBLAS) calls; n launches per CUDA graph, CUDA events."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps  # noqa: E402


def timeit(fn, n=10):
    """n launches captured in one CUDA graph (no host launch overhead), replayed 5 times"""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=st):
        for _ in range(n):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * n)


ops = CudaOps("cuda")
V, T, C = (2368, 1024), (2304, 768), (2304, 768)      # (rows, hidden) of the image / text / two-text streams
GROUPS = {
    # name: (layout, [(M, N, K)], epilogue)
    "conn qkv fwd": ("nt", [(2368, 3072, 1024), (2304, 3072, 768)] * 2 + [(2304, 2304, 768)] * 2, ""),
    "conn out fwd": ("nt", [(2368, 1024, 1024), (2304, 768, 1024)] * 2 + [(2304, 768, 768)] * 2, "drop_add"),
    "conn ffn1 fwd": ("nt", [(2368, 1024, 1024), (2304, 3072, 768)] * 2 + [(2304, 3072, 768)] * 2, "gelu"),
    "conn ffn2 fwd": ("nt", [(2368, 1024, 1024), (2304, 768, 3072)] * 2 + [(2304, 768, 3072)] * 2, "drop_add"),
    "conn dffn2": ("nn", [(2368, 1024, 1024), (2304, 3072, 768)] * 2 + [(2304, 3072, 768)] * 2, "dgelu"),
    "conn dffn1": ("nn", [(2368, 1024, 1024), (2304, 768, 3072)] * 2 + [(2304, 768, 3072)] * 2, "add"),
    "conn dqkv": ("nn", [(2368, 1024, 3072), (2304, 768, 3072)] * 2 + [(2304, 768, 2304)] * 2, "add"),
    "conn wffn": ("tn", [(1024, 1024, 2368), (3072, 768, 2304)] * 2 + [(3072, 768, 2304)] * 2, "acc"),
    "conn wqkv": ("tn", [(3072, 1024, 2368), (3072, 768, 2304)] * 2 + [(2304, 768, 2304)] * 2, "acc"),
    "conn wout": ("tn", [(1024, 1024, 2368), (768, 1024, 2304)] * 2 + [(768, 768, 2304)] * 2, "acc"),
    "layer qkv": ("nt", [(9216, 2304, 768), (4736, 3072, 1024)], ""),
    "layer out": ("nt", [(9216, 768, 768), (4736, 1024, 1024)], "drop_add"),
    "layer ffn1": ("nt", [(9216, 3072, 768), (4736, 1024, 1024)], "gelu"),
    "layer ffn2": ("nt", [(9216, 768, 3072), (4736, 1024, 1024)], "drop_add"),
    "layer wffn": ("tn", [(3072, 768, 9216), (1024, 1024, 4736)], "acc"),
}


def make(layout, M, N, K, epi):
    dev = "cuda"
    rng = torch.tensor([1, 2], dtype=torch.int64, device=dev)
    if layout == "nt":
        a, b = torch.randn(M, K, device=dev).bfloat16(), torch.randn(N, K, device=dev).bfloat16()
        ref = lambda: torch.matmul(a, b.t())                      # noqa: E731
    elif layout == "nn":
        a, b = torch.randn(M, K, device=dev).bfloat16(), torch.randn(K, N, device=dev).bfloat16()
        ref = lambda: torch.matmul(a, b)                          # noqa: E731
    else:
        a, b = torch.randn(K, M, device=dev).bfloat16(), torch.randn(K, N, device=dev).bfloat16()
        ref = lambda: torch.matmul(a.t(), b)                      # noqa: E731
    kw = dict(a=a, b=b, layout=layout)
    if epi == "acc":
        kw.update(out=torch.zeros(M, N, device=dev), accumulate=True)
    else:
        kw["out"] = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        if epi in ("drop_add", "gelu", ""):
            kw["bias"] = torch.randn(N, device=dev).bfloat16()
        if epi == "drop_add":
            kw.update(drop=(0.1, 3), rng=rng, addend=torch.randn(M, N, device=dev).bfloat16())
        if epi == "add":
            kw.update(addend=torch.randn(M, N, device=dev).bfloat16())
        if epi == "gelu":
            kw.update(act=1, aux_out=torch.empty(M, N, device=dev, dtype=torch.bfloat16))
        if epi == "dgelu":
            kw.update(act=2, aux_in=torch.randn(M, N, device=dev).bfloat16())
    return kw, ref


only = sys.argv[1] if len(sys.argv) > 1 else None
tg = ts = tc = 0.0
for name, (layout, probs, epi) in GROUPS.items():
    if only and only not in name:
        continue
    built = [make(layout, M, N, K, epi) for M, N, K in probs]
    calls = [b[0] for b in built]
    refs = [b[1] for b in built]
    fl = sum(2.0 * M * N * K for M, N, K in probs)
    t_group = timeit(lambda: ops.gemm_group(calls), n=10)
    var = ops.gemm_last_variant()
    t_single = timeit(lambda: [ops.gemm(**c) for c in calls], n=10)
    t_cublas = timeit(lambda: [r() for r in refs], n=10)
    tg, ts, tc = tg + t_group, ts + t_single, tc + t_cublas
    print("%-14s %s x%d %6.1f GF  grouped %7.1f us %7.1f TF/s | singles %7.1f us %7.1f TF/s | cublas %7.1f us %7.1f TF/s  [%s]" % (
        name, layout, len(probs), fl / 1e9, t_group * 1e3, fl / t_group / 1e9, t_single * 1e3, fl / t_single / 1e9,
        t_cublas * 1e3, fl / t_cublas / 1e9, var))
print("sum grouped %.3f ms  singles %.3f ms  cublas %.3f ms" % (tg, ts, tc))
