"""Microbenchmark of k3m_gemm_bf16 on the GEMM shapes of the K3M step vs torch.matmul (cuBLAS), CUDA events.
K3M_GEMM_BN=128|256 forces the tile width."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps  # noqa: E402

ops = CudaOps("cuda")
SHAPES = [  # (layout, M, N, K, note)
    ("nt", 9216, 2304, 768, "text qkv"), ("nt", 9216, 768, 768, "text attn-out"), ("nt", 9216, 3072, 768, "text ffn1"),
    ("nt", 9216, 768, 3072, "text ffn2"), ("nt", 4736, 3072, 1024, "img qkv"), ("nt", 4736, 1024, 1024, "img out/ffn"),
    ("nt", 2304, 3072, 768, "co qkv2"), ("nt", 2368, 1024, 1024, "co out"), ("nt", 4608, 21128, 768, "vocab"),
    ("nn", 9216, 768, 2304, "text dqkv"), ("nn", 9216, 768, 768, "text dout"), ("nn", 9216, 768, 3072, "text dffn1"),
    ("nn", 9216, 3072, 768, "text dffn2"), ("nn", 4608, 768, 21128, "vocab dgrad"),
    ("tn", 2304, 768, 9216, "text wqkv"), ("tn", 768, 768, 9216, "text wout"), ("tn", 3072, 768, 9216, "text wffn1"),
    ("tn", 768, 3072, 9216, "text wffn2"), ("tn", 1024, 1024, 4736, "img w"), ("tn", 21128, 768, 4608, "vocab wgrad"),
    ("nt", 2304, 768, 768, "co small"), ("nt", 2304, 768, 3072, "co ffn2"), ("nn", 2304, 3072, 768, "co dffn2"),
    ("tn", 768, 768, 2304, "co w small"), ("tn", 3072, 768, 2304, "co wffn1"), ("nt", 64, 768, 768, "tiny"),
]
EPI = os.environ.get("K3M_BENCH_EPI", "")     # "", "drop_add", "gelu", "dgelu"


def timeit(fn, n=20):
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


def timeit_eager(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


only = sys.argv[1] if len(sys.argv) > 1 else None
tot_k = tot_c = 0.0
for layout, M, N, K, note in SHAPES:
    if only and only != note:
        continue
    if layout == "nt":
        a, b = torch.randn(M, K, device="cuda").bfloat16(), torch.randn(N, K, device="cuda").bfloat16()
        ref = lambda: torch.matmul(a, b.t())
    elif layout == "nn":
        a, b = torch.randn(M, K, device="cuda").bfloat16(), torch.randn(K, N, device="cuda").bfloat16()
        ref = lambda: torch.matmul(a, b)
    else:
        a, b = torch.randn(K, M, device="cuda").bfloat16(), torch.randn(K, N, device="cuda").bfloat16()
        ref = lambda: torch.matmul(a.t(), b)
    if layout == "tn":
        out = torch.zeros(M, N, device="cuda")
        fn = lambda: ops.gemm(a, b, layout, out=out, accumulate=True)
    else:
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        bias = torch.randn(N, device="cuda").bfloat16()
        add = torch.randn(M, N, device="cuda").bfloat16()
        aux = torch.randn(M, N, device="cuda").bfloat16()
        rng = torch.tensor([1, 2], dtype=torch.int64, device="cuda")
        if EPI == "drop_add":
            fn = lambda: ops.gemm(a, b, layout, out=out, bias=bias, drop=(0.1, 3), rng=rng, addend=add)
        elif EPI == "gelu":
            fn = lambda: ops.gemm(a, b, layout, out=out, bias=bias, act=1, aux_out=aux)
        elif EPI == "dgelu":
            fn = lambda: ops.gemm(a, b, layout, out=out, act=2, aux_in=aux)
        else:
            fn = lambda: ops.gemm(a, b, layout, out=out)
    tk, tc = timeit(fn), timeit(ref)
    fl = 2.0 * M * N * K
    tot_k += tk
    tot_c += tc
    print("%-2s %6d %6d %6d %-12s  k3m %7.1f us %7.1f TF/s   cublas %7.1f us %7.1f TF/s" % (
        layout, M, N, K, note, tk * 1e3, fl / tk / 1e9, tc * 1e3, fl / tc / 1e9))
print("sum k3m %.2f ms  cublas %.2f ms" % (tot_k, tot_c))
