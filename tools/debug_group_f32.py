"""Reproducer of a (fixed) defect of `k3m_gemm_grouped_bf16`: fp32 outputs STORED (not accumulated) by a grouped launch
of small problems (128 x 128 single-CTA tiles, 16 epilogue warps) with a bias were wrong
(tests/test_kernels_gpu.py::test_gemm_grouped_fp32_outputs_and_in_place_addends).  Prints, per problem, how the result
differs from the fp32 reference.  What it showed (profiles/r03last_group_f32.txt): only WITH a bias, elements 3, 5, 7
of every 8-column group, every row, garbage up to 1e38 — the generic epilogue's bias unpack (`ld8` through a punned
pointer) compiled, in that one instantiation, to uniform-register code reading registers the load never wrote.
Fixed by `ld8_bits` (gemm_sm100.cu); profiles/r03fix_group_f32.txt is this script on the fixed library.
Usage: python tools/debug_group_f32.py [case ...]"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from k3m_b200.ops import CudaOps  # noqa: E402

CASES = {
    "tiny": ([(36, 384, 384), (40, 384, 384), (30, 768, 768)], True),
    "tiny_nobias": ([(36, 384, 384), (40, 384, 384), (30, 768, 768)], False),
    "two_equal": ([(128, 128, 64), (128, 128, 64)], True),
    "two_rows": ([(256, 128, 128), (128, 256, 128)], True),
    "one_plus_one": ([(36, 384, 384), (36, 384, 384)], True),
    "mixed_big": ([(300, 776, 264), (2304, 2304, 768)], True),
}


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).cuda()


def describe(i, got, want, others):
    err = (got - want).abs()
    bad = err > 2e-3 + 2e-3 * want.abs()
    nb = int(bad.sum())
    line = "  [%d] %dx%d: %d / %d bad, max err %.3e, nan %d" % (i, got.shape[0], got.shape[1], nb, bad.numel(),
                                                                float(err.nan_to_num(1e9).max()), int(got.isnan().sum()))
    if nb:
        r, c = bad.nonzero(as_tuple=True)
        line += "\n      bad rows %d..%d (mod 32: %s)  bad cols %d..%d (mod 32: %s; /32: %s)" % (
            int(r.min()), int(r.max()), sorted(set((r % 32).tolist()))[:12], int(c.min()), int(c.max()),
            sorted(set((c % 32).tolist()))[:12], sorted(set((c // 32).tolist()))[:16])
        line += "\n      zero where bad: %d; first bad: %s" % (
            int((got[bad] == 0).sum()),
            [(int(a), int(b), round(float(got[a, b]), 4), round(float(want[a, b]), 4)) for a, b in zip(r[:4], c[:4])])
        for j, o in others:
            if o.shape == want.shape and float((got - o).abs().max()) < 2e-3:
                line += "\n      == the expected output of problem %d" % j
    print(line)


def main():
    ops = CudaOps("cuda")
    for name in (sys.argv[1:] or list(CASES)):
        shapes, with_bias = CASES[name]
        calls, wants = [], []
        for i, (M, N, K) in enumerate(shapes):
            a, b = rnd(M, K, seed=3 * i + 1), rnd(N, K, seed=3 * i + 2, scale=1.0 / math.sqrt(K))
            bias = rnd(N, seed=3 * i + 3) if with_bias else None
            calls.append(dict(a=a, b=b, layout="nt", bias=bias, out_dtype=torch.float32))
            w = a.float() @ b.float().t()
            wants.append(w + bias.float() if with_bias else w)
        singles = [ops.gemm(**c) for c in calls]
        outs = ops.gemm_group(calls, group_f32_store=True)
        torch.cuda.synchronize()
        print("%s: %s" % (name, ops.gemm_last_variant()))
        for i, (g, s, w) in enumerate(zip(outs, singles, wants)):
            assert float((s - w).abs().max()) < 5e-2, "single launch wrong?"
            describe(i, g, w, [(j, x) for j, x in enumerate(wants) if j != i])


if __name__ == "__main__":
    main()
