"""Times k3m_attn_fwd / k3m_attn_bwd alone (CUDA events, L2-sized rotation of inputs) for a list of shapes.
Usage: python tools/bench_attn.py [--iters 20] [--once]      (--once: one launch per shape, for ncu)"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

SHAPES = [  # B, heads, Lq, Lk, d
    (64, 12, 256, 256, 64), (64, 8, 256, 37, 128), (64, 8, 37, 256, 128), (64, 8, 50, 256, 96),
    (16, 12, 128, 128, 64), (64, 12, 36, 36, 64), (64, 8, 37, 37, 128)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--shapes", default="")
    args = ap.parse_args()
    from k3m_b200.ops import CudaOps
    ops = CudaOps("cuda")
    rng = torch.tensor([1234, 7], dtype=torch.int64, device="cuda")
    shapes = SHAPES if not args.shapes else [SHAPES[int(i)] for i in args.shapes.split(",")]
    for B, H, Lq, Lk, d in shapes:
        g = torch.Generator(device="cuda").manual_seed(0)
        W = 3 * H * d
        nbuf = 1 if args.once else 4
        qs = [torch.randn(B * Lq, W, device="cuda", generator=g).bfloat16() for _ in range(nbuf)]
        ks = [torch.randn(B * Lk, W, device="cuda", generator=g).bfloat16() for _ in range(nbuf)]
        do = torch.randn(B * Lq, H * d, device="cuda", generator=g).bfloat16()
        mask = torch.zeros(B, Lk, device="cuda")
        dq, dkv = torch.zeros_like(qs[0]), torch.zeros_like(ks[0])
        for drop in (None, (0.1, 5)):
            n = 1 if args.once else args.iters
            ctx, lse = ops.attn_fwd(qs[0], 0, ks[0], H * d, 2 * H * d, mask, B, H, Lq, Lk, d, drop=drop, rng=rng)
            ops.attn_bwd(do, ctx, qs[0], 0, ks[0], H * d, 2 * H * d, mask, lse, dq, dkv, B, H, Lq, Lk, d, drop=drop, rng=rng)
            if args.once:
                continue
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            torch.cuda.synchronize()
            ev[0].record()
            for i in range(n):
                ops.attn_fwd(qs[i % nbuf], 0, ks[i % nbuf], H * d, 2 * H * d, mask, B, H, Lq, Lk, d, drop=drop, rng=rng, out=ctx)
            ev[1].record()
            for i in range(n):
                ops.attn_bwd(do, ctx, qs[i % nbuf], 0, ks[i % nbuf], H * d, 2 * H * d, mask, lse, dq, dkv, B, H, Lq, Lk, d,
                             drop=drop, rng=rng)
            ev[2].record()
            torch.cuda.synchronize()
            fl = 4.0 * B * H * Lq * Lk * d
            tf, tb = ev[0].elapsed_time(ev[1]) / n, ev[1].elapsed_time(ev[2]) / n
            print("B%d h%d %dx%dx%d drop=%s  fwd %.1f us (%.0f TF/s)  bwd %.1f us (%.0f TF/s)" % (
                B, H, Lq, Lk, d, "on" if drop else "off", tf * 1e3, fl / tf / 1e9, tb * 1e3, 2.5 * fl / tb / 1e9), flush=True)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
