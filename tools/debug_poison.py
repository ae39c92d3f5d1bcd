"""Uninitialised-read hunt: every torch.empty / empty_like of a floating tensor is filled with NaN (and integer tensors
with a large value) before the engine runs; any kernel that reads memory it (or a predecessor) did not write then
shows up as NaN in the losses / gradients, or as a difference from the unpoisoned run."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps  # noqa: E402
from tests.util import case_setup, load_golden, run_engine  # noqa: E402

_empty, _empty_like = torch.empty, torch.empty_like
POISON = [float("nan")]


def poison(t):
    if t.is_floating_point():
        t.fill_(POISON[0])
    elif t.dtype in (torch.int32, torch.int64) and t.numel():
        t.fill_(1 << 20)
    return t


def empty(*a, **k):
    return poison(_empty(*a, **k))


def empty_like(*a, **k):
    return poison(_empty_like(*a, **k))


ops = CudaOps("cuda")
for name, training in (("tiny_mean", True), ("tiny_hard", True), ("base_hard", True), ("base_mean", False)):
    fx = load_golden(name)
    cfg, sd, batch, noise = case_setup(fx)
    out0, L, g0 = run_engine(cfg, sd, batch, noise, ops, "cuda", torch.bfloat16, 1, training=training)
    torch.cuda.synchronize()
    res = []
    for val in (float("nan"), 1.0e4, -3.0):
        POISON[0] = val
        torch.empty, torch.empty_like = empty, empty_like
        try:
            out1, _, g1 = run_engine(cfg, sd, batch, noise, ops, "cuda", torch.bfloat16, 1, training=training)
            torch.cuda.synchronize()
        finally:
            torch.empty, torch.empty_like = _empty, _empty_like
        bad = [n for n, (off, num, _) in L.offsets.items() if not bool(torch.isfinite(g1[off:off + num]).all())]
        diff = float((g1 - g0).abs().max()) / float(g0.abs().max()) if not bad else float("nan")
        res.append((val, len(bad), bad[:3], diff, {k: float(out1[k]) for k in ("loss_t", "loss_v", "loss_pv", "loss_lpm")}))
    print(name, "training" if training else "eval")
    for r in res:
        print("   poison %-8s non-finite gradient tensors %d %s  max |g - g_clean| / max|g| = %.3e  losses %s" % r)
