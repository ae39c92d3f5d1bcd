"""Uninitialised-read hunt on the GRAPH step (`PretrainStep`), the scenario of
tests/test_step_gpu.py::test_checkpoint_resume_continues_the_run_on_gpu: every torch.empty / empty_like is filled with a
poison value before use — also inside the captured graph, where the fill becomes a node that re-poisons the buffer on
every replay — then run a (three steps) and b (two steps) -> checkpoint -> c (one step) are compared.  A kernel that
reads memory nobody wrote shows up as NaN or as a difference between the poison values / from the clean run.
(tools/debug_poison.py does the same for the eager engine; tools/debug_resume.py shows that fresh runs in a clean
process agree to 1e-7, which a stale-memory read would not disturb there: a re-used block holds the previous, identical
run's data.)"""
import os
import random
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from k3m_b200 import checkpoint as ckpt  # noqa: E402
from k3m_b200.config import K3M_TINY, BertConfig  # noqa: E402
from k3m_b200.synth import synth_batch  # noqa: E402
from k3m_b200.trainer import PretrainStep  # noqa: E402

_empty, _empty_like = torch.empty, torch.empty_like
POISON = [None]


def poison(t):
    if POISON[0] is None or not t.is_cuda:
        return t
    if t.is_floating_point():
        t.fill_(POISON[0])
    elif t.dtype in (torch.int32, torch.int64) and t.numel():
        t.fill_(3)
    elif t.dtype == torch.uint8 and t.numel():
        t.fill_(0xAB)
    return t


def empty(*a, **k):
    return poison(_empty(*a, **k))


def empty_like(*a, **k):
    return poison(_empty_like(*a, **k))


def scenario(tmp):
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 1, 1.0, 4
    kw = dict(device="cuda", init_seed=2, lr=1e-3, warmup_steps=2, total_steps=10, training=True, rng_seed=99,
              use_graph=True, len_t=12, len_v=9, len_pv=20, max_num_pv=4)
    batches = [synth_batch(cfg, 4, 12, 9, 20, max_num_pv=4, seed=60 + i) for i in range(3)]
    a = PretrainStep(cfg, 4, **kw)
    for i in range(3):
        random.seed(i)
        a.step(batches[i])
    b = PretrainStep(cfg, 4, **kw)
    for i in range(2):
        random.seed(i)
        b.step(batches[i])
    path = os.path.join(tmp, "resume.tar")
    ckpt.save_checkpoint(b, path)
    c = PretrainStep(cfg, 4, **dict(kw, init_seed=5))
    assert ckpt.load_checkpoint(c, path) == 2
    random.seed(2)
    c.step(batches[2])
    torch.cuda.synchronize()
    scenario.layout = a.layout
    return a.p32.clone(), c.p32.clone(), dict(a.read_losses()), dict(c.read_losses())


def repeat(n):
    """the clean scenario n times: how often, and where, does c leave a by more than the usual 1e-7?"""
    events = 0
    for r in range(n):
        with tempfile.TemporaryDirectory() as tmp:
            pa, pc, la, lc = scenario(tmp)
        rel = float((pc - pa).norm()) / float(pa.norm())
        if rel < 1e-6:
            continue
        events += 1
        rows = []
        for name, (off, num, _) in scenario.layout.offsets.items():
            d = (pc[off:off + num] - pa[off:off + num]).abs()
            if float(d.max()) > 0:
                rows.append((float(d.norm()), float(d.max()), int((d > 1e-5).sum()), num, name))
        rows.sort(reverse=True)
        print("repeat %d: |c - a| / |a| = %.3e, %d tensors differ; loss differences %s" % (
            r, rel, len(rows), {k: "%.1e" % abs(lc[k] - la[k]) for k in la if lc[k] != la[k]}))
        for dn, dx, cnt, num, name in rows[:12]:
            print("    %-66s |dp| norm %.2e max %.2e (%d of %d > 1e-5)" % (name, dn, dx, cnt, num))
    print("%d repeats of the clean scenario: %d with |c - a| / |a| >= 1e-6" % (n, events))


def main():
    # stale data first: the caching allocator's free blocks then hold garbage, as they do late in a pytest session
    junk = [torch.randn(1 << 22, device="cuda") * 50 for _ in range(24)] + [torch.full((1 << 20,), float("nan"), device="cuda")]
    del junk
    base = None
    for val in (None, float("nan"), 1.0e4, -3.0):
        POISON[0] = val
        torch.empty, torch.empty_like = empty, empty_like
        try:
            with tempfile.TemporaryDirectory() as tmp:
                pa, pc, la, lc = scenario(tmp)
        finally:
            torch.empty, torch.empty_like = _empty, _empty_like
        if base is None:
            base = pa
        print("poison %-6s: non-finite a %d c %d; |c - a| / |a| = %.3e; |a - a_clean| / |a| = %.3e; losses a %s c %s" % (
            val, int((~torch.isfinite(pa)).sum()), int((~torch.isfinite(pc)).sum()),
            float((pc - pa).norm()) / float(pa.norm()), float((pa - base).norm()) / float(base.norm()),
            {k: round(v, 5) for k, v in la.items()}, {k: round(v, 5) for k, v in lc.items()}))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "repeat":
        repeat(int(sys.argv[2]))
    else:
        main()
