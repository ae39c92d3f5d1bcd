"""Run-to-run reproducibility of the graph-replayed training step on the tiny config (the scenario of
tests/test_step_gpu.py::test_checkpoint_resume_continues_the_run_on_gpu, which failed ONCE in ~10 validation runs with
a relative parameter difference of 3.3e-4): N fresh `PretrainStep` objects run the same three steps; every run is
compared with run 0 — per parameter TENSOR, so that "Adam amplifies the order noise of gradients that are ~0 by
construction" (key biases: softmax is shift-invariant) can be told from a race that touches ordinary weights.
Usage: python tools/debug_resume.py [runs]   (env switches K3M_PARALLEL=0 / K3M_WGRAD_STREAM=0 / K3M_GEMM_GROUP=0 bisect)"""
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from k3m_b200.config import K3M_TINY, BertConfig  # noqa: E402
from k3m_b200.synth import synth_batch  # noqa: E402
from k3m_b200.trainer import PretrainStep  # noqa: E402


def main():
    runs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 1, 1.0, 4
    kw = dict(device=os.environ.get("K3M_DEBUG_DEVICE", "cuda"), init_seed=2, lr=1e-3, warmup_steps=2, total_steps=10, training=True, rng_seed=99,
              use_graph=os.environ.get("K3M_DEBUG_GRAPH", "1") != "0", len_t=12, len_v=9, len_pv=20, max_num_pv=4)
    batches = [synth_batch(cfg, 4, 12, 9, 20, max_num_pv=4, seed=60 + i) for i in range(3)]
    ref = None
    detail_above = float(os.environ.get("K3M_DEBUG_DETAIL_ABOVE", "0"))     # per-tensor table only for larger differences
    worst, quiet = [0.0, 0.0, 0.0], [0, 0, 0]
    for r in range(runs):
        s = PretrainStep(cfg, 4, **kw)
        snaps = []
        for i in range(3):
            random.seed(i)
            s.step(batches[i])
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            snaps.append((s.p32.clone(), s.m.clone(), s.v.clone(), dict(s.read_losses())))
        if ref is None:
            ref, L = snaps, s.layout
            print("run 0: losses %s" % {k: round(v, 6) for k, v in snaps[-1][3].items()})
            continue
        for i in range(3):
            p, m, v, lo = snaps[i]
            p0, m0, v0, lo0 = ref[i]
            if torch.equal(p, p0) and torch.equal(m, m0):
                print("run %d step %d: bit-identical parameters and first moments" % (r, i + 1))
                continue
            rows = []
            for n, (off, num, _) in L.offsets.items():
                d = (p[off:off + num] - p0[off:off + num]).abs()
                if float(d.max()) > 0:
                    dm = (m[off:off + num] - m0[off:off + num]).abs()
                    rows.append((float(d.norm()), float(d.max()), int((d > 1e-4).sum()), num,
                                 float(dm.max()), float(m0[off:off + num].abs().max()), n))
            rows.sort(reverse=True)
            tot = float((p - p0).norm()) / float(p0.norm())
            worst[i] = max(worst[i], tot)
            if tot < detail_above:
                quiet[i] += 1
                continue
            print("run %d step %d: rel param diff %.3e, %d tensors differ, loss diff %s; top by norm:" % (
                r, i + 1, tot, len(rows), {k: "%.1e" % abs(lo[k] - lo0[k]) for k in lo if lo[k] != lo0[k]}))
            for dn, dx, cnt, num, dmx, mmax, n in rows[:10]:
                print("    %-70s |dp| norm %.2e max %.2e (%d of %d > 1e-4)  |dm| max %.2e (|m| max %.2e)" % (
                    n, dn, dx, cnt, num, dmx, mmax))
            kb = sum(x[0] ** 2 for x in rows if ".key.bias" in x[6] or "key" in x[6] and "bias" in x[6]) ** 0.5
            print("    share of the difference in key biases: %.3f" % (kb / max(float((p - p0).norm()), 1e-30)))
        del s
    print("%d runs: largest relative parameter difference to run 0 after step 1 / 2 / 3: %.3e / %.3e / %.3e; "
          "run-steps below the detail threshold %.1e: %s" % (runs, worst[0], worst[1], worst[2], detail_above, quiet))


if __name__ == "__main__":
    main()
