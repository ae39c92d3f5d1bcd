import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from k3m_b200.synth import synth_batch
from k3m_b200.trainer import PretrainStep
cfg = bench.make_cfg()
B = 64
for use_graph, warm in ((False, 100), (True, 100), (True, 0)):
    step = PretrainStep(cfg, B, device="cuda", init_seed=0, training=True, use_graph=use_graph, lr=1e-4,
                        warmup_steps=warm, total_steps=100000, **bench.DIMS)
    host = [{k: v.pin_memory() for k, v in synth_batch(cfg, B, 36, 36, 36, 8, seed=1234 + i).items()} for i in range(3)]
    for it in range(4):
        step.step(host[it % 3])
        l = step.read_losses()
        def bad(t):
            return int((~torch.isfinite(t.float())).sum())
        print("graph", use_graph, "warm", warm, "it", it, {k: round(v, 4) for k, v in l.items()},
              "bad p32/w16/m/v/g", bad(step.p32), bad(step.w16), bad(step.m), bad(step.v), bad(step.g32),
              "hyper", step.hyper[0].tolist()[:1], step.hyper[0].tolist()[5:7], "rng", step.rng.tolist())
        if bad(step.p32):
            L = step.layout
            names = [n for n, (o, num, _) in L.offsets.items() if bad(step.p32[o:o + num])]
            print("   bad params:", len(names), names[:8])
            break
    del step
    torch.cuda.empty_cache()
