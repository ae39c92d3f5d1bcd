"""Multi-GPU check of the data-parallel paths (launch: torchrun --nproc-per-node N tools/check_dp.py [--base]).
Same seeds, three PretrainStep objects per rank: fused peer-memory optimizer (k3m_adamw_dp), NCCL bf16 all-reduce,
NCCL fp32 all-reduce.  After a few captured steps: every rank holds the same bf16 weights; fused == NCCL-bf16 up to the
rounding of the gradient sum; both within bf16 tolerance of the fp32 all-reduce; losses agree; sync_shards() completes
the fp32 masters."""
import os
import random
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT, K3M_TINY  # noqa: E402
from k3m_b200.synth import synth_batch  # noqa: E402
from k3m_b200.trainer import PretrainStep  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
base = "--base" in sys.argv
cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT if base else K3M_TINY)
cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 0, 1.0, 4
dims = dict(len_t=36, len_v=36, len_pv=36, max_num_pv=8) if base else dict(len_t=12, len_v=9, len_pv=20, max_num_pv=4)
B = 8 if base else 4
steps = 3


def run(**kw):
    st = PretrainStep(cfg, B, device=dev, init_seed=5, lr=1e-3, training=False, use_graph=True,
                      process_group=dist.group.WORLD, bucket_mb=16 if base else 0.25, **dict(dims, **kw))
    losses = []
    for it in range(steps):
        random.seed(100 + it)
        st.step(synth_batch(cfg, B, dims["len_t"], dims["len_v"], dims["len_pv"], dims["max_num_pv"], seed=50 + 10 * it + rank))
        losses.append(st.read_losses()["loss"])
    torch.cuda.synchronize()
    return st, losses


fused, lf = run(dp_fused=True)
assert fused.dp_fused, "peer-memory path did not come up"
nccl16, l16 = run(dp_fused=False, comm_bf16=True)
nccl32, l32 = run(dp_fused=False, comm_bf16=False)
T = fused.layout.trained_end
ok = True


def report(name, cond, detail=""):
    global ok
    ok = ok and bool(cond)
    if rank == 0:
        print("%-58s %s %s" % (name, "ok" if cond else "FAIL", detail))


# 1. every rank has the same bf16 weights
w = fused.w16[:T].clone()
ws = [torch.empty_like(w) for _ in range(world)]
dist.all_gather(ws, w)
report("fused: all ranks hold identical bf16 weights", all(torch.equal(ws[0], x) for x in ws))
# 2. fused vs NCCL bf16 / fp32: the UPDATE (w - w0) relative L2
w0 = PretrainStep(cfg, B, device=dev, init_seed=5, use_graph=False, **dims).w16[:T].float()
uf, u16, u32 = fused.w16[:T].float() - w0, nccl16.w16[:T].float() - w0, nccl32.w16[:T].float() - w0
e16 = float((uf - u16).norm() / u16.norm())
e32 = float((uf - u32).norm() / u32.norm())
report("fused vs NCCL bf16 all-reduce: update rel-L2", e16 <= 3e-2, "%.3e" % e16)
report("fused vs NCCL fp32 all-reduce: update rel-L2", e32 <= 6e-2, "%.3e" % e32)
for a, b, c in zip(lf, l16, l32):
    report("losses fused / bf16 / fp32", abs(a - c) <= 2e-3 * abs(c) and abs(b - c) <= 2e-3 * abs(c), "%.5f %.5f %.5f" % (a, b, c))
# 3. masters after sync_shards: identical across ranks, and their bf16 rounding is the weight arena
fused.sync_shards(moments=True)
p = fused.p32[:T].clone()
ps = [torch.empty_like(p) for _ in range(world)]
dist.all_gather(ps, p)
report("fused: fp32 masters identical across ranks after sync_shards", all(torch.equal(ps[0], x) for x in ps))
report("fused: bf16(master) == weight arena", torch.equal(p.to(torch.bfloat16), fused.w16[:T]))
report("fused: fp32 gradient arena zeroed", float(fused.g32.abs().max()) == 0.0)
ct = fused.comm_timing(synth_batch(cfg, B, dims["len_t"], dims["len_v"], dims["len_pv"], dims["max_num_pv"], seed=999 + rank))
if rank == 0:
    print("comm_timing", {k: v for k, v in ct.items() if k != "note"})
    print("ALL OK" if ok else "FAILED")
dist.barrier()
fused.graph = nccl16.graph = nccl32.graph = None
torch.cuda.synchronize()
os._exit(0 if ok else 1)
