"""CPU: host logic of the training step (k3m_b200.trainer.PretrainStep) with the fp32 PyTorch
restatement of the kernels injected: arena regions, AdamW groups, LR schedule, bucketed gradient
all-reduce (world_size 2 over gloo) — against the oracle + torch.optim.AdamW."""
import os
import random

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from k3m_b200.config import BertConfig, K3M_TINY
from k3m_b200.layout import ArenaLayout, is_no_decay, param_spec
from k3m_b200.optim import warmup_linear_factor
from k3m_b200.synth import synth_batch, synth_state_dict
from k3m_b200.trainer import PretrainStep
from oracle.k3m_oracle import Oracle, total_loss, warmup_linear_lambda
from oracle.ops_ref import RefOps

DIMS = dict(len_t=12, len_v=9, len_pv=20, max_num_pv=4)


def tiny_cfg():
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 0, 1.0, 4
    return cfg


def oracle_grads(cfg, sd, batch, seed):
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    random.seed(seed)
    out = Oracle(cfg, sdo).forward(batch, None)
    total_loss(out).backward()
    return out, {k: v.grad for k, v in sdo.items()}


def reference_update(sd, grads, steps_lr, wd=0.01):
    """torch.optim.AdamW with the two groups of pretrain.py:835-851; params without grad are skipped."""
    params = {k: torch.nn.Parameter(v.clone()) for k, v in sd.items()}
    groups = [dict(params=[p for n, p in params.items() if not is_no_decay(n)], weight_decay=wd),
              dict(params=[p for n, p in params.items() if is_no_decay(n)], weight_decay=0.0)]
    opt = torch.optim.AdamW(groups, lr=steps_lr[0], betas=(0.9, 0.98), eps=1e-8)
    for lr, g in zip(steps_lr, grads):
        for grp in opt.param_groups:
            grp["lr"] = lr
        for n, p in params.items():
            p.grad = None if g[n] is None else g[n].clone()
        opt.step()
    return {k: v.data for k, v in params.items()}


def check_params(step, want, tol=1e-4):
    """Adam's first steps are ~ lr * g / (|g| + eps): for gradient elements near fp32 round-off the
    update itself is noise-sensitive, hence 0.1 * lr rather than 1e-6."""
    got = step.state_dict()
    for n, w in want.items():
        err = float((got[n] - w).abs().max())
        assert err <= tol * max(1.0, float(w.abs().max())), (n, err)


def test_arena_regions_and_groups():
    cfg = tiny_cfg()
    L = ArenaLayout(cfg)
    assert 0 < L.decay_end < L.trained_end < L.total
    for n, (off, num, _) in L.offsets.items():
        if n in L.frozen:
            assert off >= L.trained_end, n
        elif is_no_decay(n):
            assert L.decay_end <= off and off + num <= L.trained_end, n
        else:
            assert off + num <= L.decay_end, n
    # the quirk of the substring rule: co-attention LayerNorm1/2 WEIGHTS decay, their biases do not
    assert not is_no_decay("encoder.c_layer.0.biOutput.LayerNorm1.weight")
    assert is_no_decay("encoder.c_layer.0.biOutput.LayerNorm1.bias")
    assert is_no_decay("encoder.layer.0.output.LayerNorm.weight")
    # fused views are contiguous
    q = L.offsets["encoder.layer.0.attention.self.query.weight"]
    k = L.offsets["encoder.layer.0.attention.self.key.weight"]
    assert q[0] + q[1] == k[0] and L.fused["encoder.layer.0.attention.self.qkv.weight"][0] == q[0]


def test_schedule_restatements_agree():
    for s in range(0, 130, 7):
        assert warmup_linear_factor(s, 10, 100) == warmup_linear_lambda(s, 10, 100)


def test_single_rank_two_steps_match_oracle_plus_torch_adamw():
    cfg = tiny_cfg()
    sd = synth_state_dict(param_spec(cfg), seed=3)
    B = 3
    step = PretrainStep(cfg, B, device="cpu", state_dict=sd, ops=RefOps(torch.float32), training=False,
                        use_graph=False, lr=1e-3, warmup_steps=2, total_steps=10, **DIMS)
    cur, grads, lrs = sd, [], []
    for it in range(2):
        batch = synth_batch(cfg, B, DIMS["len_t"], DIMS["len_v"], DIMS["len_pv"], max_num_pv=4, seed=20 + it)
        out, g = oracle_grads(cfg, cur, batch, seed=5 + it)
        grads.append(g)
        lrs.append(1e-3 * warmup_linear_lambda(it, 2, 10))
        random.seed(5 + it)
        step.step(batch)
        got = step.read_losses()
        for name, idx in (("loss_t", 0), ("loss_v", 1), ("loss_pv", 3), ("loss_itm", 6), ("loss_lpm", 9)):
            assert got[name] == pytest.approx(float(out[idx]), rel=2e-5), name
        assert got["loss"] == pytest.approx(float(total_loss(out)), rel=2e-5)
        cur = reference_update(sd, grads, lrs)
        check_params(step, cur)
    assert float(step.g32.abs().max()) == 0.0        # zeroed by the optimizer kernel
    # never-trained parameters are untouched (torch skips params whose grad is None: no decay either)
    for n in step.layout.frozen:
        assert torch.equal(step.state_dict()[n], sd[n]), n


def _dp_worker(rank, world, port, q, comm_bf16=False):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        torch.set_num_threads(2)
        cfg = tiny_cfg()
        sd = synth_state_dict(param_spec(cfg), seed=3)
        B = 3
        step = PretrainStep(cfg, B, device="cpu", state_dict=sd, ops=RefOps(torch.float32), training=False,
                            use_graph=False, lr=1e-3, process_group=dist.group.WORLD, bucket_mb=0.05,
                            comm_bf16=comm_bf16, **DIMS)
        assert len(step._buckets) > 3 and step.comm_bf16 == comm_bf16
        shards = [synth_batch(cfg, B, DIMS["len_t"], DIMS["len_v"], DIMS["len_pv"], max_num_pv=4, seed=40 + r)
                  for r in range(world)]
        random.seed(70 + rank)
        step.step(shards[rank])
        # DDP semantics: mean over ranks of the rank-local gradients
        per_rank = [oracle_grads(cfg, sd, shards[r], seed=70 + r)[1] for r in range(world)]
        mean = {n: (None if per_rank[0][n] is None else sum(g[n] for g in per_rank) / world) for n in sd}
        if comm_bf16:
            # gradients cross the wire rounded to bf16 and are summed in bf16.  Adam's first step is lr * g / (|g| + eps):
            # it normalises the gradient, so an element whose two rank gradients nearly cancel can flip the sign of its
            # +-lr update — the per-element bound of the fp32 check does not apply.  What must hold: the UPDATE as a
            # whole matches (relative L2 over all trained parameters), and no element moves by more than 2 lr.
            ref = reference_update(sd, [mean], [1e-3])
            num = den = 0.0
            for n in step.layout.spec:
                if n in step.layout.frozen:
                    continue
                got_u, ref_u = step.state_dict()[n] - sd[n], ref[n] - sd[n]
                num += float((got_u - ref_u).pow(2).sum())
                den += float(ref_u.pow(2).sum())
                assert float((got_u - ref_u).abs().max()) <= 2.01e-3, n
            assert (num / den) ** 0.5 <= 5e-2, (num / den) ** 0.5
        else:
            check_params(step, reference_update(sd, [mean], [1e-3]))
        assert step._next_bucket == len(step._buckets)
        dist.barrier()
        q.put((rank, "ok"))
    except Exception as e:      # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("comm_bf16", [False, True])
def test_two_rank_gloo_bucketed_allreduce_matches_mean_of_shard_gradients(comm_bf16):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + (977 if comm_bf16 else 0)) % 2000
    procs = [ctx.Process(target=_dp_worker, args=(r, world, port, q, comm_bf16)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), res


def _traj_worker(rank, world, port, q, comm_bf16):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        torch.set_num_threads(2)
        cfg = tiny_cfg()
        sd = synth_state_dict(param_spec(cfg), seed=3)
        step = PretrainStep(cfg, 3, device="cpu", state_dict=sd, ops=RefOps(torch.float32), training=False,
                            use_graph=False, lr=2e-3, process_group=dist.group.WORLD, bucket_mb=0.05,
                            comm_bf16=comm_bf16, **DIMS)
        losses = []
        for it in range(20):
            random.seed(500 + it)
            step.step(synth_batch(cfg, 3, DIMS["len_t"], DIMS["len_v"], DIMS["len_pv"], max_num_pv=4, seed=900 + 2 * it + rank))
            losses.append(step.read_losses()["loss"])
        q.put((rank, losses))
    except Exception:      # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def test_bf16_gradient_allreduce_tracks_fp32_allreduce_over_20_steps():
    """K3M_COMM_BF16 (the data-parallel default): 20 optimizer steps on two ranks with gradients all-reduced in bf16
    against the same run with the fp32 all-reduce — every step's training loss within 1e-3 relative."""
    world, ctx = 2, mp.get_context("spawn")
    traj = {}
    for mode in (False, True):
        q = ctx.Queue()
        port = 31500 + (os.getpid() + (313 if mode else 0)) % 2000
        procs = [ctx.Process(target=_traj_worker, args=(r, world, port, q, mode)) for r in range(world)]
        for p in procs:
            p.start()
        res = dict(q.get(timeout=900) for _ in range(world))
        for p in procs:
            p.join(60)
        assert all(isinstance(v, list) for v in res.values()), res
        traj[mode] = res
    for r in range(world):
        a, b = traj[False][r], traj[True][r]
        assert a[-1] < a[0]                       # the run trains
        for i, (x, y) in enumerate(zip(a, b)):
            assert abs(x - y) <= 1e-3 * abs(x), (r, i, x, y)
