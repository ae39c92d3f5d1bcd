"""CPU: reference-layout checkpoints (k3m_b200.checkpoint; SURVEY §8 f row N4) against a real
torch.optim.AdamW + LambdaLR built the way pretrain.py:835-851,919-931 builds them.

* what `save_checkpoint` writes has the structure and the values of `optimizer.state_dict()` /
  `scheduler.state_dict()` of that optimizer after the same steps;
* a `.tar` written from the torch side (the reference's `torch.save` call, pretrain.py:1243-1251, with a
  DDP `module.` prefix) resumes here and the next step matches torch's next step;
* our own `.tar` loads into the torch optimizer (`optimizer.load_state_dict`) and into a fresh step."""
import random

import pytest
import torch

from k3m_b200 import checkpoint as ckpt
from k3m_b200.layout import is_no_decay, param_spec
from k3m_b200.synth import synth_batch, synth_state_dict
from k3m_b200.trainer import PretrainStep
from oracle.k3m_oracle import warmup_linear_lambda
from oracle.ops_ref import RefOps
from tests.test_trainer_cpu import DIMS, oracle_grads, tiny_cfg

LR, WARM, TOTAL, B = 1e-3, 2, 10, 3


def torch_side(sd):
    """model parameters + AdamW (two groups) + the LambdaLR of WarmupLinearSchedule, as the driver does"""
    params = {k: torch.nn.Parameter(v.clone()) for k, v in sd.items()}
    groups = [dict(params=[p for n, p in params.items() if not is_no_decay(n)], weight_decay=0.01),
              dict(params=[p for n, p in params.items() if is_no_decay(n)], weight_decay=0.0)]
    opt = torch.optim.AdamW(groups, lr=LR, betas=(0.9, 0.98), eps=1e-8)
    sch = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: warmup_linear_lambda(s, WARM, TOTAL))
    return params, opt, sch


def torch_step(cfg, params, opt, sch, batch, seed):
    cur = {k: v.data.clone() for k, v in params.items()}
    _, g = oracle_grads(cfg, cur, batch, seed)
    for n, p in params.items():
        p.grad = None if g[n] is None else g[n].clone()
    opt.step()
    sch.step()


def make_step(cfg, sd):
    return PretrainStep(cfg, B, device="cpu", state_dict=sd, ops=RefOps(torch.float32), training=False,
                        use_graph=False, lr=LR, warmup_steps=WARM, total_steps=TOTAL, **DIMS)


def batch_of(cfg, it):
    return synth_batch(cfg, B, DIMS["len_t"], DIMS["len_v"], DIMS["len_pv"], max_num_pv=4, seed=40 + it)


def run_ours(step, cfg, it):
    random.seed(9 + it)
    step.step(batch_of(cfg, it))


def close(a, b, tol, what):
    err = float((a - b).abs().max())
    assert err <= tol * max(1.0, float(b.abs().max())), (what, err)


@pytest.fixture(scope="module")
def two_steps():
    cfg = tiny_cfg()
    sd = synth_state_dict(param_spec(cfg), seed=3)
    step = make_step(cfg, sd)
    params, opt, sch = torch_side(sd)
    for it in range(2):
        run_ours(step, cfg, it)
        torch_step(cfg, params, opt, sch, batch_of(cfg, it), 9 + it)
    return cfg, sd, step, params, opt, sch


def test_saved_checkpoint_has_torch_layout_and_values(two_steps, tmp_path):
    cfg, sd, step, params, opt, sch = two_steps
    ck = ckpt.save_checkpoint(step, str(tmp_path / "e0.tar"), str(tmp_path / "e0.bin"))
    ck = torch.load(str(tmp_path / "e0.tar"), weights_only=False)
    assert set(ck) == {"model_state_dict", "optimizer_state_dict", "scheduler_state_dict", "global_step"}
    assert ck["global_step"] == 2
    want, got = opt.state_dict(), ck["optimizer_state_dict"]
    assert [pg["params"] for pg in got["param_groups"]] == [pg["params"] for pg in want["param_groups"]]
    for pg_g, pg_w in zip(got["param_groups"], want["param_groups"]):
        for k in ("weight_decay", "betas", "eps", "amsgrad", "initial_lr"):
            assert pg_g[k] == pg_w[k], k
        assert pg_g["lr"] == pytest.approx(pg_w["lr"], rel=1e-12)
    # state exists exactly for the parameters torch has state for (those that received a gradient)
    assert set(got["state"]) == set(want["state"])
    for i, st in want["state"].items():
        assert float(got["state"][i]["step"]) == float(st["step"]) == 2.0
        close(got["state"][i]["exp_avg"], st["exp_avg"], 2e-5, ("exp_avg", i))
        close(got["state"][i]["exp_avg_sq"], st["exp_avg_sq"], 2e-5, ("exp_avg_sq", i))
    s_w, s_g = sch.state_dict(), ck["scheduler_state_dict"]
    for k in ("last_epoch", "_step_count", "base_lrs"):
        assert s_g[k] == s_w[k], k
    assert s_g["_last_lr"] == pytest.approx(s_w["_last_lr"], rel=1e-12)
    # the weights-only file is the state_dict, tied decoder alias included
    w = torch.load(str(tmp_path / "e0.bin"))
    assert set(w) == set(ck["model_state_dict"]) and "cls.predictions.decoder.weight" in w
    for n, p in params.items():
        close(w[n], p.data, 1e-4, n)
    # and torch's own optimizer accepts it
    _, opt2, _ = torch_side(sd)
    opt2.load_state_dict(got)
    assert len(opt2.state_dict()["state"]) == len(want["state"])


def test_resume_from_a_torch_written_checkpoint(two_steps, tmp_path):
    cfg, sd, step, params, opt, sch = two_steps
    # what the reference writes after two optimizer steps, from a DDP-wrapped model
    path = str(tmp_path / "ref.tar")
    torch.save({"model_state_dict": {"module." + k: v.data for k, v in params.items()},
                "optimizer_state_dict": opt.state_dict(), "scheduler_state_dict": sch.state_dict(),
                "global_step": 2}, path)
    fresh = make_step(cfg, synth_state_dict(param_spec(cfg), seed=77))     # different weights, no history
    fresh.warmup_steps, fresh.total_steps = WARM, TOTAL
    assert ckpt.load_checkpoint(fresh, path) == 2
    # the third step continues both runs identically
    p3, o3, s3 = torch_side(sd)
    for n, p in p3.items():
        p.data.copy_(params[n].data)
    o3.load_state_dict(opt.state_dict())
    s3.load_state_dict(sch.state_dict())
    torch_step(cfg, p3, o3, s3, batch_of(cfg, 2), 11)
    run_ours(fresh, cfg, 2)
    got = fresh.state_dict()
    for n, p in p3.items():
        close(got[n], p.data, 1e-4, n)
    assert fresh.global_step == 3


def test_own_checkpoint_round_trip(two_steps, tmp_path):
    cfg, sd, step, params, opt, sch = two_steps
    path = str(tmp_path / "own.tar")
    ckpt.save_checkpoint(step, path)
    fresh = make_step(cfg, synth_state_dict(param_spec(cfg), seed=78))
    assert ckpt.load_checkpoint(fresh, path) == 2
    assert torch.equal(fresh.p32, step.p32) and torch.equal(fresh.m, step.m) and torch.equal(fresh.v, step.v)
    assert (fresh.lr, fresh.warmup_steps, fresh.total_steps) == (step.lr, step.warmup_steps, step.total_steps)


def test_rejects_mismatched_optimizer_state(two_steps):
    cfg, sd, step, params, opt, sch = two_steps
    ck = ckpt.checkpoint_dict(step)
    bad = dict(ck, optimizer_state_dict=dict(ck["optimizer_state_dict"]))
    bad["optimizer_state_dict"]["param_groups"] = ck["optimizer_state_dict"]["param_groups"][:1]
    with pytest.raises(ValueError):
        ckpt.load_checkpoint(make_step(cfg, sd), bad)
