"""Checkpoints in the reference's on-disk layout, and the resume path the reference left commented out
(SURVEY §8 f, row N4).

The reference writes, per epoch (`pretrain.py:1238-1251`):
  * `<name>.bin` — `model.state_dict()` (999 keys for the pretraining model, `module.`-prefixed when the
    model was wrapped in DDP and saved without unwrapping; both spellings are accepted on load,
    `pretrain.py:745-762`);
  * `<name>.tar` — `{"model_state_dict", "optimizer_state_dict", "scheduler_state_dict", "global_step"}`,
    the optimizer being `torch.optim.AdamW` over the two groups of `pretrain.py:835-851` (decay, then
    no-decay, each in `named_parameters()` order) and the scheduler `WarmupLinearSchedule` (a `LambdaLR`).
Its resume code (`pretrain.py:764-779`) is commented out; `load_checkpoint` is that code, working.

`torch.optim.AdamW.state_dict()` numbers parameters 0..n-1 across the groups in order and holds, for every
parameter that has received a gradient, `{"step": fp32 scalar, "exp_avg", "exp_avg_sq"}`.  Here the moments
live in two flat fp32 arenas parallel to the parameter arena (`PretrainStep.m`, `.v`) and the step count is
`PretrainStep.global_step` for every trained parameter, so the translation is a set of views.  Parameters that
never receive a gradient (`layout.no_grad_names`, 86 tensors) have no optimizer state in either direction.
"""
from collections import OrderedDict

import torch

from . import optim


def _param_order(step):
    """Parameter names in torch.optim.AdamW index order: decay group, then no-decay group."""
    names = list(step.layout.spec)
    groups = optim.param_groups(names, step.weight_decay)
    return groups, [n for g, _ in groups for n in g]


def _trained(step, name):
    off, num, _ = step.layout.offsets[name]
    return off < step.layout.trained_end


def _current_lr(step):
    f = 1.0
    if step.total_steps > 0:
        f = optim.warmup_linear_factor(step.global_step, step.warmup_steps, step.total_steps)
    return step.lr * f


def optimizer_state_dict(step):
    """`torch.optim.AdamW.state_dict()` of the run so far (tensors are copies on the CPU)."""
    groups, order = _param_order(step)
    state = {}
    if step.global_step > 0:
        for i, n in enumerate(order):
            if not _trained(step, n):
                continue
            off, num, shape = step.layout.offsets[n]
            state[i] = {"step": torch.tensor(float(step.global_step)),
                        "exp_avg": step.m[off:off + num].detach().view(*shape).cpu().clone(),
                        "exp_avg_sq": step.v[off:off + num].detach().view(*shape).cpu().clone()}
    pgs, base = [], 0
    for g, wd in groups:
        pgs.append({"weight_decay": wd, "lr": _current_lr(step), "betas": tuple(step.betas), "eps": step.eps,
                    "amsgrad": False, "maximize": False, "foreach": None, "capturable": False,
                    "differentiable": False, "fused": None, "decoupled_weight_decay": True,
                    "initial_lr": step.lr, "params": list(range(base, base + len(g)))})
        base += len(g)
    return {"state": state, "param_groups": pgs}


def load_optimizer_state_dict(step, osd):
    """Fills the moment arenas from a `torch.optim.AdamW.state_dict()` whose parameter numbering follows
    the reference's grouping.  Returns the step count found in the state (0 when the state is empty)."""
    groups, order = _param_order(step)
    pgs = osd["param_groups"]
    if [len(pg["params"]) for pg in pgs] != [len(g) for g, _ in groups]:
        raise ValueError("optimizer_state_dict groups %s do not match this model's %s"
                         % ([len(pg["params"]) for pg in pgs], [len(g) for g, _ in groups]))
    ids = [i for pg in pgs for i in pg["params"]]
    step.m.zero_()
    step.v.zero_()
    steps = set()
    for pos, n in enumerate(order):
        st = osd["state"].get(ids[pos])
        if st is None:
            continue
        off, num, shape = step.layout.offsets[n]
        if tuple(st["exp_avg"].shape) != tuple(shape):
            raise ValueError("%s: optimizer state shape %s, expected %s" % (n, tuple(st["exp_avg"].shape), shape))
        if not _trained(step, n):
            continue        # a state entry for a parameter this step never updates: nothing to restore
        step.m[off:off + num].copy_(st["exp_avg"].reshape(-1).to(torch.float32))
        step.v[off:off + num].copy_(st["exp_avg_sq"].reshape(-1).to(torch.float32))
        steps.add(int(float(st["step"])))
    if len(steps) > 1:
        raise ValueError("optimizer state holds different step counts %s: one fused AdamW step per bucket "
                         "needs a common count" % sorted(steps))
    pg = pgs[0]
    step.betas, step.eps = tuple(pg["betas"]), pg["eps"]
    step.lr = pg.get("initial_lr", pg["lr"])
    step.weight_decay = pgs[0]["weight_decay"]
    return steps.pop() if steps else 0


def scheduler_state_dict(step):
    """`LambdaLR.state_dict()` of `WarmupLinearSchedule` after `global_step` scheduler steps."""
    lr = _current_lr(step)
    return {"warmup_steps": step.warmup_steps, "t_total": step.total_steps, "base_lrs": [step.lr, step.lr],
            "last_epoch": step.global_step, "verbose": False, "_step_count": step.global_step + 1,
            "_get_lr_called_within_step": False, "_last_lr": [lr, lr], "lr_lambdas": [None, None]}


def checkpoint_dict(step):
    sd = OrderedDict((k, v.detach().cpu().clone()) for k, v in step.state_dict().items())
    return {"model_state_dict": sd, "optimizer_state_dict": optimizer_state_dict(step),
            "scheduler_state_dict": scheduler_state_dict(step), "global_step": step.global_step}


def save_checkpoint(step, tar_path, bin_path=None):
    """Writes the reference's `.tar` checkpoint (and the weights-only `.bin` when asked).  In fused data-parallel
    mode this is a collective (every rank calls it: the owners of the optimizer shards broadcast them first)."""
    if hasattr(step, "sync_shards"):
        step.sync_shards(moments=True)
    ck = checkpoint_dict(step)
    torch.save(ck, tar_path)
    if bin_path is not None:
        torch.save(ck["model_state_dict"], bin_path)
    return ck


def load_checkpoint(step, path_or_dict, map_location="cpu"):
    """Resume: weights, AdamW moments, schedule position and global step from a reference-layout `.tar`
    (written by the reference's drivers or by `save_checkpoint`).  Returns the global step."""
    ck = path_or_dict
    if not isinstance(ck, dict):
        ck = torch.load(path_or_dict, map_location=map_location, weights_only=False)
    step.load_state_dict(ck["model_state_dict"])
    n = load_optimizer_state_dict(step, ck["optimizer_state_dict"])
    gs = int(ck.get("global_step", n))
    sch = ck.get("scheduler_state_dict") or {}
    if "t_total" in sch:
        step.warmup_steps, step.total_steps = sch["warmup_steps"], sch["t_total"]
    if "base_lrs" in sch:
        step.lr = sch["base_lrs"][0]
    # AdamW's bias correction uses the per-parameter step count, the schedule the scheduler's epoch; the
    # reference advances both once per optimizer step, so they agree unless gradient accumulation was used
    step.global_step = n if n else gs
    if sch.get("last_epoch", step.global_step) != step.global_step:
        raise ValueError("scheduler epoch %s and optimizer step %s differ: not a checkpoint of this driver"
                         % (sch.get("last_epoch"), step.global_step))
    # the dropout / Gumbel counter stream advances once per step (trainer._device_step): continue it, so a
    # resumed run with the same seed draws the masks the uninterrupted run would have drawn
    step.rng[1] = step.global_step
    return step.global_step
