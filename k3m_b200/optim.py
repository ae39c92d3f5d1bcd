"""Optimizer-side host logic of the K3M pretraining step.

`warmup_linear_factor` restates `pytorch_transformers==1.1.0` `WarmupLinearSchedule` (third-party,
not vendored in the reference; call sites `pretrain.py:380-384,927-931`): a LambdaLR whose factor is
`step / max(1, warmup)` during warm-up and `max(0, (T - step) / max(1, T - warmup))` afterwards.
The 1.1.0 source is absent here (the reference holds no test for it); the factor is held to the renamed successor
`transformers.get_linear_schedule_with_warmup` in `tests/test_oracle.py`.

`param_groups` reproduces the from-scratch grouping rule of `pretrain.py:835-851`.
"""
from .layout import is_no_decay


def warmup_linear_factor(step, warmup_steps, t_total):
    if step < warmup_steps:
        return float(step) / float(max(1, warmup_steps))
    return max(0.0, float(t_total - step) / float(max(1.0, t_total - warmup_steps)))


def param_groups(names, weight_decay=0.01):
    """-> [(names with decay, weight_decay), (names without, 0.0)] — substring rule on the name."""
    decay = [n for n in names if not is_no_decay(n)]
    no_decay = [n for n in names if is_no_decay(n)]
    return [(decay, weight_decay), (no_decay, 0.0)]
