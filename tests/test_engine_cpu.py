"""CPU: the hand-written forward/backward schedule (k3m_b200.engine) driven by the PyTorch
restatement of the kernel contracts (oracle/ops_ref.py, fp32) must reproduce the reference's
losses and all gradients (golden fixtures from the real reference)."""
import pytest
import torch

from oracle.ops_ref import RefOps
from tests.util import case_setup, compare_taps, compare_with_golden, load_golden, run_engine


@pytest.mark.parametrize("name", ["tiny_hard", "tiny_mean", "tiny_inter", "tiny_soft", "tiny_nopv", "base_hard",
                                  "base_mean", "long_hard"])
def test_schedule_matches_reference_fp32(name):
    fx = load_golden(name)
    cfg, sd, batch, noise = case_setup(fx)
    taps = {}
    out, L, g = run_engine(cfg, sd, batch, noise, RefOps(torch.float32), "cpu", torch.float32,
                           fx["seeds"]["python"], taps=taps)
    errs, worst = compare_with_golden(fx, out, L, g, loss_rtol=1e-5, grad_tol=2e-4)
    assert not errs, errs[:10]
    # every BertLayer / BertImageLayer / BertConnectionLayer output of the unmodified reference (forward hooks)
    terrs, stats = compare_taps(fx, taps, 1e-5)
    assert not terrs, terrs[:10]
    assert len(stats) == len(fx["taps"]) > 0


def test_dropout_mask_is_regenerated_consistently():
    """training mode: two runs with the same (seed, step) give identical gradients, a different
    step gives different ones (dropout masks are a pure function of the RNG state)."""
    fx = load_golden("tiny_mean")
    cfg, sd, batch, noise = case_setup(fx)
    ops = RefOps(torch.float32)
    o1, L, g1 = run_engine(cfg, sd, batch, noise, ops, "cpu", torch.float32, 1, training=True)
    o2, _, g2 = run_engine(cfg, sd, batch, noise, ops, "cpu", torch.float32, 1, training=True)
    assert torch.equal(g1, g2)
    o3, _, g3 = run_engine(cfg, sd, batch, noise, ops, "cpu", torch.float32, 1, training=False)
    assert not torch.equal(g1, g3)
    assert abs(float(o1["loss_t"]) - float(o3["loss_t"])) > 1e-6


def test_dense_and_gathered_mlm_head_agree():
    """The masked-LM head on the labelled rows only (`build_mlm_plan`, the product path) gives the losses and gradients
    of the reference's all-position decoder + ignore_index."""
    fx = load_golden("tiny_hard")
    cfg, sd, batch, noise = case_setup(fx)
    o1, L, g1 = run_engine(cfg, sd, batch, noise, RefOps(torch.float32), "cpu", torch.float32, fx["seeds"]["python"])
    o2, _, g2 = run_engine(cfg, sd, batch, noise, RefOps(torch.float32), "cpu", torch.float32, fx["seeds"]["python"],
                           gather_mlm=False)
    for k in ("loss_t", "loss_pv"):
        assert float(o1[k]) == pytest.approx(float(o2[k]), rel=1e-6)
    assert float((g1 - g2).abs().max()) <= 1e-6 * float(g2.abs().max())
