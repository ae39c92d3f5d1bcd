"""CPU: the item-alignment path (`K3MForItemAlignment`, BASELINE configs[4]) — oracle restatement and
the engine schedule on the fp32 restatement of the kernels, against fixtures made from the real
reference (`tools/make_golden.py`, `ALIGN_CASES`)."""
import pytest
import torch

from oracle.k3m_oracle import Oracle
from oracle.ops_ref import RefOps
from tests.util import (ITEM_KEYS, align_case_setup, compare_alignment_with_golden, load_golden,
                        run_engine_alignment, sample_index)

CASES = ["align_tiny_mean", "align_tiny_hard"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_alignment_matches_reference(name):
    fx = load_golden(name)
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    B = labels.shape[0]
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    it1 = {k: batch[k][:B] for k in ITEM_KEYS}
    it2 = {k: batch[k][B:] for k in ITEM_KEYS}
    n1 = n2 = None
    if noise is not None:
        n1 = {m: x[:B] for m, x in noise.items()}
        n2 = {m: x[B:] for m, x in noise.items()}
    probs, loss, f1, f2 = Oracle(cfg, sdo).forward_alignment(labels, it1, it2, n1, n2)
    loss.backward()
    assert abs(float(loss) - fx["loss"]) <= 1e-5 * abs(fx["loss"])
    assert float((probs.detach() - fx["probs"]).abs().max()) <= 1e-6
    assert torch.allclose(torch.cat((f1, f2)).detach(), fx["c_final"], rtol=1e-4, atol=1e-5)
    for n, ref in fx["grads"].items():
        g = sdo[n].grad
        if ref is None:
            assert g is None or float(g.abs().max()) == 0.0, n
            continue
        if ref["maxabs"] < 1e-9:
            continue
        got = g.reshape(-1)[sample_index(g.numel())]
        assert float((got - ref["sample"]).norm()) <= 2e-4 * (float(ref["sample"].norm()) + 1e-3 * ref["maxabs"]), n


@pytest.mark.parametrize("name", CASES)
def test_alignment_schedule_matches_reference_fp32(name):
    fx = load_golden(name)
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    out, L, g = run_engine_alignment(cfg, sd, batch, labels, noise, RefOps(torch.float32), "cpu", torch.float32)
    errs = compare_alignment_with_golden(fx, out, L, g, loss_rtol=1e-5, grad_tol=2e-4, prob_atol=1e-6)
    assert not errs, errs[:10]


def test_alignment_class_api_matches_reference():
    """`K3MForItemAlignment(config)(labels, <11 tensors of item 1>, <11 of item 2>)` -> the reference's
    (probs[:, 0], probs[:, 1], probs[:, 1], loss), parameters in the reference's order."""
    from k3m_b200.vilbert_k3m import K3MForItemAlignment
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    B = labels.shape[0]
    model = K3MForItemAlignment(cfg, ops=RefOps(torch.float32))
    assert [n for n, _ in model.named_parameters()] == list(fx["grads"].keys())
    assert list(model.state_dict().keys()) == list(fx["grads"].keys())
    model.load_state_dict(sd)
    model.eval()
    args = [labels]
    for half in (slice(0, B), slice(B, 2 * B)):
        args += [batch[k][half] for k in ITEM_KEYS]
    e1, e2, probs, loss = model(*args)
    assert abs(float(loss) - fx["loss"]) <= 1e-5 * abs(fx["loss"])
    assert float((probs.detach() - fx["probs"][:, 1]).abs().max()) <= 1e-6
    assert torch.equal(e2, probs) and float((e1.detach() - fx["probs"][:, 0]).abs().max()) <= 1e-6
    (2.0 * loss).backward()
    for n, p in model.named_parameters():
        ref = fx["grads"][n]
        if ref is None:
            assert p.grad is None, n
        elif ref["maxabs"] >= 1e-9:
            got = p.grad.reshape(-1)[sample_index(p.numel())]
            assert float((got - 2.0 * ref["sample"]).norm()) <= 4e-4 * (float(ref["sample"].norm()) + 1e-3 * ref["maxabs"]), n
    with torch.no_grad():
        again = model(*args)
    assert again[3].grad_fn is None and abs(float(again[3]) - fx["loss"]) <= 1e-5 * abs(fx["loss"])


def test_alignment_step_object_matches_oracle_and_adamw():
    """`AlignmentStep` (the fused finetune step: stacked towers, pair loss, AdamW groups) on the fp32
    restatement: two optimizer steps against oracle gradients + torch.optim.AdamW."""
    from k3m_b200.trainer import AlignmentStep
    from tests.test_trainer_cpu import check_params, reference_update
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    B, Lt, Lv, Lpv, mpv = fx["dims"]
    step = AlignmentStep(cfg, B, len_t=Lt, len_v=Lv, len_pv=Lpv, max_num_pv=mpv, device="cpu", state_dict=sd,
                         training=False, lr=1e-3, ops=RefOps(torch.float32), use_graph=False)
    assert set(step.state_dict().keys()) == set(sd.keys())
    hb = {k: batch[k] for k in ITEM_KEYS}
    hb["labels"] = labels
    cur, grads, want_losses = sd, [], []
    for it in range(2):
        sdo = {k: v.clone().requires_grad_(True) for k, v in cur.items()}
        it1 = {k: batch[k][:B] for k in ITEM_KEYS}
        it2 = {k: batch[k][B:] for k in ITEM_KEYS}
        probs, loss, _, _ = Oracle(cfg, sdo).forward_alignment(labels, it1, it2)
        loss.backward()
        grads.append({k: v.grad for k, v in sdo.items()})
        want_losses.append(float(loss))
        cur = reference_update(sd, grads, [1e-3] * len(grads))
        step.step(hb)
        got = step.read_losses()
        assert abs(got["loss"] - want_losses[-1]) <= 1e-5 * abs(want_losses[-1]), (it, got, want_losses)
        assert float((step.probs - probs.detach()).abs().max()) <= 1e-5
        check_params(step, cur)
