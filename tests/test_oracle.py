"""CPU: the oracle restatement against (a) the committed golden fixtures made from the real
reference and (b) the live reference when /root/reference is mounted (build container only)."""
import random

import pytest
import torch

from oracle import ref_harness
from oracle.k3m_oracle import Oracle, total_loss, warmup_linear_lambda
from tests.util import case_setup, load_golden, sample_index

CASES = ["tiny_hard", "tiny_mean", "tiny_inter", "tiny_soft", "tiny_nopv"]


def _run_oracle(fx):
    cfg, sd, batch, noise = case_setup(fx)
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    random.seed(fx["seeds"]["python"])
    orc = Oracle(cfg, sdo)
    orc.taps = {}
    out = orc.forward(batch, noise)
    total_loss(out).backward()
    return out, sdo, orc.taps


@pytest.mark.parametrize("name", CASES + ["base_hard", "base_mean"])
def test_oracle_matches_golden(name):
    fx = load_golden(name)
    out, sdo, taps = _run_oracle(fx)
    # per-layer activations: the forward hooks on the unmodified reference's layer modules
    assert len(fx["taps"]) > 0
    for key, ref in fx["taps"].items():
        x = taps[key].detach().reshape(-1)
        assert abs(float(x.norm()) - ref["norm"]) <= 1e-5 * ref["norm"], key
        assert float((x[sample_index(x.numel())] - ref["sample"]).abs().max()) <= 1e-5 * float(ref["sample"].abs().max()) + 1e-6, key
    got = dict(loss_t=out[0], loss_v=out[1], loss_pv=out[3], loss_itm=out[6], loss_lpm=out[9])
    for k, ref in fx["losses"].items():
        assert abs(float(got[k]) - ref) <= 1e-5 * max(1.0, abs(ref)), (k, float(got[k]), ref)
    assert torch.allclose(out[7], fx["c_initial"], atol=1e-5)
    assert torch.allclose(out[8], fx["c_final"], atol=1e-5)
    n_none = 0
    for n, ref in fx["grads"].items():
        g = sdo[n].grad
        if ref is None:
            n_none += 1
            assert g is None or float(g.abs().max()) == 0.0, n
            continue
        g = g.reshape(-1)
        tol = 1e-5 * ref["maxabs"] + 1e-7
        assert float((g[sample_index(g.numel())] - ref["sample"]).abs().max()) <= tol, n
        assert abs(float(g.norm()) - ref["norm"]) <= 1e-4 * ref["norm"] + 1e-7, n
    if name in ("base_hard", "base_mean"):
        # 86 never-trained tensors; the 18 score_* gate tensors join them unless if_pre_sampling == 1
        assert n_none == (86 if name == "base_hard" else 104) and len(fx["grads"]) == 998
    if name == "tiny_soft":
        assert fx["grads"]["soft_t.weight"] is not None and fx["grads"]["score_self_v.weight"] is not None


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")
def test_oracle_matches_live_reference():
    import make_golden
    from k3m_b200.synth import FORWARD_ORDER
    fx = load_golden("tiny_hard")
    cfg, sd, batch, noise = case_setup(fx)
    make_golden.run_case  # the generator and this test share CASES
    ref, rcfg, model = ref_harness.make_reference_model(fx["config"], fx["overrides"])
    full = dict(sd)
    full["cls.predictions.decoder.weight"] = sd["embeddings.word_embeddings.weight"]
    model.load_state_dict(full)
    model.eval()
    # mode 0 avoids patching gumbel; weights / inputs identical
    model.if_pre_sampling = 0
    cfg.if_pre_sampling = 0
    random.seed(11)
    r = model(*[batch[k] for k in FORWARD_ORDER], input_ids_pv=batch["input_ids_pv"],
              token_type_ids_pv=batch["token_type_ids_pv"], attention_mask_pv=batch["attention_mask_pv"],
              masked_lm_labels_pv=batch["masked_lm_labels_pv"],
              next_sentence_label_pv_v=batch["next_sentence_label_pv_v"],
              next_sentence_label_pv_t=batch["next_sentence_label_pv_t"],
              index_p=batch["index_p"], index_v=batch["index_v"], device="cpu")
    random.seed(11)
    o = Oracle(cfg, sd).forward(batch, None)
    for i in (0, 1, 3, 6, 9, 7, 8):
        assert torch.allclose(r[i], o[i], atol=1e-6), i


def test_warmup_linear_schedule():
    assert warmup_linear_lambda(0, 10, 100) == 0.0
    assert warmup_linear_lambda(5, 10, 100) == 0.5
    assert warmup_linear_lambda(10, 10, 100) == 1.0
    assert abs(warmup_linear_lambda(55, 10, 100) - 0.5) < 1e-12
    assert warmup_linear_lambda(100, 10, 100) == 0.0
    assert warmup_linear_lambda(150, 10, 100) == 0.0


def test_warmup_linear_schedule_against_the_library_successor():
    """`WarmupLinearSchedule` of pytorch_transformers 1.1.0 (`pretrain.py:380-384,927-931`) is not in this image; its
    renamed successor in the same library line, `transformers.get_linear_schedule_with_warmup` (transformers 5.5 here),
    is.  Both the oracle's and the product's factor are held to the successor's LambdaLR over whole runs — learning
    rates of a real torch optimizer, step by step, incl. the degenerate warm-ups (0, 1, == total, > total).  This pins
    the schedule to the published successor, not to the 1.1.0 source (DESIGN.md section 4 says which)."""
    transformers = pytest.importorskip("transformers")
    from k3m_b200.optim import warmup_linear_factor
    for warm, total in [(0, 7), (1, 9), (10, 100), (37, 1000), (50, 50), (60, 50), (3, 4)]:
        p = torch.nn.Parameter(torch.zeros(1))
        opt = torch.optim.SGD([p], lr=1e-4)
        sch = transformers.get_linear_schedule_with_warmup(opt, num_warmup_steps=warm, num_training_steps=total)
        for step in range(total + 5):
            want = sch.get_last_lr()[0]
            assert want == 1e-4 * warmup_linear_lambda(step, warm, total), (warm, total, step)
            assert want == 1e-4 * warmup_linear_factor(step, warm, total), (warm, total, step)
            opt.step()
            sch.step()
