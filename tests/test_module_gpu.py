"""GPU: the drop-in class API on the CUDA op table — `.cuda()` moves the parameter arena, forward returns
the reference's tuples, `loss.backward()` fills `.grad` views of the fp32 gradient arena, a stock
`torch.optim.AdamW` step changes what the kernels read.  Fixtures come from the unmodified reference."""
import random

import pytest
import torch

from tests.test_module_cpu import call
from tests.util import (ITEM_KEYS, align_case_setup, case_setup, compare_alignment_with_golden,
                        compare_with_golden, load_golden)

pytestmark = pytest.mark.gpu


def test_pretrain_class_on_cuda_matches_reference():
    from k3m_b200.vilbert_k3m import BertForMultiModalPreTraining_tri_stru
    fx = load_golden("tiny_mean")
    cfg, sd, batch, noise = case_setup(fx)
    model = BertForMultiModalPreTraining_tri_stru(cfg)
    model.load_state_dict(sd, strict=False)
    model = model.cuda()
    assert all(p.is_cuda for p in model.parameters())
    model.eval()
    random.seed(fx["seeds"]["python"])
    b = {k: v.cuda() for k, v in batch.items()}
    out = call(model, b)
    (out[0] + out[1] + out[3] + out[9]).backward()
    torch.cuda.synchronize()
    got = dict(loss_t=out[0], loss_v=out[1], loss_pv=out[3], loss_itm=out[6], loss_lpm=out[9],
               c_initial=out[7], c_final=out[8])
    errs, _ = compare_with_golden(fx, got, model._layout, model._g32, loss_rtol=1e-3, grad_tol=8e-2,
                                  key_bias_atol=1e-3, itm_rtol=1e-2)
    assert not errs, errs[:10]
    # a stock optimizer over the arena-backed parameters moves the loss of the next forward
    model.train()
    opt = torch.optim.AdamW([p for p in model.parameters()], lr=1e-3)
    first = None
    for it in range(3):
        opt.zero_grad()
        random.seed(5)
        o = call(model, b)
        loss = o[0] + o[1] + o[3] + o[9]
        loss.backward()
        opt.step()
        first = float(loss) if first is None else first
    assert torch.isfinite(loss) and float(loss) < first
    frozen = [n for n, p in model.named_parameters() if p.grad is None]
    assert len(frozen) == sum(1 for v in fx["grads"].values() if v is None)


def test_alignment_class_on_cuda_matches_reference():
    from k3m_b200.vilbert_k3m import K3MForItemAlignment
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    B = labels.shape[0]
    model = K3MForItemAlignment(cfg)
    model.load_state_dict(sd)
    model = model.cuda().eval()
    args = [labels.cuda()]
    for half in (slice(0, B), slice(B, 2 * B)):
        args += [batch[k][half].cuda() for k in ITEM_KEYS]
    e1, e2, probs, loss = model(*args)
    loss.backward()
    torch.cuda.synchronize()
    out = dict(loss=loss, probs=torch.stack((e1, probs), dim=1), c_initial=fx["c_initial"], c_final=fx["c_final"])
    errs = compare_alignment_with_golden(fx, out, model._layout, model._g32, loss_rtol=1e-3, grad_tol=0.1,
                                         prob_atol=2e-3)
    assert not errs, errs[:10]


@pytest.mark.parametrize("lens", [(128, 100, 128), (50, 36, 256), (200, 140, 256)])
def test_long_context_dims_run_and_match_restatement(lens):
    """BASELINE configs[3] geometry (128-token titles, 100 regions + 1, 128 PV tokens), the finetune scripts'
    256 PV tokens, and every stream beyond 128 tokens (the blocked attention kernels in all nine attention
    sites) on the tiny model: CUDA kernels against the bf16-storage restatement on the same device."""
    from k3m_b200.config import BertConfig, K3M_TINY
    from k3m_b200.layout import param_spec
    from k3m_b200.ops import CudaOps
    from k3m_b200.synth import synth_batch, synth_state_dict
    from oracle.ops_ref import RefOps
    from tests.util import run_engine
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 0, 1.0, 4
    cfg.max_position_embeddings = 256
    sd = synth_state_dict(param_spec(cfg), seed=5)
    batch = synth_batch(cfg, 3, lens[0], lens[1], lens[2], max_num_pv=8, seed=31)
    out, L, g = run_engine(cfg, sd, batch, None, CudaOps("cuda"), "cuda", torch.bfloat16, 7)
    outr, _, gr = run_engine(cfg, sd, batch, None, RefOps(torch.bfloat16), "cuda", torch.bfloat16, 7)
    for k in ("loss_t", "loss_v", "loss_pv", "loss_lpm"):
        assert float(out[k]) == pytest.approx(float(outr[k]), rel=1e-3), k
    rel = []
    for n, (off, num, _) in L.offsets.items():
        a, b = gr[off:off + num], g[off:off + num]
        if float(a.norm()) > 1e-7:
            rel.append(float((a - b).norm() / a.norm()))
    rel.sort()
    assert rel[len(rel) // 2] <= 2e-2, rel[len(rel) // 2]


def test_class_api_graph_replay_matches_eager_launches():
    """The drop-in class replays two CUDA graphs from the third call of a geometry on (call 1 eager, call 2 captures):
    losses and every gradient of a replayed step equal the eagerly launched step (eval mode: no dropout)."""
    from k3m_b200.config import BertConfig, K3M_TINY
    from k3m_b200.synth import FORWARD_ORDER, synth_batch
    from k3m_b200.vilbert_k3m import BertForMultiModalPreTraining_tri_stru, _GraphedStep
    import random
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 0, 1.0, 4
    torch.manual_seed(1)
    model = BertForMultiModalPreTraining_tri_stru(cfg).cuda()
    model.eval()
    batches = [synth_batch(cfg, 4, 12, 9, 20, max_num_pv=4, seed=31 + i) for i in range(2)]

    def step(b, seed):
        random.seed(seed)
        model.zero_grad(set_to_none=True)
        out = model(*[b[k] for k in FORWARD_ORDER], input_ids_pv=b["input_ids_pv"], token_type_ids_pv=b["token_type_ids_pv"],
                    attention_mask_pv=b["attention_mask_pv"], masked_lm_labels_pv=b["masked_lm_labels_pv"],
                    next_sentence_label_pv_v=b["next_sentence_label_pv_v"], next_sentence_label_pv_t=b["next_sentence_label_pv_t"],
                    index_p=b["index_p"], index_v=b["index_v"])
        (out[0] + 0.5 * out[1] + out[3] + out[9]).backward()
        torch.cuda.synchronize()
        return [float(out[i]) for i in (0, 1, 3, 6, 9)], model._g32.clone(), out[8].clone()

    l0, g0, c0 = step(batches[0], 5)            # eager
    step(batches[1], 6)                         # captures
    assert any(isinstance(v, _GraphedStep) for v in model._graphs.values())
    l2, g2, c2 = step(batches[0], 5)            # replay, same batch and negatives as the eager step
    assert l2 == pytest.approx(l0, rel=1e-6)
    assert torch.equal(c0, c2)
    assert float((g0 - g2).abs().max()) <= 1e-5 * float(g0.abs().max())
    assert float(g0.abs().max()) > 0
