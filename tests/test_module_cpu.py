"""CPU: the drop-in class API (`k3m_b200.vilbert_k3m.BertForMultiModalPreTraining_tri_stru`) — the
reference's constructor / forward signature / 10-tuple / parameter names / state_dict / checkpoint
entry points — with the fp32 restatement of the kernels injected, against the golden fixtures made
from the real reference (`tools/make_golden.py`)."""
import os
import random
import sys

import pytest
import torch

from k3m_b200.vilbert_k3m import BertConfig, BertForMultiModalPreTraining_tri_stru, install_as_reference
from oracle.ops_ref import RefOps
from tests.util import case_setup, compare_with_golden, load_golden


def call(model, b):
    """positional / keyword split of the reference's driver, `pretrain.py:990-1011`"""
    return model(b["input_ids"], b["image_feat"], b["image_loc"], b["token_type_ids"], b["attention_mask"],
                 b["image_attention_mask"], b["masked_lm_labels"], b["image_label"], b["image_target"],
                 b["next_sentence_label"], False, b["input_ids_pv"], b["token_type_ids_pv"], b["attention_mask_pv"],
                 b["masked_lm_labels_pv"], b["next_sentence_label_pv_v"], b["next_sentence_label_pv_t"],
                 b["index_p"], b["index_v"], device="cpu")


def make(name):
    fx = load_golden(name)
    cfg, sd, batch, noise = case_setup(fx)
    model = BertForMultiModalPreTraining_tri_stru(cfg, ops=RefOps(torch.float32))
    res = model.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys and res.missing_keys == ["cls.predictions.decoder.weight"]
    return fx, cfg, sd, batch, model


def test_names_and_state_dict_follow_the_reference():
    fx, cfg, sd, batch, model = make("tiny_mean")
    names = [n for n, _ in model.named_parameters()]
    assert names == list(fx["grads"].keys())            # the reference's named_parameters() order
    keys = list(model.state_dict().keys())
    assert len(keys) == len(names) + 1 and "cls.predictions.decoder.weight" in keys
    assert model.state_dict()["cls.predictions.decoder.weight"].data_ptr() == \
        model.state_dict()["embeddings.word_embeddings.weight"].data_ptr()
    with pytest.raises(ValueError):
        BertForMultiModalPreTraining_tri_stru({"hidden_size": 8})


def test_forward_backward_match_reference_through_class_api():
    fx, cfg, sd, batch, model = make("tiny_mean")
    model.eval()                                         # dropout off, like the fixture
    random.seed(fx["seeds"]["python"])
    out = call(model, batch)
    assert len(out) == 10 and out[2] == 0 and out[4] == 0 and out[5] == 0
    assert all(tuple(out[i].shape) == (1,) for i in (0, 1, 3, 6, 9))
    loss = out[0] + out[1] + out[3] + out[9]             # `pretrain.py:1055` with loss_img_weight 1
    loss.backward()
    got = dict(loss_t=out[0], loss_v=out[1], loss_pv=out[3], loss_itm=out[6], loss_lpm=out[9],
               c_initial=out[7], c_final=out[8])
    errs, _ = compare_with_golden(fx, got, model._layout, model._g32, loss_rtol=1e-5, grad_tol=2e-4)
    assert not errs, errs[:10]
    for n, p in model.named_parameters():
        if fx["grads"][n] is None:
            assert p.grad is None, n                     # the 86 never-trained tensors
        else:
            assert p.grad is not None and p.grad.shape == p.shape, n


def test_upstream_gradients_scale_each_loss_and_grads_accumulate():
    fx, cfg, sd, batch, model = make("tiny_mean")
    model.eval()
    name = "encoder.layer.0.output.dense.weight"
    p = dict(model.named_parameters())[name]

    def grad_of(weights):
        model.zero_grad(set_to_none=True)
        random.seed(3)
        out = call(model, batch)
        (weights[0] * out[0] + weights[1] * out[1] + weights[2] * out[3] + weights[3] * out[9]).backward()
        return p.grad.clone()

    g_all = grad_of((1.0, 1.0, 1.0, 1.0))
    parts = [grad_of(tuple(1.0 if i == j else 0.0 for i in range(4))) for j in range(4)]
    assert torch.allclose(sum(parts), g_all, rtol=1e-4, atol=1e-7 * float(g_all.abs().max()) + 1e-9)
    g_half = grad_of((1.0, 0.5, 1.0, 1.0))
    assert torch.allclose(g_half, g_all - 0.5 * parts[1], rtol=1e-4, atol=1e-6 * float(g_all.abs().max()))
    # no zero_grad in between: gradients accumulate like autograd's
    random.seed(3)
    out = call(model, batch)
    (out[0] + 0.5 * out[1] + out[3] + out[9]).backward()
    assert torch.allclose(p.grad, 2 * g_half, rtol=1e-4, atol=1e-6 * float(g_all.abs().max()))


def test_torch_optimizer_and_no_grad_path():
    fx, cfg, sd, batch, model = make("tiny_mean")
    model.train()
    decay = [p for n, p in model.named_parameters() if "bias" not in n]
    opt = torch.optim.AdamW(decay, lr=1e-3)
    before = model.state_dict()["struc_w1.weight"].clone()
    random.seed(1)
    out = call(model, batch)
    (out[0] + out[1] + out[3] + out[9]).backward()
    opt.step()
    assert not torch.equal(before, model.state_dict()["struc_w1.weight"])
    with torch.no_grad():
        out2 = call(model, batch)
    assert out2[0].grad_fn is None and model._engine.saved is None
    assert float(out2[0]) != float(out[0])               # the update reached the kernels' shadow copy


def test_checkpoint_round_trip(tmp_path):
    fx, cfg, sd, batch, model = make("tiny_mean")
    model.save_pretrained(str(tmp_path))
    loaded = BertForMultiModalPreTraining_tri_stru.from_pretrained(str(tmp_path), config=cfg, default_gpu=True)
    assert loaded is not None and not loaded.training
    for (n1, a), (n2, b) in zip(model.state_dict().items(), loaded.state_dict().items()):
        assert n1 == n2 and torch.equal(a, b)
    # DDP-saved checkpoints carry `module.` (pretrain.py:1238-1251)
    torch.save({"module." + k: v for k, v in sd.items()}, os.path.join(str(tmp_path), "ddp.bin"))
    again = BertForMultiModalPreTraining_tri_stru.from_pretrained(os.path.join(str(tmp_path), "ddp.bin"), config=cfg)
    assert torch.equal(again.state_dict()["struc_w3.bias"], sd["struc_w3.bias"])
    assert BertForMultiModalPreTraining_tri_stru.from_pretrained(os.path.join(str(tmp_path), "nope"), config=cfg) is None


def test_install_as_reference_module_path():
    install_as_reference()
    from vilbert_k3m.vilbert_k3m import BertConfig as C2, BertForMultiModalPreTraining_tri_stru as M2
    assert C2 is BertConfig and M2 is BertForMultiModalPreTraining_tri_stru
    for k in ("vilbert_k3m", "vilbert_k3m.vilbert_k3m"):
        sys.modules.pop(k, None)


def test_product_path_refuses_cpu():
    cfg = BertConfig.from_dict(load_golden("tiny_mean")["config"])
    model = BertForMultiModalPreTraining_tri_stru(cfg)
    fx, _, _, batch, _ = make("tiny_mean")
    with pytest.raises(RuntimeError):
        call(model, batch)
