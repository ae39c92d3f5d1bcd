"""CPU: the item-alignment evaluation sweep and prediction dump (k3m_b200.finetune; SURVEY §8 f row N1)
against sklearn's metrics (what finetune.py:1119-1121 calls) and, through the class API on the fp32
restatement of the kernels, against the probabilities of the real reference stored in the golden fixture."""
import json

import numpy as np
import pytest
import torch

from k3m_b200 import finetune as ft
from oracle.ops_ref import RefOps
from tests.util import align_case_setup, load_golden


def test_metrics_match_sklearn():
    sk = pytest.importorskip("sklearn.metrics")
    rng = np.random.RandomState(0)
    for n in (1, 7, 200):
        for trial in range(5):
            labels = rng.randint(0, 2, n)
            probs = rng.rand(n).astype(np.float32)
            for th, p, r, f1 in ft.threshold_sweep(labels, probs):
                pred = probs >= th
                assert p == pytest.approx(sk.precision_score(labels, pred, zero_division=0), abs=1e-12)
                assert r == pytest.approx(sk.recall_score(labels, pred, zero_division=0), abs=1e-12)
                assert f1 == pytest.approx(sk.f1_score(labels, pred, zero_division=0), abs=1e-12)
    # degenerate: no positives predicted / no positive labels -> 0.0, as sklearn (with its warning)
    assert ft.precision_recall_f1([1, 0, 1], [0, 0, 0]) == (0.0, 0.0, 0.0)
    assert ft.precision_recall_f1([0, 0], [1, 0]) == (0.0, 0.0, 0.0)
    assert [round(r[0], 1) for r in ft.threshold_sweep([1], [0.5])] == [0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9]


def loader_batch(batch, labels, B):
    """the fixture's stacked items -> the 27-field tuple of the reference's pair loader"""
    out = [labels]
    for half in (slice(0, B), slice(B, 2 * B)):
        g = lambda k: batch[k][half]
        Lv = g("image_feat").shape[1] - 1
        out += [g("input_ids"), g("attention_mask"), g("token_type_ids"), g("input_ids_pv"), g("attention_mask_pv"),
                g("token_type_ids_pv"), g("image_feat"), g("image_loc"), torch.zeros(B, Lv, 1601),
                g("image_attention_mask")]
    out += [["src%d" % i for i in range(B)], batch["index_p"][:B].numpy(), batch["index_v"][:B].numpy(),
            ["tgt%d" % i for i in range(B)], batch["index_p"][B:].numpy(), batch["index_v"][B:].numpy()]
    return tuple(out)


@pytest.fixture(scope="module")
def model_and_batch():
    from k3m_b200.vilbert_k3m import K3MForItemAlignment
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, _ = align_case_setup(fx)
    model = K3MForItemAlignment(cfg, ops=RefOps(torch.float32))
    model.load_state_dict(sd)
    return fx, model, loader_batch(batch, labels, labels.shape[0]), labels


def test_evaluate_reproduces_reference_probabilities(model_and_batch):
    fx, model, lb, labels = model_and_batch
    model.train()
    lines = []
    probs, lab, rows = ft.evaluate(model, [lb, lb], log=lines.append)
    assert model.training                      # mode restored
    B = labels.shape[0]
    assert probs.shape == (2 * B,) and np.array_equal(lab, np.concatenate([labels.numpy()] * 2))
    assert np.abs(probs[:B] - fx["probs"][:, 1].numpy()).max() <= 1e-6
    assert np.array_equal(probs[:B], probs[B:])
    assert len(rows) == 9 and len(lines) == 9 and lines[0].startswith("threshold=0.1, precision=")
    for th, p, r, f1 in rows:
        assert (p, r, f1) == ft.precision_recall_f1(lab, probs >= th)


def test_predict_writes_the_reference_jsonl(model_and_batch, tmp_path):
    fx, model, lb, labels = model_and_batch
    path = str(tmp_path / "deepAI_result_threshold=0.5.jsonl")
    n = ft.predict(model, [lb], path, 0.5)
    B = labels.shape[0]
    assert n == B
    rows = [json.loads(l) for l in open(path, encoding="utf-8")]
    assert [r["src_item_id"] for r in rows] == ["src%d" % i for i in range(B)]
    assert [r["tgt_item_id"] for r in rows] == ["tgt%d" % i for i in range(B)]
    for i, r in enumerate(rows):
        assert set(r) == {"src_item_id", "src_item_emb", "tgt_item_id", "tgt_item_emb", "threshold"}
        assert r["threshold"] == 0.5
        # "ce" head: the two per-item slots are the class probabilities (vilbert_k3m.py:3454), printed as
        # str(np.float32) inside brackets
        assert abs(float(r["src_item_emb"][1:-1]) - float(fx["probs"][i, 0])) <= 1e-6
        assert abs(float(r["tgt_item_emb"][1:-1]) - float(fx["probs"][i, 1])) <= 1e-6
    vec = ft.prediction_record("a", "b", np.array([0.5, 1.25], np.float32), np.float32(2.0), 0.3)
    assert vec["src_item_emb"] == "[0.5,1.25]" and vec["tgt_item_emb"] == "[2.0]"


def test_split_rejects_wrong_layout():
    with pytest.raises(ValueError):
        ft.split_pair_batch(tuple([torch.zeros(1)] * 20) + ([], [], [], [], [], []))
