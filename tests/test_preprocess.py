"""Batch preprocessing (SURVEY §8 f row N2).  CPU: the numpy restatement (oracle/preprocess_ref.py) against the
golden outputs of the real reference (`tools/make_golden_preprocess.py`: BertPreprocessBatch + the loader's
prepend, with recorded random draws), bit for bit.  GPU: the CUDA kernels through the C ABI against the same
golden outputs and, on larger random batches, against the restatement — bit-exact for ids / masks / labels /
span indices AND for the float32 features and locations (same IEEE operations in the same order)."""
import numpy as np
import pytest
import torch

from oracle.preprocess_ref import preprocess_batch
from tests.util import load_golden

CASES = ("preprocess_small", "preprocess_script")


def run_oracle(fx):
    d = fx["dims"]
    return preprocess_batch(fx["raws"], fx["u_tok"], fx["rnd_tok"], fx["u_reg"], d["Lt"], d["Lpv"], d["R"],
                            d["max_num_pv"])


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_bit_for_bit(name):
    fx = load_golden(name)
    out = run_oracle(fx)
    assert len(out) == 15
    for k, v in out.items():
        w = fx["want"][k].numpy()
        assert v.shape == w.shape and v.dtype == w.dtype, (k, v.dtype, w.dtype)
        assert np.array_equal(v, w), k
    # the fixture exercises what it should: all three title outcomes, masked and kept regions, malformed PV rows
    lab, ids = out["masked_lm_labels"], out["input_ids"]
    assert (lab >= 0).any() and ((lab >= 0) & (ids == 103)).any() and ((lab >= 0) & (ids == lab)).any()
    assert ((lab >= 0) & (ids != 103) & (ids != lab)).any()
    assert (out["image_label"] == 1).any() and (out["index_p"].reshape(len(lab), -1).sum(1) == 0).any()


def synth_raws(rng, B, Lt, Lpv, R, F, T, max_num_pv):
    from tools.make_golden_preprocess import pv_string
    raws = []
    for b in range(B):
        n = int(rng.randint(1, R + 1))
        w, h = int(rng.randint(200, 900)), int(rng.randint(200, 900))
        x1, y1 = rng.rand(n) * w * 0.6, rng.rand(n) * h * 0.6
        boxes = np.stack([x1, y1, x1 + rng.rand(n) * w * 0.4, y1 + rng.rand(n) * h * 0.4], 1).astype(np.float32)
        raws.append(dict(tokens=[int(t) for t in rng.randint(1000, 21128, rng.randint(0, Lt - 1))],
                         tokens_pv=pv_string(rng, int(rng.randint(0, max_num_pv + 3)), b % 7)[:Lpv - 2],
                         boxes=boxes, image_w=w, image_h=h, features=rng.randn(n, F).astype(np.float32),
                         targets=rng.rand(n, T).astype(np.float32)))
    draw = lambda *s: np.where(rng.rand(*s) < 0.4, rng.rand(*s) * 0.15, 0.15 + rng.rand(*s) * 0.85)   # noqa: E731
    return raws, draw(B, Lt), rng.randint(0, 21128, (B, Lt)), draw(B, R)


def run_gpu(raws, u_tok, rnd_tok, u_reg, Lt, Lpv, R, max_num_pv, fused=True):
    from k3m_b200.preprocess import GpuBatchPreprocessor, pack_raw_batch
    raw = pack_raw_batch(raws, Lt, Lpv, R, device="cuda")
    draws = dict(u_tok=torch.from_numpy(np.ascontiguousarray(u_tok, np.float64)).cuda(),
                 rnd_tok=torch.from_numpy(np.ascontiguousarray(rnd_tok, np.int64)).cuda(),
                 u_reg=torch.from_numpy(np.ascontiguousarray(u_reg, np.float64)).cuda())
    out = GpuBatchPreprocessor(max_num_pv)(raw, draws, fused=fused)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", CASES)
def test_kernels_match_reference_bit_for_bit(name, fused):
    fx = load_golden(name)
    d = fx["dims"]
    got = run_gpu(fx["raws"], fx["u_tok"], fx["rnd_tok"], fx["u_reg"], d["Lt"], d["Lpv"], d["R"], d["max_num_pv"],
                  fused=fused)
    for k, w in fx["want"].items():
        if k == "item_id":
            continue
        w = w.numpy()
        assert got[k].shape == w.shape and got[k].dtype == w.dtype, (k, got[k].dtype, w.dtype)
        assert np.array_equal(got[k], w), k


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [dict(B=64, Lt=36, Lpv=36, R=36, F=2048, T=1601, max_num_pv=8),
                                  dict(B=16, Lt=128, Lpv=256, R=100, F=2048, T=64, max_num_pv=30)])
def test_kernels_match_restatement_at_training_sizes(dims):
    rng = np.random.RandomState(dims["B"])
    d = dims
    raws, u_tok, rnd_tok, u_reg = synth_raws(rng, d["B"], d["Lt"], d["Lpv"], d["R"], d["F"], d["T"], d["max_num_pv"])
    want = preprocess_batch(raws, u_tok, rnd_tok, u_reg, d["Lt"], d["Lpv"], d["R"], d["max_num_pv"])
    got = run_gpu(raws, u_tok, rnd_tok, u_reg, d["Lt"], d["Lpv"], d["R"], d["max_num_pv"])
    for k, w in want.items():
        assert got[k].dtype == w.dtype and np.array_equal(got[k], w), k


@pytest.mark.gpu
def test_preprocessed_batch_feeds_the_training_step():
    """raw records -> GPU preprocessing with device-drawn randomness -> one PretrainStep step."""
    from k3m_b200.config import BertConfig, K3M_TINY
    from k3m_b200.preprocess import GpuBatchPreprocessor, pack_raw_batch
    from k3m_b200.trainer import PretrainStep
    cfg = BertConfig.from_dict(K3M_TINY)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 1, 1.0, 4
    B, Lt, Lpv, R, P = 4, 12, 20, 9, 4
    rng = np.random.RandomState(3)
    raws, _, _, _ = synth_raws(rng, B, Lt, Lpv, R, cfg.v_feature_size, cfg.v_target_size, P)
    for r in raws:
        r["tokens"] = [t % cfg.vocab_size for t in r["tokens"]] or [5]
        r["tokens_pv"] = [t if t in (131, 132) else 200 + t % (cfg.vocab_size - 200) for t in r["tokens_pv"]]
    pre = GpuBatchPreprocessor(P, vocab_size=cfg.vocab_size, seed=1)
    batch = pre(pack_raw_batch(raws, Lt, Lpv, R))
    assert int(batch["attention_mask"].sum()) == sum(len(r["tokens"]) + 2 for r in raws)
    step = PretrainStep(cfg, B, len_t=Lt, len_v=R, len_pv=Lpv, max_num_pv=P, device="cuda", init_seed=1,
                        training=True, use_graph=False)
    step.step({k: v.cpu() for k, v in batch.items()})
    out = step.read_losses()
    assert all(v == v for v in out.values()), out
