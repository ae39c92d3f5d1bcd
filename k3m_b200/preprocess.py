"""GPU-side batch preprocessing (SURVEY §8 f row N2): the reference builds every training sample in Python
worker processes (`BertPreprocessBatch`, `vilbert_k3m/datasets/concept_cap_dataset_struc.py:532-933`) and
prepends the global image feature in the loader (`:372-413`); at thousands of samples per second per GPU that
pipeline is the bottleneck.  Here the padded RAW batch (token ids, pixel boxes, detector features) is copied to
the device once and four kernels of `libk3m_sm100a.so` (`csrc/preprocess.cu`) produce the tensors
`PretrainStep.step` / the model's `forward` take, bit-identical to the reference for the same random draws.

`raw` (device tensors): tokens int64 [B, Lt], n_tok int32 [B] (ids without specials, truncated to Lt - 2),
tokens_pv int64 [B, Lpv], n_tok_pv int32 [B], boxes f32 [B, R, 4], num_boxes int32 [B], whp f32 [B, 3]
(`pack_raw_batch` builds all of it from per-sample records), features f32 [B, R, F], targets f32 [B, R, T]."""
import ctypes

import numpy as np
import torch

from . import _lib

CLS, SEP, MASK = 101, 102, 103          # bert-base-chinese ids of [CLS] / [SEP] / [MASK]
_vp, _i, _ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong


def _p(t):
    return _vp(t.data_ptr())


def pack_raw_batch(samples, Lt, Lpv, R, device="cuda"):
    """Per-sample records (dict: tokens, tokens_pv, boxes [n, 4], image_w, image_h, features [n, F], targets
    [n, T]) -> the padded raw batch on `device`.  Truncation to L - 2 tokens happens here (`:660-661`)."""
    B = len(samples)
    F, T = samples[0]["features"].shape[1], samples[0]["targets"].shape[1]
    tok, tok_pv = np.zeros((B, Lt), np.int64), np.zeros((B, Lpv), np.int64)
    n_tok, n_pv, nb = np.zeros(B, np.int32), np.zeros(B, np.int32), np.zeros(B, np.int32)
    boxes, whp = np.zeros((B, R, 4), np.float32), np.zeros((B, 3), np.float32)
    feat, targ = np.zeros((B, R, F), np.float32), np.zeros((B, R, T), np.float32)
    for b, s in enumerate(samples):
        t, p = list(s["tokens"])[:Lt - 2], list(s["tokens_pv"])[:Lpv - 2]
        tok[b, :len(t)], tok_pv[b, :len(p)] = t, p
        n_tok[b], n_pv[b] = len(t), len(p)
        n = min(int(s["boxes"].shape[0]), R)
        nb[b] = n
        boxes[b, :n], feat[b, :n], targ[b, :n] = s["boxes"][:n], s["features"][:n], s["targets"][:n]
        w, h = float(s["image_w"]), float(s["image_h"])
        whp[b] = (w, h, w * h)                      # the product in double, rounded once (`:589`)
    mk = lambda a: torch.from_numpy(a).to(device)   # noqa: E731
    return dict(tokens=mk(tok), n_tok=mk(n_tok), tokens_pv=mk(tok_pv), n_tok_pv=mk(n_pv), boxes=mk(boxes),
                num_boxes=mk(nb), whp=mk(whp), features=mk(feat), targets=mk(targ))


class GpuBatchPreprocessor(object):
    """`prep(raw, draws=None) -> dict` named like the model's forward arguments.  `draws` = dict(u_tok f64
    [B, Lt], rnd_tok int64 [B, Lt], u_reg f64 [B, R]) to reproduce a given run of the reference; by default
    they are drawn on the device from `generator`."""

    def __init__(self, max_num_pv, device="cuda", vocab_size=21128, seed=0):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.K3MError("GpuBatchPreprocessor needs a CUDA device (there is no CPU path)")
        self.max_num_pv, self.vocab = max_num_pv, vocab_size
        self.gen = torch.Generator(device=self.device).manual_seed(seed)

    def __call__(self, raw, draws=None, fused=True):
        """`fused=False` goes through the three per-job entry points (one launch each) instead of
        `k3m_prep_batch`; same results."""
        dev, i64 = self.device, torch.int64
        B, Lt = raw["tokens"].shape
        Lpv = raw["tokens_pv"].shape[1]
        R, F = raw["features"].shape[1], raw["features"].shape[2]
        if draws is None:
            draws = dict(u_tok=torch.rand(B, Lt, dtype=torch.float64, device=dev, generator=self.gen),
                         rnd_tok=torch.randint(0, self.vocab, (B, Lt), device=dev, generator=self.gen),
                         u_reg=torch.rand(B, R, dtype=torch.float64, device=dev, generator=self.gen))
        st = _vp(torch.cuda.current_stream().cuda_stream)
        e = lambda *shape, dtype=i64: torch.empty(*shape, dtype=dtype, device=dev)      # noqa: E731
        out = dict(input_ids=e(B, Lt), attention_mask=e(B, Lt), token_type_ids=e(B, Lt), masked_lm_labels=e(B, Lt),
                   input_ids_pv=e(B, Lpv), attention_mask_pv=e(B, Lpv), token_type_ids_pv=e(B, Lpv),
                   masked_lm_labels_pv=e(B, Lpv), index_p=e(B, self.max_num_pv, 2), index_v=e(B, self.max_num_pv, 2),
                   image_feat=e(B, R + 1, F, dtype=torch.float32), image_loc=e(B, R + 1, 5, dtype=torch.float32),
                   image_attention_mask=e(B, R + 1), image_label=e(B, R), image_target=raw["targets"])
        zero_flag, count0 = e(B, R, dtype=torch.uint8), e(B, dtype=torch.int32)
        if fused:
            _lib.check(self.lib.k3m_prep_batch(
                _p(raw["tokens"]), _p(raw["n_tok"]), _p(draws["u_tok"]), _p(draws["rnd_tok"]), _p(raw["tokens_pv"]),
                _p(raw["n_tok_pv"]), _p(raw["boxes"]), _p(raw["num_boxes"]), _p(raw["whp"]), _p(draws["u_reg"]),
                _p(raw["features"]), _i(B), _i(Lt), _i(Lpv), _i(R), _i(F), _i(self.max_num_pv), _ll(CLS), _ll(SEP),
                _ll(MASK), _p(out["input_ids"]), _p(out["attention_mask"]), _p(out["token_type_ids"]),
                _p(out["masked_lm_labels"]), _p(out["input_ids_pv"]), _p(out["attention_mask_pv"]),
                _p(out["token_type_ids_pv"]), _p(out["masked_lm_labels_pv"]), _p(out["index_p"]), _p(out["index_v"]),
                _p(out["image_feat"]), _p(out["image_loc"]), _p(out["image_attention_mask"]), _p(out["image_label"]),
                _p(zero_flag), _p(count0), st), "k3m_prep_batch")
            return self._finish(out, B)
        _lib.check(self.lib.k3m_prep_title(_p(raw["tokens"]), _p(raw["n_tok"]), _p(draws["u_tok"]), _p(draws["rnd_tok"]),
                                           _i(B), _i(Lt), _ll(CLS), _ll(SEP), _ll(MASK), _p(out["input_ids"]),
                                           _p(out["attention_mask"]), _p(out["token_type_ids"]),
                                           _p(out["masked_lm_labels"]), st), "k3m_prep_title")
        _lib.check(self.lib.k3m_prep_pv(_p(raw["tokens_pv"]), _p(raw["n_tok_pv"]), _i(B), _i(Lpv), _i(self.max_num_pv),
                                        _ll(CLS), _ll(SEP), _ll(MASK), _p(out["input_ids_pv"]),
                                        _p(out["attention_mask_pv"]), _p(out["token_type_ids_pv"]),
                                        _p(out["masked_lm_labels_pv"]), _p(out["index_p"]), _p(out["index_v"]), st),
                   "k3m_prep_pv")
        _lib.check(self.lib.k3m_prep_regions(_p(raw["boxes"]), _p(raw["num_boxes"]), _p(raw["whp"]), _p(draws["u_reg"]),
                                             _p(raw["features"]), _i(B), _i(R), _i(F), _p(out["image_feat"]),
                                             _p(out["image_loc"]), _p(out["image_attention_mask"]),
                                             _p(out["image_label"]), _p(zero_flag), _p(count0), st), "k3m_prep_regions")
        return self._finish(out, B)

    def _finish(self, out, B):
        # the three ITM labels are constant 0 in this loader (`:606`)
        z = torch.zeros(B, dtype=torch.int64, device=self.device)
        out.update(next_sentence_label=z, next_sentence_label_pv_v=z, next_sentence_label_pv_t=z)
        return out
