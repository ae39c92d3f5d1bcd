"""Golden vectors for the batch preprocessing (SURVEY §8 f row N2) from the UNMODIFIED reference:
imports `/root/reference/vilbert_k3m/datasets/concept_cap_dataset_struc.py` in this container (stub modules for
the absent `lmdb`, `tensorpack`, `msgpack_numpy`; nothing is copied), runs the real `BertPreprocessBatch.__call__`
per sample with `random.random` / `np.random.randint` replaced by recorded pre-drawn sequences, then the real
`ConceptCapLoaderTrain_struc.__iter__` (global-feature prepend) over the stacked batch, and stores inputs, draws
and outputs in `tests/golden/preprocess_<case>.pt`.

Run here (needs /root/reference): python tools/make_golden_preprocess.py"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/vilbert_k3m/datasets/concept_cap_dataset_struc.py"
CLS, SEP, MASK, VOCAB = 101, 102, 103, 21128


def load_reference_module():
    for name in ("lmdb", "tensorpack", "tensorpack.dataflow", "msgpack_numpy"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.patch = lambda: None
            m.__getattr__ = lambda attr: type(attr, (object,), {})      # class-level defaults like td.LMDBSerializer
            sys.modules[name] = m
    sys.modules["tensorpack"].dataflow = sys.modules["tensorpack.dataflow"]
    spec = importlib.util.spec_from_file_location("ref_concept_cap_dataset_struc", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class Tokenizer(object):
    """ids in, ids out: the samples are pre-tokenised, the reference only needs these five members"""
    mask_token = "[MASK]"

    def encode(self, ids):
        return list(ids)

    def convert_tokens_to_ids(self, tok):
        assert tok == self.mask_token
        return MASK

    def add_special_tokens_single_sentence(self, ids):
        return [CLS] + ids + [SEP]

    def __len__(self):
        return VOCAB


class Draws(object):
    """stands in for the `random` module inside the reference: hands out the pre-drawn doubles in call order"""

    def __init__(self, seq):
        self.seq, self.i = list(seq), 0

    def random(self):
        v = self.seq[self.i]
        self.i += 1
        return v


def pv_string(rng, n_pairs, malformed):
    toks = []
    for k in range(n_pairs):
        toks += list(rng.randint(1000, VOCAB, rng.randint(1, 4))) + [131]
        toks += list(rng.randint(1000, VOCAB, rng.randint(1, 5))) + [132]
    if malformed == 1 and toks:
        toks = toks[:-1]                       # last pair left open
    elif malformed == 2:
        toks = [131] + toks                    # a stray ':' in front
    elif malformed == 3:
        toks = [t for t in toks if t != 132]   # no ';' at all
    elif malformed == 4:
        toks = [131, 2000, 131, 2001, 132, 2002, 132] + toks   # a span that swallows a ';'
    return [int(t) for t in toks]


def make_case(mod, name, B, Lt, Lpv, R, max_num_pv, F, T, seed):
    rng = np.random.RandomState(seed)
    pre = mod.BertPreprocessBatch(Tokenizer(), max_seq_len=Lt, max_seq_len_pv=Lpv, max_num_pv=max_num_pv,
                                  max_region_len=R, v_feature_size=F, v_target_size=T)
    raws, rows, u_tok, rnd_tok, u_reg = [], [], [], [], []
    for b in range(B):
        n_t = int(rng.randint(0, Lt + 3))                         # some longer than Lt - 2: truncated
        tokens = [int(t) for t in rng.randint(1000, VOCAB, n_t)]
        tokens_pv = pv_string(rng, int(rng.randint(0, max_num_pv + 3)), b % 5)
        n = int(rng.randint(1, R + 1))
        w, h = int(rng.randint(200, 900)), int(rng.randint(200, 900))
        x1, y1 = rng.rand(n) * w * 0.6, rng.rand(n) * h * 0.6
        boxes = np.stack([x1, y1, x1 + rng.rand(n) * w * 0.4, y1 + rng.rand(n) * h * 0.4], 1).astype(np.float32)
        if n > 2:
            boxes[1] = boxes[0] + np.float32(1.5)                  # heavy overlap -> propagation
        feats = rng.randn(n, F).astype(np.float32)
        targs = rng.rand(n, T).astype(np.float32)
        draw = lambda k: np.where(rng.rand(k) < 0.5, rng.rand(k) * 0.15, 0.15 + rng.rand(k) * 0.85)
        ut, ur = draw(Lt), draw(R)
        rt = rng.randint(0, VOCAB, Lt)
        # run the real thing: draws are consumed title tokens first, then regions (:670, :714)
        tt = tokens[:Lt - 2]
        used = []
        mod.random = Draws(list(ut[:len(tt)]) + list(ur[:n]))
        real_randint = np.random.randint
        order = [i for i in range(len(tt)) if ut[i] < 0.15 and 0.8 <= ut[i] / 0.15 < 0.9]
        np.random.randint = lambda hi, _o=iter(order): (used.append(1), int(rt[next(_o)]))[1]
        try:
            row = pre((b, list(tokens), list(tokens_pv), "cat", h, w, n, boxes.copy(), feats.copy(), targs.copy()))
        finally:
            np.random.randint = real_randint
        assert mod.random.i == len(tt) + n and len(used) == len(order)
        rows.append(row)
        raws.append(dict(tokens=tt, tokens_pv=tokens_pv[:Lpv - 2], boxes=boxes, image_w=w, image_h=h,
                         features=feats, targets=targs))
        u_tok.append(ut)
        rnd_tok.append(rt)
        u_reg.append(ur)
    # the loader's __iter__ over the stacked batch (global feature, whole-image box, mask)
    stacked = tuple(np.stack([np.asarray(r[k]) for r in rows]) for k in range(len(rows[0])))
    loader = object.__new__(mod.ConceptCapLoaderTrain_struc)
    loader.ds = types.SimpleNamespace(get_data=lambda: iter([stacked]))
    out = next(iter(loader))
    names = ("input_ids", "attention_mask", "token_type_ids", "masked_lm_labels", "next_sentence_label",
             "input_ids_pv", "attention_mask_pv", "token_type_ids_pv", "masked_lm_labels_pv",
             "next_sentence_label_pv_v", "next_sentence_label_pv_t", "image_feat", "image_loc", "image_target",
             "image_label", "image_attention_mask", "index_p", "index_v", "item_id")
    want = {k: torch.as_tensor(np.asarray(v)) for k, v in zip(names, out)}
    fx = dict(dims=dict(B=B, Lt=Lt, Lpv=Lpv, R=R, max_num_pv=max_num_pv, F=F, T=T), raws=raws,
              u_tok=np.stack(u_tok), rnd_tok=np.stack(rnd_tok), u_reg=np.stack(u_reg), want=want)
    path = os.path.join(ROOT, "tests", "golden", "preprocess_%s.pt" % name)
    torch.save(fx, path)
    print(name, "->", path, os.path.getsize(path) // 1024, "KB")


def main():
    mod = load_reference_module()
    make_case(mod, "small", B=10, Lt=16, Lpv=24, R=9, max_num_pv=4, F=16, T=8, seed=1)
    make_case(mod, "script", B=6, Lt=36, Lpv=64, R=36, max_num_pv=30, F=32, T=12, seed=2)


if __name__ == "__main__":
    main()
