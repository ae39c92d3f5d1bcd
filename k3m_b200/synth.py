"""Synthetic K3M pretraining batches and deterministic weights.

The reference's loader (`vilbert_k3m/datasets/concept_cap_dataset_struc.py:372-413`) yields 16
CPU tensors plus `index_p` / `index_v`.  There is no dataset here, so `synth_batch` builds a
batch with the same names, shapes, dtypes and structural invariants:

  * ids start with [CLS]=101 and end (last valid position) with [SEP]=102;
  * PV ids are `p... : v... ;` groups (':'=131, ';'=132) and `index_p[i,j] = [begin, colon]`,
    `index_v[i,j] = [colon+1, semicolon]` exactly as `BertPreprocessBatch` emits them
    (`concept_cap_dataset_struc.py:785-813`); unused rows are `[0, 0]`;
  * 15 % MLM labels on valid positions (>= 1 per batch), -1 elsewhere;
  * image row 0 is the mean feature with box [0,0,1,1,1] (`:382-397`); `image_target` is a
    soft distribution over `v_target_size` classes; `image_label` in {-1, 1} (>= 1 per batch);
  * `ragged=True` (default): every item has its own `num_boxes` in [len_v/2, len_v]; rows beyond it are what the
    loader pads with (`:580-586`, `:716-724`): zero feature / box / target rows, `image_label` -1 and
    `image_attention_mask` 0, and the global feature is the mean over the valid boxes (`:382-397`);
  * next-sentence labels: the loader only ever emits 0 (`:604`), but `forward` accepts any 0/1 triple and
    turns it into the ITM target `1 - ((l1 + l2 + l3) == 0)` (`vilbert_k3m.py:2823-2824`), so with
    `ragged=True` the three labels are random 0/1 draws (both ITM classes are exercised).

Everything is generated with numpy's PCG64 from the given seed, so the GPU box and this
container produce bit-identical batches.
"""
import numpy as np
import torch

CLS, SEP, COLON, SEMI = 101, 102, 131, 132


def synth_batch(cfg, batch, len_t=36, len_v=36, len_pv=36, max_num_pv=8, seed=1234,
                ragged=True, no_pv=()):
    """`no_pv`: items that get no property/value group at all (`[CLS] [SEP]` PV row, all-zero index rows) —
    the case the reference's aggregator handles through its bare `except` (`vilbert_k3m.py:2452-2456`)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    V = cfg.vocab_size
    lo = min(1000, V // 2)
    B = batch

    def text(L):
        ids = rng.integers(lo, V, size=(B, L), dtype=np.int64)
        lens = rng.integers(max(4, L // 2), L + 1, size=B) if ragged else np.full(B, L)
        mask = (np.arange(L)[None, :] < lens[:, None]).astype(np.int64)
        ids[:, 0] = CLS
        ids[np.arange(B), lens - 1] = SEP
        ids = ids * mask
        return ids, mask, lens

    input_ids, attention_mask, lens_t = text(len_t)
    labels = np.full((B, len_t), -1, dtype=np.int64)
    pick = (rng.random((B, len_t)) < 0.15) & (attention_mask == 1)
    pick[:, 0] = False
    if not pick.any():
        pick[0, 1] = True
    labels[pick] = input_ids[pick]

    # PV stream: [CLS] p p : v v ; p : v ; ... [SEP]
    ids_pv = np.zeros((B, len_pv), dtype=np.int64)
    mask_pv = np.zeros((B, len_pv), dtype=np.int64)
    labels_pv = np.full((B, len_pv), -1, dtype=np.int64)
    index_p = np.zeros((B, max_num_pv, 2), dtype=np.int64)
    index_v = np.zeros((B, max_num_pv, 2), dtype=np.int64)
    for i in range(B):
        pos = 1
        ids_pv[i, 0] = CLS
        n_target = int(rng.integers(1, max_num_pv + 1))
        if i in no_pv:
            n_target = 0
        j = 0
        while j < n_target:
            lp = int(rng.integers(1, 3))
            lv = int(rng.integers(1, 3))
            need = lp + 1 + lv + 1
            if pos + need + 1 > len_pv:
                break
            begin = pos
            ids_pv[i, pos:pos + lp] = rng.integers(lo, V, size=lp)
            pos += lp
            colon = pos
            ids_pv[i, pos] = COLON
            pos += 1
            vals = rng.integers(lo, V, size=lv)
            ids_pv[i, pos:pos + lv] = vals
            # value tokens of groups 2.. are masked-PV targets with prob 0.5
            if j >= 1 and rng.random() < 0.5:
                labels_pv[i, pos:pos + lv] = vals
            pos += lv
            semi = pos
            ids_pv[i, pos] = SEMI
            pos += 1
            index_p[i, j] = (begin, colon)
            index_v[i, j] = (colon + 1, semi)
            j += 1
        ids_pv[i, pos] = SEP
        mask_pv[i, :pos + 1] = 1
    if (labels_pv != -1).sum() == 0:
        c = int(index_v[0, 0, 0])
        labels_pv[0, c] = ids_pv[0, c]

    Fv, Tv = cfg.v_feature_size, cfg.v_target_size
    feat = np.abs(rng.standard_normal((B, len_v, Fv), dtype=np.float32))
    zero = rng.random((B, len_v)) < 0.15
    feat[zero] = 0.0
    nbox = rng.integers(max(1, len_v // 2), len_v + 1, size=B) if ragged else np.full(B, len_v)
    box_ok = np.arange(len_v)[None, :] < nbox[:, None]
    feat[~box_ok] = 0.0
    g = feat.sum(axis=1, keepdims=True) / nbox[:, None, None].astype(np.float32)
    image_feat = np.concatenate([g, feat], axis=1).astype(np.float32)
    xy = np.sort(rng.random((B, len_v, 2, 2), dtype=np.float32), axis=2)
    loc = np.zeros((B, len_v, 5), dtype=np.float32)
    loc[..., 0] = xy[:, :, 0, 0]
    loc[..., 1] = xy[:, :, 0, 1]
    loc[..., 2] = xy[:, :, 1, 0]
    loc[..., 3] = xy[:, :, 1, 1]
    loc[..., 4] = (loc[..., 2] - loc[..., 0]) * (loc[..., 3] - loc[..., 1])
    loc[~box_ok] = 0.0
    gl = np.tile(np.array([0, 0, 1, 1, 1], dtype=np.float32), (B, 1, 1))
    image_loc = np.concatenate([gl, loc], axis=1)
    logits = rng.standard_normal((B, len_v, Tv), dtype=np.float32) * 2.0
    logits -= logits.max(axis=-1, keepdims=True)
    e = np.exp(logits)
    image_target = (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)
    image_target[~box_ok] = 0.0
    image_label = np.where((rng.random((B, len_v)) < 0.15) & box_ok, 1, -1).astype(np.int64)
    if (image_label == 1).sum() == 0:
        image_label[0, 0] = 1
    image_mask = np.concatenate([np.ones((B, 1), dtype=np.int64), box_ok.astype(np.int64)], axis=1)

    z = np.zeros(B, dtype=np.int64)
    nsl = [rng.integers(0, 2, size=B).astype(np.int64) if ragged else z.copy() for _ in range(3)]
    t = torch.from_numpy
    return dict(
        input_ids=t(input_ids), image_feat=t(image_feat), image_loc=t(image_loc),
        token_type_ids=t(np.zeros_like(input_ids)), attention_mask=t(attention_mask),
        image_attention_mask=t(image_mask), masked_lm_labels=t(labels),
        image_label=t(image_label), image_target=t(image_target),
        next_sentence_label=t(nsl[0]),
        input_ids_pv=t(ids_pv), token_type_ids_pv=t(np.zeros_like(ids_pv)),
        attention_mask_pv=t(mask_pv), masked_lm_labels_pv=t(labels_pv),
        next_sentence_label_pv_v=t(nsl[1]), next_sentence_label_pv_t=t(nsl[2]),
        index_p=t(index_p), index_v=t(index_v),
    )


FORWARD_ORDER = (
    "input_ids", "image_feat", "image_loc", "token_type_ids", "attention_mask",
    "image_attention_mask", "masked_lm_labels", "image_label", "image_target",
    "next_sentence_label",
)


def synth_state_dict(spec, seed=0, scale=1.0):
    """Deterministic fp32 weights for a `param_spec`: N(0, 0.02*scale) matrices, small non-zero
    biases, LayerNorm weights around 1 — every parameter non-trivial so that every gradient
    path is exercised.  CPU torch generator -> identical on every machine."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in spec.items():
        if "LayerNorm" in name and name.endswith("weight"):
            sd[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif len(shape) == 1:
            sd[name] = 0.02 * torch.randn(shape, generator=g)
        else:
            sd[name] = 0.02 * scale * torch.randn(shape, generator=g)
    return sd
