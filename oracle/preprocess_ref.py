"""TEST INFRASTRUCTURE ONLY — CPU restatement (numpy, per sample) of the reference's batch preprocessing, the
checker of `k3m_b200/csrc/preprocess.cu` (SURVEY §8 f row N2).  Never imported by the product path.

Follows `vilbert_k3m/datasets/concept_cap_dataset_struc.py`:
  `iou` :40-78, `BertPreprocessBatch.__call__` :564-652 (image tensors, location normalisation),
  `convert_example_to_features` :654-747, `mask_word` :763-783, `index_pv` :785-813, `mask_word_pv` :815-839,
  `mask_region` :887-933, and the loader's global-feature prepend `ConceptCapLoaderTrain_struc.__iter__` :372-413.
The reference draws `random.random()` once per title token and once per region, and `np.random.randint` when a
token is replaced; here the draws are arguments (`u_tok[i]`, `rnd_tok[i]`, `u_reg[r]`), which is what makes the
GPU path checkable bit for bit.  Pinned against the real reference by `tools/make_golden_preprocess.py`
(`tests/golden/preprocess_*.pt`).

Raw sample = dict(tokens, tokens_pv: lists of ids without specials, already truncated to L - 2;
boxes f32 [n, 4] pixel corners; image_w, image_h; features f32 [n, F]; targets f32 [n, T])."""
import numpy as np

CLS, SEP, MASK = 101, 102, 103
COLON, SEMI = 131, 132


def pairwise_iou(boxes):
    """:40-78 with anchors == gt_boxes, float32 throughout (the "+ 1" pixel convention)."""
    b = np.asarray(boxes, np.float32)
    area = (b[:, 2] - b[:, 0] + np.float32(1)) * (b[:, 3] - b[:, 1] + np.float32(1))
    iw = np.minimum(b[:, None, 2], b[None, :, 2]) - np.maximum(b[:, None, 0], b[None, :, 0]) + np.float32(1)
    ih = np.minimum(b[:, None, 3], b[None, :, 3]) - np.maximum(b[:, None, 1], b[None, :, 1]) + np.float32(1)
    iw = np.where(iw < 0, np.float32(0), iw)
    ih = np.where(ih < 0, np.float32(0), ih)
    inter = iw * ih
    return inter / (area[:, None] + area[None, :] - inter)


def mask_title(tokens, u, rnd):
    """:763-783 — 15 % of the tokens are selected; 80 % of those become [MASK], 10 % a random id, 10 % stay."""
    out, labels = list(tokens), []
    for i, t in enumerate(tokens):
        prob = float(u[i])
        if prob < 0.15:
            prob /= 0.15
            if prob < 0.8:
                out[i] = MASK
            elif prob < 0.9:
                out[i] = int(rnd[i])
            labels.append(t)
        else:
            labels.append(-1)
    return out, labels


def mask_pv_values(tokens):
    """:815-839 — every value span (between ':' and ';') except the first pair's is masked; an open last pair
    is closed at the end of the list; the lists are zipped to the shorter one."""
    out = list(tokens)
    colons = [i for i, t in enumerate(tokens) if t == COLON]
    semis = [i for i, t in enumerate(tokens) if t == SEMI]
    if len(semis) == len(colons) - 1:
        semis.append(len(tokens))
    if len(semis) > 1:
        semis, colons = semis[1:], colons[1:]
    labels = [-1] * len(tokens)
    for a, b in zip(colons, semis):
        for i in range(a + 1, b):
            labels[i] = out[i]
            out[i] = MASK
    return out, labels


def pv_spans(row, max_num_pv):
    """:785-813 on the row WITH specials: [begin, colon] property spans and [colon + 1, semicolon] value spans."""
    colons = [i for i, t in enumerate(row) if t == COLON]
    semis = [i for i, t in enumerate(row) if t == SEMI]
    if len(semis) == len(colons):
        pass
    elif len(semis) == len(colons) - 1:
        colons = colons[:-1]
    else:
        colons, semis = [], []
    p, v, begin = [], [], 1
    for a, b in zip(colons, semis):
        p.append([begin, a])
        v.append([a + 1, b])
        begin = b + 1
        if len(p) >= max_num_pv:
            break
    return p, v


def pad_row(row, labels, L):
    """:664-690 — specials, attention mask, zero segments, padding to L."""
    ids = [CLS] + list(row) + [SEP]
    lab = [-1] + list(labels) + [-1]
    att = [1] * len(ids)
    pad = L - len(ids)
    return ids + [0] * pad, att + [0] * pad, [0] * L, lab + [-1] * pad


def mask_regions(features, n, R, overlaps, u):
    """:887-933 — a selected region (15 %) is zeroed 90 % of the time; every region overlapping a selected one
    by IoU > 0.4 is flagged in `masked_label` (which only feeds the global-feature count)."""
    feat = np.array(features, np.float32, copy=True)
    masked = np.zeros(R, bool)
    ov = np.zeros((n, R), np.float32)
    ov[:, :n] = overlaps
    label = []
    for i in range(n):
        prob = float(u[i])
        if prob < 0.15:
            prob /= 0.15
            if prob < 0.9:
                feat[i] = 0
            masked = np.logical_or(masked, ov[i] > 0.4)
            label.append(1)
        else:
            label.append(-1)
    return feat, label + [-1] * (R - n), masked


def preprocess_sample(raw, u_tok, rnd_tok, u_reg, Lt, Lpv, R, max_num_pv):
    """One sample -> the reference's per-sample tensors (before the loader's global-feature prepend)."""
    n = int(raw["boxes"].shape[0])
    F, T = raw["features"].shape[1], raw["targets"].shape[1]
    feat = np.zeros((R, F), np.float32)
    targ = np.zeros((R, T), np.float32)
    loc = np.zeros((R, 5), np.float32)
    feat[:n], targ[:n], loc[:n, :4] = raw["features"], raw["targets"], raw["boxes"]
    w, h = float(raw["image_w"]), float(raw["image_h"])
    loc[:, 4] = (loc[:, 3] - loc[:, 1]) * (loc[:, 2] - loc[:, 0]) / np.float32(w * h)     # :586-590
    loc[:, 0] /= np.float32(w)
    loc[:, 1] /= np.float32(h)
    loc[:, 2] /= np.float32(w)
    loc[:, 3] /= np.float32(h)
    t_ids, t_lab = mask_title(raw["tokens"], u_tok, rnd_tok)
    p_ids, p_lab = mask_pv_values(raw["tokens_pv"])
    ids, att, seg, lab = pad_row(t_ids, t_lab, Lt)
    ids_pv, att_pv, seg_pv, lab_pv = pad_row(p_ids, p_lab, Lpv)
    ip, iv = pv_spans([CLS] + p_ids + [SEP], max_num_pv)
    ip += [[0, 0]] * (max_num_pv - len(ip))
    iv += [[0, 0]] * (max_num_pv - len(iv))
    feat, label, masked = mask_regions(feat, n, R, pairwise_iou(raw["boxes"]), u_reg)
    return dict(input_ids=ids, attention_mask=att, token_type_ids=seg, masked_lm_labels=lab,
                input_ids_pv=ids_pv, attention_mask_pv=att_pv, token_type_ids_pv=seg_pv, masked_lm_labels_pv=lab_pv,
                index_p=ip, index_v=iv, image_feat=feat, image_loc=loc, image_target=targ, image_label=label,
                image_attention_mask=[1] * n + [0] * (R - n), masked_label=masked)


def preprocess_batch(raws, u_tok, rnd_tok, u_reg, Lt, Lpv, R, max_num_pv):
    """Batch of samples + the loader's prepend (:372-413): a global feature (sum over the R rows divided by the
    number of rows NOT flagged in masked_label — padding rows count, as in the reference; float32 sum in row
    order, float64 division, back to float32), the whole-image box (0, 0, 1, 1, 1) and a 1 in the mask."""
    per = [preprocess_sample(r, u_tok[i], rnd_tok[i], u_reg[i], Lt, Lpv, R, max_num_pv) for i, r in enumerate(raws)]
    out = {}
    for k in per[0]:
        out[k] = np.stack([np.asarray(p[k]) for p in per])
    masked = out.pop("masked_label")
    count = np.sum(masked == 0, axis=1, keepdims=True)
    count[count == 0] = 1
    s = np.zeros(out["image_feat"][:, 0].shape, np.float32)
    for r in range(R):                                   # numpy reduces axis 1 slice by slice
        s = s + out["image_feat"][:, r]
    g = (s.astype(np.float64) / count).astype(np.float32)
    B = len(raws)
    out["image_feat"] = np.concatenate([g[:, None], out["image_feat"]], axis=1).astype(np.float32)
    gl = np.tile(np.array([[0, 0, 1, 1, 1]], np.float32), (B, 1))
    out["image_loc"] = np.concatenate([gl[:, None], out["image_loc"]], axis=1).astype(np.float32)
    out["image_attention_mask"] = np.concatenate([np.ones((B, 1), np.int64), out["image_attention_mask"]], axis=1)
    for k, v in out.items():
        if v.dtype.kind in "iu" or v.dtype == bool:
            out[k] = v.astype(np.int64)
    return out
