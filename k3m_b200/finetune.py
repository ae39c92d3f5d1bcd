"""Host side of the item-alignment task around `K3MForItemAlignment` (SURVEY §8 f, row N1): the evaluation
sweep and the prediction dump of `finetune.py`, written against the model's forward only — the model is
`k3m_b200.vilbert_k3m.K3MForItemAlignment` (CUDA kernels) in production; the training step itself is
`k3m_b200.trainer.AlignmentStep`.

Batch layout (what the reference's `ConceptCapLoaderTrain_struc` pair loader yields, `finetune.py:1032-1050`):
a sequence whose first 21 entries are
  labels, then for item 1 and item 2:  input_ids, input_mask, segment_ids, input_ids_pv, input_mask_pv,
  segment_ids_pv, image_feat, image_loc, image_target, image_mask
and whose last 6 are  src_item_ids, index_p_1, index_v_1, tgt_item_ids, index_p_2, index_v_2.
"""
import json

import numpy as np
import torch


def split_pair_batch(batch, device=None):
    """-> (labels, model positional arguments after `labels`, src_item_ids, tgt_item_ids), the arguments in
    the order `K3MForItemAlignment.forward` takes them (`finetune.py:1075-1100`)."""
    src_ids, tgt_ids = batch[-6], batch[-3]
    idx = [torch.as_tensor(batch[i]) for i in (-5, -4, -2, -1)]
    t = [torch.as_tensor(x) for x in batch[:-6]]
    if len(t) != 21:
        raise ValueError("pair batch holds %d tensors before the six trailing fields, expected 21" % len(t))
    if device is not None:
        idx = [x.to(device, non_blocking=True) for x in idx]
        t = [x.to(device, non_blocking=True) for x in t]
    labels = t[0]
    args = []
    for item, (ip, iv) in enumerate(((idx[0], idx[1]), (idx[2], idx[3]))):
        (input_ids, input_mask, segment_ids, input_ids_pv, input_mask_pv, segment_ids_pv, image_feat, image_loc,
         _image_target, image_mask) = t[1 + 10 * item:11 + 10 * item]
        args += [input_ids, segment_ids, input_mask, input_ids_pv, segment_ids_pv, input_mask_pv, ip, iv,
                 image_feat, image_loc, image_mask]
    return labels, args, src_ids, tgt_ids


def precision_recall_f1(labels, pred):
    """Binary precision / recall / F1 of the positive class with sklearn's conventions
    (`precision_score`, `recall_score`, `f1_score`, `finetune.py:1119-1121`): 0.0 when a denominator is 0."""
    labels = np.asarray(labels).astype(bool)
    pred = np.asarray(pred).astype(bool)
    tp = float(np.sum(labels & pred))
    fp = float(np.sum(~labels & pred))
    fn = float(np.sum(labels & ~pred))
    p = tp / (tp + fp) if tp + fp > 0 else 0.0
    r = tp / (tp + fn) if tp + fn > 0 else 0.0
    f1 = 2.0 * tp / (2.0 * tp + fp + fn) if 2.0 * tp + fp + fn > 0 else 0.0
    return p, r, f1


def threshold_sweep(labels, probs, thresholds=None):
    """`finetune.py:1118-1122`: metrics of `probs >= threshold` for threshold in arange(0.1, 1.0, 0.1)."""
    thresholds = np.arange(0.1, 1.0, 0.1) if thresholds is None else thresholds
    probs = np.asarray(probs)
    return [(float(th),) + precision_recall_f1(labels, probs >= th) for th in thresholds]


def evaluate(model, batches, device=None, log=None):
    """The per-epoch evaluation of `finetune.py:1028-1124`: eval mode, no grad, positive-class probabilities
    of every pair, then the threshold sweep.  -> (probs, labels, sweep rows)."""
    was_training = model.training
    model.eval()
    all_p, all_l = [], []
    with torch.no_grad():
        for batch in batches:
            labels, args, _, _ = split_pair_batch(batch, device)
            _, _, probs, _ = model(labels, *args, output_all_attention_masks=False)
            all_p.append(probs.detach().float().cpu().numpy().reshape(-1))
            all_l.append(labels.detach().cpu().numpy().reshape(-1))
    model.train(was_training)
    probs = np.concatenate(all_p) if all_p else np.zeros(0, np.float32)
    labels = np.concatenate(all_l) if all_l else np.zeros(0, np.int64)
    rows = threshold_sweep(labels, probs)
    if log is not None:
        for th, p, r, f1 in rows:
            log("threshold=%s, precision=%s, recall=%s, f1=%s" % (th, p, r, f1))
    return probs, labels, rows


def _emb_text(e):
    """`finetune.py:1196-1197`: a vector prints as its comma-joined elements, a scalar as itself."""
    return ",".join(str(x) for x in e) if isinstance(e, np.ndarray) and e.ndim > 0 else str(e)


def prediction_record(src_item_id, tgt_item_id, src_embed, tgt_embed, threshold):
    """One line of `deepAI_result_threshold=<t>.jsonl` (`finetune.py:1198-1201`)."""
    return {"src_item_id": src_item_id, "src_item_emb": "[%s]" % _emb_text(src_embed),
            "tgt_item_id": tgt_item_id, "tgt_item_emb": "[%s]" % _emb_text(tgt_embed), "threshold": threshold}


def predict(model, batches, out_path, threshold, device=None):
    """`--do_pred` (`finetune.py:1126-1210`): the two per-item outputs of the model for every pair, one JSON
    object per line.  Returns the number of pairs written."""
    model.eval()
    n = 0
    with torch.no_grad(), open(out_path, "w", encoding="utf-8") as w:
        for batch in batches:
            labels, args, src_ids, tgt_ids = split_pair_batch(batch, device)
            e1, e2, _, _ = model(labels, *args, output_all_attention_masks=False)
            e1 = e1.detach().float().cpu().numpy()
            e2 = e2.detach().float().cpu().numpy()
            for s, t, a, b in zip(src_ids, tgt_ids, e1, e2):
                w.write(json.dumps(prediction_record(s, t, a, b, threshold)) + "\n")
                n += 1
    return n
