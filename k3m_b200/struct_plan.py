"""Host-side index plan for the structure aggregator and the link-prediction (LPM) loss.

The reference walks `index_p` / `index_v` in Python, one (item, PV) pair at a time, syncing the
device for every `index_p[i, j, 0] == 0` test (`vilbert_k3m/vilbert_k3m.py:2437-2446`), and
draws LPM negatives with `random.sample` (`:2480`, `:2492`).  Both depend only on the integer
index tensors, which arrive from the host loader — so the whole plan is built on the host,
once per batch, *before* any kernel runs, and the device side becomes four batched kernels.

The `random.sample` call sequence (order, population, k) is reproduced exactly, so seeding
Python's `random` the same way yields the same negatives as the reference.
"""
import random

import numpy as np


def build_struct_plan(index_p, index_v, L_pv, num_negative_pv, rng=random, halves=1):
    """index_p / index_v: integer arrays [B, max_num_pv, 2] (numpy or nested lists).
    `halves=2`: the batch stacks the two towers of the item-alignment model (`item_embedding` is called
    once per item of the pair, `vilbert_k3m.py:3408-3433`): the stale-`t` fallback of an item without
    PVs does not cross from one call into the next, and there are no LPM terms (`:3105-3161`).

    Returns a dict of int32 numpy arrays:
      pair_item [P]      item of each valid (item, PV) pair, in reference loop order
      pair_rows [P, 4]   rows of the stacked PV sequence [B*L_pv, H]: (p0, p1, v0, v1)
      seg       [B, 2]   (start, end) pair range whose `t` rows item i attends over
      terms     [T, 4]   LPM terms (item, pair, negative item, negative value pair)
      fallback  [halves, 2]  per aggregator call (row of `t`, item) of the pseudo pair that stands for the
                         reference's `t = c_initial[i]` fallback (`:2452-2456`: leading items without any
                         pair), or (-1, -1).  The pseudo pair gathers row 0 of its item (value unused: its `t`
                         row is overwritten), takes part in no LPM term and gets no struc_w1 gradient.
    """
    ip = np.asarray(index_p).astype(np.int64)
    iv = np.asarray(index_v).astype(np.int64)
    B, J = ip.shape[0], ip.shape[1]
    ip, iv = ip.tolist(), iv.tolist()     # plain ints: numpy scalar indexing dominated this loop (first step of a run
    #                                       waits for it: 4.5 -> 2.7 ms at B = 64; the rest is the random.sample calls)
    pair_item, pair_rows, seg, first = [], [], [], []
    last = None
    fallback = [(-1, -1)] * halves
    counts = [0] * B
    for i in range(B):
        if halves > 1 and i % (B // halves) == 0:
            last = None                   # a new `structure_aggregator` call: `t` is unbound again
        start = len(pair_item)
        for j in range(J):
            if ip[i][j][0] == 0:          # loop stops at the first [0, *] row (`:2441`)
                break
            base = i * L_pv
            pair_item.append(i)
            pair_rows.append((base + ip[i][j][0], base + ip[i][j][1],
                              base + iv[i][j][0], base + iv[i][j][1]))
            counts[i] += 1
        end = len(pair_item)
        first.append(start)
        if end > start:
            last = (start, end)
        elif last is None:
            # reference: NameError -> bare except -> t = c_initial[i] (`:2452-2456`); only for
            # leading items without PVs: a pseudo pair whose `t` row the device overwrites with c_initial[i]
            fallback[i // max(1, B // halves)] = (start, i)
            pair_item.append(i)
            pair_rows.append((i * L_pv,) * 4)
            last = (start, start + 1)
        # an item without PVs silently reuses the previous item's `t` (stale variable)
        seg.append(last)
    terms = []
    if halves > 1:
        counts = [0] * B                  # alignment: no link-prediction terms, no random draws
    n_ent = num_negative_pv // 2
    n_val = num_negative_pv - n_ent
    for i in range(B):
        cand_items = [k for k in range(B) if k != i]      # same population for every pair of the item
        for j in range(counts[i]):
            pj = first[i] + j
            cand = cand_items
            if cand:
                for k in rng.sample(cand, min(len(cand), n_ent)):
                    terms.append((i, pj, k, pj))
            cand = [k for k in range(counts[i]) if k != j]
            if cand:
                for k in rng.sample(cand, min(len(cand), n_val)):
                    terms.append((i, pj, i, first[i] + k))
    return dict(
        pair_item=np.asarray(pair_item, dtype=np.int32).reshape(-1),
        pair_rows=np.asarray(pair_rows, dtype=np.int32).reshape(-1, 4),
        seg=np.asarray(seg, dtype=np.int32).reshape(-1, 2),
        terms=np.asarray(terms, dtype=np.int32).reshape(-1, 4),
        fallback=np.asarray(fallback, dtype=np.int32).reshape(-1, 2),
    )


def plan_capacity(B, max_num_pv, num_negative_pv):
    """(pairs, terms, maxn) upper bounds for a batch of B items with at most `max_num_pv` pairs."""
    P = B * max_num_pv
    return P, max(1, P * max(1, num_negative_pv)), max_num_pv


def pad_struct_plan(plan, B, max_num_pv, num_negative_pv):
    """Fixed-shape version of a plan, for CUDA-graph replay: pairs are padded with (item 0, row 0)
    entries that no `seg` range and no term refers to (their `t` rows get a zero gradient, so they
    contribute nothing to any weight gradient), terms are padded with zeros and the real count is
    returned as `n_terms` (int32 [1])."""
    P_cap, T_cap, _ = plan_capacity(B, max_num_pv, num_negative_pv)
    P, T = plan["pair_item"].shape[0], plan["terms"].shape[0]
    if P > P_cap or T > T_cap:
        raise ValueError("struct plan exceeds its capacity (P=%d/%d, T=%d/%d)" % (P, P_cap, T, T_cap))
    out = dict(pair_item=np.zeros(P_cap, dtype=np.int32), pair_rows=np.zeros((P_cap, 4), dtype=np.int32),
               seg=plan["seg"].copy(), terms=np.zeros((T_cap, 4), dtype=np.int32),
               fallback=plan["fallback"].copy(),
               n_terms=np.asarray([T], dtype=np.int32))
    out["pair_item"][:P] = plan["pair_item"]
    out["pair_rows"][:P] = plan["pair_rows"]
    out["terms"][:T] = plan["terms"]
    return out


def mlm_capacity(n_title_rows, n_pv_rows, frac=(0.25, 0.4)):
    """Static capacities (multiples of 128 rows) of the labelled-row lists of the title / PV masked-LM losses: the
    loader masks 15 % of the title tokens and a part of the PV value tokens (`concept_cap_dataset_struc.py:785-860`)."""
    r = lambda x: max(128, (int(x) + 127) // 128 * 128)      # noqa: E731
    return min(r(n_title_rows * frac[0]), r(n_title_rows)), min(r(n_pv_rows * frac[1]), r(n_pv_rows))


def build_mlm_plan(labels_t, labels_pv, cap=None):
    """Rows of the stacked fused sequence S = [title rows | PV rows] that carry a masked-LM label, for the gathered MLM
    head (`Engine.forward`).  labels_*: integer arrays [B, L] (-1 = no label).  cap = (cap_t, cap_pv) pads both lists to
    fixed sizes (CUDA graphs) and raises when a batch has more labels than that; None sizes them to the batch (eager).
    -> dict(mlm_rows int32 [cap_t + cap_pv] (-1 = padding), mlm_labels int64 [cap_t + cap_pv], cap_t int)"""
    lt = np.asarray(labels_t).reshape(-1)
    lp = np.asarray(labels_pv).reshape(-1)
    it, ip = np.nonzero(lt != -1)[0], np.nonzero(lp != -1)[0]
    if cap is None:
        cap = (max(8, (len(it) + 7) // 8 * 8), max(8, (len(ip) + 7) // 8 * 8))
    if len(it) > cap[0] or len(ip) > cap[1]:
        raise ValueError("batch has %d / %d masked-LM labels, the step was built for at most %d / %d "
                         "(raise mlm_cap_frac)" % (len(it), len(ip), cap[0], cap[1]))
    rows = np.full(cap[0] + cap[1], -1, dtype=np.int32)
    labs = np.full(cap[0] + cap[1], -1, dtype=np.int64)
    rows[:len(it)] = it
    labs[:len(it)] = lt[it]
    rows[cap[0]:cap[0] + len(ip)] = len(lt) + ip
    labs[cap[0]:cap[0] + len(ip)] = lp[ip]
    return dict(mlm_rows=rows, mlm_labels=labs, cap_t=int(cap[0]))
