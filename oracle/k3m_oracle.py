"""TEST INFRASTRUCTURE — CPU oracle for the K3M pretraining step.  Not shipped, not measured as
product: only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / reference arm
may import this file.

It is a plain fp32 PyTorch *restatement* (functional, state_dict in / losses out; autograd
provides the backward) of the reference's
`BertForMultiModalPreTraining_tri_stru.forward` (`vilbert_k3m/vilbert_k3m.py:2673-2859`).
Every function cites the reference lines it follows.

Pinning: the reference has no tests / golden vectors of its own for this path, so
the oracle is pinned against the reference itself, imported unmodified in the build container
(`oracle/ref_harness.py`): `tests/test_oracle.py::test_oracle_matches_live_reference` compares
losses, c_initial / c_final and the gradients on identical weights and inputs where `/root/reference`
is mounted, and `tools/make_golden.py` writes reference outputs (losses, per-layer taps, gradient
samples) to `tests/golden/*.pt`, which `tests/test_oracle.py::test_oracle_matches_golden` (and
`tests/test_alignment_cpu.py::test_oracle_alignment_matches_reference` for the alignment model)
re-check anywhere, including the GPU box.
"""
import math
import random

import torch
import torch.nn.functional as F


def gelu(x):
    # `vilbert_k3m.py:119-125`: exact erf GELU
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def layer_norm(x, w, b, eps=1e-12):
    # `vilbert_k3m.py:328-332`: TF-style, biased variance, eps inside the sqrt
    u = x.mean(-1, keepdim=True)
    s = (x - u).pow(2).mean(-1, keepdim=True)
    return w * ((x - u) / torch.sqrt(s + eps)) + b


class Oracle(object):
    """Holds a state_dict (name -> fp32 tensor, reference names) and evaluates the step."""

    def __init__(self, cfg, state_dict, training=False):
        self.cfg = cfg
        self.sd = state_dict
        self.training = training
        self.taps = None  # optional dict: name -> activation (for per-layer parity)
        self._calls = {}

    # ---- small helpers -------------------------------------------------------------------
    def lin(self, name, x):
        return F.linear(x, self.sd[name + ".weight"], self.sd.get(name + ".bias"))

    def ln(self, name, x):
        return layer_norm(x, self.sd[name + ".weight"], self.sd[name + ".bias"])

    def drop(self, x, p):
        return F.dropout(x, p, self.training)

    def tap(self, name, x):
        if self.taps is not None:
            self.taps[name] = x

    def tap_call(self, module, outs):
        """Same keys as the forward hooks `tools/make_golden.py` puts on the unmodified reference's
        BertLayer / BertImageLayer / BertConnectionLayer modules: "<module>#<call index>[/<output>]"."""
        if self.taps is None:
            return
        k = self._calls.get(module, 0)
        self._calls[module] = k + 1
        for j, x in enumerate(outs):
            self.taps["%s#%d" % (module, k) + ("/%d" % j if len(outs) > 1 else "")] = x

    # ---- embeddings ------------------------------------------------------------------------
    def text_embeddings(self, ids, type_ids):
        # `vilbert_k3m.py:361-382`; position ids are always arange(L) (`:364`)
        L = ids.shape[1]
        pos = torch.arange(L, device=ids.device)
        # nn.Embedding(padding_idx=0) (`:343-345`): row 0 gets no gradient from this path
        e = (F.embedding(ids, self.sd["embeddings.word_embeddings.weight"], padding_idx=0)
             + self.sd["embeddings.position_embeddings.weight"][pos][None]
             + self.sd["embeddings.token_type_embeddings.weight"][type_ids])
        return self.drop(self.ln("embeddings.LayerNorm", e), self.cfg.hidden_dropout_prob)

    def image_embeddings(self, feat, loc):
        # `vilbert_k3m.py:2153-2161`
        e = self.lin("v_embeddings.image_embeddings", feat) + self.lin(
            "v_embeddings.image_location_embeddings", loc)
        return self.drop(self.ln("v_embeddings.LayerNorm", e), self.cfg.hidden_dropout_prob)

    # ---- attention -------------------------------------------------------------------------
    def _attend(self, q, k, v, add_mask, heads, p_drop):
        # `vilbert_k3m.py:449-464`: scores/sqrt(d) + mask -> softmax -> dropout -> @V
        B, Lq, A = q.shape
        d = A // heads
        q = q.view(B, Lq, heads, d).permute(0, 2, 1, 3)
        k = k.view(B, -1, heads, d).permute(0, 2, 1, 3)
        v = v.view(B, -1, heads, d).permute(0, 2, 1, 3)
        s = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d) + add_mask
        p = self.drop(torch.softmax(s, dim=-1), p_drop)
        return torch.matmul(p, v).permute(0, 2, 1, 3).reshape(B, Lq, A)

    def bert_layer(self, p, x, add_mask, image):
        # BertLayer `:535-548` / BertImageLayer `:696-709`
        c = self.cfg
        heads = c.v_num_attention_heads if image else c.num_attention_heads
        pa = c.v_attention_probs_dropout_prob if image else c.attention_probs_dropout_prob
        ph = c.v_hidden_dropout_prob if image else c.hidden_dropout_prob
        ctx = self._attend(self.lin(p + ".attention.self.query", x),
                           self.lin(p + ".attention.self.key", x),
                           self.lin(p + ".attention.self.value", x), add_mask, heads, pa)
        a = self.ln(p + ".attention.output.LayerNorm",
                    self.drop(self.lin(p + ".attention.output.dense", ctx), ph) + x)
        h = gelu(self.lin(p + ".intermediate.dense", a))
        y = self.ln(p + ".output.LayerNorm", self.drop(self.lin(p + ".output.dense", h), ph) + a)
        self.tap_call(p, (y,))
        return y

    def connection(self, p, x1, m1, x2, m2):
        # BertConnectionLayer `:1030-1071` / _two_text `:1074-1111`.  Stream 1 = image (or PV),
        # stream 2 = text.  ctx1 = queries of 2 over keys/values of 1 (`:786-800`), and the
        # image stream receives ctx2 through dense1 (`:1061-1063`, Appendix A.2).
        c = self.cfg
        heads = c.bi_num_attention_heads
        b = p + ".biattention."
        q1, k1, v1 = (self.lin(b + n, x1) for n in ("query1", "key1", "value1"))
        q2, k2, v2 = (self.lin(b + n, x2) for n in ("query2", "key2", "value2"))
        ctx1 = self._attend(q2, k1, v1, m1, heads, c.v_attention_probs_dropout_prob)
        ctx2 = self._attend(q1, k2, v2, m2, heads, c.attention_probs_dropout_prob)
        o = p + ".biOutput."
        a1 = self.ln(o + "LayerNorm1",
                     self.drop(self.lin(o + "dense1", ctx2), c.v_hidden_dropout_prob) + x1)
        a2 = self.ln(o + "LayerNorm2",
                     self.drop(self.lin(o + "dense2", ctx1), c.hidden_dropout_prob) + x2)
        # both FFN output dropouts use the class of the sub-module: v_output of c_layer is a
        # BertImageOutput (v_hidden_dropout_prob, `:685`); in _two_text it is a BertOutput.
        two_text = "c_layer_pv_t" in p
        p1 = c.hidden_dropout_prob if two_text else c.v_hidden_dropout_prob
        h1 = gelu(self.lin(p + ".v_intermediate.dense", a1))
        y1 = self.ln(p + ".v_output.LayerNorm", self.drop(self.lin(p + ".v_output.dense", h1), p1) + a1)
        h2 = gelu(self.lin(p + ".t_intermediate.dense", a2))
        y2 = self.ln(p + ".t_output.LayerNorm",
                     self.drop(self.lin(p + ".t_output.dense", h2), c.hidden_dropout_prob) + a2)
        self.tap_call(p, (y1, y2))
        return y1, y2

    def encoder_pass(self, tag, x_txt, m_txt, x_1, m_1, conn, image):
        """One pairwise pass (`calculate_for_text_img` `:1154-1330`, `_pv_img` `:1332`,
        `_two_text` `:1510`): stream "1" (image, or PV in the two-text pass) runs
        `v_layer` (or `layer`), stream "2" (text) runs `layer`; connection `conn[count]` after
        text layer t_biattention_id[count]-1 / stream-1 layer ids[count]-1."""
        c = self.cfg
        ids1 = c.v_biattention_id if image else c.t_biattention_id   # zip quirk `:1534`
        n1 = c.v_num_hidden_layers if image else c.num_hidden_layers
        l1 = "encoder.v_layer.%d" if image else "encoder.layer.%d"
        s1 = s2 = 0
        for count, (e1, e2) in enumerate(zip(ids1, c.t_biattention_id)):
            for i in range(s2, e2):
                x_txt = self.bert_layer("encoder.layer.%d" % i, x_txt, m_txt, False)
            for i in range(s1, e1):
                x_1 = self.bert_layer(l1 % i, x_1, m_1, image)
            x_1, x_txt = self.connection("encoder.%s.%d" % (conn, count), x_1, m_1, x_txt, m_txt)
            self.tap("%s.c%d.s1" % (tag, count), x_1)
            self.tap("%s.c%d.s2" % (tag, count), x_txt)
            s1, s2 = e1, e2
        for i in range(s1, n1):
            x_1 = self.bert_layer(l1 % i, x_1, m_1, image)
        for i in range(s2, c.num_hidden_layers):
            x_txt = self.bert_layer("encoder.layer.%d" % i, x_txt, m_txt, False)
        return x_txt, x_1

    # ---- fusion (modal selection) ----------------------------------------------------------
    def fuse(self, mod, indiv, c1, c2, noise):
        """`get_sequence_pooled_output_final` `:2376-2411`.  if_pre_sampling: 1 = hard Gumbel
        select per (token, channel) (`pre_sampling_sequence` `:2331-2374`), 0 = mean of three,
        2 = soft gates (`:2300-2329`), else mean of the two interactive sequences."""
        mode = self.cfg.if_pre_sampling
        if mode == 0:
            return (indiv + c1 + c2) / 3
        if mode not in (1, 2):
            return (c1 + c2) / 2
        s = [F.relu(indiv), F.relu(c1), F.relu(c2)]
        cat = torch.cat(s, 2)
        alphas = [torch.sigmoid(self.lin("score_%s_%s" % (n, mod), cat))
                  for n in ("self", "cross1", "cross2")]
        if mode == 2:
            return self.lin("soft_" + mod, torch.cat([a * b for a, b in zip(s, alphas)], 2))
        ak = torch.stack(alphas, 2)                               # [B, L, 3, H]
        if noise is None:
            g = -torch.empty_like(ak).exponential_().log()        # F.gumbel_softmax's draw
        else:
            g = noise
        y_soft = torch.softmax((ak + g) / 1.0, dim=2)
        idx = y_soft.max(2, keepdim=True)[1]
        y_hard = torch.zeros_like(ak).scatter_(2, idx, 1.0)
        a = y_hard - y_soft.detach() + y_soft                     # straight-through
        return s[0] * a[:, :, 0] + s[1] * a[:, :, 1] + s[2] * a[:, :, 2]

    # ---- structure aggregation + LPM -------------------------------------------------------
    def structure_aggregator(self, c_initial, seq_pv, index_p, index_v, lpm=True):
        """`structure_aggregator` `:2413-2505`.  Same loop order and the same `random.sample`
        call sequence as the reference, so seeding `random` identically gives identical
        negatives.  Items without any PV reuse the previous item's `t` (the reference's bare
        `except`, `:2452-2456`, only fires for item 0)."""
        B = seq_pv.shape[0]
        ip = index_p.tolist()
        iv = index_v.tolist()
        W1, b1 = self.sd["struc_w1.weight"], self.sd["struc_w1.bias"]
        props, vals, finals = [], [], []
        t = None
        for i in range(B):
            props.append([])
            vals.append([])
            rows = []
            for j in range(len(ip[i])):
                if ip[i][j][0] == 0:
                    break
                p = seq_pv[i, ip[i][j]].mean(0)        # mean of exactly two rows
                v = seq_pv[i, iv[i][j]].mean(0)
                props[i].append(p)
                vals[i].append(v)
                rows.append(F.linear(torch.cat((c_initial[i], p, v)), W1, b1))
            if rows:
                t = torch.stack(rows)
            elif t is None:
                t = c_initial[i][None]
            att = torch.softmax(self.lin("struc_w2", F.leaky_relu(t)), dim=0)
            finals.append(c_initial[i] + self.lin("struc_w3", (att * t).sum(0)))
        c_final = torch.stack(finals)
        if not lpm:          # the alignment model's aggregator (`:3105-3161`) stops here
            return c_final, None
        nneg = self.cfg.num_negative_pv
        pos, neg = [], []
        for i in range(B):
            e = c_final[i]
            for j, (p, v) in enumerate(zip(props[i], vals[i])):
                n_ent = nneg // 2
                cand = [k for k in range(B) if k != i]
                if cand:
                    pn = torch.norm(e + p - v)
                    for k in random.sample(cand, min(len(cand), n_ent)):
                        pos.append(pn)
                        neg.append(torch.norm(c_final[k] + p - v))
                n_val = nneg - n_ent
                cand = [k for k in range(len(props[i])) if k != j]
                if cand:
                    pn = torch.norm(e + p - v)
                    for k in random.sample(cand, min(len(cand), n_val)):
                        pos.append(pn)
                        neg.append(torch.norm(e + p - vals[i][k]))
        if pos:
            pos_t, neg_t = torch.stack(pos), torch.stack(neg)
            # MarginRankingLoss(pos, neg, y=+1) = mean(max(0, -(pos - neg) + margin)) (`:2502`)
            loss = F.margin_ranking_loss(pos_t, neg_t, torch.ones_like(pos_t),
                                         margin=self.cfg.margin)
        else:
            loss = c_final.sum() * 0 + float("nan")  # mean over an empty set, as torch gives
        return c_final, loss

    # ---- heads -----------------------------------------------------------------------------
    def mlm_scores(self, x):
        # `:1795-1799` + `:1836-1839`; decoder weight tied to the word embeddings (`:2266`)
        h = self.ln("cls.predictions.transform.LayerNorm",
                    gelu(self.lin("cls.predictions.transform.dense", x)))
        return F.linear(h, self.sd["embeddings.word_embeddings.weight"]) + self.sd["cls.predictions.bias"]

    def image_scores(self, x):
        # `:1814-1818` + `:1921-1924`
        h = self.ln("cls.imagePredictions.transform.LayerNorm",
                    gelu(self.lin("cls.imagePredictions.transform.dense", x)))
        return self.lin("cls.imagePredictions.decoder", h)

    # ---- the step --------------------------------------------------------------------------
    def forward(self, batch, gumbel_noise=None):
        """Returns the reference's 10-tuple (`:2834-2846`)."""
        c = self.cfg
        f32 = torch.float32
        seq_v, seq_t, seq_pv, pool_v, pool_t, pool_pv, c_initial = self.tower(batch, gumbel_noise)
        c_final, loss_lpm = self.structure_aggregator(c_initial, seq_pv, batch["index_p"],
                                                      batch["index_v"])
        # heads `:1875-1909`
        itm = self.lin("cls.seq_relationship", self.drop(pool_t + pool_pv + pool_v, 0.1))
        sc_t = self.mlm_scores(seq_t)
        sc_pv = self.mlm_scores(seq_pv)
        sc_v = self.image_scores(seq_v)[:, 1:]                                # `:2745`
        # losses `:2754-2832`
        lab = (batch["image_label"] == 1)
        kl = F.kl_div(F.log_softmax(sc_v, dim=2), batch["image_target"], reduction="none")
        # reference divides by max(count, 0) (`:2760`): NaN/inf when nothing is masked
        loss_v = (kl * lab.unsqueeze(2).to(f32)).sum() / max(int(lab.sum()), 0) \
            if int(lab.sum()) > 0 else (kl * 0).sum() / 0.0
        V = c.vocab_size
        loss_t = F.cross_entropy(sc_t.reshape(-1, V), batch["masked_lm_labels"].reshape(-1),
                                 ignore_index=-1)
        loss_pv = F.cross_entropy(sc_pv.reshape(-1, V), batch["masked_lm_labels_pv"].reshape(-1),
                                  ignore_index=-1)
        nsl = 1 - 1 * ((batch["next_sentence_label"] + batch["next_sentence_label_pv_v"]
                        + batch["next_sentence_label_pv_t"]) == 0)
        loss_itm = F.cross_entropy(itm.view(-1, 2), nsl.view(-1).long(), ignore_index=-1)
        return (loss_t[None], loss_v[None], 0, loss_pv[None], 0, 0, loss_itm[None],
                c_initial, c_final, loss_lpm[None])

    def forward_alignment(self, labels, item1, item2, noise1=None, noise2=None):
        """`K3MForItemAlignment.forward`, "ce" branch (`:3379-3454`): two `item_embedding` calls with
        shared weights (`:3329-3377`), ClassificationHead (`:2164-2183`) on cat(c_final_1, c_final_2),
        cross-entropy.  Returns (probs [B, 2], loss [1], c_final_1, c_final_2)."""
        finals = []
        for item, gn in ((item1, noise1), (item2, noise2)):
            seq_v, seq_t, seq_pv, pool_v, pool_t, pool_pv, c_initial = self.tower(item, gn)
            finals.append(self.structure_aggregator(c_initial, seq_pv, item["index_p"], item["index_v"],
                                                    lpm=False)[0])
        x = self.drop(torch.cat(finals, dim=1), self.cfg.hidden_dropout_prob)
        x = torch.tanh(self.lin("classifier.dense", x))
        logits = self.lin("classifier.out_proj", self.drop(x, self.cfg.hidden_dropout_prob))
        loss = F.cross_entropy(logits.view(-1, 2), labels.view(-1).long())
        return torch.softmax(logits, dim=1), loss[None], finals[0], finals[1]

    def tower(self, batch, gumbel_noise=None):
        """`bert_tri` `:2507-2671` + fusion / pooling `:2376-2411` + `c_initial` `:2723` (`:3368-3373`)."""
        f32 = torch.float32

        def ext(m):  # `:2547-2580`
            return ((1.0 - m.to(f32)) * -10000.0)[:, None, None, :]

        m_t, m_pv, m_v = ext(batch["attention_mask"]), ext(batch["attention_mask_pv"]), ext(
            batch["image_attention_mask"])
        e_v = self.image_embeddings(batch["image_feat"], batch["image_loc"])
        e_t = self.text_embeddings(batch["input_ids"], batch["token_type_ids"])
        e_pv = self.text_embeddings(batch["input_ids_pv"], batch["token_type_ids_pv"])
        self.tap("emb.t", e_t), self.tap("emb.pv", e_pv), self.tap("emb.v", e_v)
        # three pairwise passes, `BertEncoder.forward` `:1686-1750`
        t_v, v_t = self.encoder_pass("tv", e_t, m_t, e_v, m_v, "c_layer", True)
        pv_v, v_pv = self.encoder_pass("pvv", e_pv, m_pv, e_v, m_v, "c_layer_pv_v", True)
        t_pv, pv_t = self.encoder_pass("tpv", e_t, m_t, e_pv, m_pv, "c_layer_pv_t", False)
        for n, x in (("t_v", t_v), ("v_t", v_t), ("pv_v", pv_v), ("v_pv", v_pv),
                     ("t_pv", t_pv), ("pv_t", pv_t)):
            self.tap("enc." + n, x)
        gn = gumbel_noise or {}
        seq_v = self.fuse("v", e_v, v_t, v_pv, gn.get("v"))
        seq_t = self.fuse("t", e_t, t_v, t_pv, gn.get("t"))
        seq_pv = self.fuse("pv", e_pv, pv_v, pv_t, gn.get("pv"))
        self.tap("fuse.v", seq_v), self.tap("fuse.t", seq_t), self.tap("fuse.pv", seq_pv)
        pool_v = self.lin("map_bi_to_individual", seq_v[:, 1:].mean(1))   # `:2405`
        pool_t = seq_t[:, 1:].mean(1)
        pool_pv = seq_pv[:, 1:].mean(1)
        c_initial = (pool_v + pool_t + pool_pv) / 3                           # `:2723`
        return seq_v, seq_t, seq_pv, pool_v, pool_t, pool_pv, c_initial


def total_loss(out, loss_img_weight=1.0):
    """`pretrain.py:1055`: mlm_t + w*mrm + mlm_pv + lpm (ITM is returned but never added)."""
    return out[0] + out[1] * loss_img_weight + out[3] + out[9]


def warmup_linear_lambda(step, warmup_steps, t_total):
    """pytorch_transformers==1.1.0 `WarmupLinearSchedule.lr_lambda` (third-party, absent here;
    restated from its published definition — parity unpinned)."""
    if step < warmup_steps:
        return float(step) / float(max(1, warmup_steps))
    return max(0.0, float(t_total - step) / float(max(1.0, t_total - warmup_steps)))


def param_groups(named_params, weight_decay=0.01):
    """Two groups by the substring rule of `pretrain.py:835-851`."""
    nd = ("bias", "LayerNorm.bias", "LayerNorm.weight")
    decay = [p for n, p in named_params if not any(s in n for s in nd)]
    no_decay = [p for n, p in named_params if any(s in n for s in nd)]
    return [{"params": decay, "weight_decay": weight_decay},
            {"params": no_decay, "weight_decay": 0.0}]
