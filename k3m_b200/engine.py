"""Host-side schedule of the K3M pretraining step: forward, hand-written backward, in terms of
the kernel contracts of `include/k3m.h` (op table `k3m_b200.ops.CudaOps`).

What the schedule follows in the reference (all in `vilbert_k3m/vilbert_k3m.py`):
`bert_tri` `:2507-2671` -> `BertEncoder.forward` `:1686-1750` (three pairwise passes) ->
`get_sequence_pooled_output_final` `:2376-2411` -> `structure_aggregator` `:2413-2505` ->
`BertPreTrainingHeads.forward` `:1875-1909` -> loss block `:2739-2846`.

B200-first differences from the reference's eager graph (results are the same):
  * The three passes run in lock-step.  The four text-side streams (title@t<->v, title@t<->pv,
    PV@pv<->v, PV@t<->pv) all apply the same `encoder.layer[i]` weights at the same point of
    their pass, so they are stacked along M into one [4*B*L, H] GEMM per projection; the two
    image streams are stacked the same way for `encoder.v_layer[i]`.
  * q/k/v (and the three fusion gates) are single fused GEMMs over adjacent arena weights.
  * bias, GELU, dropout, residual add live in GEMM epilogues; dropout masks are regenerated
    from a counter-based RNG in backward, never stored.
  * There is no autograd graph: every block has an explicit backward that writes weight
    gradients straight into one flat fp32 arena (the all-reduce / AdamW buffer).
  * Static shapes, no host synchronisation inside -> the whole step is CUDA-graph capturable.
"""
import os

import torch

from .struct_plan import build_struct_plan

ACT_NONE, ACT_GELU, ACT_DGELU, ACT_SIGMOID, ACT_TANH, ACT_DTANH, ACT_GELU_GRAD, ACT_MULAUX = 0, 1, 2, 3, 4, 5, 6, 7


class _Seg(object):
    """A run of `B` sequences of length `L` starting at row `r0` of a stacked activation."""
    __slots__ = ("r0", "B", "L", "mask")

    def __init__(self, r0, B, L, mask):
        self.r0, self.B, self.L, self.mask = r0, B, L, mask

    @property
    def r1(self):
        return self.r0 + self.B * self.L


def _merge_segs(segs):
    """Adjacent segments with equal L become one attention launch (masks concatenated)."""
    out = []
    for s in segs:
        if out and out[-1].L == s.L and out[-1].r1 == s.r0:
            p = out[-1]
            out[-1] = _Seg(p.r0, p.B + s.B, p.L, torch.cat([p.mask, s.mask], 0))
        else:
            out.append(s)
    return out


class Engine(object):
    def __init__(self, cfg, layout, ops, weights, grads, rng, act_dtype):
        """weights: flat arena in `act_dtype` (bf16 shadow of the fp32 masters);
        grads: flat fp32 arena of the same layout; rng: int64 [2] device tensor (seed, step)."""
        self.cfg, self.layout, self.ops = cfg, layout, ops
        self.w, self.g, self.rng, self.adt = weights, grads, rng, act_dtype
        self.training = False
        self.tape = []
        self._site = 0
        self._wv, self._gv = {}, {}
        self.on_ready = None     # callback(prefixes | None): weight gradients of these blocks are final
        # Weight / bias gradients of one layer group (or one connection stage) are collected while its backward runs
        # and leave as grouped launches at its end (`_flush_grads`): a text + image layer has 4 x 2 weight gradients
        # of one layout -> ONE persistent launch of 8 problems instead of four launches of two (each of which drains
        # the GPU behind its own last wave).  None = not collecting (every caller outside those two blocks).
        self._pend_w = self._pend_b = None
        self.defer_grads = os.environ.get("K3M_DEFER_GRADS", "1") != "0"
        self.ready_per_layer = os.environ.get("K3M_READY_PER_LAYER", "0") != "0"
        # BertIntermediate (`:515-518`): the forward epilogue stores gelu'(pre) instead of the pre-activation — Phi(pre) is
        # already in registers there, the derivative costs three more instructions — and the backward epilogue of the
        # FFN-down dgrad is a plain multiply instead of ~18 instructions per element of erf arithmetic in the K = 768
        # GEMMs whose tile time IS the epilogue (K3M_GELU_GRAD=0: pre-activation + ACT_DGELU as before)
        if os.environ.get("K3M_GELU_GRAD", "1") != "0":
            self.act_ffn_fwd, self.act_ffn_bwd = ACT_GELU_GRAD, ACT_MULAUX
        else:
            self.act_ffn_fwd, self.act_ffn_bwd = ACT_GELU, ACT_DGELU
        # per-layer parity taps (tests only; never set inside a captured step): when a dict, `tower_fwd` stores the
        # output of every BertLayer / BertImageLayer / BertConnectionLayer application under the key the reference's
        # module call would have: "<module name>#<call index>[/<output index>]" (tools/make_golden.py install_taps)
        self.taps = None

    # ---- views / small helpers -------------------------------------------------------------
    def W(self, name):
        v = self._wv.get(name)
        if v is None:
            v = self._wv[name] = self.layout.view(self.w, name)
        return v

    def G(self, name):
        v = self._gv.get(name)
        if v is None:
            v = self._gv[name] = self.layout.view(self.g, name)
        return v

    def site(self, p):
        """A dropout application: (p, unique site id) in training mode, else None."""
        self._site += 1
        if self.training and p > 0.0:
            return (float(p), self._site)
        return None

    def empty(self, rows, cols, dtype=None):
        return torch.empty(rows, cols, dtype=dtype or self.adt, device=self.w.device)

    def zeros(self, rows, cols, dtype=None):
        return torch.zeros(rows, cols, dtype=dtype or self.adt, device=self.w.device)

    def wgrad(self, dy, x, wname):
        """G[w] += dy^T x    (dy [M,out], x [M,in] -> [out,in])."""
        g = self.G(wname)
        self._aux(lambda: self.ops.gemm(dy, x, "tn", out=g, accumulate=True), (dy, x))

    def bgrad(self, dy, bname):
        """G[b] += column sums of dy."""
        g = self.G(bname)
        self._aux(lambda: self.ops.colsum(dy, g), (dy,))

    def bgrads(self, items):
        """G[b] += column sums of dy for every (dy, bname): one grouped launch on the auxiliary stream."""
        if self._pend_b is not None:
            self._pend_b.extend(items)
            return
        grp = getattr(self.ops, "colsum_group", None)
        if grp is None or len(items) == 1:
            for dy, b in items:
                self.bgrad(dy, b)
            return
        pairs = [(dy, self.G(b)) for dy, b in items]
        self._aux(lambda: grp(pairs), tuple(dy for dy, _ in items))

    def _aux(self, fn, keep):
        """Weight / bias gradients feed only the optimizer: they leave the dgrad chain (the critical
        path of backward) for an auxiliary stream when the op table has one."""
        aux = getattr(self.ops, "aux", None)
        if aux is None:
            fn()
        else:
            aux(fn, keep)

    def _join_aux(self):
        j = getattr(self.ops, "join_aux", None)
        if j is not None:
            j()

    def dgrad(self, dy, wname, **kw):
        """dy [M,out] @ W[out,in] -> [M,in]."""
        return self.ops.gemm(dy, self.W(wname), "nn", rng=self.rng, **kw)

    # ---- blocks ------------------------------------------------------------------------------
    # Every block works on a LIST of independent streams in lock-step (the three connection layers of a stage;
    # the text and the image layer of a stage): the same projection of all of them is ONE grouped GEMM launch
    # (`ops.gemm_group` -> k3m_gemm_grouped_bf16), so M ~ 2304-row problems fill whole waves of CTA pairs
    # together.  Attention and LayerNorm launches of the streams run on concurrent CUDA streams (`ops.parallel`),
    # weight / bias gradients leave for the auxiliary stream, grouped as well.
    def gemms(self, calls):
        grp = getattr(self.ops, "gemm_group", None)
        if grp is None or len(calls) == 1:
            return [self.ops.gemm(**c) for c in calls]
        return grp(calls)

    def wgrads(self, items):
        """G[w] += dy^T x for every (dy, x, wname): one grouped launch on the auxiliary stream."""
        if self._pend_w is not None:
            self._pend_w.extend(items)
            return
        calls = [dict(a=dy, b=x, layout="tn", out=self.G(w), accumulate=True) for dy, x, w in items]
        keep = tuple(t for dy, x, _ in items for t in (dy, x))
        self._aux(lambda: self.gemms(calls), keep)

    def _collect_grads(self):
        if self.defer_grads:
            self._pend_w, self._pend_b = [], []

    def _flush_grads(self):
        """Launch what `wgrads` / `bgrads` collected since `_collect_grads`: operands are fresh outputs of this
        block's backward or saved forward activations, none of which is written again before the flush."""
        pw, pb = self._pend_w, self._pend_b
        self._pend_w = self._pend_b = None
        if pb:
            for i in range(0, len(pb), 8):
                self.bgrads(pb[i:i + 8])
        if pw:
            # same reduction length (rows of the stream) first, so that a launch's tiles have equal k extents
            pw = sorted(pw, key=lambda it: -it[0].shape[0])
            for i in range(0, len(pw), 8):
                self.wgrads(pw[i:i + 8])

    def lns_fwd(self, items):
        """LayerNorm of every stream of a group: one grouped launch when the op table has one -> [(y, mean, rstd)]"""
        grp = getattr(self.ops, "ln_fwd_group", None)
        if grp is None or len(items) == 1:
            return [self.ops.ln_fwd(**it) for it in items]
        return grp(items)

    def lns_bwd(self, items):
        grp = getattr(self.ops, "ln_bwd_group", None)
        if grp is None or len(items) == 1:
            return [self.ops.ln_bwd(**it) for it in items]
        return grp(items)

    def _each(self, fns):
        """independent non-GEMM launches of the streams of a group: concurrent CUDA streams"""
        if len(fns) == 1:
            fns[0]()
        else:
            self.ops.parallel(fns)

    def attn_out_fwd_m(self, items):
        """LN(dropout(ctx W^T + b) + x)  — BertSelfOutput `:485-489` and the BertBiOutput halves `:986-996`.
        items: (ctx, x, p_dense, p_ln, p_drop) -> [(a, saved)]."""
        ops = self.ops
        sites = [self.site(it[4]) for it in items]
        zs = self.gemms([dict(a=ctx, b=self.W(pd + ".weight"), layout="nt", bias=self.W(pd + ".bias"), drop=s,
                              rng=self.rng, addend=x) for (ctx, x, pd, pl, _), s in zip(items, sites)])
        lns = self.lns_fwd([dict(x=zs[i], gamma=self.W(it[3] + ".weight"), beta=self.W(it[3] + ".bias"))
                            for i, it in enumerate(items)])
        return [(lns[i][0], (it[0], zs[i], lns[i][1], lns[i][2], sites[i], it[2], it[3])) for i, it in enumerate(items)]

    def attn_out_bwd_m(self, items):
        """items: (da, saved) -> [(dx residual grad, dctx)]."""
        lns = self.lns_bwd([dict(dy=da, x=z, gamma=self.W(pl + ".weight"), mean=mean, rstd=rstd,
                                 dgamma=self.G(pl + ".weight"), dbeta=self.G(pl + ".bias"), drop_in=s, rng=self.rng,
                                 dbias=self.G(pd + ".bias")) for da, (ctx, z, mean, rstd, s, pd, pl) in items])
        self.wgrads([(lns[i][1], it[1][0], it[1][5] + ".weight") for i, it in enumerate(items)])
        dctxs = self.gemms([dict(a=lns[i][1], b=self.W(it[1][5] + ".weight"), layout="nn", rng=self.rng)
                            for i, it in enumerate(items)])
        return [(lns[i][0], dctxs[i]) for i in range(len(items))]

    def ffn_fwd_m(self, items):
        """LN(dropout(gelu(a W1^T + b1) W2^T + b2) + a) — BertIntermediate `:515-518` + BertOutput `:528-532`
        (and the image twins).  items: (a, p_int, p_out, p_drop, out) -> [(y, saved)]."""
        ops = self.ops
        pres = [self.empty(a.shape[0], self.W(pi + ".dense.weight").shape[0]) for a, pi, _, _, _ in items]
        hs = self.gemms([dict(a=a, b=self.W(pi + ".dense.weight"), layout="nt", bias=self.W(pi + ".dense.bias"),
                              act=self.act_ffn_fwd, aux_out=pre) for (a, pi, _, _, _), pre in zip(items, pres)])
        sites = [self.site(it[3]) for it in items]
        zs = self.gemms([dict(a=h, b=self.W(po + ".dense.weight"), layout="nt", bias=self.W(po + ".dense.bias"),
                              drop=s, rng=self.rng, addend=a) for (a, _, po, _, _), h, s in zip(items, hs, sites)])
        lns = self.lns_fwd([dict(x=zs[i], gamma=self.W(it[2] + ".LayerNorm.weight"), beta=self.W(it[2] + ".LayerNorm.bias"),
                                 out=it[4]) for i, it in enumerate(items)])
        return [(lns[i][0], (it[0], pres[i], hs[i], zs[i], lns[i][1], lns[i][2], sites[i], it[1], it[2]))
                for i, it in enumerate(items)]

    def ffn_bwd_m(self, items):
        """items: (dy, saved) -> [da]."""
        lns = self.lns_bwd([dict(dy=dy, x=z, gamma=self.W(po + ".LayerNorm.weight"), mean=mean, rstd=rstd,
                                 dgamma=self.G(po + ".LayerNorm.weight"), dbeta=self.G(po + ".LayerNorm.bias"),
                                 drop_in=s, rng=self.rng, dbias=self.G(po + ".dense.bias"))
                            for dy, (a, pre, h, z, mean, rstd, s, pi, po) in items])
        self.wgrads([(lns[i][1], it[1][2], it[1][8] + ".dense.weight") for i, it in enumerate(items)])
        dpres = self.gemms([dict(a=lns[i][1], b=self.W(it[1][8] + ".dense.weight"), layout="nn", rng=self.rng,
                                 act=self.act_ffn_bwd, aux_in=it[1][1]) for i, it in enumerate(items)])
        self.bgrads([(dpres[i], it[1][7] + ".dense.bias") for i, it in enumerate(items)])
        self.wgrads([(dpres[i], it[1][0], it[1][7] + ".dense.weight") for i, it in enumerate(items)])
        return self.gemms([dict(a=dpres[i], b=self.W(it[1][7] + ".dense.weight"), layout="nn", rng=self.rng,
                                addend=lns[i][0]) for i, it in enumerate(items)])

    def layers_fwd_m(self, items):
        """BertLayer `:535-548` / BertImageLayer `:696-709` on stacked activations.
        items: (p, x, segs, image, out) -> [(y, saved)]."""
        c, ops = self.cfg, self.ops
        qkvs = self.gemms([dict(a=x, b=self.W(p + ".attention.self.qkv.weight"), layout="nt",
                                bias=self.W(p + ".attention.self.qkv.bias")) for p, x, _, _, _ in items])
        ctxs = [self.empty(x.shape[0], x.shape[1]) for _, x, _, _, _ in items]
        atts = [[] for _ in items]
        meta = []
        for p, x, segs, image, _ in items:
            H = x.shape[1]
            heads = c.v_num_attention_heads if image else c.num_attention_heads
            meta.append((H, heads, H // heads, c.v_attention_probs_dropout_prob if image else c.attention_probs_dropout_prob,
                         c.v_hidden_dropout_prob if image else c.hidden_dropout_prob))

        def attn(i):
            def run():
                H, heads, d, p_att, _ = meta[i]
                for sg in items[i][2]:
                    s = self.site(p_att)
                    v = qkvs[i][sg.r0:sg.r1]
                    _, lse = ops.attn_fwd(v, 0, v, H, 2 * H, sg.mask, sg.B, heads, sg.L, sg.L, d, drop=s, rng=self.rng,
                                          out=ctxs[i][sg.r0:sg.r1])
                    atts[i].append((sg, s, lse))
            return run
        self._each([attn(i) for i in range(len(items))])
        ao = self.attn_out_fwd_m([(ctxs[i], it[1], it[0] + ".attention.output.dense", it[0] + ".attention.output.LayerNorm",
                                   meta[i][4]) for i, it in enumerate(items)])
        ff = self.ffn_fwd_m([(ao[i][0], it[0] + ".intermediate", it[0] + ".output", meta[i][4], it[4])
                             for i, it in enumerate(items)])
        return [(ff[i][0], (it[0], it[1], qkvs[i], atts[i], meta[i][1], meta[i][2], ao[i][1], ff[i][1]))
                for i, it in enumerate(items)]

    def layers_bwd_m(self, items):
        """items: (dy, saved) -> [dx]."""
        ops = self.ops
        self._collect_grads()
        das = self.ffn_bwd_m([(dy, sv[7]) for dy, sv in items])
        ao = self.attn_out_bwd_m([(das[i], sv[6]) for i, (_, sv) in enumerate(items)])
        dqkvs = [self.empty(sv[1].shape[0], 3 * sv[1].shape[1]) for _, sv in items]

        def attn(i):
            def run():
                p, x, qkv, att, heads, d, sv_o, sv_f = items[i][1]
                H, ctx, dctx = x.shape[1], sv_o[0], ao[i][1]
                for sg, s, lse in att:
                    v, dv = qkv[sg.r0:sg.r1], dqkvs[i][sg.r0:sg.r1]
                    ops.attn_bwd(dctx[sg.r0:sg.r1], ctx[sg.r0:sg.r1], v, 0, v, H, 2 * H, sg.mask, lse, dv, dv,
                                 sg.B, heads, sg.L, sg.L, d, drop=s, rng=self.rng)
            return run
        self._each([attn(i) for i in range(len(items))])
        self.bgrads([(dqkvs[i], sv[0] + ".attention.self.qkv.bias") for i, (_, sv) in enumerate(items)])
        self.wgrads([(dqkvs[i], sv[1], sv[0] + ".attention.self.qkv.weight") for i, (_, sv) in enumerate(items)])
        self._flush_grads()
        return self.gemms([dict(a=dqkvs[i], b=self.W(sv[0] + ".attention.self.qkv.weight"), layout="nn", rng=self.rng,
                                addend=ao[i][0]) for i, (_, sv) in enumerate(items)])

    def connections_fwd_m(self, specs):
        """BertConnectionLayer `:1030-1071` / `_two_text` `:1074-1111`, the independent connection layers of one
        stage in lock-step.  specs: (p, x1, sg1, x2, sg2, two_text, out1, out2).  Stream 1 = image (PV in the
        two-text pass), stream 2 = text.  ctx1 (text-length) = queries of 2 over K/V of 1, masked by mask1
        (`:786-800`); ctx2 = queries of 1 over K/V of 2.  The image stream gets ctx2 through dense1, the text
        stream ctx1 through dense2 (`:1061-1063`).  -> [saved]."""
        c, ops = self.cfg, self.ops
        heads = c.bi_num_attention_heads
        calls = []
        for p, x1, sg1, x2, sg2, two_text, _, _ in specs:
            b = p + ".biattention."
            calls.append(dict(a=x1, b=self.W(b + "qkv1.weight"), layout="nt", bias=self.W(b + "qkv1.bias")))
            calls.append(dict(a=x2, b=self.W(b + "qkv2.weight"), layout="nt", bias=self.W(b + "qkv2.bias")))
        qk = self.gemms(calls)
        att = [None] * len(specs)

        def attn(i):
            def run():
                p, x1, sg1, x2, sg2, two_text, _, _ = specs[i]
                Hb = c.hidden_size if two_text else c.bi_hidden_size
                d = Hb // heads
                qkv1, qkv2 = qk[2 * i], qk[2 * i + 1]
                s1 = self.site(c.v_attention_probs_dropout_prob)
                ctx1, lse1 = ops.attn_fwd(qkv2, 0, qkv1, Hb, 2 * Hb, sg1.mask, sg1.B, heads, sg2.L, sg1.L, d,
                                          drop=s1, rng=self.rng)
                s2 = self.site(c.attention_probs_dropout_prob)
                ctx2, lse2 = ops.attn_fwd(qkv1, 0, qkv2, Hb, 2 * Hb, sg2.mask, sg1.B, heads, sg1.L, sg2.L, d,
                                          drop=s2, rng=self.rng)
                att[i] = (Hb, d, ctx1, ctx2, (s1, lse1), (s2, lse2))
            return run
        self._each([attn(i) for i in range(len(specs))])
        items = []
        for i, (p, x1, sg1, x2, sg2, two_text, _, _) in enumerate(specs):
            o = p + ".biOutput."
            items.append((att[i][3], x1, o + "dense1", o + "LayerNorm1", c.v_hidden_dropout_prob))
            items.append((att[i][2], x2, o + "dense2", o + "LayerNorm2", c.hidden_dropout_prob))
        ao = self.attn_out_fwd_m(items)
        items = []
        for i, (p, x1, sg1, x2, sg2, two_text, out1, out2) in enumerate(specs):
            p1 = c.hidden_dropout_prob if two_text else c.v_hidden_dropout_prob
            items.append((ao[2 * i][0], p + ".v_intermediate", p + ".v_output", p1, out1))
            items.append((ao[2 * i + 1][0], p + ".t_intermediate", p + ".t_output", c.hidden_dropout_prob, out2))
        ff = self.ffn_fwd_m(items)
        saved = []
        for i, (p, x1, sg1, x2, sg2, two_text, _, _) in enumerate(specs):
            Hb, d, ctx1, ctx2, a1, a2 = att[i]
            saved.append((p, x1, sg1, x2, sg2, Hb, heads, d, qk[2 * i], qk[2 * i + 1], a1, a2,
                          ao[2 * i][1], ao[2 * i + 1][1], ff[2 * i][1], ff[2 * i + 1][1]))
        return saved

    def connections_bwd_m(self, items):
        """items: (dy1, dy2, saved, dx1_out, dx2_out)."""
        ops = self.ops
        self._collect_grads()
        ffi = []
        for dy1, dy2, sv, _, _ in items:
            ffi += [(dy1, sv[14]), (dy2, sv[15])]
        das = self.ffn_bwd_m(ffi)
        aoi = []
        for i, (_, _, sv, _, _) in enumerate(items):
            aoi += [(das[2 * i], sv[12]), (das[2 * i + 1], sv[13])]
        ao = self.attn_out_bwd_m(aoi)         # [2i] -> (dz1 residual grad of x1, dctx2), [2i+1] -> (dz2, dctx1)
        dq = []
        for _, _, sv, _, _ in items:
            dq.append((self.empty(sv[1].shape[0], 3 * sv[5]), self.empty(sv[3].shape[0], 3 * sv[5])))

        def attn(i):
            def run():
                (p, x1, sg1, x2, sg2, Hb, heads, d, qkv1, qkv2, (s1, lse1), (s2, lse2), sv_o1, sv_o2, _, _) = items[i][2]
                dqkv1, dqkv2 = dq[i]
                dctx2, dctx1 = ao[2 * i][1], ao[2 * i + 1][1]
                ctx2, ctx1 = sv_o1[0], sv_o2[0]
                ops.attn_bwd(dctx1, ctx1, qkv2, 0, qkv1, Hb, 2 * Hb, sg1.mask, lse1, dqkv2, dqkv1,
                             sg1.B, heads, sg2.L, sg1.L, d, drop=s1, rng=self.rng)
                ops.attn_bwd(dctx2, ctx2, qkv1, 0, qkv2, Hb, 2 * Hb, sg2.mask, lse2, dqkv1, dqkv2,
                             sg1.B, heads, sg1.L, sg2.L, d, drop=s2, rng=self.rng)
            return run
        self._each([attn(i) for i in range(len(items))])
        wg, bg, calls = [], [], []
        for i, (_, _, sv, dx1_out, dx2_out) in enumerate(items):
            b = sv[0] + ".biattention."
            bg += [(dq[i][0], b + "qkv1.bias"), (dq[i][1], b + "qkv2.bias")]
            wg += [(dq[i][0], sv[1], b + "qkv1.weight"), (dq[i][1], sv[3], b + "qkv2.weight")]
            calls.append(dict(a=dq[i][0], b=self.W(b + "qkv1.weight"), layout="nn", rng=self.rng, addend=ao[2 * i][0],
                              out=dx1_out))
            calls.append(dict(a=dq[i][1], b=self.W(b + "qkv2.weight"), layout="nn", rng=self.rng, addend=ao[2 * i + 1][0],
                              out=dx2_out))
        self.bgrads(bg)
        self.wgrads(wg)
        self._flush_grads()
        self.gemms(calls)

    # ---- embeddings ------------------------------------------------------------------------
    def text_embed_fwd(self, ids, type_ids, L):
        """BertEmbeddings `:361-382` (position ids are always arange(L), `:364`)."""
        ops = self.ops
        e = "embeddings."
        x = ops.embed_fwd(ids, type_ids, self.W(e + "word_embeddings.weight"),
                          self.W(e + "position_embeddings.weight"),
                          self.W(e + "token_type_embeddings.weight"), L)
        s = self.site(self.cfg.hidden_dropout_prob)
        y, mean, rstd = ops.ln_fwd(x, self.W(e + "LayerNorm.weight"), self.W(e + "LayerNorm.bias"),
                                   drop=s, rng=self.rng)
        return y, (ids, type_ids, L, x, mean, rstd, s)

    def text_embed_bwd(self, dy, saved):
        ids, type_ids, L, x, mean, rstd, s = saved
        e = "embeddings."
        dx, _ = self.ops.ln_bwd(dy, x, self.W(e + "LayerNorm.weight"), mean, rstd,
                                self.G(e + "LayerNorm.weight"), self.G(e + "LayerNorm.bias"),
                                drop_out=s, rng=self.rng)
        self.ops.embed_bwd(dx, ids, type_ids, L, self.G(e + "word_embeddings.weight"),
                           self.G(e + "position_embeddings.weight"),
                           self.G(e + "token_type_embeddings.weight"))

    def image_embed_fwd(self, feat, loc):
        """BertImageEmbeddings `:2153-2161`: the 5-d box projection is the GEMM's addend."""
        ops = self.ops
        e = "v_embeddings."
        fa = ops.cast(feat, self.adt)
        le = ops.loc_embed_fwd(loc, self.W(e + "image_location_embeddings.weight"),
                               self.W(e + "image_location_embeddings.bias"))
        z = ops.gemm(fa, self.W(e + "image_embeddings.weight"), "nt",
                     bias=self.W(e + "image_embeddings.bias"), addend=le)
        s = self.site(self.cfg.hidden_dropout_prob)
        y, mean, rstd = ops.ln_fwd(z, self.W(e + "LayerNorm.weight"), self.W(e + "LayerNorm.bias"),
                                   drop=s, rng=self.rng)
        return y, (fa, loc, z, mean, rstd, s)

    def image_embed_bwd(self, dy, saved):
        fa, loc, z, mean, rstd, s = saved
        e = "v_embeddings."
        _, dz = self.ops.ln_bwd(dy, z, self.W(e + "LayerNorm.weight"), mean, rstd,
                                self.G(e + "LayerNorm.weight"), self.G(e + "LayerNorm.bias"),
                                drop_out=s, rng=self.rng, dbias=self.G(e + "image_embeddings.bias"),
                                want_masked=True)
        self.wgrad(dz, fa, e + "image_embeddings.weight")
        self.ops.loc_embed_bwd(dz, loc, self.G(e + "image_location_embeddings.weight"),
                               self.G(e + "image_location_embeddings.bias"))

    # ---- fusion ----------------------------------------------------------------------------
    def fuse_fwd(self, mod, indiv, c1, c2, noise, out):
        """`get_sequence_pooled_output_final` `:2376-2411`; mode 1 = `pre_sampling_sequence`
        `:2331-2374` (hard Gumbel pick per token and channel, straight-through backward)."""
        ops, mode = self.ops, self.cfg.if_pre_sampling
        if mode == 0:
            ops.lincomb3(indiv, c1, c2, 1.0 / 3, 1.0 / 3, 1.0 / 3, out=out)
            return (mode,)
        if mode == 2:
            # `pre_sampling_sequence_soft` `:2300-2329`: soft_m(cat(relu(.) * sigmoid(score_*(cat))))
            cat = ops.relu_cat3(indiv, c1, c2)
            alpha = ops.gemm(cat, self.W("score_%s.weight" % mod), "nt", bias=self.W("score_%s.bias" % mod),
                             act=ACT_SIGMOID)
            gated = ops.gate_mul(cat, alpha)
            ops.gemm(gated, self.W("soft_%s.weight" % mod), "nt", bias=self.W("soft_%s.bias" % mod), out=out)
            return (mode, mod, cat, alpha, gated)
        if mode != 1:
            ops.lincomb3(c1, c2, None, 0.5, 0.5, 0.0, out=out)
            return (mode,)
        cat = ops.relu_cat3(indiv, c1, c2)
        gates = ops.gemm(cat, self.W("score_%s.weight" % mod), "nt",
                         bias=self.W("score_%s.bias" % mod), out_dtype=torch.float32)
        self._site += 1
        site = self._site
        ops.fuse_select_fwd(cat, gates, noise=noise, site=site, rng=self.rng, out=out)
        return (mode, mod, cat, gates, noise, site)

    def fuse_bwd(self, dout, saved, d_indiv, d_c1, d_c2):
        """Accumulates into d_indiv / d_c1 / d_c2 (all pre-initialised)."""
        ops, mode = self.ops, saved[0]
        if mode == 0:
            for t in (d_indiv, d_c1, d_c2):
                ops.axpy(dout, t, 1.0 / 3)
            return
        if mode == 2:
            _, mod, cat, alpha, gated = saved
            self.bgrad(dout, "soft_%s.bias" % mod)
            self.wgrad(dout, gated, "soft_%s.weight" % mod)
            dgated = self.dgrad(dout, "soft_%s.weight" % mod)
            dcat, dscore = ops.gate_bwd(dgated, cat, alpha)
            self.bgrad(dscore, "score_%s.bias" % mod)
            self.wgrad(dscore, cat, "score_%s.weight" % mod)
            self.dgrad(dscore, "score_%s.weight" % mod, addend=dcat, out=dcat)
            ops.relu_bwd_split3(dcat, cat, d_indiv, d_c1, d_c2)
            return
        if mode != 1:
            ops.axpy(dout, d_c1, 0.5)
            ops.axpy(dout, d_c2, 0.5)
            return
        _, mod, cat, gates, noise, site = saved
        dgates, dcat = ops.fuse_select_bwd(dout, cat, gates, noise=noise, site=site, rng=self.rng)
        self.bgrad(dgates, "score_%s.bias" % mod)
        self.wgrad(dgates, cat, "score_%s.weight" % mod)
        self.dgrad(dgates, "score_%s.weight" % mod, addend=dcat, out=dcat)
        ops.relu_bwd_split3(dcat, cat, d_indiv, d_c1, d_c2)

    # ---- heads -----------------------------------------------------------------------------
    def head_fwd(self, p, x, p_dec_w, p_dec_b, logits_dtype, n_cols):
        """transform (dense -> GELU -> LN, `:1795-1799` / `:1814-1818`) + decoder."""
        ops = self.ops
        M, H = x.shape
        pre = self.empty(M, H)
        h = ops.gemm(x, self.W(p + ".transform.dense.weight"), "nt",
                     bias=self.W(p + ".transform.dense.bias"), act=ACT_GELU, aux_out=pre)
        hn, mean, rstd = ops.ln_fwd(h, self.W(p + ".transform.LayerNorm.weight"),
                                    self.W(p + ".transform.LayerNorm.bias"))
        ld = (n_cols + 7) // 8 * 8
        logits = self.zeros(M, ld, logits_dtype) if ld != n_cols else self.empty(M, ld, logits_dtype)
        ops.gemm(hn, self.W(p_dec_w), "nt", bias=self.W(p_dec_b), out=logits[:, :n_cols])
        return logits, (p, x, pre, h, hn, mean, rstd, p_dec_w, p_dec_b, n_cols)

    def head_bwd(self, dlogits, saved, dx):
        """dlogits [M, ld] in act dtype (columns >= n_cols are zero); dx += grad of x."""
        p, x, pre, h, hn, mean, rstd, p_dec_w, p_dec_b, n_cols = saved
        ops = self.ops
        dl = dlogits[:, :n_cols]
        self.bgrad(dl, p_dec_b)
        self.wgrad(dl, hn, p_dec_w)
        if n_cols >= 8192:
            # vocabulary decoder: [rows, 21128] x [21128, H] has 331 k-blocks but only 21 output tiles (labelled rows
            # x H) — a bf16 output cannot be split along k, so it ran on 21 of 74 CTA pairs (170 us, 317 TF/s).
            # Accumulating into a zeroed fp32 buffer lets the split-K planner fill the machine; one rounding at the end.
            acc = self.zeros(dl.shape[0], self.W(p_dec_w).shape[1], torch.float32)
            ops.gemm(dl, self.W(p_dec_w), "nn", out=acc, accumulate=True)
            dhn = acc if self.adt == torch.float32 else ops.cast(acc, self.adt)
        else:
            dhn = self.dgrad(dl, p_dec_w)
        dh, _ = ops.ln_bwd(dhn, h, self.W(p + ".transform.LayerNorm.weight"), mean, rstd,
                           self.G(p + ".transform.LayerNorm.weight"),
                           self.G(p + ".transform.LayerNorm.bias"))
        dpre = ops.act_bwd(dh, pre, ACT_DGELU)
        self.bgrad(dpre, p + ".transform.dense.bias")
        self.wgrad(dpre, x, p + ".transform.dense.weight")
        self.dgrad(dpre, p + ".transform.dense.weight", addend=dx, out=dx)

    # ---- structure aggregator + LPM ----------------------------------------------------------
    def struct_fwd(self, c_init, seq_pv, plan, lpm=True):
        """`structure_aggregator` `:2413-2505` as batched kernels over the host-built plan
        (`lpm=False`: the alignment model's aggregator, `:3105-3161`, which has no LPM loss)."""
        ops, f32 = self.ops, torch.float32
        B, H = c_init.shape
        X, PV = ops.pair_gather_fwd(c_init, seq_pv, plan["pair_item"], plan["pair_rows"])
        t = ops.gemm(X, self.W("struc_w1.weight"), "nt", bias=self.W("struc_w1.bias"), out_dtype=f32)
        ops.struct_fallback_fwd(t, c_init, plan["fallback"])      # leading items without PVs (`:2452-2456`)
        s, att = ops.struct_attn_fwd(t, self.W("struc_w2.weight"), self.W("struc_w2.bias"), plan["seg"])
        s_a = ops.cast(s, self.adt)
        c_final = ops.gemm(s_a, self.W("struc_w3.weight"), "nt", bias=self.W("struc_w3.bias"),
                           addend=c_init, out_dtype=f32)
        dc_final = self.zeros(B, H, f32)
        dPV = self.zeros(PV.shape[0], 2 * H, f32)
        loss = None
        if lpm:
            loss = ops.lpm_fwd_bwd(c_final, PV, plan["terms"], float(self.cfg.margin), dc_final, dPV,
                                   n_valid=plan.get("n_terms"))
        return c_final, loss, (plan, X, t, att, s_a, dc_final, dPV)

    def struct_bwd(self, saved, w_lpm, dc_init, dseq_pv):
        """dc_init [B,H] fp32 += ; dseq_pv [B*L,H] += ."""
        plan, X, t, att, s_a, dc_final, dPV = saved
        ops, f32 = self.ops, torch.float32
        if w_lpm != 1.0:
            ops.scale_(dc_final, w_lpm)
            ops.scale_(dPV, w_lpm)
        ops.axpy(dc_final, dc_init, 1.0)
        self.bgrad(dc_final, "struc_w3.bias")
        dcf = ops.cast(dc_final, self.adt)
        self.wgrad(dcf, s_a, "struc_w3.weight")
        ds = self.dgrad(dcf, "struc_w3.weight", out_dtype=f32)
        dt = self.zeros(t.shape[0], t.shape[1], f32)
        ops.struct_attn_bwd(ds, t, att, self.W("struc_w2.weight"), plan["seg"], dt,
                            self.G("struc_w2.weight"), self.G("struc_w2.bias"))
        ops.struct_fallback_bwd(dt, dc_init, plan["fallback"])
        self.bgrad(dt, "struc_w1.bias")
        dta = ops.cast(dt, self.adt)
        self.wgrad(dta, X, "struc_w1.weight")
        dX = self.dgrad(dta, "struc_w1.weight", out_dtype=f32)
        ops.pair_gather_bwd(dX, dPV, plan["pair_item"], plan["pair_rows"], dc_init, dseq_pv)

    # ---- the step --------------------------------------------------------------------------
    def forward(self, batch, plan, noise=None, loss_weights=(1.0, 1.0, 1.0, 1.0)):
        """batch: dict of device tensors named as the reference's forward() arguments
        (`vilbert_k3m.py:2673-2696`); plan: device tensors from `build_struct_plan`;
        noise: optional dict mod -> [B*L, 3, H] fp32 Gumbel noise (parity mode);
        loss_weights: (mlm_t, mrm, mlm_pv, lpm) weights of the summed loss (`pretrain.py:1055`).
        Returns dict(loss_t, loss_v, loss_pv, loss_itm, loss_lpm (each fp32 [1]), c_initial,
        c_final)."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        tw = self.tower_fwd(batch, noise)
        B, Lt, Lpv, Lv1, Mt, Mp, Mv = self.saved["dims"]
        S, seqV, c_init = tw["S"], tw["seqV"], tw["c_init"]
        pool_t, pool_pv, pool_v = tw["pools"]
        c_final, loss_lpm, sv_s = self.struct_fwd(c_init, S[Mt:], plan)
        loss_itm = ops.itm_loss(pool_t, pool_pv, pool_v, self.W("cls.seq_relationship.weight"),
                                self.W("cls.seq_relationship.bias"), batch["next_sentence_label"],
                                batch["next_sentence_label_pv_v"], batch["next_sentence_label_pv_t"],
                                drop=self.site(0.1), rng=self.rng)
        w_t, w_v, w_pv, w_lpm = [float(x) for x in loss_weights]
        V, Tv = c.vocab_size, c.v_target_size
        lab_t, lab_pv = batch["masked_lm_labels"].reshape(-1), batch["masked_lm_labels_pv"].reshape(-1)
        mlm_rows = plan.get("mlm_rows")
        if mlm_rows is not None:
            # The reference decodes EVERY position and lets CrossEntropyLoss ignore the unlabelled ~85 % (`:1836-1839`,
            # `:2817-2826`): those rows give neither loss nor gradient, so transform + decoder + loss run on the
            # labelled rows only, gathered by the host-built row list (`struct_plan.build_mlm_plan`; bit-exact row
            # copies, same losses and gradients).  The gathered logits [~1.2 k, 21128] are materialised on purpose:
            # recomputing them tile by tile in backward would cost two more vocabulary GEMM passes (75 GF) to save
            # ~250 MB of traffic (40 us).
            cap_t = plan["mlm_cap_t"]
            Sg = ops.gather_rows(S, mlm_rows)
            logits, sv_h = self.head_fwd("cls.predictions", Sg, "embeddings.word_embeddings.weight",
                                         "cls.predictions.bias", self.adt, V)
            lab_g = plan["mlm_labels"]
            loss_t = ops.xent_fwd_bwd(logits[:cap_t], lab_g[:cap_t], ops.count_valid(lab_t, -1), V,
                                      dlogits=logits[:cap_t], grad_scale=w_t)
            loss_pv = ops.xent_fwd_bwd(logits[cap_t:], lab_g[cap_t:], ops.count_valid(lab_pv, -1), V,
                                       dlogits=logits[cap_t:], grad_scale=w_pv)
        else:
            logits, sv_h = self.head_fwd("cls.predictions", S, "embeddings.word_embeddings.weight",
                                         "cls.predictions.bias", self.adt, V)
            loss_t = ops.xent_fwd_bwd(logits[:Mt], lab_t, ops.count_valid(lab_t, -1), V,
                                      dlogits=logits[:Mt], grad_scale=w_t)
            loss_pv = ops.xent_fwd_bwd(logits[Mt:], lab_pv, ops.count_valid(lab_pv, -1), V,
                                       dlogits=logits[Mt:], grad_scale=w_pv)
        logits_v, sv_hv = self.head_fwd("cls.imagePredictions", seqV,
                                        "cls.imagePredictions.decoder.weight",
                                        "cls.imagePredictions.decoder.bias", f32, Tv)
        dlv = self.zeros(Mv, logits_v.shape[1])
        il = batch["image_label"].reshape(-1)
        loss_v = ops.kl_fwd_bwd(logits_v, batch["image_target"].reshape(-1, Tv), il,
                                ops.count_eq(il, 1), Tv, Lv1, dlogits=dlv, grad_scale=w_v)
        self.saved.update(struct=sv_s, w_lpm=w_lpm, head=(logits, sv_h), head_v=(dlv, sv_hv), mlm_rows=mlm_rows,
                          mlm_cap_t=plan.get("mlm_cap_t"))
        return dict(loss_t=loss_t, loss_v=loss_v, loss_pv=loss_pv, loss_itm=loss_itm,
                    loss_lpm=loss_lpm, c_initial=c_init, c_final=c_final)

    def tower_fwd(self, batch, noise=None):
        """Everything up to the initial item embedding (`bert_tri` `:2507-2671`, the three-pass
        encoder, fusion `:2376-2411`, pooling, `c_initial`): shared by the pretraining model and the
        item-alignment towers (`item_embedding`, `:3329-3377`).  Fills `self.saved` for
        `tower_bwd`; returns dict(S [title | PV fused sequences], seqV, c_init, pools)."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        self.tape, self._site = [], 0
        noise = noise or {}
        ids_t, ids_pv = batch["input_ids"], batch["input_ids_pv"]
        B, Lt = ids_t.shape
        Lpv = ids_pv.shape[1]
        Lv1 = batch["image_feat"].shape[1]
        H, Hv = c.hidden_size, c.v_hidden_size
        Mt, Mp, Mv = B * Lt, B * Lpv, B * Lv1
        mask_t = ops.mask_additive(batch["attention_mask"])          # `:2561`
        mask_pv = ops.mask_additive(batch["attention_mask_pv"])      # `:2570`
        mask_v = ops.mask_additive(batch["image_attention_mask"])    # `:2580`
        e_t, sv_et = self.text_embed_fwd(ids_t.reshape(-1), batch["token_type_ids"].reshape(-1), Lt)
        e_pv, sv_epv = self.text_embed_fwd(ids_pv.reshape(-1), batch["token_type_ids_pv"].reshape(-1), Lpv)
        e_v, sv_ev = self.image_embed_fwd(batch["image_feat"].reshape(Mv, -1),
                                          batch["image_loc"].reshape(Mv, -1))
        # stacked streams: text [T1 | T3 | P2 | P3], image [V1 | V2]
        T1, T3 = slice(0, Mt), slice(Mt, 2 * Mt)
        P2, P3 = slice(2 * Mt, 2 * Mt + Mp), slice(2 * Mt + Mp, 2 * Mt + 2 * Mp)
        V1, V2 = slice(0, Mv), slice(Mv, 2 * Mv)
        X = self.empty(2 * Mt + 2 * Mp, H)
        Y = self.empty(2 * Mv, Hv)
        for sl in (T1, T3):
            X[sl].copy_(e_t)
        for sl in (P2, P3):
            X[sl].copy_(e_pv)
        for sl in (V1, V2):
            Y[sl].copy_(e_v)
        sgT, sgP, sgV = _Seg(0, B, Lt, mask_t), _Seg(0, B, Lpv, mask_pv), _Seg(0, B, Lv1, mask_v)
        segs_x = _merge_segs([_Seg(0, B, Lt, mask_t), _Seg(Mt, B, Lt, mask_t),
                              _Seg(2 * Mt, B, Lpv, mask_pv), _Seg(2 * Mt + Mp, B, Lpv, mask_pv)])
        segs_y = _merge_segs([_Seg(0, B, Lv1, mask_v), _Seg(Mv, B, Lv1, mask_v)])
        # stages: ("tv", [text layer saves], [image layer saves]) | ("c", save_tv, save_pvv, save_tpv).
        # The text and image stacks between two connection stages are independent, and so are the three
        # connection layers of a stage: they are handed to ops.parallel (concurrent streams).
        stages = []
        s1 = s2 = 0
        box = {"X": X, "Y": Y}

        taps = self.taps
        t_last = c.t_biattention_id[-1]

        def tap_text(i):
            if taps is not None:
                # call order of the shared module over the three passes (`:1686-1750`): title (t<->v), PV
                # (pv<->v), then the two-text pass — title before PV inside its loop (`:1552-1576`), PV
                # before title in the closing loops (`:1655-1672`)
                order = (T1, P2, P3, T3) if i >= t_last else (T1, P2, T3, P3)
                for k, sl in enumerate(order):
                    taps["encoder.layer.%d#%d" % (i, k)] = box["X"][sl]

        def tap_image(i):
            if taps is not None:
                for k, sl in enumerate((V1, V2)):
                    taps["encoder.v_layer.%d#%d" % (i, k)] = box["Y"][sl]

        def run_layers(t_range, v_range):
            """text layers `t_range` and image layers `v_range` (independent stacks): pairs of one text and one image
            layer in lock-step (grouped GEMM launches), the rest of the longer stack alone"""
            tsv, vsv = [], []
            t_list, v_list = list(t_range), list(v_range)
            for k in range(max(len(t_list), len(v_list))):
                items = []
                if k < len(t_list):
                    items.append(("encoder.layer.%d" % t_list[k], box["X"], segs_x, False, None))
                if k < len(v_list):
                    items.append(("encoder.v_layer.%d" % v_list[k], box["Y"], segs_y, True, None))
                res = self.layers_fwd_m(items)
                j = 0
                if k < len(t_list):
                    box["X"], sv = res[j]
                    tsv.append(sv)
                    tap_text(t_list[k])
                    j += 1
                if k < len(v_list):
                    box["Y"], sv = res[j]
                    vsv.append(sv)
                    tap_image(v_list[k])
            stages.append(("tv", tsv, vsv))

        for count, (e1, e2) in enumerate(zip(c.v_biattention_id, c.t_biattention_id)):
            run_layers(range(s2, e2), range(s1, e1))
            X, Y = box["X"], box["Y"]
            Xn, Yn = torch.empty_like(X), torch.empty_like(Y)
            sv_c = self.connections_fwd_m([
                ("encoder.c_layer.%d" % count, Y[V1], sgV, X[T1], sgT, False, Yn[V1], Xn[T1]),
                ("encoder.c_layer_pv_v.%d" % count, Y[V2], sgV, X[P2], sgP, False, Yn[V2], Xn[P2]),
                ("encoder.c_layer_pv_t.%d" % count, X[P3], sgP, X[T3], sgT, True, Xn[P3], Xn[T3])])
            stages.append(("c", sv_c[0], sv_c[1], sv_c[2]))
            if taps is not None:
                for nm, o0, o1 in (("c_layer", Yn[V1], Xn[T1]), ("c_layer_pv_v", Yn[V2], Xn[P2]),
                                   ("c_layer_pv_t", Xn[P3], Xn[T3])):
                    taps["encoder.%s.%d#0/0" % (nm, count)] = o0
                    taps["encoder.%s.%d#0/1" % (nm, count)] = o1
            box["X"], box["Y"] = Xn, Yn
            s1, s2 = e1, e2
        run_layers(range(s2, c.num_hidden_layers), range(s1, c.v_num_hidden_layers))
        X, Y = box["X"], box["Y"]
        # fusion: title / PV sequences land in one stacked buffer S for the shared MLM head
        S = self.empty(Mt + Mp, H)
        seqV = self.empty(Mv, Hv)
        sv_fv = self.fuse_fwd("v", e_v, Y[V1], Y[V2], noise.get("v"), seqV)
        sv_ft = self.fuse_fwd("t", e_t, X[T1], X[T3], noise.get("t"), S[:Mt])
        sv_fp = self.fuse_fwd("pv", e_pv, X[P2], X[P3], noise.get("pv"), S[Mt:])
        pool_t = ops.mean_pool_fwd(S[:Mt], B, Lt, f32)
        pool_pv = ops.mean_pool_fwd(S[Mt:], B, Lpv, f32)
        pool_vr = ops.mean_pool_fwd(seqV, B, Lv1, self.adt)
        pool_v = ops.gemm(pool_vr, self.W("map_bi_to_individual.weight"), "nt",
                          bias=self.W("map_bi_to_individual.bias"), out_dtype=f32)
        c_init = ops.lincomb3(pool_v, pool_t, pool_pv, 1.0 / 3, 1.0 / 3, 1.0 / 3, out_dtype=f32)
        self.saved = dict(dims=(B, Lt, Lpv, Lv1, Mt, Mp, Mv), stages=stages, emb=(sv_et, sv_epv, sv_ev),
                          fuse=(sv_fv, sv_ft, sv_fp), pool_vr=pool_vr)
        return dict(S=S, seqV=seqV, c_init=c_init, pools=(pool_t, pool_pv, pool_v))

    def backward(self, grad_scales=None):
        """Gradient of w_t*loss_t + w_v*loss_v + w_pv*loss_pv + w_lpm*loss_lpm, accumulated into
        the fp32 gradient arena (which the caller zeroes once per optimizer step).
        `grad_scales`: optional (g_t, g_v, g_pv, g_lpm) fp32 DEVICE scalars (autograd's upstream
        gradients of the four losses, `k3m_b200.vilbert_k3m`); each multiplies the stored
        d(loss)/d(logits) of its loss before anything is propagated."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        sv = self.saved
        B, Lt, Lpv, Lv1, Mt, Mp, Mv = sv["dims"]
        if grad_scales is not None:
            g_t, g_v, g_pv, g_lpm = grad_scales
            dlog = sv["head"][0]
            cut = sv["mlm_cap_t"] if sv.get("mlm_rows") is not None else Mt
            ops.scale_dev_(dlog[:cut], g_t)
            ops.scale_dev_(dlog[cut:], g_pv)
            ops.scale_dev_(sv["head_v"][0], g_v)
            ops.scale_dev_(sv["struct"][5], g_lpm)      # dc_final
            ops.scale_dev_(sv["struct"][6], g_lpm)      # dPV
        H, Hv = c.hidden_size, c.v_hidden_size
        dS = self.zeros(Mt + Mp, H)
        dseqV = self.zeros(Mv, Hv)
        if sv.get("mlm_rows") is not None:
            dSg = self.zeros(sv["mlm_rows"].numel(), H)
            self.head_bwd(sv["head"][0], sv["head"][1], dSg)
            ops.scatter_add_rows(dSg, sv["mlm_rows"], dS)
        else:
            self.head_bwd(sv["head"][0], sv["head"][1], dS)
        self.head_bwd(sv["head_v"][0], sv["head_v"][1], dseqV)
        dc_init = self.zeros(B, H, f32)
        self.struct_bwd(sv["struct"], sv["w_lpm"], dc_init, dS[Mt:])
        self.tower_bwd(dS, dseqV, dc_init)

    def tower_bwd(self, dS, dseqV, dc_init):
        """Backward of `tower_fwd`: dS [Mt + Mp, H] / dseqV [Mv, Hv] (act dtype) and dc_init [B, H] fp32
        are the gradients of its outputs (consumed / modified in place)."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        sv = self.saved
        B, Lt, Lpv, Lv1, Mt, Mp, Mv = sv["dims"]
        H, Hv = c.hidden_size, c.v_hidden_size
        T1, T3 = slice(0, Mt), slice(Mt, 2 * Mt)
        P2, P3 = slice(2 * Mt, 2 * Mt + Mp), slice(2 * Mt + Mp, 2 * Mt + 2 * Mp)
        V1, V2 = slice(0, Mv), slice(Mv, 2 * Mv)
        ops.scale_(dc_init, 1.0 / 3)                       # c_initial = (v + t + pv) / 3
        ops.mean_pool_bwd(dc_init, dS[:Mt], B, Lt)
        ops.mean_pool_bwd(dc_init, dS[Mt:], B, Lpv)
        self.bgrad(dc_init, "map_bi_to_individual.bias")
        dpa = ops.cast(dc_init, self.adt)
        self.wgrad(dpa, sv["pool_vr"], "map_bi_to_individual.weight")
        dpr = self.dgrad(dpa, "map_bi_to_individual.weight", out_dtype=f32)
        ops.mean_pool_bwd(dpr, dseqV, B, Lv1)
        dX = self.zeros(2 * Mt + 2 * Mp, H)
        dY = self.zeros(2 * Mv, Hv)
        d_et, d_epv, d_ev = self.zeros(Mt, H), self.zeros(Mp, H), self.zeros(Mv, Hv)
        sv_fv, sv_ft, sv_fp = sv["fuse"]
        self.fuse_bwd(dseqV, sv_fv, d_ev, dY[V1], dY[V2])
        self.fuse_bwd(dS[:Mt], sv_ft, d_et, dX[T1], dX[T3])
        self.fuse_bwd(dS[Mt:], sv_fp, d_epv, dX[P2], dX[P3])
        on_ready = self.on_ready or (lambda prefixes: None)

        def ready(prefixes):
            self._join_aux()
            on_ready(prefixes)

        ready(None)
        box = {"dX": dX, "dY": dY}
        for st in reversed(sv["stages"]):
            if st[0] == "tv":
                tsv, vsv = st[1], st[2]
                for k in reversed(range(max(len(tsv), len(vsv)))):
                    items = []
                    if k < len(tsv):
                        items.append((box["dX"], tsv[k]))
                    if k < len(vsv):
                        items.append((box["dY"], vsv[k]))
                    res = self.layers_bwd_m(items)
                    j = 0
                    if k < len(tsv):
                        box["dX"] = res[j]
                        j += 1
                    if k < len(vsv):
                        box["dY"] = res[j]
                    if self.ready_per_layer and not (tsv and vsv):
                        # (single-stack stage only: the arena lists a stage's text layers before its image layers,
                        # so in a lock-step stage "text layer k done" says nothing about image layers < k.)  The
                        # layer's weight gradients were launched a moment ago: an optimizer bucket that they
                        # complete reads them while they still sit in the L2
                        ready([sv_[k][0] for sv_ in (tsv, vsv) if k < len(sv_)])
                ready([s_[0] for s_ in tsv] + [s_[0] for s_ in vsv])
            else:
                dX, dY = box["dX"], box["dY"]
                dXn, dYn = torch.empty_like(dX), torch.empty_like(dY)
                self.connections_bwd_m([(dY[V1], dX[T1], st[1], dYn[V1], dXn[T1]),
                                        (dY[V2], dX[P2], st[2], dYn[V2], dXn[P2]),
                                        (dX[P3], dX[T3], st[3], dXn[P3], dXn[T3])])
                box["dX"], box["dY"] = dXn, dYn
                ready([st[1][0], st[2][0], st[3][0]])
        dX, dY = box["dX"], box["dY"]
        for sl in (T1, T3):
            ops.axpy(dX[sl], d_et, 1.0)
        for sl in (P2, P3):
            ops.axpy(dX[sl], d_epv, 1.0)
        for sl in (V1, V2):
            ops.axpy(dY[sl], d_ev, 1.0)
        sv_et, sv_epv, sv_ev = sv["emb"]
        self.text_embed_bwd(d_et, sv_et)
        self.text_embed_bwd(d_epv, sv_epv)
        self.image_embed_bwd(d_ev, sv_ev)
        self._join_aux()
        self.saved = None

    # ---- item alignment (K3MForItemAlignment, `vilbert_k3m.py:2862-3454`) ----------------------
    def forward_alignment(self, batch, plan, labels, noise=None):
        """Both items of every pair run through ONE tower pass: `batch` holds item 1 of the B pairs
        followed by item 2 (rows [0, B) and [B, 2B) of every tensor; the reference calls
        `item_embedding` twice with the same weights, `:3408-3433`), `plan` is the aggregator plan
        of the stacked 2B items, `labels` int64 [B].  "ce" head (`:3443-3450`, ClassificationHead
        `:2164-2183`): logits = out_proj(dropout(tanh(dense(dropout(cat(c_final_1, c_final_2)))))).
        Returns dict(loss [1], probs [B, 2], c_initial [2B, H], c_final [2B, H])."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        tw = self.tower_fwd(batch, noise)
        B2, Lt, Lpv, Lv1, Mt, Mp, Mv = self.saved["dims"]
        B, H = B2 // 2, c.hidden_size
        c_final, _, sv_s = self.struct_fwd(tw["c_init"], tw["S"][Mt:], plan, lpm=False)
        feat = self.empty(B, 2 * H)
        ops.dropout(c_final[:B], None, self.rng, out=feat[:, :H])
        ops.dropout(c_final[B:], None, self.rng, out=feat[:, H:])
        s_in = self.site(c.hidden_dropout_prob)
        feat_d = ops.dropout(feat, s_in, self.rng)
        h = ops.gemm(feat_d, self.W("classifier.dense.weight"), "nt", bias=self.W("classifier.dense.bias"),
                     act=ACT_TANH)
        s_h = self.site(c.hidden_dropout_prob)
        h_d = ops.dropout(h, s_h, self.rng)
        logits = ops.gemm(h_d, self.W("classifier.out_proj.weight"), "nt",
                          bias=self.W("classifier.out_proj.bias"), out_dtype=f32)
        dlog = self.zeros(B, 8)                        # bf16, columns 0..1 used (K = 2 operand, 16 B rows)
        loss, probs = ops.pair_cls_loss(logits, labels, dlogits=dlog)
        self.saved.update(struct=sv_s, cls=(feat_d, h, h_d, dlog, s_in, s_h))
        return dict(loss=loss, probs=probs, c_initial=tw["c_init"], c_final=c_final)

    def backward_alignment(self, grad_scale=None):
        """Gradient of the pair loss into the fp32 gradient arena; `grad_scale`: optional fp32 device
        scalar (autograd's upstream gradient)."""
        c, ops, f32 = self.cfg, self.ops, torch.float32
        sv = self.saved
        B2, Lt, Lpv, Lv1, Mt, Mp, Mv = sv["dims"]
        B, H, Hv = B2 // 2, c.hidden_size, c.v_hidden_size
        feat_d, h, h_d, dlog, s_in, s_h = sv["cls"]
        if grad_scale is not None:
            ops.scale_dev_(dlog, grad_scale)
        dl = dlog[:, :2]
        self.bgrad(dl, "classifier.out_proj.bias")
        self.wgrad(dl, h_d, "classifier.out_proj.weight")
        # dropout mask of h and tanh' in the dgrad epilogue: dz = (dl W_out) * keep/(1-p) * (1 - h^2)
        dz = self.dgrad(dl, "classifier.out_proj.weight", act=ACT_DTANH, aux_in=h, drop=s_h)
        self.bgrad(dz, "classifier.dense.bias")
        self.wgrad(dz, feat_d, "classifier.dense.weight")
        dfeat = self.dgrad(dz, "classifier.dense.weight", drop=s_in)
        dc_final = sv["struct"][5]
        ops.dropout(dfeat[:, :H], None, self.rng, out=dc_final[:B])
        ops.dropout(dfeat[:, H:], None, self.rng, out=dc_final[B:])
        dS = self.zeros(Mt + Mp, H)
        dseqV = self.zeros(Mv, Hv)
        dc_init = self.zeros(B2, H, f32)
        self.struct_bwd(sv["struct"], 1.0, dc_init, dS[Mt:])
        self.tower_bwd(dS, dseqV, dc_init)

