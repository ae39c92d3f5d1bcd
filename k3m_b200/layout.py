"""Parameter inventory and flat-arena layout of the K3M pretraining model.

`param_spec(config)` lists every parameter of the reference's
`BertForMultiModalPreTraining_tri_stru` (`vilbert_k3m/vilbert_k3m.py:2190-2264`) with the
reference's state_dict name and shape, in the reference's registration order (998 tensors,
461 727 692 elements for `bert_base_6layer_6conect.json`).  `ArenaLayout` places them in ONE
contiguous buffer so that

  * query/key/value weights (and biases) of every attention block are adjacent -> one fused
    [3H, H] projection GEMM reads them as a single matrix,
  * the three fusion gates score_self/score_cross1/score_cross2 of a modality are adjacent ->
    one [3H, 3H] GEMM,
  * multi-tensor AdamW and the gradient all-reduce run over plain contiguous ranges.
"""
from collections import OrderedDict


def _linear(spec, name, out_f, in_f, bias=True):
    spec[name + ".weight"] = (out_f, in_f)
    if bias:
        spec[name + ".bias"] = (out_f,)


def _ln(spec, name, h):
    spec[name + ".weight"] = (h,)
    spec[name + ".bias"] = (h,)


def _bert_layer(spec, p, h, inter):
    # BertLayer / BertImageLayer, `vilbert_k3m.py:535-548,696-709`
    for n in ("query", "key", "value"):
        _linear(spec, p + ".attention.self." + n, h, h)
    _linear(spec, p + ".attention.output.dense", h, h)
    _ln(spec, p + ".attention.output.LayerNorm", h)
    _linear(spec, p + ".intermediate.dense", inter, h)
    _linear(spec, p + ".output.dense", h, inter)
    _ln(spec, p + ".output.LayerNorm", h)


def _connection(spec, p, h1, i1, h2, i2, bi):
    # BertConnectionLayer(_two_text), `vilbert_k3m.py:1030-1111`; stream 1 = image (or PV)
    for n in ("query1", "key1", "value1"):
        _linear(spec, p + ".biattention." + n, bi, h1)
    for n in ("query2", "key2", "value2"):
        _linear(spec, p + ".biattention." + n, bi, h2)
    _linear(spec, p + ".biOutput.dense1", h1, bi)
    _ln(spec, p + ".biOutput.LayerNorm1", h1)
    _linear(spec, p + ".biOutput.q_dense1", h1, bi)
    _linear(spec, p + ".biOutput.dense2", h2, bi)
    _ln(spec, p + ".biOutput.LayerNorm2", h2)
    _linear(spec, p + ".biOutput.q_dense2", h2, bi)
    _linear(spec, p + ".v_intermediate.dense", i1, h1)
    _linear(spec, p + ".v_output.dense", h1, i1)
    _ln(spec, p + ".v_output.LayerNorm", h1)
    _linear(spec, p + ".t_intermediate.dense", i2, h2)
    _linear(spec, p + ".t_output.dense", h2, i2)
    _ln(spec, p + ".t_output.LayerNorm", h2)


def _encoder(s, cfg):
    H, Hv, Hb = cfg.hidden_size, cfg.v_hidden_size, cfg.bi_hidden_size
    I, Iv = cfg.intermediate_size, cfg.v_intermediate_size
    nc = len(cfg.v_biattention_id)
    for i in range(cfg.num_hidden_layers):
        _bert_layer(s, "encoder.layer.%d" % i, H, I)
    for i in range(cfg.v_num_hidden_layers):
        _bert_layer(s, "encoder.v_layer.%d" % i, Hv, Iv)
    for i in range(nc):
        _connection(s, "encoder.c_layer.%d" % i, Hv, Iv, H, I, Hb)
    for i in range(nc):
        _connection(s, "encoder.c_layer_pv_v.%d" % i, Hv, Iv, H, I, Hb)
    for i in range(nc):
        _connection(s, "encoder.c_layer_pv_t.%d" % i, H, I, H, I, H)


def _fusion_and_aggregator_head(s, cfg):
    H, Hb = cfg.hidden_size, cfg.bi_hidden_size
    _linear(s, "map_individual_to_bi", Hb, H)
    _linear(s, "map_bi_to_individual", H, Hb)
    for n in ("score_self_v", "score_cross1_v", "score_cross2_v", "soft_v"):
        _linear(s, n, Hb, Hb * 3)
    for m in ("t", "pv"):
        for n in ("score_self_", "score_cross1_", "score_cross2_", "soft_"):
            _linear(s, n + m, H, H * 3)


def param_spec(cfg, kind="pretrain"):
    """OrderedDict name -> shape, in the reference's `named_parameters()` order.
    kind "pretrain": `BertForMultiModalPreTraining_tri_stru` (`vilbert_k3m.py:2190-2264`);
    kind "alignment": `K3MForItemAlignment` (`:2866-2943`; registration order checked against the
    unmodified reference by `tools/make_golden.py`)."""
    if not (cfg.use_image and cfg.with_coattention):
        raise ValueError("k3m_b200 implements the tri-modal path (use_image and with_coattention)")
    if cfg.task_specific_tokens or cfg.dynamic_attention:
        raise ValueError("task_specific_tokens / dynamic_attention are off on the K3M path")
    # driver flags that change the computation (`pretrain.py:1350-1366` -> `vilbert_k3m.py:1185-1195` frozen
    # leading layers, `:1228-1260` in-batch pairs, `:1262-1273` fast mode): not built — refuse instead of
    # silently computing the default
    for flag, default in (("fixed_t_layer", 0), ("fixed_v_layer", 0), ("in_batch_pairs", False), ("fast_mode", False)):
        if getattr(cfg, flag, default) not in (default, None):
            raise ValueError("%s=%r is not supported by k3m_b200 (only the default %r)" % (flag, getattr(cfg, flag), default))
    if getattr(cfg, "visual_target", 0) != 0:
        raise ValueError("visual_target=%r is not supported (only 0: KL against detector class distributions)" % cfg.visual_target)
    H, Hv, Hb = cfg.hidden_size, cfg.v_hidden_size, cfg.bi_hidden_size
    s = OrderedDict()
    s["embeddings.word_embeddings.weight"] = (cfg.vocab_size, H)
    s["embeddings.position_embeddings.weight"] = (cfg.max_position_embeddings, H)
    s["embeddings.token_type_embeddings.weight"] = (cfg.type_vocab_size, H)
    _ln(s, "embeddings.LayerNorm", H)
    if kind == "alignment":
        _linear(s, "v_embeddings.image_embeddings", Hv, cfg.v_feature_size)
        _linear(s, "v_embeddings.image_location_embeddings", Hv, 5)
        _ln(s, "v_embeddings.LayerNorm", Hv)
        _linear(s, "v_pooler.dense", Hb, Hv)
        _encoder(s, cfg)
        _linear(s, "t_pooler.dense", Hb, H)
        _fusion_and_aggregator_head(s, cfg)
        if getattr(cfg, "loss_type", "ce") == "ce":
            _linear(s, "classifier.dense", H, H * 2)          # ClassificationHead `:2164-2183`
            _linear(s, "classifier.out_proj", 2, H)
    elif kind == "pretrain":
        _encoder(s, cfg)
        _linear(s, "t_pooler.dense", Hb, H)
        _linear(s, "v_embeddings.image_embeddings", Hv, cfg.v_feature_size)
        _linear(s, "v_embeddings.image_location_embeddings", Hv, 5)
        _ln(s, "v_embeddings.LayerNorm", Hv)
        _linear(s, "v_pooler.dense", Hb, Hv)
        s["cls.predictions.bias"] = (cfg.vocab_size,)
        _linear(s, "cls.predictions.transform.dense", H, H)
        _ln(s, "cls.predictions.transform.LayerNorm", H)
        _linear(s, "cls.seq_relationship", 2, H)
        _linear(s, "cls.imagePredictions.transform.dense", Hv, Hv)
        _ln(s, "cls.imagePredictions.transform.LayerNorm", Hv)
        _linear(s, "cls.imagePredictions.decoder", cfg.v_target_size, Hv)
        _fusion_and_aggregator_head(s, cfg)
    else:
        raise ValueError("kind must be 'pretrain' or 'alignment'")
    _linear(s, "struc_w1", H, H * 3)
    _linear(s, "struc_w2", 1, H)
    _linear(s, "struc_w3", H, H)
    return s


TIED_ALIAS = ("cls.predictions.decoder.weight", "embeddings.word_embeddings.weight")

NO_DECAY_SUBSTRINGS = ("bias", "LayerNorm.bias", "LayerNorm.weight")  # `pretrain.py:835`


def is_no_decay(name):
    """Substring rule of `pretrain.py:835-851` (note: `biOutput.LayerNorm1.weight` decays)."""
    return any(nd in name for nd in NO_DECAY_SUBSTRINGS)


def _round_up(x, m):
    return (x + m - 1) // m * m


def no_grad_names(cfg, names):
    """Parameters that never receive a gradient in the K3M pretraining step:
    `q_dense1/2` are never called (`vilbert_k3m.py:976,983`), the pooler outputs are discarded by
    `forward` (`:2620-2631`), `map_individual_to_bi` is never called on this path, `soft_*` are only used by
    `if_pre_sampling == 2`, the `score_*` gates only by `if_pre_sampling` 1 and 2, and the ITM loss is
    never added to the training loss (`pretrain.py:1055`) so `cls.seq_relationship` is value-only.
    torch.optim.AdamW skips such parameters entirely (no decay, no state)."""
    out = set()
    for n in names:
        if (".q_dense" in n or n.startswith("t_pooler.") or n.startswith("v_pooler.")
                or n.startswith("map_individual_to_bi.")
                or n.startswith("cls.seq_relationship.")):
            out.add(n)
        elif n.startswith("soft_") and getattr(cfg, "if_pre_sampling", 1) != 2:
            out.add(n)
        elif n.startswith("score_") and getattr(cfg, "if_pre_sampling", 1) not in (1, 2):
            out.add(n)
    return out


def exec_order(cfg, names):
    """Parameter names sorted by the point of the forward schedule that first uses them
    (`BertEncoder.forward` `vilbert_k3m.py:1686-1750`: text layers up to the next connection,
    image layers up to it, the three connection layers, ...; then fusion, aggregator, heads).
    Backward finishes weight gradients in the reverse of this order, so contiguous arena ranges
    taken from the END become ready first: gradient buckets are plain slices."""
    stage = {}

    def put(prefix, k):
        stage[prefix] = k

    k = 0
    put("embeddings.", k)
    put("v_embeddings.", k)
    k += 1
    s1 = s2 = 0
    for count, (e1, e2) in enumerate(zip(cfg.v_biattention_id, cfg.t_biattention_id)):
        for i in range(s2, e2):
            put("encoder.layer.%d." % i, k)
            k += 1
        for i in range(s1, e1):
            put("encoder.v_layer.%d." % i, k)
            k += 1
        for c in ("c_layer", "c_layer_pv_v", "c_layer_pv_t"):
            put("encoder.%s.%d." % (c, count), k)
            k += 1
        s1, s2 = e1, e2
    for i in range(s1, cfg.v_num_hidden_layers):
        put("encoder.v_layer.%d." % i, k)
        k += 1
    for i in range(s2, cfg.num_hidden_layers):
        put("encoder.layer.%d." % i, k)
        k += 1
    prefixes = sorted(stage.keys(), key=len, reverse=True)

    def key(item):
        idx, n = item
        for p in prefixes:
            if n.startswith(p):
                return (stage[p], idx)
        return (k + 1, idx)           # fusion gates, aggregator, heads, poolers: after the encoder

    return [n for _, n in sorted(enumerate(names), key=key)]


class ArenaLayout(object):
    """name -> (offset, numel, shape) in one flat buffer; every tensor starts 64-element aligned
    (128 B in bf16, 256 B in fp32) so TMA descriptors can point straight into the arena.

    The arena has three contiguous regions, in this order:
      [0, decay_end)            trained, weight decay 0.01   (`pretrain.py:835-851` substring rule)
      [decay_end, trained_end)  trained, no weight decay     (biases, `LayerNorm.{weight,bias}`)
      [trained_end, total)      parameters that never get a gradient (`no_grad_names`)
    so AdamW is two launches over plain ranges and the gradient all-reduce covers exactly
    [0, trained_end) in contiguous buckets.

    Fused groups are emitted back to back *without* padding in between (their row sizes are
    multiples of 64 elements for every supported config), so that e.g.
    `fused("encoder.layer.0.attention.self.qkv.weight")` is a plain [3H, H] view.
    """

    ALIGN = 64

    def __init__(self, cfg, kind="pretrain"):
        self.kind = kind
        self.spec = param_spec(cfg, kind)
        self.offsets = OrderedDict()
        self.fused = OrderedDict()      # fused name -> (offset, shape)
        names = list(self.spec.keys())
        frozen = no_grad_names(cfg, names)
        self.frozen = frozen
        state = {"off": 0}
        placed = set()

        def place(name):
            shape = self.spec[name]
            n = 1
            for d in shape:
                n *= d
            self.offsets[name] = (state["off"], n, shape)
            state["off"] += n
            placed.add(name)

        def place_group(fused_name, members):
            state["off"] = _round_up(state["off"], self.ALIGN)
            start = state["off"]
            for m in members:
                place(m)
            shapes = [self.spec[m] for m in members]
            rows = sum(sh[0] for sh in shapes)
            self.fused[fused_name] = (start, (rows,) + tuple(shapes[0][1:]))

        def region_of(name):
            if name in frozen:
                return 2
            return 1 if is_no_decay(name) else 0

        def group_of(name):
            """-> (fused name, members) if `name` opens a fused group, else None."""
            for q, grp, tag in (
                (".attention.self.query", ("query", "key", "value"), "qkv"),
                (".biattention.query1", ("query1", "key1", "value1"), "qkv1"),
                (".biattention.query2", ("query2", "key2", "value2"), "qkv2"),
            ):
                for kind in (".weight", ".bias"):
                    if name.endswith(q + kind):
                        prefix = name[: -len(q + kind)] + q[: q.rfind(".") + 1]
                        return prefix + tag + kind, [prefix + g + kind for g in grp]
            if name.startswith("score_self_"):
                mod, kind = name[len("score_self_"):].split(".")
                return ("score_%s.%s" % (mod, kind),
                        ["score_%s_%s.%s" % (g, mod, kind) for g in ("self", "cross1", "cross2")])
            return None

        bounds = []
        ordered = exec_order(cfg, names)
        for region in (0, 1, 2):
            for name in ordered:
                if name in placed or region_of(name) != region:
                    continue
                grp = group_of(name)
                if grp is not None:
                    assert all(region_of(m) == region for m in grp[1])
                    place_group(grp[0], grp[1])
                    continue
                state["off"] = _round_up(state["off"], self.ALIGN)
                place(name)
            state["off"] = _round_up(state["off"], 1024)
            bounds.append(state["off"])
        self.decay_end, self.trained_end, self.total = bounds
        assert placed == set(names)

    def view(self, flat, name):
        if name in self.fused:
            off, shape = self.fused[name]
        else:
            off, _, shape = self.offsets[name]
        n = 1
        for d in shape:
            n *= d
        return flat[off:off + n].view(*shape)

    def segments(self, group_fn=None):
        """Merged (start, end, key) runs over the trained region; key = group_fn(name) (default:
        the weight-decay flag).  Alignment gaps are absorbed into the preceding run (they hold
        zeros with zero gradients, so any update leaves them at zero... except weight decay of 0)."""
        group_fn = group_fn or (lambda n: not is_no_decay(n))
        items = sorted(((off, num, n) for n, (off, num, _) in self.offsets.items() if n not in self.frozen))
        runs = []
        for off, num, n in items:
            k = group_fn(n)
            if runs and runs[-1][2] == k:
                runs[-1][1] = off + num
            else:
                if runs:
                    runs[-1][1] = off
                runs.append([off, off + num, k])
        if runs:
            runs[-1][1] = self.trained_end
        return [tuple(r) for r in runs]
