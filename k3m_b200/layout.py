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


def param_spec(cfg):
    """OrderedDict name -> shape, in the reference's `named_parameters()` order."""
    if not (cfg.use_image and cfg.with_coattention):
        raise ValueError("k3m_b200 implements the tri-modal path (use_image and with_coattention)")
    if cfg.task_specific_tokens or cfg.dynamic_attention:
        raise ValueError("task_specific_tokens / dynamic_attention are off on the K3M path")
    H, Hv, Hb = cfg.hidden_size, cfg.v_hidden_size, cfg.bi_hidden_size
    I, Iv = cfg.intermediate_size, cfg.v_intermediate_size
    nc = len(cfg.v_biattention_id)
    s = OrderedDict()
    s["embeddings.word_embeddings.weight"] = (cfg.vocab_size, H)
    s["embeddings.position_embeddings.weight"] = (cfg.max_position_embeddings, H)
    s["embeddings.token_type_embeddings.weight"] = (cfg.type_vocab_size, H)
    _ln(s, "embeddings.LayerNorm", H)
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
    _linear(s, "map_individual_to_bi", Hb, H)
    _linear(s, "map_bi_to_individual", H, Hb)
    for n in ("score_self_v", "score_cross1_v", "score_cross2_v", "soft_v"):
        _linear(s, n, Hb, Hb * 3)
    for m in ("t", "pv"):
        for n in ("score_self_", "score_cross1_", "score_cross2_", "soft_"):
            _linear(s, n + m, H, H * 3)
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


class ArenaLayout(object):
    """name -> (offset, numel, shape) in one flat buffer; every tensor starts 64-element aligned
    (128 B in bf16, 256 B in fp32) so TMA descriptors can point straight into the arena.

    Fused groups are emitted back to back *without* padding in between (their row sizes are
    multiples of 64 elements for every supported config), so that e.g.
    `fused("encoder.layer.0.attention.self.qkv.weight")` is a plain [3H, H] view.
    """

    ALIGN = 64

    def __init__(self, cfg):
        self.spec = param_spec(cfg)
        self.offsets = OrderedDict()
        self.fused = OrderedDict()      # fused name -> (offset, shape)
        self.decay_ranges = []          # (start, end) element ranges with weight decay
        off = 0
        placed = set()
        names = list(self.spec.keys())

        def place(name):
            nonlocal off
            shape = self.spec[name]
            n = 1
            for d in shape:
                n *= d
            self.offsets[name] = (off, n, shape)
            off += n
            placed.add(name)

        def place_group(fused_name, members):
            nonlocal off
            off = _round_up(off, self.ALIGN)
            start = off
            for m in members:
                place(m)
            shapes = [self.spec[m] for m in members]
            rows = sum(sh[0] for sh in shapes)
            self.fused[fused_name] = (start, (rows,) + tuple(shapes[0][1:]))

        for name in names:
            if name in placed:
                continue
            base = None
            for q, grp, tag in (
                (".attention.self.query", ("query", "key", "value"), "qkv"),
                (".biattention.query1", ("query1", "key1", "value1"), "qkv1"),
                (".biattention.query2", ("query2", "key2", "value2"), "qkv2"),
            ):
                for kind in (".weight", ".bias"):
                    if name.endswith(q + kind):
                        base = name[: -len(q + kind)]
                        prefix = base + q[: q.rfind(".") + 1]
                        members = [prefix + g + kind for g in grp]
                        place_group(prefix + tag + kind, members)
            if base is not None:
                continue
            if name.startswith("score_self_"):
                m = name[len("score_self_"):]           # e.g. "v.weight"
                mod, kind = m.split(".")
                members = ["score_%s_%s.%s" % (g, mod, kind) for g in ("self", "cross1", "cross2")]
                place_group("score_%s.%s" % (mod, kind), members)
                continue
            off = _round_up(off, self.ALIGN)
            place(name)
        self.total = _round_up(off, self.ALIGN)
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
