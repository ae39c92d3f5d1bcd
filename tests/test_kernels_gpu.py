"""GPU: every entry point of libk3m_sm100a.so (through k3m_b200.ops.CudaOps -> the C ABI) against
the PyTorch restatement of its contract (oracle/ops_ref.py) on the same seeded inputs.

Tolerances: integer / index work (gathers, masks, counts, dropout keep-masks) is bit-exact;
bf16 outputs are compared to the fp32 result rounded to bf16 with a few ulp of slack; fp32
outputs (statistics, losses, parameter gradients) to 1e-4 relative unless stated."""
import math

import pytest
import torch

from oracle.ops_ref import RefOps, dropout_keep

pytestmark = pytest.mark.gpu
BF, F32 = torch.bfloat16, torch.float32
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    from k3m_b200.ops import CudaOps
    return CudaOps(DEV)


@pytest.fixture(scope="module")
def ref():
    return RefOps(F32)


def rnd(*shape, seed=0, scale=1.0, dtype=BF):
    g = torch.Generator().manual_seed(seed + 7919 * len(shape) + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(DEV)


def close(got, want, rtol, atol, what=""):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = err > tol
    assert not bool(bad.any()), "%s: %d/%d bad, max err %.4e (want max %.4e)" % (
        what, int(bad.sum()), bad.numel(), float(err.max()), float(want.abs().max()))


def close_bf16(got, want, what="", ulps=2.0, atol=1e-6):
    close(got, want, ulps * 2.0 ** -8, atol, what)


RNG = None


def rng_state():
    return torch.tensor([0x1234ABCD5678, 3], dtype=torch.int64, device=DEV)


# ---- GEMM ----------------------------------------------------------------------------------------
GEMM_SHAPES = [(128, 128, 64), (256, 768, 768), (2304, 768, 3072), (300, 200, 136), (37, 1601, 1024),
               (2368, 1024, 2048), (64, 2, 768), (1, 768, 768), (130, 64, 8), (4608, 2304, 768)]


@pytest.mark.parametrize("layout", ["nt", "nn", "tn"])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_plain(ops, ref, layout, M, N, K):
    if layout != "nt" and (N % 8 or M % 8):
        Np, Mp = (N + 7) // 8 * 8, (M + 7) // 8 * 8
    else:
        Np, Mp = N, M
    s = 1.0 / math.sqrt(K)
    if layout == "nt":
        a, b = rnd(M, K), rnd(N, K, seed=1, scale=s)
    elif layout == "nn":
        a, b = rnd(M, K), rnd(K, Np, seed=1, scale=s)[:, :N]
    else:
        a, b = rnd(K, Mp)[:, :M], rnd(K, Np, seed=1, scale=s)[:, :N]
    got = ops.gemm(a, b, layout, out_dtype=F32)
    want = ref.gemm(a, b, layout, out_dtype=F32)
    close(got, want, 1e-4, 2e-4 * math.sqrt(K) * s * 4, "gemm %s" % layout)
    got16 = ops.gemm(a, b, layout)
    close_bf16(got16, want, "gemm bf16 %s" % layout, atol=1e-3)


@pytest.mark.parametrize("act", [0, 1, 3, 4, 6])
def test_gemm_epilogue_bias_act_addend(ops, ref, act):
    M, N, K = 300, 776, 264
    a, w = rnd(M, K), rnd(N, K, seed=1, scale=0.06)
    bias, add = rnd(N, seed=2), rnd(M, N, seed=3)
    aux_g = torch.zeros(M, N, dtype=BF, device=DEV)
    aux_r = torch.zeros(M, N, dtype=F32, device=DEV)
    got = ops.gemm(a, w, "nt", bias=bias, act=act, aux_out=aux_g if act in (1, 6) else None, addend=add)
    want = ref.gemm(a, w, "nt", bias=bias, act=act, aux_out=aux_r if act in (1, 6) else None, addend=add)
    close_bf16(got, want, "epilogue act=%d" % act, atol=2e-3)
    if act in (1, 6):          # 1: the pre-activation, 6 (ACT_GELU_GRAD): gelu'(pre-activation)
        close_bf16(aux_g, aux_r, "aux_out", atol=2e-3)


@pytest.mark.parametrize("act", [2, 5, 7])
def test_gemm_epilogue_derivative(ops, ref, act):
    M, N, K = 260, 384, 200
    a, w = rnd(M, K), rnd(K, N, seed=1, scale=0.07)
    aux = rnd(M, N, seed=4, scale=0.8 if act == 5 else 1.5)
    add = rnd(M, N, seed=5, dtype=F32)
    got = ops.gemm(a, w, "nn", act=act, aux_in=aux, addend=add, out_dtype=F32)
    want = ref.gemm(a, w, "nn", act=act, aux_in=aux, addend=add, out_dtype=F32)
    close(got, want, 1e-4, 1e-4, "derivative epilogue")


def test_gemm_dropout_mask_bit_exact(ops, ref):
    M, N, K = 200, 328, 128
    a, w = rnd(M, K), rnd(N, K, seed=1, scale=0.1)
    r = rng_state()
    got = ops.gemm(a, w, "nt", drop=(0.1, 17), rng=r, out_dtype=F32)
    plain = ops.gemm(a, w, "nt", out_dtype=F32)
    keep = dropout_keep(r, 17, M, N, 0.1)
    nz = plain != 0
    assert torch.equal((got != 0) & nz, keep & nz)          # the keep mask itself: bit-exact
    close(got, torch.where(keep, plain / 0.9, torch.zeros_like(plain)), 1e-5, 1e-6, "dropout scale")
    frac = float(keep.float().mean())
    assert 0.88 < frac < 0.92


def test_gemm_accumulate_split_k(ops, ref):
    # weight-gradient shape: small M,N, long K -> split-K with fp32 atomics onto a non-zero output
    K, M, N = 4608, 768, 776
    a, b = rnd(K, M, scale=0.05), rnd(K, N, seed=1, scale=0.05)[:, :770]
    base = rnd(M, 770, seed=2, dtype=F32)
    got, want = base.clone(), base.clone()
    ops.gemm(a, b, "tn", out=got, accumulate=True)
    ref.gemm(a, b, "tn", out=want, accumulate=True)
    close(got, want, 1e-4, 1e-4, "split-K accumulate")


# ---- the benchmark's shapes (B = 64: M = 9216 text rows, 4736 image rows, 2304 / 2368 connection rows) ----------
# The CTA-pair kernel (cta_group::2, 256 x 256 tile per pair) is only selected once a problem has >= 45 pair
# tiles, so every fused epilogue is also checked at the sizes the timed step runs, and the test asserts which
# variant the library dispatched.
BENCH_EPILOGUES = [
    # name, layout, M, N, K, kwargs builder, expected variant substrings
    ("attn_out_text", "nt", 9216, 768, 768, "bias_drop_add", ("ctas=2", "epi=F_ADD", "tin=2")),
    ("ffn_down_text", "nt", 9216, 768, 3072, "bias_drop_add", ("ctas=2", "epi=F_ADD", "tin=2")),
    ("ffn_up_text", "nt", 9216, 3072, 768, "bias_gelu", ("ctas=2", "epi=F_GELU")),
    ("ffn_up_text_grad", "nt", 9216, 3072, 768, "bias_gelu_grad", ("ctas=2", "epi=F_GELU")),
    ("ffn_up_image_grad", "nt", 4736, 1024, 1024, "bias_gelu_grad", ("ctas=2", "epi=F_GELU")),
    ("qkv_text", "nt", 9216, 2304, 768, "bias", ("ctas=2", "epi=F_PLAIN")),
    ("ffn_down_dgrad_text", "nn", 9216, 3072, 768, "dgelu", ("ctas=2", "epi=F_DGELU", "tin=1")),
    ("ffn_down_dgrad_text_mulaux", "nn", 9216, 3072, 768, "mulaux", ("ctas=2", "epi=F_DGELU", "tin=1")),
    ("ffn_up_dgrad_text", "nn", 9216, 768, 3072, "add", ("ctas=2", "epi=F_ADD", "tin=2")),
    ("out_image", "nt", 4736, 1024, 1024, "bias_drop_add", ("ctas=2", "epi=F_ADD")),
    ("ffn_up_conn", "nt", 2304, 3072, 768, "bias_gelu", ("ctas=2", "epi=F_GELU")),
    ("ffn_up_conn_grad", "nt", 2304, 3072, 768, "bias_gelu_grad", ("ctas=2", "epi=F_GELU")),
    ("ffn_up_ragged_grad", "nt", 2300, 3048, 768, "bias_gelu_grad", ("ctas=2", "epi=F_GELU")),
    ("ffn_down_conn", "nt", 2304, 768, 3072, "bias_drop_add", ("epi=F_ADD",)),
    ("out_conn_image", "nt", 2368, 1024, 1024, "bias_drop_add", ("epi=F_ADD",)),
    ("dgrad_conn", "nn", 2304, 3072, 768, "dgelu", ("ctas=2", "epi=F_DGELU")),
    ("vocab", "nt", 4608, 21128, 768, "bias", ("ctas=2", "epi=F_PLAIN")),
    ("vocab_dgrad", "nn", 4608, 768, 21128, "plain", ("ctas=2", "epi=F_PLAIN")),
]


@pytest.mark.parametrize("name,layout,M,N,K,kind,variant", BENCH_EPILOGUES, ids=[c[0] for c in BENCH_EPILOGUES])
def test_gemm_bench_shapes_fused_epilogues(ops, ref, name, layout, M, N, K, kind, variant):
    s = 1.0 / math.sqrt(K)
    a = rnd(M, K)
    b = rnd(N, K, seed=1, scale=s) if layout == "nt" else rnd(K, N, seed=1, scale=s)
    r = rng_state()
    kw = {}
    if kind.startswith("bias"):
        kw["bias"] = rnd(N, seed=2)
    if kind == "bias_drop_add":
        kw.update(drop=(0.1, 23), rng=r, addend=rnd(M, N, seed=3))
    if kind == "add":
        kw.update(addend=rnd(M, N, seed=3))
    if kind == "dgelu":
        kw.update(act=2, aux_in=rnd(M, N, seed=4, scale=1.5))
    if kind == "mulaux":
        kw.update(act=7, aux_in=rnd(M, N, seed=4, scale=0.6))
    aux_g = aux_r = None
    gelu_act = {"bias_gelu": 1, "bias_gelu_grad": 6}.get(kind)
    if gelu_act:
        aux_g = torch.zeros(M, N, dtype=BF, device=DEV)
        aux_r = torch.zeros(M, N, dtype=F32, device=DEV)
    got = ops.gemm(a, b, layout, act=gelu_act, aux_out=aux_g, **kw) if gelu_act else ops.gemm(a, b, layout, **kw)
    v = ops.gemm_last_variant()
    print("%s %s %dx%dx%d: %s" % (name, layout, M, N, K, v))
    for sub in variant:
        assert sub in v, (name, v, sub)
    want = (ref.gemm(a, b, layout, act=gelu_act, aux_out=aux_r, out_dtype=F32, **kw) if gelu_act
            else ref.gemm(a, b, layout, out_dtype=F32, **kw))
    close_bf16(got, want, name, atol=4e-3)
    if aux_g is not None:
        close_bf16(aux_g, aux_r, name + " aux_out", atol=4e-3)
    # (with dropout the comparison above also pins the keep mask: a flipped element of |value| ~ 1 is far outside
    # the tolerance; the mask alone is checked bit for bit in test_gemm_dropout_mask_bit_exact)


BENCH_WGRADS = [("wffn2_text", 768, 3072, 9216), ("wffn1_text", 3072, 768, 9216), ("wqkv_text", 2304, 768, 9216),
                ("wout_text", 768, 768, 9216), ("w_image", 1024, 1024, 4736), ("w_conn", 1024, 1024, 2368),
                ("wffn1_conn", 3072, 768, 2304), ("w_small_conn", 768, 768, 2304), ("w_vocab", 21128, 768, 4608)]


@pytest.mark.parametrize("name,M,N,K", BENCH_WGRADS, ids=[c[0] for c in BENCH_WGRADS])
def test_gemm_bench_shapes_weight_gradients(ops, ref, name, M, N, K):
    """tn, fp32 reduce-add onto a non-zero gradient (TMA bulk reduce, split-K where the rule picks it)."""
    a, b = rnd(K, M, scale=0.05), rnd(K, N, seed=1, scale=0.05)
    base = rnd(M, N, seed=2, dtype=F32)
    got, want = base.clone(), base.clone()
    ops.gemm(a, b, "tn", out=got, accumulate=True)
    v = ops.gemm_last_variant()
    print("%s tn %dx%dx%d: %s" % (name, M, N, K, v))
    assert "acc=1" in v and "tma_out=1" in v
    ref.gemm(a, b, "tn", out=want, accumulate=True)
    close(got, want, 2e-4, 2e-4 * math.sqrt(K) * 0.05 * 0.05 * 8, name)


# ---- grouped launches (k3m_gemm_grouped_bf16) ---------------------------------------------------------------
def _group_case(kind, M, N, K, seed):
    """one problem of a group with epilogue `kind`; returns (kwargs for ops.gemm / gemm_group, extra outputs)"""
    s = 1.0 / math.sqrt(K)
    layout = {"bias": "nt", "bias_drop_add": "nt", "bias_gelu": "nt", "bias_gelu_grad": "nt", "dgelu": "nn", "mulaux": "nn",
              "add": "nn", "acc": "tn"}[kind]
    if layout == "nt":
        a, b = rnd(M, K, seed=seed), rnd(N, K, seed=seed + 1, scale=s)
    elif layout == "nn":
        a, b = rnd(M, K, seed=seed), rnd(K, N, seed=seed + 1, scale=s)
    else:
        a, b = rnd(K, M, seed=seed, scale=0.05), rnd(K, N, seed=seed + 1, scale=0.05)
    kw = dict(a=a, b=b, layout=layout)
    if kind.startswith("bias"):
        kw["bias"] = rnd(N, seed=seed + 2)
    if kind == "bias_drop_add":
        kw.update(drop=(0.1, 40 + seed), rng=rng_state(), addend=rnd(M, N, seed=seed + 3))
    if kind == "add":
        kw.update(addend=rnd(M, N, seed=seed + 3))
    if kind == "dgelu":
        kw.update(act=2, aux_in=rnd(M, N, seed=seed + 4, scale=1.5))
    if kind == "mulaux":
        kw.update(act=7, aux_in=rnd(M, N, seed=seed + 4, scale=0.6))
    if kind == "bias_gelu":
        kw.update(act=1)
    if kind == "bias_gelu_grad":
        kw.update(act=6)
    return kw


GROUPS = {
    # the six problems of a connection stage (three connection layers x two streams), B = 64
    "conn_qkv": [("bias", 2368, 3072, 1024), ("bias", 2304, 3072, 768), ("bias", 2368, 3072, 1024),
                 ("bias", 2304, 3072, 768), ("bias", 2304, 2304, 768), ("bias", 2304, 2304, 768)],
    "conn_out": [("bias_drop_add", 2368, 1024, 1024), ("bias_drop_add", 2304, 768, 1024), ("bias_drop_add", 2368, 1024, 1024),
                 ("bias_drop_add", 2304, 768, 1024), ("bias_drop_add", 2304, 768, 768), ("bias_drop_add", 2304, 768, 768)],
    "conn_ffn_up": [("bias_gelu", 2368, 1024, 1024), ("bias_gelu", 2304, 3072, 768), ("bias_gelu", 2368, 1024, 1024),
                    ("bias_gelu", 2304, 3072, 768), ("bias_gelu", 2304, 3072, 768), ("bias_gelu", 2304, 3072, 768)],
    "conn_dgrad_mixed": [("dgelu", 2368, 1024, 1024), ("dgelu", 2304, 3072, 768), ("add", 2304, 768, 3072),
                         ("add", 2368, 1024, 3072)],
    "conn_ffn_up_grad": [("bias_gelu_grad", 2368, 1024, 1024), ("bias_gelu_grad", 2304, 3072, 768), ("bias_gelu", 2368, 1024, 1024),
                         ("bias_gelu_grad", 2304, 3072, 768), ("bias_gelu_grad", 2304, 3072, 768), ("bias_gelu", 2304, 3072, 768)],
    "conn_dffn_mulaux": [("mulaux", 2368, 1024, 1024), ("mulaux", 2304, 3072, 768), ("dgelu", 2368, 1024, 1024),
                         ("mulaux", 2304, 3072, 768), ("add", 2304, 768, 3072), ("mulaux", 2304, 3072, 768)],
    "conn_wgrad": [("acc", 1024, 1024, 2368), ("acc", 3072, 768, 2304), ("acc", 768, 3072, 2304), ("acc", 768, 768, 2304),
                   ("acc", 1024, 1024, 2368), ("acc", 768, 1024, 2304)],
    # a text and an image layer in lock-step
    "text_image": [("bias_drop_add", 9216, 768, 3072), ("bias_drop_add", 4736, 1024, 1024)],
    # small problems (tiny configs): 128 x 128 single-CTA tiles, ragged edges
    "small": [("bias_gelu", 100, 256, 128), ("bias_gelu", 48, 128, 256), ("bias_gelu", 300, 776, 264)],
    "small_grad": [("bias_gelu_grad", 100, 256, 128), ("bias_gelu_grad", 48, 128, 256), ("bias_gelu_grad", 300, 776, 264)],
    "small_wgrad": [("acc", 256, 128, 100), ("acc", 128, 256, 48), ("acc", 384, 384, 300)],
    # nine problems: more than one launch's worth (8 per launch)
    "nine": [("bias", 256, 256, 128)] * 9,
}


@pytest.mark.parametrize("name", list(GROUPS))
def test_gemm_grouped_matches_single_problem_restatement(ops, ref, name):
    probs = [_group_case(kind, M, N, K, 10 * i) for i, (kind, M, N, K) in enumerate(GROUPS[name])]
    calls, wants, auxs = [], [], []
    for kw in probs:
        c = dict(kw)
        M = kw["a"].shape[1] if kw["layout"] == "tn" else kw["a"].shape[0]
        N = kw["b"].shape[0] if kw["layout"] == "nt" else kw["b"].shape[1]
        r = dict(kw)
        aux_g = aux_r = None
        if kw.get("act") in (1, 6):
            aux_g, aux_r = torch.zeros(M, N, dtype=BF, device=DEV), torch.zeros(M, N, dtype=F32, device=DEV)
            c["aux_out"], r["aux_out"] = aux_g, aux_r
        if kw["layout"] == "tn":
            base = rnd(M, N, seed=99, dtype=F32)
            c.update(out=base.clone(), accumulate=True)
            r.update(out=base.clone(), accumulate=True)
            wants.append(ref.gemm(**r))
        else:
            wants.append(ref.gemm(out_dtype=F32, **r))
        calls.append(c)
        auxs.append((aux_g, aux_r))
    outs = ops.gemm_group(calls)
    v = ops.gemm_last_variant()
    print("group %s: %s" % (name, v))
    big = all(M >= 256 and N >= 256 for _, M, N, _ in GROUPS[name]) and name != "nine"
    assert ("ctas=2" in v) == big, v
    assert "group=%d" % (len(probs) if len(probs) <= 8 else len(probs) - 8) in v or len(probs) > 8, v
    for i, (got, want, (aux_g, aux_r)) in enumerate(zip(outs, wants, auxs)):
        kind, M, N, K = GROUPS[name][i]
        if kind == "acc":
            close(got, want, 2e-4, 2e-4 * math.sqrt(K) * 0.05 * 0.05 * 8, "%s[%d]" % (name, i))
        else:
            close_bf16(got, want, "%s[%d] %s" % (name, i, kind), atol=4e-3)
        if aux_g is not None:
            close_bf16(aux_g, aux_r, "%s[%d] aux_out" % (name, i), atol=4e-3)


@pytest.mark.parametrize("shapes", [[(36, 384, 384), (40, 384, 384), (30, 768, 768)], [(300, 776, 264), (2304, 2304, 768)]],
                         ids=["tiny", "mixed"])
def test_gemm_grouped_fp32_outputs_and_in_place_addends(ops, ref, shapes):
    """`gemm_group` with the epilogues the step itself only issues as single launches: fp32 STORE outputs with a bias
    (the GENERIC epilogue; the op table issues them singly by default, `group_f32_store=True` sends them through
    k3m_gemm_grouped_bf16) and a residual addend that IS the output buffer (the fusion gates' input gradient).
    The grouped generic epilogue with a bias was wrong once in the 128 x 128 variant (bias elements 3, 5, 7 of every 8
    read registers the load never wrote — `ld8_bits` in gemm_sm100.cu, tools/debug_group_f32.py); the grouped arm
    below and the sigmoid / tanh groups are its regression test."""
    calls, wants = [], []
    for i, (M, N, K) in enumerate(shapes):
        a, b, bias = rnd(M, K, seed=3 * i), rnd(N, K, seed=3 * i + 1, scale=1.0 / math.sqrt(K)), rnd(N, seed=3 * i + 2)
        calls.append(dict(a=a, b=b, layout="nt", bias=bias, out_dtype=F32))
        wants.append(ref.gemm(a, b, "nt", bias=bias, out_dtype=F32))
    for grouped in (False, True):
        outs = ops.gemm_group(calls, group_f32_store=grouped)
        v = ops.gemm_last_variant()
        print("group fp32 out (grouped launch: %s): %s" % (grouped, v))
        assert ("group=%d" % len(shapes) in v) == grouped and "epi=generic" in v, v
        for i, (got, want) in enumerate(zip(outs, wants)):
            assert got.dtype == F32
            close(got, want, 2e-3, 2e-3, "fp32 out [%d] grouped=%s" % (i, grouped))
    # bf16 outputs through the generic epilogue of a grouped launch: bias + sigmoid / tanh (no specialised epilogue)
    for act in (3, 4):
        calls, wants = [], []
        for i, (M, N, K) in enumerate(shapes):
            a, b, bias = rnd(M, K, seed=50 + 3 * i), rnd(N, K, seed=51 + 3 * i, scale=1.0 / math.sqrt(K)), rnd(N, seed=52 + 3 * i)
            calls.append(dict(a=a, b=b, layout="nt", bias=bias, act=act))
            wants.append(ref.gemm(a, b, "nt", bias=bias, act=act, out_dtype=F32))
        outs = ops.gemm_group(calls)
        v = ops.gemm_last_variant()
        assert "group=%d" % len(shapes) in v and "epi=generic" in v, v
        for i, (got, want) in enumerate(zip(outs, wants)):
            close_bf16(got, want, "generic act %d [%d]" % (act, i), atol=4e-3)
    calls, wants = [], []
    for i, (M, N, K) in enumerate(shapes):
        a, b = rnd(M, K, seed=30 + 3 * i), rnd(K, N, seed=31 + 3 * i, scale=1.0 / math.sqrt(K))
        buf = rnd(M, N, seed=32 + 3 * i)
        wants.append(ref.gemm(a, b, "nn", addend=buf.clone(), out_dtype=F32))
        calls.append(dict(a=a, b=b, layout="nn", addend=buf, out=buf))
    outs = ops.gemm_group(calls)
    print("group in-place addend: %s" % ops.gemm_last_variant())
    for i, (got, want) in enumerate(zip(outs, wants)):
        close_bf16(got, want, "in-place addend [%d]" % i, atol=4e-3)


def test_gemm_strided_views(ops, ref):
    # operands / outputs that are column slices of wider buffers (qkv, logits[:, :V])
    M, K, N = 192, 256, 200
    A = rnd(M, 3 * K)
    a = A[:, K:2 * K]
    w = rnd(N, K, seed=1, scale=0.06)
    out_g = torch.full((M, 208), 7.0, dtype=BF, device=DEV)
    out_r = torch.full((M, 208), 7.0, dtype=F32, device=DEV)
    ops.gemm(a, w, "nt", out=out_g[:, :N])
    ref.gemm(a, w, "nt", out=out_r[:, :N])
    close_bf16(out_g, out_r, "strided", atol=1e-3)
    assert bool((out_g[:, N:] == 7.0).all())


def test_gemm_rejects_bad_arguments(ops):
    from k3m_b200._lib import K3MError
    a, w = rnd(64, 68), rnd(64, 68, seed=1)
    with pytest.raises(K3MError):
        ops.gemm(a[:, :60][:, 1:], w[:, 1:60], "nt")       # misaligned base pointer
    with pytest.raises(K3MError):
        ops.gemm(rnd(64, 64), rnd(64, 64), "nt", out=torch.zeros(64, 64, dtype=BF, device=DEV), accumulate=True)


# ---- LayerNorm -------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,H", [(2304, 768), (2368, 1024), (37, 64), (1, 768)])
@pytest.mark.parametrize("drop", [None, (0.1, 5)])
def test_ln_fwd(ops, ref, M, H, drop):
    x, g, b = rnd(M, H, scale=2.0), rnd(H, seed=1), rnd(H, seed=2)
    r = rng_state()
    y, mean, rstd = ops.ln_fwd(x, g, b, drop=drop, rng=r)
    yr, mr, rr = ref.ln_fwd(x, g, b, drop=drop, rng=r)
    close(mean, mr, 1e-5, 1e-5, "mean")
    close(rstd, rr, 1e-4, 1e-6, "rstd")
    close_bf16(y, yr, "ln y", atol=1e-3)
    if drop is not None:
        assert torch.equal(y == 0, (yr == 0) | (y == 0) & (yr.abs() < 1e-3))


# ragged row counts, fewer rows than CTAs, the text stack's 9216 rows (with K3M_LN_BWD_TILE=1 the H > 512 cases run the
# tile kernel — cp.async double buffering, column sums in registers — which passes the same bounds)
@pytest.mark.parametrize("M,H", [(2304, 768), (2368, 1024), (37, 64), (9216, 768), (2301, 1024), (5, 768), (300, 512)])
@pytest.mark.parametrize("mode", ["plain", "drop_in", "drop_out"])
def test_ln_bwd(ops, ref, M, H, mode):
    x, g, dy = rnd(M, H, scale=2.0), rnd(H, seed=1), rnd(M, H, seed=3)
    r = rng_state()
    _, mean, rstd = ref.ln_fwd(x, g, g)
    kw = dict(drop_in=(0.1, 9)) if mode == "drop_in" else dict(drop_out=(0.1, 9), want_masked=True) if mode == "drop_out" else {}
    res = []
    for o in (ops, ref):
        dg = torch.zeros(H, dtype=F32, device=DEV) + 1.0
        db = torch.zeros(H, dtype=F32, device=DEV) - 2.0
        dbias = torch.zeros(H, dtype=F32, device=DEV) + 0.5
        dx, dxm = o.ln_bwd(dy, x, g, mean, rstd, dg, db, rng=r, dbias=dbias, **kw)
        res.append((dx, dxm, dg, db, dbias))
    (dx, dxm, dg, db, dbias), (dxr, dxmr, dgr, dbr, dbiasr) = res
    close_bf16(dx, dxr, "dx", atol=2e-3, ulps=3)
    close_bf16(dxm, dxmr, "dxm", atol=2e-3, ulps=3)
    s = math.sqrt(M)
    close(dg, dgr, 2e-3, 2e-3 * s, "dgamma")
    close(db, dbr, 2e-3, 2e-3 * s, "dbeta")
    close(dbias, dbiasr, 2e-2, 2e-2 * s, "dbias")   # sums of bf16-rounded dx in the ref vs fp32 in the kernel


def test_ln_grouped_mixed_hidden_sizes(ops, ref):
    """k3m_ln_fwd_grouped / k3m_ln_bwd_grouped: the six LayerNorms of a connection stage (hidden 1024 and 768, 2368 /
    2304 rows) in one launch each, against the single-problem restatement."""
    shapes = [(2368, 1024), (2304, 768), (2368, 1024), (2304, 768), (2304, 768), (37, 768)]
    r = rng_state()
    items, items_r = [], []
    for i, (M, H) in enumerate(shapes):
        it = dict(x=rnd(M, H, seed=i, scale=2.0), gamma=rnd(H, seed=50 + i), beta=rnd(H, seed=60 + i))
        if i == 1:
            it.update(drop=(0.1, 7), rng=r)
        items.append(it)
    got = ops.ln_fwd_group(items)
    want = [ref.ln_fwd(**it) for it in items]
    for (y, mean, rstd), (yr, mr, rr) in zip(got, want):
        close(mean, mr, 1e-5, 1e-5, "mean")
        close(rstd, rr, 1e-4, 1e-6, "rstd")
        close_bf16(y, yr, "ln y", atol=1e-3)
    bw, res = [], []
    for i, (M, H) in enumerate(shapes):
        bw.append(dict(dy=rnd(M, H, seed=70 + i), x=items[i]["x"], gamma=items[i]["gamma"], mean=want[i][1], rstd=want[i][2],
                       rng=r, drop_in=(0.1, 9 + i) if i % 2 == 0 else None))
    for o, grouped in ((ops, True), (ref, False)):
        accs = [(torch.zeros(H, dtype=F32, device=DEV) + 1.0, torch.zeros(H, dtype=F32, device=DEV) - 2.0,
                 torch.zeros(H, dtype=F32, device=DEV) + 0.5) for _, H in shapes]
        its = [dict(it, dgamma=a[0], dbeta=a[1], dbias=a[2]) for it, a in zip(bw, accs)]
        outs = o.ln_bwd_group(its) if grouped else [o.ln_bwd(**it) for it in its]
        res.append((outs, accs))
    for i, (M, H) in enumerate(shapes):
        (dx, dxm), (dxr, dxmr) = res[0][0][i], res[1][0][i]
        close_bf16(dx, dxr, "dx[%d]" % i, atol=2e-3, ulps=3)
        close_bf16(dxm, dxmr, "dxm[%d]" % i, atol=2e-3, ulps=3)
        sc = math.sqrt(M)
        for a, b, n, t in zip(res[0][1][i], res[1][1][i], ("dgamma", "dbeta", "dbias"), (2e-3, 2e-3, 2e-2)):
            close(a, b, t, t * sc, "%s[%d]" % (n, i))


def test_colsum_cast_axpy_lincomb_scale_mask(ops, ref):
    x = rnd(2304, 776)[:, :770]
    o1, o2 = torch.ones(770, dtype=F32, device=DEV), torch.ones(770, dtype=F32, device=DEV)
    ops.colsum(x, o1)
    ref.colsum(x, o2)
    close(o1, o2, 1e-4, 1e-3, "colsum bf16")
    xf = rnd(100, 2304, dtype=F32)
    o1, o2 = torch.zeros(2304, dtype=F32, device=DEV), torch.zeros(2304, dtype=F32, device=DEV)
    ops.colsum(xf, o1)
    ref.colsum(xf, o2)
    close(o1, o2, 1e-5, 1e-4, "colsum f32")
    assert torch.equal(ops.cast(xf, BF), xf.to(BF))
    assert torch.equal(ops.cast(xf.to(BF), F32), xf.to(BF).float())
    a, b, c = rnd(64, 768), rnd(64, 768, seed=1), rnd(64, 768, seed=2)
    close(ops.lincomb3(a, b, c, 1 / 3, 1 / 3, 1 / 3, out_dtype=F32), ref.lincomb3(a, b, c, 1 / 3, 1 / 3, 1 / 3, out_dtype=F32),
          1e-6, 1e-6, "lincomb3")
    close_bf16(ops.lincomb3(a, b, None, 0.5, 0.5, 0.0), ref.lincomb3(a, b, None, 0.5, 0.5, 0.0), "lincomb2")
    y1 = rnd(64, 1536, seed=5)
    y2 = y1.clone().float()
    ops.axpy(a, y1[:, 768:], 0.25)
    ref.axpy(a, y2[:, 768:], 0.25)
    close_bf16(y1, y2, "axpy")
    z = rnd(1000, dtype=F32)
    z2 = z.clone()
    ops.scale_(z, 1 / 3)
    assert torch.allclose(z, z2 / 3, rtol=1e-6)
    pre, dyy = rnd(128, 768, scale=1.5), rnd(128, 768, seed=1)
    close_bf16(ops.act_bwd(dyy, pre, 2), ref.act_bwd(dyy, pre, 2), "dgelu", atol=1e-3)
    mask = (torch.rand(64, 36) > 0.3).long().to(DEV)
    assert torch.equal(ops.mask_additive(mask), ref.mask_additive(mask))


# ---- attention -------------------------------------------------------------------------------------
ATTN = [  # B, heads, Lq, Lk, d
    (8, 12, 36, 36, 64), (8, 8, 37, 37, 128), (8, 8, 36, 37, 128), (8, 8, 37, 36, 128), (4, 8, 36, 36, 96),
    (2, 12, 128, 128, 64), (2, 8, 101, 128, 128), (2, 8, 128, 101, 128), (3, 2, 5, 7, 16),
    # more than 128 tokens on one side: the blocked (flash-style) kernels; the item-alignment finetune
    # runs 256 PV tokens (run_finetune_item_alignment.sh:40-43) against 50 title tokens / 37 regions
    (2, 12, 256, 256, 64), (2, 8, 256, 37, 128), (2, 8, 37, 256, 128), (2, 8, 50, 256, 96), (2, 8, 256, 50, 96),
    (1, 2, 200, 131, 16), (1, 12, 512, 300, 64)]


def _attn_inputs(B, H, Lq, Lk, d):
    qbuf = rnd(B * Lq, 3 * H * d, scale=1.0)
    kvbuf = rnd(B * Lk, 3 * H * d, seed=1, scale=1.0)
    lens = torch.randint(max(1, Lk // 2), Lk + 1, (B,), generator=torch.Generator().manual_seed(B * Lk))
    mask = (torch.arange(Lk)[None, :] < lens[:, None]).long().to(DEV)
    return qbuf, kvbuf, (1.0 - mask.float()) * -10000.0


@pytest.mark.parametrize("B,H,Lq,Lk,d", ATTN)
@pytest.mark.parametrize("drop", [None, (0.1, 21)])
def test_attn_fwd(ops, ref, B, H, Lq, Lk, d, drop):
    qbuf, kvbuf, mask = _attn_inputs(B, H, Lq, Lk, d)
    r = rng_state()
    ctx, lse = ops.attn_fwd(qbuf, 0, kvbuf, H * d, 2 * H * d, mask, B, H, Lq, Lk, d, drop=drop, rng=r)
    ctxr, lser = ref.attn_fwd(qbuf, 0, kvbuf, H * d, 2 * H * d, mask, B, H, Lq, Lk, d, drop=drop, rng=r)
    close(lse[0], lser, 1e-5, 1e-4, "lse")
    # P is rounded to bf16 before the P V tensor-core product (2^-9 relative per term, like flash-attention)
    close_bf16(ctx, ctxr, "ctx", atol=8e-3)


@pytest.mark.parametrize("B,H,Lq,Lk,d", ATTN)
@pytest.mark.parametrize("drop", [None, (0.1, 21)])
def test_attn_bwd(ops, ref, B, H, Lq, Lk, d, drop):
    qbuf, kvbuf, mask = _attn_inputs(B, H, Lq, Lk, d)
    r = rng_state()
    dctx = rnd(B * Lq, H * d, seed=4)
    ctxr, lser = ref.attn_fwd(qbuf, 0, kvbuf, H * d, 2 * H * d, mask, B, H, Lq, Lk, d, drop=drop, rng=r)
    ctx = ctxr.to(BF)
    W = 3 * H * d
    dq_g, dkv_g = torch.zeros(B * Lq, W, dtype=BF, device=DEV), torch.zeros(B * Lk, W, dtype=BF, device=DEV)
    dq_r, dkv_r = torch.zeros(B * Lq, W, dtype=F32, device=DEV), torch.zeros(B * Lk, W, dtype=F32, device=DEV)
    lse2 = torch.stack([lser, torch.zeros_like(lser)])      # row 1: scratch of the blocked backward
    ops.attn_bwd(dctx, ctx, qbuf, 0, kvbuf, H * d, 2 * H * d, mask, lse2, dq_g, dkv_g, B, H, Lq, Lk, d, drop=drop, rng=r)
    ref.attn_bwd(dctx, ctx, qbuf, 0, kvbuf, H * d, 2 * H * d, mask, lser, dq_r, dkv_r, B, H, Lq, Lk, d, drop=drop, rng=r)
    sc = float(dq_r.abs().max())
    close_bf16(dq_g, dq_r, "dq", atol=4e-3 * sc, ulps=3)
    close_bf16(dkv_g, dkv_r, "dkv", atol=4e-3 * float(dkv_r.abs().max()), ulps=3)


def test_attn_key_mask_with_trailing_padding(ops, ref):
    """Image self-attention / co-attention over padded regions (num_boxes < max_region_len): the key mask has
    trailing zeros for Lk = 37; masked keys get ~0 probability (the reference adds -10000, `:2580`, `:449`)."""
    B, H, Lq, Lk, d = 16, 8, 36, 37, 128
    qbuf = rnd(B * Lq, 3 * H * d)
    kvbuf = rnd(B * Lk, 3 * H * d, seed=1)
    nbox = torch.randint(18, 37, (B,), generator=torch.Generator().manual_seed(9))
    mask = torch.cat([torch.ones(B, 1, dtype=torch.long), (torch.arange(36)[None, :] < nbox[:, None]).long()], 1).to(DEV)
    madd = ops.mask_additive(mask)
    assert torch.equal(madd, ref.mask_additive(mask))
    ctx, lse = ops.attn_fwd(qbuf, 0, kvbuf, H * d, 2 * H * d, madd, B, H, Lq, Lk, d)
    ctxr, lser = ref.attn_fwd(qbuf, 0, kvbuf, H * d, 2 * H * d, madd, B, H, Lq, Lk, d)
    close(lse[0], lser, 1e-5, 1e-4, "lse")
    close_bf16(ctx, ctxr, "ctx", atol=8e-3)
    # changing the VALUES of masked keys must not change the output beyond exp(-10000) = 0
    kv2 = kvbuf.clone().reshape(B, Lk, -1)
    kv2[~mask.bool()] = 7.0
    ctx2, _ = ops.attn_fwd(qbuf, 0, kv2.reshape(B * Lk, -1), H * d, 2 * H * d, madd, B, H, Lq, Lk, d)
    assert torch.equal(ctx, ctx2)


# ---- embeddings ------------------------------------------------------------------------------------
def test_embed_fwd_bwd_bit_exact_gather(ops, ref):
    V, L, H, B = 21128, 36, 768, 16
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(0, V, (B * L,), generator=g)
    ids[::7] = 0                       # padding rows
    ids[1::9] = ids[1]                 # collisions in the scatter-add
    ids, typ_ids = ids.to(DEV), torch.randint(0, 2, (B * L,), generator=g).to(DEV)
    word, pos, typ = rnd(V, H, scale=0.02), rnd(512, H, seed=1, scale=0.02), rnd(2, H, seed=2, scale=0.02)
    out = ops.embed_fwd(ids, typ_ids, word, pos, typ, L)
    want = ref.embed_fwd(ids, typ_ids, word, pos, typ, L)
    assert torch.equal(out, want.to(BF))          # fp32 sum of three bf16 rows, one rounding
    dx = rnd(B * L, H, seed=3)
    res = []
    for o in (ops, ref):
        dw, dp, dt = (torch.zeros(V, H, dtype=F32, device=DEV), torch.zeros(512, H, dtype=F32, device=DEV),
                      torch.zeros(2, H, dtype=F32, device=DEV))
        o.embed_bwd(dx, ids, typ_ids, L, dw, dp, dt)
        res.append((dw, dp, dt))
    assert float(res[0][0][0].abs().max()) == 0.0   # padding_idx row gets nothing from this path
    for a, b, n in zip(res[0], res[1], ("dword", "dpos", "dtype")):
        close(a, b, 1e-5, 1e-4, n)
    assert torch.equal(res[0][0] != 0, res[1][0] != 0)


def test_loc_embed(ops, ref):
    M, N = 2368, 1024
    loc = torch.rand(M, 5, generator=torch.Generator().manual_seed(1)).to(DEV)
    w, b = rnd(N, 5), rnd(N, seed=1)
    close_bf16(ops.loc_embed_fwd(loc, w, b), ref.loc_embed_fwd(loc, w, b), "loc fwd", atol=1e-3)
    dy = rnd(M, N, seed=2)
    res = []
    for o in (ops, ref):
        dw, db = torch.zeros(N, 5, dtype=F32, device=DEV), torch.zeros(N, dtype=F32, device=DEV)
        o.loc_embed_bwd(dy, loc, dw, db)
        res.append((dw, db))
    close(res[0][0], res[1][0], 1e-4, 1e-3, "loc dw")
    close(res[0][1], res[1][1], 1e-4, 1e-3, "loc db")


def test_gather_and_scatter_rows_bit_exact(ops, ref):
    """Labelled-row gather of the MLM head: row copies are bit-exact, padding (-1) gives zero rows / is skipped."""
    M, H, n = 4608, 768, 1152
    src = rnd(M, 3 * H)[:, H:2 * H]                       # a strided view
    g = torch.Generator().manual_seed(3)
    idx = torch.randperm(M, generator=g)[:n].to(torch.int32)
    idx[-100:] = -1
    idx = idx.to(DEV)
    got, want = ops.gather_rows(src, idx), ref.gather_rows(src, idx)
    assert torch.equal(got, want) and float(got[-100:].abs().max()) == 0.0
    d = rnd(n, H, seed=5)
    a, b = rnd(M, H, seed=6), None
    b = a.clone()
    ops.scatter_add_rows(d, idx, a)
    ref.scatter_add_rows(d, idx, b)
    assert torch.equal(a, b)


def test_soft_fusion_gate_kernels(ops, ref):
    """if_pre_sampling = 2 (`vilbert_k3m.py:2300-2329`): gated = cat * alpha and its backward."""
    M, H = 333, 768
    cat = ops.relu_cat3(rnd(M, H), rnd(M, H, seed=1), rnd(M, H, seed=2))
    alpha = torch.sigmoid(rnd(M, 3 * H, seed=3).float()).to(BF)
    assert torch.equal(ops.gate_mul(cat, alpha), ref.gate_mul(cat, alpha).to(BF))
    dg = rnd(M, 3 * H, seed=4)
    (dc, dsc), (dcr, dscr) = ops.gate_bwd(dg, cat, alpha), ref.gate_bwd(dg, cat, alpha)
    close_bf16(dc, dcr.float(), "dcat", ulps=1.01)
    close_bf16(dsc, dscr.float(), "dscore", ulps=1.01)


def test_struct_fallback_rows(ops, ref):
    """Leading items without a PV pair (`:2452-2456`): t[row] = c_init[item]; backward moves the row's gradient."""
    P, B, H = 9, 4, 768
    t, c = rnd(P, H, dtype=F32), rnd(B, H, seed=1, dtype=F32)
    fb = torch.tensor([[2, 1], [-1, -1]], dtype=torch.int32, device=DEV)
    t1, t2 = t.clone(), t.clone()
    ops.struct_fallback_fwd(t1, c, fb)
    ref.struct_fallback_fwd(t2, c, fb)
    assert torch.equal(t1, t2) and torch.equal(t1[2], c[1]) and torch.equal(t1[3], t[3])
    dt1, dc1 = rnd(P, H, seed=2, dtype=F32), rnd(B, H, seed=3, dtype=F32)
    dt2, dc2 = dt1.clone(), dc1.clone()
    ops.struct_fallback_bwd(dt1, dc1, fb)
    ref.struct_fallback_bwd(dt2, dc2, fb)
    assert torch.equal(dt1, dt2) and torch.equal(dc1, dc2) and float(dt1[2].abs().max()) == 0.0


# ---- fusion (hard Gumbel pick) ----------------------------------------------------------------------
@pytest.mark.parametrize("use_noise", [True, False])
def test_fuse_select(ops, ref, use_noise):
    M, H = 333, 768
    a, b, c = rnd(M, H), rnd(M, H, seed=1), rnd(M, H, seed=2)
    cat = ops.relu_cat3(a, b, c)
    assert torch.equal(cat, ref.relu_cat3(a, b, c).to(BF))
    gates = rnd(M, 3 * H, seed=3, dtype=F32)
    r = rng_state()
    noise = None
    if use_noise:
        noise = -torch.empty(M, 3, H).exponential_(generator=torch.Generator().manual_seed(3)).log().to(DEV)
    out = ops.fuse_select_fwd(cat, gates, noise=noise, site=4, rng=r)
    want = ref.fuse_select_fwd(cat, gates, noise=noise, site=4, rng=r)
    mism = (out.float() != want.to(BF).float())
    # the pick is an argmax over fp32 values computed with slightly different exp/log: allow a
    # vanishing fraction of near-ties with injected noise, a small one with on-device noise
    assert float(mism.float().mean()) <= (1e-5 if use_noise else 2e-3), float(mism.float().mean())
    dout = rnd(M, H, seed=5)
    dg, dc = ops.fuse_select_bwd(dout, cat, gates, noise=noise, site=4, rng=r)
    dgr, dcr = ref.fuse_select_bwd(dout, cat, gates, noise=noise, site=4, rng=r)
    ok = ~mism
    ok3 = torch.cat([ok, ok, ok], 1)
    close_bf16(torch.where(ok3, dc.float(), 0 * dcr), torch.where(ok3, dcr, 0 * dcr), "dcat")
    if use_noise:
        close_bf16(dg, dgr, "dgates", atol=2e-3 * float(dgr.abs().max()), ulps=3)
    d1 = [torch.zeros(M, H, dtype=BF, device=DEV) + 1 for _ in range(3)]
    d2 = [torch.zeros(M, H, dtype=F32, device=DEV) + 1 for _ in range(3)]
    ops.relu_bwd_split3(dcr.to(BF), cat, *d1)
    ref.relu_bwd_split3(dcr.to(BF), cat, *d2)
    for x, y in zip(d1, d2):
        close_bf16(x, y, "relu split")


def test_gumbel_noise_distribution(ops, ref):
    """on-device noise path: selection frequencies with zero gates must be ~1/3 each"""
    M, H = 512, 768
    cat = torch.cat([torch.full((M, H), v, dtype=BF, device=DEV) for v in (1.0, 2.0, 3.0)], 1)
    gates = torch.zeros(M, 3 * H, dtype=F32, device=DEV)
    out = ops.fuse_select_fwd(cat, gates, site=9, rng=rng_state())
    for v in (1.0, 2.0, 3.0):
        assert abs(float((out == v).float().mean()) - 1 / 3) < 0.01


def test_mean_pool(ops, ref):
    B, L, H = 64, 37, 1024
    seq = rnd(B * L, 2 * H)[:, H:]
    close(ops.mean_pool_fwd(seq, B, L, F32), ref.mean_pool_fwd(seq, B, L, F32), 1e-5, 1e-5, "pool f32")
    close_bf16(ops.mean_pool_fwd(seq, B, L, BF), ref.mean_pool_fwd(seq, B, L, F32), "pool bf16")
    dp = rnd(B, H, seed=1, dtype=F32)
    d1 = rnd(B * L, H, seed=2)
    d2 = d1.clone().float()
    ops.mean_pool_bwd(dp, d1, B, L)
    ref.mean_pool_bwd(dp, d2, B, L)
    close_bf16(d1, d2, "pool bwd")


# ---- structure aggregator + LPM ------------------------------------------------------------------
def _plan(B, L, mpv, seed):
    import random
    import numpy as np
    from k3m_b200.struct_plan import build_struct_plan
    rs = np.random.RandomState(seed)
    ip = np.zeros((B, mpv, 2), dtype=np.int64)
    iv = np.zeros((B, mpv, 2), dtype=np.int64)
    for i in range(B):
        n = rs.randint(1, mpv + 1)
        for j in range(n):
            p = np.sort(rs.choice(np.arange(1, L), 4, replace=False))
            ip[i, j] = p[:2]
            iv[i, j] = p[2:]
    random.seed(seed)
    return build_struct_plan(ip, iv, L, 4)


def test_struct_aggregator_kernels(ops, ref):
    B, L, H = 16, 36, 768
    plan_np = _plan(B, L, 6, 3)
    plan = ops.plan_to_device(plan_np, DEV)
    c_init, seq = rnd(B, H, dtype=F32), rnd(B * L, H, seed=1)
    X, PV = ops.pair_gather_fwd(c_init, seq, plan["pair_item"], plan["pair_rows"])
    Xr, PVr = ref.pair_gather_fwd(c_init, seq, plan["pair_item"], plan["pair_rows"])
    assert torch.equal(PV, PVr)                     # 2-row means of bf16 rows: exact in fp32
    assert torch.equal(X, Xr.to(BF))
    P = X.shape[0]
    t = rnd(P, H, seed=2, dtype=F32)
    w2, b2 = rnd(1, H, seed=3, scale=0.05), rnd(1, seed=4)
    s, att = ops.struct_attn_fwd(t, w2, b2, plan["seg"])
    sr, attr = ref.struct_attn_fwd(t, w2, b2, plan["seg"])
    close(att, attr, 1e-4, 1e-6, "att")
    close(s, sr, 1e-4, 1e-5, "s")
    ds = rnd(B, H, seed=5, dtype=F32)
    res = []
    for o, a in ((ops, att), (ref, attr)):
        dt = torch.zeros(P, H, dtype=F32, device=DEV)
        dw2, db2 = torch.zeros(1, H, dtype=F32, device=DEV), torch.zeros(1, dtype=F32, device=DEV)
        o.struct_attn_bwd(ds, t, a, w2, plan["seg"], dt, dw2, db2)
        res.append((dt, dw2, db2))
    for a, b, n in zip(res[0], res[1], ("dt", "dw2", "db2")):
        close(a, b, 1e-3, 1e-4 * float(b.abs().max()) + 1e-5, n)
    c_final = rnd(B, H, seed=6, dtype=F32)
    res = []
    for o in (ops, ref):
        dc, dpv = torch.zeros(B, H, dtype=F32, device=DEV), torch.zeros(P, 2 * H, dtype=F32, device=DEV)
        loss = o.lpm_fwd_bwd(c_final, PV, plan["terms"], 60.0, dc, dpv)
        res.append((loss, dc, dpv))
    assert float(res[1][0]) > 0
    for a, b, n in zip(res[0], res[1], ("lpm loss", "dc_final", "dPV")):
        close(a, b, 1e-4, 1e-6, n)
    res = []
    dX, dPV = rnd(P, 3 * H, seed=7, dtype=F32), rnd(P, 2 * H, seed=8, dtype=F32)
    for o, dt_ in ((ops, BF), (ref, F32)):
        dci = torch.zeros(B, H, dtype=F32, device=DEV)
        dseq = torch.zeros(B * L, H, dtype=dt_, device=DEV)
        o.pair_gather_bwd(dX, dPV, plan["pair_item"], plan["pair_rows"], dci, dseq)
        res.append((dci, dseq))
    close(res[0][0], res[1][0], 1e-5, 1e-5, "dc_init")
    close_bf16(res[0][1], res[1][1], "dseq", atol=1e-3)


# ---- losses ----------------------------------------------------------------------------------------
# one-pass row kernel (row packed in registers): V % 8 == 0 up to 8192 (4 vectors per thread) / 24576 (12); everything
# else (ragged V, larger vocabularies) takes the three-sweep scalar kernel
@pytest.mark.parametrize("M,V", [(512, 21128), (40, 1000), (3, 21128), (9, 1003), (6, 8192), (4, 24576), (5, 30000)])
def test_xent(ops, ref, M, V):
    ld = (V + 7) // 8 * 8
    logits = rnd(M, ld, scale=2.0)
    g = torch.Generator().manual_seed(M)
    labels = torch.randint(0, V, (M,), generator=g)
    labels[torch.rand(M, generator=g) > 0.2] = -1
    labels[0] = V - 1
    labels = labels.to(DEV)
    cnt = ops.count_valid(labels, -1)
    assert float(cnt) == float((labels != -1).sum())
    assert float(ops.count_eq(labels, -1)) == float((labels == -1).sum())
    dl_r = torch.zeros(M, ld, dtype=F32, device=DEV)
    want = ref.xent_fwd_bwd(logits, labels, cnt, V, dlogits=dl_r, grad_scale=0.7)
    lg = logits.clone()
    got = ops.xent_fwd_bwd(lg, labels, cnt, V, dlogits=lg, grad_scale=0.7)
    close(got, want, 1e-5, 1e-6, "xent loss")
    close_bf16(lg, dl_r, "dlogits", atol=1e-6)
    assert bool((lg[labels == -1] == 0).all())
    lg2 = logits.clone()
    got2 = ops.xent_fwd_bwd(lg2, labels, cnt, V)          # value only: logits untouched
    assert torch.equal(lg2, logits) and float(got2) == pytest.approx(float(got), rel=1e-6)


def test_kl(ops, ref):
    B, Lv1, T = 16, 37, 1601
    ld = (T + 7) // 8 * 8
    logits = rnd(B * Lv1, ld, dtype=F32, scale=2.0)
    target = torch.softmax(rnd(B * (Lv1 - 1), T, dtype=F32, seed=1) * 3, 1)
    target[5] = 0.0                                        # xlogy(0, 0) = 0
    il = (torch.rand(B * (Lv1 - 1), generator=torch.Generator().manual_seed(2)) < 0.15).long() * 2 - 1
    il = il.to(DEV)
    il[5] = 1
    cnt = ops.count_eq(il, 1)
    dl_g = torch.full((B * Lv1, ld), 3.0, dtype=BF, device=DEV)
    dl_r = torch.zeros(B * Lv1, ld, dtype=F32, device=DEV)
    got = ops.kl_fwd_bwd(logits, target, il, cnt, T, Lv1, dlogits=dl_g, grad_scale=1.3)
    want = ref.kl_fwd_bwd(logits, target, il, cnt, T, Lv1, dlogits=dl_r, grad_scale=1.3)
    close(got, want, 1e-5, 1e-6, "kl loss")
    close_bf16(dl_g, dl_r, "kl dlogits", atol=1e-7)


def test_itm(ops, ref):
    B, H = 64, 768
    pt, pp, pv = rnd(B, H, dtype=F32), rnd(B, H, seed=1, dtype=F32), rnd(B, H, seed=2, dtype=F32)
    w, b = rnd(2, H, seed=3, scale=0.05), rnd(2, seed=4)
    z = torch.zeros(B, dtype=torch.int64, device=DEV)
    l1 = z.clone()
    l1[::3] = 1
    r = rng_state()
    for drop in (None, (0.1, 77)):
        close(ops.itm_loss(pt, pp, pv, w, b, l1, z, z, drop=drop, rng=r),
              ref.itm_loss(pt, pp, pv, w, b, l1, z, z, drop=drop, rng=r), 1e-5, 1e-6, "itm")


def test_pair_cls_loss_and_dropout(ops, ref):
    """item-alignment head kernels (K3MForItemAlignment "ce" branch, ClassificationHead dropout)"""
    B = 37
    logits = rnd(B, 2, dtype=F32, scale=2.0)
    labels = (torch.arange(B, device=DEV) % 3 == 0).to(torch.int64)
    d1 = torch.zeros(B, 8, dtype=BF, device=DEV)
    d2 = torch.zeros(B, 8, dtype=BF, device=DEV)
    l1, p1 = ops.pair_cls_loss(logits, labels, dlogits=d1, grad_scale=0.5)
    l2, p2 = ref.pair_cls_loss(logits, labels, dlogits=d2, grad_scale=0.5)
    close(l1, l2, 1e-6, 1e-6, "pair loss")
    close(p1, p2, 1e-6, 1e-6, "pair probs")
    close_bf16(d1, d2, "pair dlogits")
    assert float(d1[:, 2:].abs().max()) == 0.0
    r = rng_state()
    x = rnd(33, 1536, seed=5)
    for drop in (None, (0.1, 9)):
        y = ops.dropout(x, drop, r)
        close_bf16(y, ref.dropout(x, drop, r), "dropout")
        if drop is not None:                         # keep-mask bit-exact against the counter RNG restatement
            assert torch.equal(y != 0, dropout_keep(r, 9, 33, 1536, 0.1) & (x != 0))
    # strided fp32 -> bf16 copy used to build cat(c_final_1, c_final_2)
    c = rnd(8, 768, dtype=F32, seed=6)
    feat = torch.zeros(8, 1536, dtype=BF, device=DEV)
    ops.dropout(c, None, r, out=feat[:, 768:])
    assert torch.equal(feat[:, 768:], c.to(BF)) and float(feat[:, :768].abs().max()) == 0.0


# ---- optimizer -------------------------------------------------------------------------------------
def test_adamw_matches_torch_optim(ops, ref):
    n = 1 << 20
    p0 = rnd(n, dtype=F32, scale=0.02)
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([p], lr=1e-4, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01)
    p32, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    p16 = torch.zeros(n, dtype=BF, device=DEV)
    for step in (1, 2, 3):
        g = rnd(n, dtype=F32, seed=step, scale=1e-3)
        p.grad = g.clone()
        opt.step()
        g32 = g.clone() * 4.0                              # grad_scale 0.25 undoes it (1/world)
        hyper = torch.tensor([1e-4, 0.9, 0.98, 1e-8, 0.01, float(step), 0.25, 0.0], dtype=F32, device=DEV)
        ops.adamw(p32, g32, m, v, p16, hyper, zero_grad=True)
        assert float(g32.abs().max()) == 0.0
        close(p32, p.data, 1e-6, 1e-8, "adamw step %d" % step)
        assert torch.equal(p16, p32.to(BF))
