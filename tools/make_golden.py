"""Generates tests/golden/*.pt by running the UNMODIFIED reference model
(/root/reference/vilbert_k3m/vilbert_k3m.py, imported through oracle/ref_harness.py) in fp32 on
CPU, deterministic mode (eval -> dropout off, Gumbel noise injected by patching
torch.nn.functional.gumbel_softmax in THIS process, random.seed before the forward for the LPM
negatives).  Only works in the build container; the fixtures are committed.

Each fixture holds: config dict + driver overrides, the seeds that regenerate weights
(`synth_state_dict`) and inputs (`synth_batch`), the Gumbel noise, the five losses, c_initial /
c_final, and for every parameter the gradient's L2 norm, max-abs and 16 sampled elements.
"""
import os
import random
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT, K3M_TINY  # noqa: E402
from k3m_b200.layout import param_spec  # noqa: E402
from k3m_b200.synth import FORWARD_ORDER, synth_batch, synth_state_dict  # noqa: E402
from oracle import ref_harness  # noqa: E402

CASES = {
    # name: (config, overrides, B, Lt, Lv, Lpv, max_num_pv, weight seed, batch seed, noise seed, py seed[, no_pv items])
    "tiny_hard": (K3M_TINY, dict(if_pre_sampling=1, margin=10.0, num_negative_pv=4), 4, 12, 9, 20, 4, 1, 7, 5, 3),
    "tiny_mean": (K3M_TINY, dict(if_pre_sampling=0, margin=1.0, num_negative_pv=4), 3, 16, 8, 24, 5, 2, 8, 0, 4),
    "tiny_inter": (K3M_TINY, dict(if_pre_sampling=3, margin=10.0, num_negative_pv=2), 5, 8, 12, 16, 3, 3, 9, 0, 5),
    "base_mean": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=0, margin=10.0, num_negative_pv=4),
                  4, 36, 36, 36, 8, 1, 7, 0, 3),
    "base_hard": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=1, margin=10.0, num_negative_pv=4),
                  4, 36, 36, 36, 8, 1, 7, 5, 3),
    # if_pre_sampling = 2: `pre_sampling_sequence_soft` (`vilbert_k3m.py:2300-2329`)
    "tiny_soft": (K3M_TINY, dict(if_pre_sampling=2, margin=10.0, num_negative_pv=4), 4, 12, 9, 20, 4, 4, 10, 0, 6),
    # items 0 and 1 without any property/value pair: the aggregator's bare `except` (`:2452-2456`) and the stale `t`
    "tiny_nopv": (K3M_TINY, dict(if_pre_sampling=0, margin=1.0, num_negative_pv=4), 5, 12, 9, 20, 4, 5, 11, 0, 7, (0, 1, 3)),
    # the benchmark's batch (BASELINE configs[1]): the CTA-pair GEMM variants and their fused epilogues only run at
    # these sizes (M = 9216 / 4736 / 2304 rows)
    "base_hard_b64": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=1, margin=10.0, num_negative_pv=4),
                      64, 36, 36, 36, 8, 1, 7, 5, 3),
    "base_mean_b64": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=0, margin=10.0, num_negative_pv=4),
                      64, 36, 36, 36, 8, 1, 7, 0, 3),
    # BASELINE configs[3] geometry (128-token titles, 100 regions, 128 PV tokens): the > 128-token attention kernels
    "long_hard": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=1, margin=10.0, num_negative_pv=4),
                  2, 128, 100, 128, 16, 6, 12, 8, 9),
}


def sample_index(numel, k=256):
    g = torch.Generator().manual_seed(numel)
    return torch.randint(0, numel, (k,), generator=g)


def make_noise(cfg, B, Lt, Lv1, Lpv, seed):
    g = torch.Generator().manual_seed(seed)
    out = {}
    for m, (L, H) in (("v", (Lv1, cfg.bi_hidden_size)), ("t", (Lt, cfg.hidden_size)), ("pv", (Lpv, cfg.hidden_size))):
        out[m] = -torch.empty(B, L, 3, H).exponential_(generator=g).log()
    return out


TAP_PREFIXES = ("encoder.layer.", "encoder.v_layer.", "encoder.c_layer.", "encoder.c_layer_pv_v.",
                "encoder.c_layer_pv_t.")


def install_taps(model, taps):
    """Forward hooks on every `BertLayer` (`vilbert_k3m.py:535-548`), `BertImageLayer` (`:696-709`) and
    `BertConnectionLayer` / `_two_text` (`:1030-1111`) of the unmodified reference: per call, the L2 norm and
    256 sampled elements of each hidden-state output.  Key: "<module name>#<call index>[/<output index>]";
    the call order of a shared module is the order of the three encoder passes (`:1686-1750`)."""
    calls = {}

    def hook_for(name, n_out):
        def hook(mod, args, out):
            k = calls.get(name, 0)
            calls[name] = k + 1
            for j in range(n_out):
                x = out[j].detach().reshape(-1)
                key = "%s#%d" % (name, k) + ("/%d" % j if n_out > 1 else "")
                taps[key] = dict(norm=float(x.norm()), sample=x[sample_index(x.numel())].clone(), shape=tuple(out[j].shape))
        return hook

    handles = []
    for name, mod in model.named_modules():
        for p in TAP_PREFIXES:
            if name.startswith(p) and name[len(p):].isdigit():
                handles.append(mod.register_forward_hook(hook_for(name, 2 if ".c_layer" in name else 1)))
    return handles


def run_case(name):
    d, ov, B, Lt, Lv, Lpv, mpv, wseed, bseed, nseed, pyseed = CASES[name][:11]
    no_pv = CASES[name][11] if len(CASES[name]) > 11 else ()
    ref, rcfg, model = ref_harness.make_reference_model(d, ov)
    cfg = BertConfig.from_dict(d)
    for k, v in ov.items():
        setattr(cfg, k, v)
    sd = synth_state_dict(param_spec(cfg), seed=wseed)
    full = dict(sd)
    full["cls.predictions.decoder.weight"] = sd["embeddings.word_embeddings.weight"]
    model.load_state_dict(full)
    model.tie_weights()
    model.eval()
    batch = synth_batch(cfg, B, Lt, Lv, Lpv, max_num_pv=mpv, seed=bseed, no_pv=no_pv)
    noise = make_noise(cfg, B, Lt, Lv + 1, Lpv, nseed) if cfg.if_pre_sampling == 1 else None
    taps = {}
    handles = install_taps(model, taps)
    orig = F.gumbel_softmax
    if noise is not None:
        order = iter(["v", "t", "pv"])   # call order in get_sequence_pooled_output_final (:2387-2392)

        def patched(logits, tau=1, hard=False, eps=1e-10, dim=-1):
            y = ((logits + noise[next(order)]) / tau).softmax(dim)
            idx = y.max(dim, keepdim=True)[1]
            yh = torch.zeros_like(logits).scatter_(dim, idx, 1.0)
            return yh - y.detach() + y
        F.gumbel_softmax = patched
    random.seed(pyseed)
    out = model(*[batch[k] for k in FORWARD_ORDER], input_ids_pv=batch["input_ids_pv"],
                token_type_ids_pv=batch["token_type_ids_pv"], attention_mask_pv=batch["attention_mask_pv"],
                masked_lm_labels_pv=batch["masked_lm_labels_pv"],
                next_sentence_label_pv_v=batch["next_sentence_label_pv_v"],
                next_sentence_label_pv_t=batch["next_sentence_label_pv_t"],
                index_p=batch["index_p"], index_v=batch["index_v"], device="cpu")
    F.gumbel_softmax = orig
    for h in handles:
        h.remove()
    (out[0] + out[1] + out[3] + out[9]).backward()      # pretrain.py:1055
    grads = {}
    for n, p in model.named_parameters():
        if p.grad is None:
            grads[n] = None
            continue
        g = p.grad.reshape(-1)
        grads[n] = dict(norm=float(g.norm()), maxabs=float(g.abs().max()), sample=g[sample_index(g.numel())].clone())
    fx = dict(case=name, config=d, overrides=ov, dims=(B, Lt, Lv, Lpv, mpv), no_pv=tuple(no_pv), taps=taps,
              seeds=dict(weights=wseed, batch=bseed, noise=nseed, python=pyseed),
              noise=noise,
              losses=dict(loss_t=float(out[0]), loss_v=float(out[1]), loss_pv=float(out[3]),
                          loss_itm=float(out[6]), loss_lpm=float(out[9])),
              c_initial=out[7].detach().clone(), c_final=out[8].detach().clone(), grads=grads)
    if noise is not None:
        fx["noise"] = None   # regenerated from the seed by make_noise (keeps the fixture small)
    path = os.path.join(ROOT, "tests", "golden", name + ".pt")
    torch.save(fx, path)
    print(name, fx["losses"], os.path.getsize(path))


# ---- item alignment (K3MForItemAlignment, finetune.py / run_finetune_item_alignment.sh) ---------------
ALIGN_CASES = {
    # name: (config, overrides, pairs, Lt, Lv, Lpv, max_num_pv, weight seed, batch seed, noise seed)
    "align_tiny_mean": (K3M_TINY, dict(if_pre_sampling=0, loss_type="ce"), 3, 12, 9, 20, 4, 11, 17, 0),
    "align_tiny_hard": (K3M_TINY, dict(if_pre_sampling=1, loss_type="ce"), 4, 10, 8, 16, 3, 12, 18, 6),
    # run_finetune_item_alignment.sh:40-43 lengths (50-token titles, 36 regions, 256 PV tokens), base config
    "align_base_hard": (K3M_BASE_6LAYER_6CONNECT, dict(if_pre_sampling=1, loss_type="ce"), 2, 50, 36, 256, 16, 13, 19, 7),
}
ITEM_KEYS = ("input_ids", "token_type_ids", "attention_mask", "input_ids_pv", "token_type_ids_pv",
             "attention_mask_pv", "index_p", "index_v", "image_feat", "image_loc", "image_attention_mask")


def split_items(batch, B):
    """stacked synthetic batch of 2B items -> (item-1 dict, item-2 dict) of B items each"""
    return ({k: batch[k][:B] for k in ITEM_KEYS}, {k: batch[k][B:] for k in ITEM_KEYS})


def align_labels(B, seed):
    return torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(seed))


def run_align_case(name):
    d, ov, B, Lt, Lv, Lpv, mpv, wseed, bseed, nseed = ALIGN_CASES[name]
    ref = ref_harness.import_reference()
    rcfg = ref.BertConfig.from_dict(dict(d))
    cfg = BertConfig.from_dict(d)
    for k, v in ov.items():
        setattr(rcfg, k, v)
        setattr(cfg, k, v)
    model = ref.K3MForItemAlignment(rcfg)
    spec = param_spec(cfg, "alignment")
    assert [n for n, _ in model.named_parameters()] == list(spec.keys()), "alignment parameter order"
    sd = synth_state_dict(spec, seed=wseed)
    model.load_state_dict(sd)
    model.eval()
    batch = synth_batch(cfg, 2 * B, Lt, Lv, Lpv, max_num_pv=mpv, seed=bseed)
    it1, it2 = split_items(batch, B)
    labels = align_labels(B, bseed)
    noise = make_noise(cfg, 2 * B, Lt, Lv + 1, Lpv, nseed) if cfg.if_pre_sampling == 1 else None
    orig = F.gumbel_softmax
    if noise is not None:
        # two item_embedding calls (:3408-3433), each fusing v, t, pv (:2387-2392)
        order = iter([("v", 0), ("t", 0), ("pv", 0), ("v", 1), ("t", 1), ("pv", 1)])

        def patched(logits, tau=1, hard=False, eps=1e-10, dim=-1):
            m, half = next(order)
            y = ((logits + noise[m][half * B:(half + 1) * B]) / tau).softmax(dim)
            idx = y.max(dim, keepdim=True)[1]
            yh = torch.zeros_like(logits).scatter_(dim, idx, 1.0)
            return yh - y.detach() + y
        F.gumbel_softmax = patched
    args = [labels]
    for it in (it1, it2):
        args += [it[k] for k in ITEM_KEYS]
    out = model(*args)
    out[3].backward()
    finals = []
    with torch.no_grad():
        for half, it in enumerate((it1, it2)):
            if noise is not None:
                order = iter([("v", half), ("t", half), ("pv", half)])
            finals.append(model.item_embedding(it["input_ids"], it["image_feat"], it["image_loc"], it["token_type_ids"],
                                               it["attention_mask"], it["image_attention_mask"], False,
                                               it["input_ids_pv"], it["token_type_ids_pv"], it["attention_mask_pv"],
                                               it["index_p"], it["index_v"]))
    F.gumbel_softmax = orig
    grads = {}
    for n, p in model.named_parameters():
        if p.grad is None:
            grads[n] = None
            continue
        g = p.grad.reshape(-1)
        grads[n] = dict(norm=float(g.norm()), maxabs=float(g.abs().max()), sample=g[sample_index(g.numel())].clone())
    fx = dict(case=name, kind="alignment", config=d, overrides=ov, dims=(B, Lt, Lv, Lpv, mpv),
              seeds=dict(weights=wseed, batch=bseed, noise=nseed), noise=None,
              loss=float(out[3]), probs=torch.stack((out[0], out[1]), dim=1).detach().clone(),
              c_initial=torch.cat([f[0] for f in finals]).clone(), c_final=torch.cat([f[1] for f in finals]).clone(),
              grads=grads)
    path = os.path.join(ROOT, "tests", "golden", name + ".pt")
    torch.save(fx, path)
    print(name, fx["loss"], fx["probs"][:, 1].tolist(), os.path.getsize(path))


if __name__ == "__main__":
    for n in (sys.argv[1:] or list(CASES) + list(ALIGN_CASES)):
        if n in ALIGN_CASES:
            run_align_case(n)
        else:
            run_case(n)
