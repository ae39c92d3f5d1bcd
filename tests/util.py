"""Shared helpers for the parity tests."""
import os
import random
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

from k3m_b200.config import BertConfig  # noqa: E402
from k3m_b200.engine import Engine  # noqa: E402
from k3m_b200.layout import ArenaLayout, param_spec  # noqa: E402
from k3m_b200.struct_plan import build_mlm_plan, build_struct_plan  # noqa: E402
from k3m_b200.synth import synth_batch, synth_state_dict  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def sample_index(numel, k=256):
    g = torch.Generator().manual_seed(numel)
    return torch.randint(0, numel, (k,), generator=g)


def make_noise(cfg, B, Lt, Lv1, Lpv, seed):
    g = torch.Generator().manual_seed(seed)
    out = {}
    for m, (L, H) in (("v", (Lv1, cfg.bi_hidden_size)), ("t", (Lt, cfg.hidden_size)), ("pv", (Lpv, cfg.hidden_size))):
        out[m] = -torch.empty(B, L, 3, H).exponential_(generator=g).log()
    return out


def case_setup(fx):
    """fixture -> (cfg, state_dict, batch, noise or None)"""
    cfg = BertConfig.from_dict(fx["config"])
    for k, v in fx["overrides"].items():
        setattr(cfg, k, v)
    B, Lt, Lv, Lpv, mpv = fx["dims"]
    sd = synth_state_dict(param_spec(cfg), seed=fx["seeds"]["weights"])
    batch = synth_batch(cfg, B, Lt, Lv, Lpv, max_num_pv=mpv, seed=fx["seeds"]["batch"], no_pv=fx.get("no_pv", ()))
    noise = make_noise(cfg, B, Lt, Lv + 1, Lpv, fx["seeds"]["noise"]) if cfg.if_pre_sampling == 1 else None
    return cfg, sd, batch, noise


def run_engine(cfg, sd, batch, noise, ops, device, adt, py_seed, training=False, taps=None, gather_mlm=True):
    """One forward+backward of the engine; returns (outputs dict, layout, fp32 grad arena).
    `taps`: optional dict that receives the per-layer outputs (see `Engine.taps`)."""
    L = ArenaLayout(cfg)
    w32 = torch.zeros(L.total)
    for n, (off, num, _) in L.offsets.items():
        w32[off:off + num] = sd[n].reshape(-1)
    w = w32.to(device=device, dtype=adt)
    g = torch.zeros(L.total, dtype=torch.float32, device=device)
    rng = torch.tensor([1234, 0], dtype=torch.int64, device=device)
    eng = Engine(cfg, L, ops, w, g, rng, adt)
    eng.training = training
    eng.taps = taps
    random.seed(py_seed)
    plan = build_struct_plan(batch["index_p"].numpy(), batch["index_v"].numpy(), batch["input_ids_pv"].shape[1],
                             cfg.num_negative_pv)
    if gather_mlm:       # the product path: masked-LM head on the labelled rows only
        plan.update(build_mlm_plan(batch["masked_lm_labels"].numpy(), batch["masked_lm_labels_pv"].numpy()))
    plan = ops.plan_to_device(plan, device)
    b = {k: v.to(device) for k, v in batch.items()}
    en = None
    if noise is not None:
        en = {m: x.reshape(-1, 3, x.shape[-1]).contiguous().to(device) for m, x in noise.items()}
    out = eng.forward(b, plan, en)
    eng.backward()
    return out, L, g


def compare_with_golden(fx, out, L, g, loss_rtol, grad_tol, key_bias_atol=None, itm_rtol=None, stats=None):
    """losses: relative error; gradients, per parameter tensor: |norm - ref| / ref and the relative
    L2 error over 256 sampled elements, both against `grad_tol`.  `stats` (a list) receives the
    per-tensor sampled relative L2 errors."""
    errs = []
    for k, ref in fx["losses"].items():
        got = float(out[k].reshape(-1)[0])
        tol = itm_rtol if (k == "loss_itm" and itm_rtol is not None) else loss_rtol
        if not abs(got - ref) <= tol * max(abs(ref), 1e-6):
            errs.append("%s: got %.6f want %.6f" % (k, got, ref))
    for k in ("c_initial", "c_final"):
        ref = fx[k]
        got = out[k].float().cpu()
        e = float((got - ref).norm()) / (float(ref.norm()) + 1e-12)
        if not e <= max(grad_tol, 10 * loss_rtol):
            errs.append("%s: rel err %.3e" % (k, e))
    gc = g.cpu()
    worst = 0.0
    for n, (off, num, shape) in L.offsets.items():
        ref = fx["grads"][n]
        got = gc[off:off + num]
        if ref is None:
            if float(got.abs().max()) != 0.0:
                errs.append("%s: reference has no gradient, got max %.3e" % (n, float(got.abs().max())))
            continue
        scale = ref["maxabs"]
        if scale < 1e-7:      # key biases / struc_w2.bias: mathematically zero gradients
            if not float(got.abs().max()) <= (key_bias_atol or 1e-5):
                errs.append("%s: expected ~0 gradient, got %.3e" % (n, float(got.abs().max())))
            continue
        e1 = abs(float(got.norm()) - ref["norm"]) / ref["norm"]
        smp = ref["sample"]
        e2 = float((got[sample_index(num)] - smp).norm()) / (float(smp.norm()) + 1e-3 * scale)
        worst = max(worst, e1, e2)
        if stats is not None:
            stats.append((e2, n))
        if not (e1 <= grad_tol and e2 <= grad_tol):
            errs.append("%s: norm err %.3e sampled rel-L2 err %.3e" % (n, e1, e2))
    return errs, worst


def compare_taps(fx, taps, tol):
    """Per-layer activations against the hooks on the unmodified reference (fixture `taps`): for every
    BertLayer / BertImageLayer / BertConnectionLayer call, |norm - ref| / ref and the relative L2 error over
    256 sampled elements.  -> (error strings, list of (sampled rel-L2, key))."""
    errs, stats = [], []
    ref_taps = fx["taps"]
    missing = sorted(set(ref_taps) - set(taps))
    if missing:
        errs.append("missing taps: %s" % missing[:4])
    for key, ref in ref_taps.items():
        if key not in taps:
            continue
        got = taps[key].float().reshape(-1).cpu()
        if got.numel() != ref["sample"].numel() and tuple(taps[key].shape) != (ref["shape"][0] * ref["shape"][1], ref["shape"][2]):
            errs.append("%s: shape %s vs reference %s" % (key, tuple(taps[key].shape), ref["shape"]))
            continue
        e1 = abs(float(got.norm()) - ref["norm"]) / ref["norm"]
        smp = ref["sample"]
        e2 = float((got[sample_index(got.numel())] - smp).norm()) / float(smp.norm())
        stats.append((e2, key))
        if not (e1 <= tol and e2 <= tol):
            errs.append("%s: norm err %.3e sampled rel-L2 err %.3e" % (key, e1, e2))
    return errs, stats


# ---- item alignment (K3MForItemAlignment) ---------------------------------------------------------
ITEM_KEYS = ("input_ids", "token_type_ids", "attention_mask", "input_ids_pv", "token_type_ids_pv",
             "attention_mask_pv", "index_p", "index_v", "image_feat", "image_loc", "image_attention_mask")


def align_case_setup(fx):
    """alignment fixture -> (cfg, state_dict, stacked batch of 2B items, labels [B], noise or None)"""
    cfg = BertConfig.from_dict(fx["config"])
    for k, v in fx["overrides"].items():
        setattr(cfg, k, v)
    B, Lt, Lv, Lpv, mpv = fx["dims"]
    sd = synth_state_dict(param_spec(cfg, "alignment"), seed=fx["seeds"]["weights"])
    batch = synth_batch(cfg, 2 * B, Lt, Lv, Lpv, max_num_pv=mpv, seed=fx["seeds"]["batch"])
    labels = torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(fx["seeds"]["batch"]))
    noise = make_noise(cfg, 2 * B, Lt, Lv + 1, Lpv, fx["seeds"]["noise"]) if cfg.if_pre_sampling == 1 else None
    return cfg, sd, batch, labels, noise


def run_engine_alignment(cfg, sd, batch, labels, noise, ops, device, adt, training=False):
    """One forward+backward of the alignment path; returns (outputs dict, layout, fp32 grad arena)."""
    L = ArenaLayout(cfg, "alignment")
    w32 = torch.zeros(L.total)
    for n, (off, num, _) in L.offsets.items():
        w32[off:off + num] = sd[n].reshape(-1)
    w = w32.to(device=device, dtype=adt)
    g = torch.zeros(L.total, dtype=torch.float32, device=device)
    rng = torch.tensor([1234, 0], dtype=torch.int64, device=device)
    eng = Engine(cfg, L, ops, w, g, rng, adt)
    eng.training = training
    plan = build_struct_plan(batch["index_p"].numpy(), batch["index_v"].numpy(), batch["input_ids_pv"].shape[1],
                             cfg.num_negative_pv, halves=2)
    plan = ops.plan_to_device(plan, device)
    b = {k: v.to(device) for k, v in batch.items()}
    en = None
    if noise is not None:
        en = {m: x.reshape(-1, 3, x.shape[-1]).contiguous().to(device) for m, x in noise.items()}
    out = eng.forward_alignment(b, plan, labels.to(device), en)
    eng.backward_alignment()
    return out, L, g


def compare_alignment_with_golden(fx, out, L, g, loss_rtol, grad_tol, prob_atol):
    errs = []
    got = float(out["loss"].reshape(-1)[0])
    if not abs(got - fx["loss"]) <= loss_rtol * abs(fx["loss"]):
        errs.append("loss: got %.6f want %.6f" % (got, fx["loss"]))
    pe = float((out["probs"].float().cpu() - fx["probs"]).abs().max())
    if not pe <= prob_atol:
        errs.append("probs: max abs err %.3e" % pe)
    for k in ("c_initial", "c_final"):
        ref = fx[k]
        e = float((out[k].float().cpu() - ref).norm()) / (float(ref.norm()) + 1e-12)
        if not e <= max(grad_tol, 10 * loss_rtol):
            errs.append("%s: rel err %.3e" % (k, e))
    gc = g.cpu()
    for n, (off, num, shape) in L.offsets.items():
        ref = fx["grads"][n]
        got = gc[off:off + num]
        if ref is None:
            if float(got.abs().max()) != 0.0:
                errs.append("%s: reference has no gradient, got max %.3e" % (n, float(got.abs().max())))
            continue
        scale = ref["maxabs"]
        if scale < 1e-9:
            if not float(got.abs().max()) <= 1e-6:
                errs.append("%s: expected ~0 gradient, got %.3e" % (n, float(got.abs().max())))
            continue
        e1 = abs(float(got.norm()) - ref["norm"]) / ref["norm"]
        smp = ref["sample"]
        e2 = float((got[sample_index(num)] - smp).norm()) / (float(smp.norm()) + 1e-3 * scale)
        if not (e1 <= grad_tol and e2 <= grad_tol):
            errs.append("%s: norm err %.3e sampled rel-L2 err %.3e" % (n, e1, e2))
    return errs


# ---- every launch of a step against the restatement, on the launch's own inputs -----------------------------
class CheckedOps(object):
    """Op table that runs every call twice: on the CUDA library and — on clones of exactly the same input
    tensors — on the PyTorch restatement of the kernel contract (oracle/ops_ref.py), then compares every tensor
    the call returned or modified in place.  The engine continues with the CUDA results, so each kernel of a
    whole step is checked in isolation at the shapes, strides, epilogue combinations and data the step really
    uses (nothing accumulates from kernel to kernel).  `stats`: entry name -> list of (relative L2 error, dtype,
    shape) per compared tensor."""
    SEQUENTIAL = ("parallel",)

    def __init__(self, cuda_ops, ref_ops):
        self._cuda, self._ref = cuda_ops, ref_ops
        self.adt = cuda_ops.adt
        self.stats = {}

    def parallel(self, fns):
        for f in fns:
            if f is not None:
                f()

    def plan_to_device(self, plan, device):
        return self._cuda.plan_to_device(plan, device)

    def __getattr__(self, name):
        if name in ("aux", "join_aux"):
            raise AttributeError(name)          # weight gradients stay on the launching stream here
        target = getattr(self._cuda, name)
        if not callable(target):
            return target
        ref_fn = getattr(self._ref, name)

        def call(*args, **kw):
            clones = {}

            def cl(x):
                if isinstance(x, (list, tuple)):
                    return type(x)(cl(v) for v in x)
                if isinstance(x, dict):
                    return {k: cl(v) for k, v in x.items()}
                if not isinstance(x, torch.Tensor):
                    return x
                key = (x.data_ptr(), tuple(x.shape), tuple(x.stride()), x.dtype)
                if key not in clones:
                    clones[key] = x.clone()
                return clones[key]

            def flat(x, out):
                if isinstance(x, (list, tuple)):
                    for v in x:
                        flat(v, out)
                elif isinstance(x, dict):
                    for k in sorted(x):
                        flat(x[k], out)
                else:
                    out.append(x)
                return out

            rargs = [cl(x) for x in args]
            rkw = {k: cl(v) for k, v in kw.items()}
            out = target(*args, **kw)
            rout = ref_fn(*rargs, **rkw)
            pairs = list(zip(flat(list(args), []), flat(rargs, []))) + list(zip(flat(kw, []), flat(rkw, [])))
            pairs += list(zip(flat(out, []), flat(rout, [])))
            seen = set()
            for a, b in pairs:
                if not isinstance(a, torch.Tensor) or not isinstance(b, torch.Tensor) or a.data_ptr() in seen:
                    continue
                seen.add(a.data_ptr())
                if name == "attn_fwd" and a.dim() == 2 and a.shape[0] == 2 and b.dim() == 1:
                    a = a[0]                    # lse: row 1 is backward scratch in the CUDA table
                if a.shape != b.shape:
                    continue
                if a.dtype in (torch.int32, torch.int64):
                    err = float((a != b).float().mean())
                else:
                    # output buffers are torch.empty: parts a launch does not write hold the same (cloned) garbage on
                    # both sides, possibly NaN — bit-identical elements count as equal
                    it = {2: torch.int16, 4: torch.int32}[a.element_size()]
                    same = a.contiguous().view(it) == b.contiguous().view(it)
                    af = torch.where(same, torch.zeros((), device=a.device), a.float())
                    bf = torch.where(same, torch.zeros((), device=a.device), b.float())
                    ref_norm = float(torch.nan_to_num(b.float(), nan=0.0, posinf=0.0, neginf=0.0).norm())
                    # floor: gradients that are mathematically zero (struc_w2.bias: a sum of softmax gradients)
                    err = float((af - bf).norm()) / (ref_norm + 1e-6 * a.numel() ** 0.5)
                self.stats.setdefault(name, []).append((err, str(a.dtype).replace("torch.", ""), tuple(a.shape)))
            return out
        return call
