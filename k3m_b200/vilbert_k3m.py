"""Drop-in model classes: the reference's Python class API over the sm_100a kernels.

`BertForMultiModalPreTraining_tri_stru(config)` mirrors the reference class of the same name
(`vilbert_k3m/vilbert_k3m.py:2190-2860`): same constructor, same `forward(...)` signature and
10-tuple return (`:2673-2696`, `:2834-2846`), same `named_parameters()` names and order (998) and
`state_dict()` keys (999, tied decoder alias), `from_pretrained` (`utils.py:884-1078`), so it drops
into `pretrain.py:990-1024` / `train_concap_struc.py:503`:

    out = model(input_ids, image_feat, image_loc, segment_ids, input_mask, image_mask, lm_label_ids,
                image_label, image_target, is_next, False, input_ids_pv, ...)
    loss = out[0] + out[1] * w + out[3] + out[9]; loss.backward(); optimizer.step()

How it differs inside (results are the reference's):
  * every parameter is an fp32 `nn.Parameter` that is a VIEW into one flat arena (`ArenaLayout`);
    `torch.optim.AdamW` / `clip_grad_norm_` / checkpoints see ordinary parameters;
  * forward casts the arena to a bf16 shadow and runs `Engine.forward` (hand-written kernels through
    the C ABI); the four training losses come back attached to ONE autograd node whose backward runs
    `Engine.backward` and points every trained parameter's `.grad` at its slice of the fp32
    gradient arena (`.grad` of the 86 never-trained tensors stays None, like in the reference);
  * data parallel: autograd hooks of `DistributedDataParallel` never fire (there is no per-parameter
    autograd graph) — call `model.set_process_group(pg)` instead and backward all-reduces the
    gradient arena itself (mean over ranks, DDP semantics).
There is no CPU path: without the CUDA library `forward` raises (`k3m_b200._lib`).  Tests inject
the fp32 restatement of the kernels through `ops=` to check THIS host logic on CPU.

The graph-captured production step (`k3m_b200.trainer.PretrainStep`: static buffers, fused AdamW,
overlapped all-reduce) is what `bench.py` measures; this class is the compatibility surface.
"""
import logging
import os
import sys

import torch
from torch import nn

from .config import BertConfig  # noqa: F401  (re-exported under the reference's module path)
from .engine import Engine
from .layout import ArenaLayout, TIED_ALIAS
from .struct_plan import build_mlm_plan, build_struct_plan, mlm_capacity, pad_struct_plan, plan_capacity

logger = logging.getLogger(__name__)

WEIGHTS_NAME = "pytorch_model.bin"      # `utils.py:44`


class _Node(nn.Module):
    """Anonymous container: gives parameters the reference's dotted names."""

    def forward(self, *a, **k):
        raise RuntimeError("k3m_b200 modules are parameter containers; call the model's forward()")


class _GraphedStep(object):
    """Forward and backward of one input geometry as two CUDA graphs (shared memory pool): the class API then costs
    two graph launches per step instead of ~640 Python -> ctypes launches (~25 ms of host time against an 18 ms
    device step).  Inputs are copied into static buffers; the aggregator plan is padded to its capacity
    (`pad_struct_plan`), the four upstream loss gradients live in static device scalars."""

    def __init__(self, owner, batch, max_num_pv):
        dev = owner._arena.device
        cfg = owner.config
        self.owner = owner
        self.inputs = {k: torch.empty_like(v) for k, v in batch.items()}
        self.B, self.max_num_pv = batch["input_ids"].shape[0], max_num_pv
        P_cap, T_cap, maxn = plan_capacity(self.B, max_num_pv, cfg.num_negative_pv)
        i32 = torch.int32
        self.plan = dict(pair_item=torch.zeros(P_cap, dtype=i32, device=dev), pair_rows=torch.zeros(P_cap, 4, dtype=i32, device=dev),
                         seg=torch.zeros(self.B, 2, dtype=i32, device=dev), terms=torch.zeros(T_cap, 4, dtype=i32, device=dev),
                         n_terms=torch.zeros(1, dtype=i32, device=dev), fallback=torch.full((1, 2), -1, dtype=i32, device=dev))
        self.plan["seg"]._k3m_maxn = maxn
        self.mlm_cap = mlm_capacity(batch["input_ids"].numel(), batch["input_ids_pv"].numel(), (0.3, 0.5))
        self.plan["mlm_rows"] = torch.full((sum(self.mlm_cap),), -1, dtype=i32, device=dev)
        self.plan["mlm_labels"] = torch.full((sum(self.mlm_cap),), -1, dtype=torch.int64, device=dev)
        self.plan["mlm_cap_t"] = self.mlm_cap[0]
        self.scales = [torch.ones(1, dtype=torch.float32, device=dev) for _ in range(4)]
        self.g_fwd = self.g_bwd = None
        self.out = None

    def load(self, batch, plan_np, labels):
        self._labels = labels
        for k, v in batch.items():
            self.inputs[k].copy_(v, non_blocking=True)
        struct = {k: v for k, v in plan_np.items() if k not in ("mlm_rows", "mlm_labels", "cap_t")}
        padded = pad_struct_plan(struct, self.B, self.max_num_pv, self.owner.config.num_negative_pv)
        mp = build_mlm_plan(self._labels[0], self._labels[1], self.mlm_cap)
        padded["mlm_rows"], padded["mlm_labels"] = mp["mlm_rows"], mp["mlm_labels"]
        for k, v in padded.items():
            self.plan[k].copy_(torch.from_numpy(v).reshape(self.plan[k].shape), non_blocking=False)

    def capture(self):
        eng = self.owner._engine
        s = torch.cuda.Stream(device=self.owner._arena.device)
        s.wait_stream(torch.cuda.current_stream())
        import gc
        gc.collect()
        was_enabled = gc.isenabled()
        gc.disable()              # see PretrainStep._capture: no finaliser may run inside an open capture
        try:
            self.g_fwd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_fwd, stream=s):
                self.out = eng.forward(self.inputs, self.plan, None, (1.0, 1.0, 1.0, 1.0))
            saved = eng.saved
            self.g_bwd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_bwd, pool=self.g_fwd.pool(), stream=s):
                eng.saved = saved
                eng.backward(tuple(self.scales))
        finally:
            if was_enabled:
                gc.enable()
        torch.cuda.current_stream().wait_stream(s)

    def forward(self):
        self.g_fwd.replay()
        o = self.out
        return tuple(o[k].clone() for k in ("loss_t", "loss_v", "loss_pv", "loss_lpm", "loss_itm", "c_initial", "c_final"))

    def backward(self, scales):
        for dst, src in zip(self.scales, scales):
            dst.copy_(src)
        self.g_bwd.replay()


class _StepFunction(torch.autograd.Function):
    """One autograd node for the whole step.  Inputs: a dummy leaf (so that autograd records the
    node) and the owning module; outputs: the four training losses (differentiable) and the
    value-only results."""

    @staticmethod
    def forward(ctx, anchor, owner, batch, plan, graphed=None):
        ctx.owner = owner
        ctx.token = owner._fwd_token
        ctx.graphed = graphed
        if graphed is not None:
            res = graphed.forward()
        else:
            out = owner._engine.forward(batch, plan, None, (1.0, 1.0, 1.0, 1.0))
            res = (out["loss_t"], out["loss_v"], out["loss_pv"], out["loss_lpm"], out["loss_itm"],
                   out["c_initial"], out["c_final"])
        ctx.mark_non_differentiable(res[4], res[5], res[6])
        return res

    @staticmethod
    def backward(ctx, g_t, g_v, g_pv, g_lpm, *unused):
        owner = ctx.owner
        if ctx.token != owner._fwd_token:
            raise RuntimeError("backward() of a K3M step whose activations were released by a later forward(); "
                               "call backward before the next forward (the reference's drivers do)")
        owner._run_backward((g_t, g_v, g_pv, g_lpm), ctx.graphed)
        return None, None, None, None, None


class K3MPreTrainedModel(nn.Module):
    """Parameter arena + the reference's `PreTrainedModel` entry points used by the drivers."""

    config_class = BertConfig
    base_model_prefix = "bert"
    _kind = "pretrain"                      # parameter inventory (`layout.param_spec`)

    def __init__(self, config, ops=None):
        super(K3MPreTrainedModel, self).__init__()
        if not isinstance(config, BertConfig):
            raise ValueError(
                "Parameter config in `{}(config)` should be an instance of class `BertConfig`. "
                "To create a model from a Google pretrained model use "
                "`model = {}.from_pretrained(PRETRAINED_MODEL_NAME)`".format(
                    self.__class__.__name__, self.__class__.__name__))
        self.config = config
        self._ops = ops
        self._layout = ArenaLayout(config, self._kind)
        self._arena = torch.zeros(self._layout.total, dtype=torch.float32)
        self._engine = None
        self._w16 = self._g32 = self._rng = None
        self._anchor = torch.zeros(1, requires_grad=True)
        self._fwd_token = 0
        self._group = None
        self._params = {}
        self._build_parameters()
        self.init_weights()

    # ---- parameters ------------------------------------------------------------------------
    def _build_parameters(self):
        L = self._layout
        for name in L.spec:
            off, num, shape = L.offsets[name]
            p = nn.Parameter(self._arena[off:off + num].view(*shape))
            self._params[name] = p
            self._node_for(name).register_parameter(name.rsplit(".", 1)[1], p)
        # the tied decoder (`tie_weights`, `vilbert_k3m.py:2266-2272`): same Parameter object under a
        # second name -> named_parameters() lists it once, state_dict() twice (999 keys)
        if self._kind == "pretrain":
            self._node_for(TIED_ALIAS[0]).register_parameter(TIED_ALIAS[0].rsplit(".", 1)[1],
                                                             self._params[TIED_ALIAS[1]])

    def _node_for(self, dotted):
        node = self
        for part in dotted.split(".")[:-1]:
            nxt = node._modules.get(part)
            if nxt is None:
                nxt = _Node()
                node.add_module(part, nxt)
            node = nxt
        return node

    def _rebind(self):
        """Point every Parameter at the (new) arena; drops device-side state."""
        L = self._layout
        with torch.no_grad():
            for name, p in self._params.items():
                off, num, shape = L.offsets[name]
                p.data = self._arena[off:off + num].view(*shape)
                p.grad = None
        self._engine = None
        self._w16 = self._g32 = self._rng = None
        self._anchor = torch.zeros(1, requires_grad=True, device=self._arena.device)

    def _apply(self, fn, recurse=True):
        """`.cuda()` / `.to(device)` / `.float()`: move the ARENA, then re-view the parameters (the
        default per-parameter move would scatter them).  Master weights stay fp32 — `.half()` /
        `.bfloat16()` are accepted and ignored: the kernels compute in bf16 from a shadow copy anyway."""
        new = fn(self._arena)
        if new.dtype != torch.float32:
            new = new.to(torch.float32)
        if new is not self._arena:
            self._arena = new.contiguous()
            self._rebind()
        return self

    def init_weights(self, seed=None):
        """`init_weights` (`vilbert_k3m.py:1940-1951`) applied by `self.apply` BEFORE `struc_w1/2/3`
        exist (`:2252-2259`): N(0, initializer_range) Linear / Embedding weights, zero biases,
        LayerNorm (1, 0); the aggregator keeps nn.Linear's default U(-1/sqrt(in), 1/sqrt(in))."""
        import math
        g = None
        if seed is not None:
            g = torch.Generator(device=self._arena.device).manual_seed(seed)
        with torch.no_grad():
            for n, p in self._params.items():
                if n.startswith("struc_w"):
                    fan_in = self._layout.spec[n.rsplit(".", 1)[0] + ".weight"][1]
                    bound = 1.0 / math.sqrt(fan_in)
                    p.uniform_(-bound, bound, generator=g)
                elif "LayerNorm" in n:
                    p.fill_(1.0 if n.endswith("weight") else 0.0)
                elif p.dim() == 1:
                    p.zero_()
                else:
                    p.normal_(0.0, self.config.initializer_range, generator=g)

    def tie_weights(self):
        """Already tied by construction (one Parameter, two names)."""

    def set_process_group(self, group, world_size=None):
        """Data parallel without DDP: backward all-reduces the fp32 gradient arena over `group` and
        divides by its size (what DDP's hooks do per bucket, `pretrain.py:166`)."""
        self._group = group
        self._world = world_size if world_size is not None else torch.distributed.get_world_size(group)

    # ---- checkpoints -----------------------------------------------------------------------
    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path, *model_args, **kwargs):
        """`PreTrainedModel.from_pretrained` (`utils.py:884-1078`) for local files: a directory holding
        `pytorch_model.bin` or a path to a `.bin`; keys may carry `module.` (saved from DDP,
        `pretrain.py:1238-1251`) or `bert.` prefixes.  Returns the model in eval mode, or None (after
        logging) when the file does not exist — the reference's convention (`utils.py:977`)."""
        config = kwargs.pop("config", None)
        state_dict = kwargs.pop("state_dict", None)
        kwargs.pop("cache_dir", None)
        kwargs.pop("from_tf", None)
        kwargs.pop("default_gpu", None)
        output_loading_info = kwargs.pop("output_loading_info", False)
        if config is None:
            cfg_path = os.path.join(pretrained_model_name_or_path, "config.json")
            config = BertConfig.from_json_file(cfg_path)
        model = cls(config, *model_args, **kwargs)
        if state_dict is None:
            path = pretrained_model_name_or_path
            if os.path.isdir(path):
                path = os.path.join(path, WEIGHTS_NAME)
            if not os.path.isfile(path):
                logger.error("Model name '%s' was not found: no file %s", pretrained_model_name_or_path, path)
                return None
            state_dict = torch.load(path, map_location="cpu")
        clean = {}
        for k, v in state_dict.items():
            for pre in ("module.", "bert."):
                if k.startswith(pre):
                    k = k[len(pre):]
            clean[k] = v
        res = model.load_state_dict(clean, strict=False)
        if res.missing_keys:
            logger.info("Weights of %s not initialized from pretrained model: %s", cls.__name__, res.missing_keys)
        if res.unexpected_keys:
            logger.info("Weights from pretrained model not used in %s: %s", cls.__name__, res.unexpected_keys)
        model.eval()
        if output_loading_info:
            return model, {"missing_keys": res.missing_keys, "unexpected_keys": res.unexpected_keys,
                           "error_msgs": []}
        return model

    def save_pretrained(self, save_directory):
        """`utils.py:847-882`: config.json + pytorch_model.bin."""
        assert os.path.isdir(save_directory), "Saving path should be a directory where the model and configuration can be saved"
        with open(os.path.join(save_directory, "config.json"), "w", encoding="utf-8") as f:
            f.write(self.config.to_json_string())
        torch.save({k: v.detach().cpu().clone() for k, v in self.state_dict().items()},
                   os.path.join(save_directory, WEIGHTS_NAME))

    # ---- device state ----------------------------------------------------------------------
    def _ensure_engine(self):
        if self._engine is not None:
            return
        dev = self._arena.device
        if self._ops is None:
            if dev.type != "cuda":
                from . import _lib
                raise _lib.K3MError("k3m_b200 runs on a CUDA device (sm_100a) only: move the model with "
                                    ".cuda() first; there is no CPU path")
            from .ops import CudaOps
            self._ops = CudaOps(dev)
        adt = getattr(self._ops, "adt", torch.bfloat16)
        L = self._layout
        self._w16 = torch.zeros(L.total, dtype=adt, device=dev)
        self._g32 = torch.zeros(L.total, dtype=torch.float32, device=dev)
        self._rng = torch.tensor([getattr(self.config, "rng_seed", 1234), 0], dtype=torch.int64, device=dev)
        self._engine = Engine(self.config, L, self._ops, self._w16, self._g32, self._rng, adt)
        self._w16_stamp = None
        self._graphs = {}

    def _param_stamp(self):
        """Changes whenever a parameter (or the arena) was written in place: the version counters torch keeps anyway."""
        return (self._arena._version, self._arena.data_ptr(), sum(p._version for p in self._params.values()))

    def _refresh_shadow(self):
        """bf16 shadow of the fp32 masters, re-cast only after the masters changed (an optimizer step, a checkpoint
        load): eval loops and gradient accumulation reuse it instead of moving 2.8 GB per forward."""
        stamp = self._param_stamp()
        if stamp != self._w16_stamp:
            self._ops.cast(self._arena, self._w16.dtype, out=self._w16)
            self._w16_stamp = stamp

    def _run_backward(self, grads, graphed=None):
        L, ops = self._layout, self._ops
        dev = self._arena.device
        first = self._params[next(n for n in L.spec if n not in L.frozen)]
        if first.grad is None:
            self._g32.zero_()          # fresh accumulation (optimizer.zero_grad(set_to_none=True))
        scales = []
        for g in grads:
            if g is None:
                g = torch.zeros(1, dtype=torch.float32, device=dev)
            scales.append(g.reshape(1).to(torch.float32).contiguous())
        if graphed is not None:
            graphed.backward(scales)
        elif self._kind == "alignment":
            self._engine.backward_alignment(scales[0])
        else:
            self._engine.backward(tuple(scales))
        if (self._group is None and torch.distributed.is_available() and torch.distributed.is_initialized()
                and torch.distributed.get_world_size() > 1 and not getattr(self, "_allow_unsynced", False)):
            # exactly the state a stock `DistributedDataParallel(model)` leaves this model in: its per-parameter
            # hooks never fire (one autograd node for the whole step), so training would silently diverge per rank
            raise RuntimeError("k3m_b200: torch.distributed is initialised with world size > 1 but this model has no "
                               "process group: wrap it with k3m_b200.parallel.DistributedDataParallel(model) (or call "
                               "model.set_process_group(group)); torch's DistributedDataParallel cannot reduce its gradients")
        if self._group is not None and self._world > 1:
            torch.distributed.all_reduce(self._g32[:L.trained_end], group=self._group)
            ops.scale_(self._g32[:L.trained_end], 1.0 / self._world)
        for n, p in self._params.items():
            if n in L.frozen:
                continue
            if p.grad is None:
                off, num, shape = L.offsets[n]
                p.grad = self._g32[off:off + num].view(*shape)


def _as_dev(t, device, dtype=None):
    if t is None:
        return None
    t = t.to(device=device, non_blocking=True)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


class BertForMultiModalPreTraining_tri_stru(K3MPreTrainedModel):
    """`vilbert_k3m/vilbert_k3m.py:2190-2860`: title / image / PV tri-stream ViLBERT with the
    structure aggregator and the MLM + MRM-KL + MLM-pv + LPM (+ ITM value) pretraining losses."""

    def __init__(self, config, ops=None):
        if getattr(config, "visual_target", 0) != 0:
            raise NotImplementedError("visual_target 1 (MSE) / 2 (NCE) are not on the K3M path "
                                      "(the reference's NCE branch reads an undefined attribute, vilbert_k3m.py:2764)")
        super(BertForMultiModalPreTraining_tri_stru, self).__init__(config, ops)

    def forward(
            self,
            input_ids,
            image_feat,
            image_loc,
            token_type_ids=None,
            attention_mask=None,
            image_attention_mask=None,
            masked_lm_labels=None,
            image_label=None,
            image_target=None,
            next_sentence_label=None,
            output_all_attention_masks=False,
            input_ids_pv=None,
            token_type_ids_pv=None,
            attention_mask_pv=None,
            masked_lm_labels_pv=None,
            next_sentence_label_pv_v=None,
            next_sentence_label_pv_t=None,
            index_p=None,
            index_v=None,
            device=None,
    ):
        """Returns the reference's 10-tuple (`:2834-2846`):
        (mlm_t[1], mrm[1], 0, mlm_pv[1], 0, 0, itm[1], c_initial[B,768], c_final[B,768], lpm[1]).
        The label-free branch of the reference (`:2853-2859`) reads undefined names and cannot run
        there either; here it raises."""
        if masked_lm_labels is None or next_sentence_label is None or image_target is None:
            raise NotImplementedError("the K3M pretraining forward needs the labels "
                                      "(the reference's label-free branch is not runnable, vilbert_k3m.py:2853-2859)")
        if input_ids_pv is None or index_p is None or index_v is None:
            raise ValueError("input_ids_pv / index_p / index_v are required (tri-stream model)")
        self._ensure_engine()
        dev = self._arena.device
        i64, f32 = torch.int64, torch.float32
        ones = lambda ref: torch.ones_like(ref)          # noqa: E731  (`bert_tri` defaults, `:2546-2556`)
        zeros = lambda ref: torch.zeros_like(ref)        # noqa: E731
        B = input_ids.shape[0]
        batch = dict(
            input_ids=_as_dev(input_ids, dev, i64),
            token_type_ids=_as_dev(token_type_ids if token_type_ids is not None else zeros(input_ids), dev, i64),
            attention_mask=_as_dev(attention_mask if attention_mask is not None else ones(input_ids), dev, i64),
            masked_lm_labels=_as_dev(masked_lm_labels, dev, i64),
            input_ids_pv=_as_dev(input_ids_pv, dev, i64),
            token_type_ids_pv=_as_dev(token_type_ids_pv if token_type_ids_pv is not None else zeros(input_ids_pv), dev, i64),
            attention_mask_pv=_as_dev(attention_mask_pv if attention_mask_pv is not None else ones(input_ids_pv), dev, i64),
            masked_lm_labels_pv=_as_dev(masked_lm_labels_pv, dev, i64),
            image_feat=_as_dev(image_feat, dev, f32),
            image_loc=_as_dev(image_loc, dev, f32),
            image_attention_mask=_as_dev(image_attention_mask if image_attention_mask is not None
                                         else torch.ones(image_feat.shape[:2], dtype=i64), dev, i64),
            image_label=_as_dev(image_label, dev, i64),
            image_target=_as_dev(image_target, dev, f32),
            next_sentence_label=_as_dev(next_sentence_label.reshape(B), dev, i64),
            next_sentence_label_pv_v=_as_dev((next_sentence_label_pv_v if next_sentence_label_pv_v is not None
                                              else zeros(next_sentence_label)).reshape(B), dev, i64),
            next_sentence_label_pv_t=_as_dev((next_sentence_label_pv_t if next_sentence_label_pv_t is not None
                                              else zeros(next_sentence_label)).reshape(B), dev, i64),
        )
        # the aggregator's pair lists and LPM negatives are drawn on the host with the reference's own
        # `random.sample` call sequence (`:2437-2502`) — one device->host copy of two small index
        # tensors replaces the reference's one sync per (item, PV) pair
        plan_np = build_struct_plan(index_p.detach().cpu().numpy(), index_v.detach().cpu().numpy(),
                                    batch["input_ids_pv"].shape[1], self.config.num_negative_pv)
        # rows that carry a masked-LM label: the head runs on those only (`Engine.forward`)
        labels_np = (masked_lm_labels.detach().cpu().numpy(), masked_lm_labels_pv.detach().cpu().numpy())
        plan_np.update(build_mlm_plan(*labels_np))
        self._refresh_shadow()
        eng = self._engine
        eng.training = self.training
        self._fwd_token += 1
        # training steps of a geometry seen before replay two CUDA graphs (K3M_CLASS_GRAPH=0: eager launches)
        graphed = None
        if (torch.is_grad_enabled() and dev.type == "cuda" and hasattr(self._ops, "lib")
                and os.environ.get("K3M_CLASS_GRAPH", "1") != "0"):
            key = (tuple(input_ids.shape), tuple(image_feat.shape), tuple(input_ids_pv.shape), tuple(index_p.shape),
                   bool(self.training))
            ent = self._graphs.get(key)
            if ent is None:
                self._graphs[key] = 0                    # first sight: eager (also the warm-up of every kernel variant)
            else:
                if ent == 0:
                    ent = self._graphs[key] = _GraphedStep(self, batch, index_p.shape[1])
                    ent.load(batch, plan_np, labels_np)
                    ent.capture()
                else:
                    ent.load(batch, plan_np, labels_np)
                graphed = ent
        plan = None if graphed is not None else self._ops.plan_to_device(plan_np, dev)
        if torch.is_grad_enabled():
            lt, lv, lpv, llpm, litm, c_init, c_fin = _StepFunction.apply(self._anchor, self, batch, plan, graphed)
        else:
            out = eng.forward(batch, plan, None, (1.0, 1.0, 1.0, 1.0))
            eng.saved = None
            lt, lv, lpv, llpm, litm = (out[k] for k in ("loss_t", "loss_v", "loss_pv", "loss_lpm", "loss_itm"))
            c_init, c_fin = out["c_initial"], out["c_final"]
        self._rng[1:].add_(1)
        return (lt.reshape(1), lv.reshape(1), 0, lpv.reshape(1), 0, 0, litm.reshape(1), c_init, c_fin,
                llpm.reshape(1))


class _PairFunction(torch.autograd.Function):
    """Autograd node of the item-alignment step: the pair loss is differentiable, the rest is value-only."""

    @staticmethod
    def forward(ctx, anchor, owner, batch, plan, labels):
        out = owner._engine.forward_alignment(batch, plan, labels, None)
        ctx.owner = owner
        ctx.token = owner._fwd_token
        res = (out["loss"], out["probs"], out["c_final"])
        ctx.mark_non_differentiable(res[1], res[2])
        return res

    @staticmethod
    def backward(ctx, g_loss, *unused):
        owner = ctx.owner
        if ctx.token != owner._fwd_token:
            raise RuntimeError("backward() of an alignment step whose activations were released by a later forward()")
        owner._run_backward((g_loss,))
        return None, None, None, None, None


class K3MForItemAlignment(K3MPreTrainedModel):
    """`vilbert_k3m/vilbert_k3m.py:2862-3454`: two item towers with shared weights -> `c_final` of each
    item -> ClassificationHead on their concatenation -> cross-entropy (`loss_type == "ce"`, the mode of
    `run_finetune_item_alignment.sh`).  Both towers run as ONE stacked pass of 2B items."""

    _kind = "alignment"

    def __init__(self, config, ops=None):
        if getattr(config, "loss_type", "ce") != "ce":
            raise NotImplementedError("loss_type %r: only the \"ce\" head of K3MForItemAlignment is built "
                                      "(\"cosine\" compares item 1 with itself in the reference, vilbert_k3m.py:3442; "
                                      "\"inner\" has no loss_fct, :2925-2931)" % (config.loss_type,))
        super(K3MForItemAlignment, self).__init__(config, ops)

    def forward(
            self,
            labels,
            # item 1
            input_ids_1, token_type_ids_1, attention_mask_1, input_ids_pv_1, token_type_ids_pv_1,
            attention_mask_pv_1, index_p_1, index_v_1, image_feat_1, image_loc_1, image_attention_mask_1,
            # item 2
            input_ids_2, token_type_ids_2, attention_mask_2, input_ids_pv_2, token_type_ids_pv_2,
            attention_mask_pv_2, index_p_2, index_v_2, image_feat_2, image_loc_2, image_attention_mask_2,
            output_all_attention_masks=False,
    ):
        """Same 23 positional arguments and return as the reference (`:3379-3454`):
        (probs[:, 0], probs[:, 1], probs[:, 1], loss) — in "ce" mode the two "embeddings" slots are
        overwritten with the class probabilities (`:3446-3450`)."""
        self._ensure_engine()
        dev = self._arena.device
        i64, f32 = torch.int64, torch.float32

        def both(a, b, dtype):
            return torch.cat((_as_dev(a, dev, dtype), _as_dev(b, dev, dtype)), dim=0)

        batch = dict(
            input_ids=both(input_ids_1, input_ids_2, i64),
            token_type_ids=both(token_type_ids_1, token_type_ids_2, i64),
            attention_mask=both(attention_mask_1, attention_mask_2, i64),
            input_ids_pv=both(input_ids_pv_1, input_ids_pv_2, i64),
            token_type_ids_pv=both(token_type_ids_pv_1, token_type_ids_pv_2, i64),
            attention_mask_pv=both(attention_mask_pv_1, attention_mask_pv_2, i64),
            image_feat=both(image_feat_1, image_feat_2, f32),
            image_loc=both(image_loc_1, image_loc_2, f32),
            image_attention_mask=both(image_attention_mask_1, image_attention_mask_2, i64),
        )
        ip = torch.cat((index_p_1.detach().cpu(), index_p_2.detach().cpu()), dim=0).numpy()
        iv = torch.cat((index_v_1.detach().cpu(), index_v_2.detach().cpu()), dim=0).numpy()
        plan = build_struct_plan(ip, iv, batch["input_ids_pv"].shape[1], self.config.num_negative_pv, halves=2)
        plan = self._ops.plan_to_device(plan, dev)
        lab = _as_dev(labels.reshape(-1), dev, i64)
        self._refresh_shadow()
        eng = self._engine
        eng.training = self.training
        self._fwd_token += 1
        if torch.is_grad_enabled():
            loss, probs, _ = _PairFunction.apply(self._anchor, self, batch, plan, lab)
        else:
            out = eng.forward_alignment(batch, plan, lab, None)
            eng.saved = None
            loss, probs = out["loss"], out["probs"]
        self._rng[1:].add_(1)
        return probs[:, 0], probs[:, 1], probs[:, 1], loss.reshape(())


def install_as_reference():
    """Make `from vilbert_k3m.vilbert_k3m import BertConfig, BertForMultiModalPreTraining_tri_stru`
    (`pretrain.py:25`, `train_concap_struc.py:26`) resolve to this package."""
    import types
    pkg = sys.modules.get("vilbert_k3m")
    if pkg is None:
        pkg = types.ModuleType("vilbert_k3m")
        pkg.__path__ = []
        sys.modules["vilbert_k3m"] = pkg
    mod = sys.modules[__name__]
    sys.modules["vilbert_k3m.vilbert_k3m"] = mod
    pkg.vilbert_k3m = mod
    return mod
