"""The K3M pretraining step as one object: forward + backward + gradient all-reduce + AdamW.

This is the part of the reference's driver loop between `optimizer.zero_grad()` and
`scheduler.step()` (`pretrain.py:986-1110`): `model(...)` -> `loss = out[0] + out[1]*w + out[3] +
out[9]` (`:1021,1055`) -> `loss.backward()` (DDP all-reduce inside, `:166`) -> `optimizer.step()`
(`torch.optim.AdamW`, `:919-924`) -> `scheduler.step()` (`WarmupLinearSchedule`, `:927-931`).

Memory (all on one B200, replicated per rank; 461.7 M parameters -> 8.3 GB):
  p32 / m / v   fp32 master weights and AdamW moments        [total]
  w16           bf16 shadow of p32 (what the kernels read)   [total]
  g32           fp32 gradient arena (wgrad GEMMs accumulate straight into it)
arranged by `ArenaLayout`: [decay | no-decay | never-trained], each in forward execution order, so
AdamW is two launches and gradient buckets are contiguous slices that become final from the end of
the arena backwards while backward is still running.

Static shapes + no host synchronisation inside -> the whole step is one CUDA graph.  Per step the
host only (1) builds the structure-aggregator index plan from `index_p/index_v` (host integers
that the loader produced on the host anyway), (2) copies the batch host->device into the graph's
static input buffers, (3) writes the learning rate, (4) replays the graph.
"""
import math
import os

import numpy as np
import torch

from . import optim
from .engine import Engine
from .layout import ArenaLayout
from .ops import CudaOps
from .struct_plan import build_mlm_plan, build_struct_plan, mlm_capacity, pad_struct_plan, plan_capacity

LOSS_NAMES = ("loss_t", "loss_v", "loss_pv", "loss_itm", "loss_lpm")


def input_spec(cfg, B, Lt, Lv1, Lpv, max_num_pv):
    """name -> (shape, dtype) of the tensors the reference's loader yields
    (`concept_cap_dataset_struc.py:372-413`) and `forward` takes (`vilbert_k3m.py:2673-2696`)."""
    i64, f32 = torch.int64, torch.float32
    return dict(
        input_ids=((B, Lt), i64), token_type_ids=((B, Lt), i64), attention_mask=((B, Lt), i64),
        masked_lm_labels=((B, Lt), i64),
        input_ids_pv=((B, Lpv), i64), token_type_ids_pv=((B, Lpv), i64), attention_mask_pv=((B, Lpv), i64),
        masked_lm_labels_pv=((B, Lpv), i64),
        image_feat=((B, Lv1, cfg.v_feature_size), f32), image_loc=((B, Lv1, 5), f32),
        image_attention_mask=((B, Lv1), i64), image_label=((B, Lv1 - 1), i64),
        image_target=((B, Lv1 - 1, cfg.v_target_size), f32),
        next_sentence_label=((B,), i64), next_sentence_label_pv_v=((B,), i64),
        next_sentence_label_pv_t=((B,), i64),
    )


def _flat_views(spec, device):
    """spec: ordered (name, shape, dtype) -> (flat uint8 buffer, {name: view}); every view starts 256-byte aligned, so
    two buffers built from the same spec have the same layout and copy into each other as ONE memcpy."""
    offs, off = [], 0
    for _, shape, dtype in spec:
        n = 1
        for d_ in shape:
            n *= int(d_)
        offs.append((off, n * torch.empty((), dtype=dtype).element_size()))
        off += (offs[-1][1] + 255) // 256 * 256
    flat = torch.zeros(max(off, 256), dtype=torch.uint8, device=device)
    return flat, {name: flat[o:o + nb].view(dtype).view(shape) for (name, shape, dtype), (o, nb) in zip(spec, offs)}


class PretrainStep(object):
    KIND = "pretrain"       # parameter inventory / engine entry points (AlignmentStep: "alignment")
    HALVES = 1              # item groups that get their own `structure_aggregator` call

    def __init__(self, cfg, batch_size, len_t=36, len_v=36, len_pv=36, max_num_pv=8, device="cuda",
                 state_dict=None, init_seed=0, lr=1e-4, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01,
                 warmup_steps=0, total_steps=0, loss_img_weight=1.0, training=True, rng_seed=1234,
                 use_graph=True, process_group=None, bucket_mb=None, overlap=True, ops=None, comm_bf16=None,
                 dp_fused=None, mlm_cap_frac=(0.25, 0.4)):
        """`ops`: kernel table; default (and the only product path) is `CudaOps`, which raises without
        a CUDA device.  Tests inject the fp32 PyTorch restatement to check this host logic on CPU."""
        self.cfg, self.device = cfg, torch.device(device)
        self.B, self.Lt, self.Lv1, self.Lpv, self.max_num_pv = batch_size, len_t, len_v + 1, len_pv, max_num_pv
        self.ops = ops if ops is not None else CudaOps(self.device)
        self._cuda = self.device.type == "cuda"
        adt = getattr(self.ops, "adt", torch.bfloat16)
        self.layout = L = ArenaLayout(cfg, self.KIND)
        self.n_items = batch_size * self.HALVES
        dev, f32 = self.device, torch.float32
        self.p32 = torch.zeros(L.total, dtype=f32, device=dev)
        self.m = torch.zeros(L.total, dtype=f32, device=dev)
        self.v = torch.zeros(L.total, dtype=f32, device=dev)
        self.g32 = torch.zeros(L.total, dtype=f32, device=dev)
        self.group = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self.rank = torch.distributed.get_rank(process_group) if process_group is not None else 0
        self.w16 = None
        self.dp_fused = False
        if dp_fused is None:
            dp_fused = os.environ.get("K3M_DP_FUSED", "1") != "0"
        if self.world > 1 and self._cuda and adt == torch.bfloat16 and dp_fused:
            self._setup_peer_memory(L, dev)
        if self.w16 is None:
            self.w16 = torch.zeros(L.total, dtype=adt, device=dev)
        self.rng = torch.tensor([rng_seed, 0], dtype=torch.int64, device=dev)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.warmup_steps, self.total_steps = warmup_steps, total_steps
        self.loss_weights = (1.0, float(loss_img_weight), 1.0, 1.0)     # `pretrain.py:1055`
        self.global_step = 0
        self.overlap = overlap
        self.engine = Engine(cfg, L, self.ops, self.w16, self.g32, self.rng, adt)
        self.engine.training = training
        if state_dict is not None:
            self.load_state_dict(state_dict)
        else:
            self.init_weights(init_seed)
        # static graph inputs
        # Inputs, index plan and hyper-parameters are views into ONE flat buffer (`_flat_views`): a step's staged
        # copy reaches the graph's static buffers as one device-to-device memcpy instead of ~25 small ones.
        self.spec = self._input_spec()
        P_cap, T_cap, maxn = plan_capacity(self.n_items, max_num_pv, cfg.num_negative_pv)
        i32 = torch.int32
        plan_spec = [("pair_item", (P_cap,), i32), ("pair_rows", (P_cap, 4), i32), ("seg", (self.n_items, 2), i32),
                     ("terms", (T_cap, 4), i32), ("n_terms", (1,), i32), ("fallback", (self.HALVES, 2), i32)]
        # masked-LM head on the labelled rows only: static row / label lists of a fixed capacity (`build_mlm_plan`)
        self.mlm_cap = None
        if self.KIND == "pretrain" and mlm_cap_frac is not None:
            self.mlm_cap = mlm_capacity(batch_size * len_t, batch_size * len_pv, mlm_cap_frac)
            n = self.mlm_cap[0] + self.mlm_cap[1]
            plan_spec += [("mlm_rows", (n,), i32), ("mlm_labels", (n,), torch.int64)]
        self._flat_spec = ([("in:" + k, tuple(s_), d) for k, (s_, d) in self.spec.items()] +
                           [("plan:" + k, s_, d) for k, s_, d in plan_spec] + [("hyper", (2, 8), f32)])
        self._static_flat, views = _flat_views(self._flat_spec, dev)
        self.inputs = {k[3:]: v for k, v in views.items() if k.startswith("in:")}
        self.plan = {k[5:]: v for k, v in views.items() if k.startswith("plan:")}
        for k in ("fallback", "mlm_rows", "mlm_labels"):
            if k in self.plan:
                self.plan[k].fill_(-1)
        self.plan["seg"]._k3m_maxn = maxn
        pin = (lambda t: t.pin_memory()) if self._cuda else (lambda t: t)
        self._plan_host = [{k: pin(torch.zeros(v.shape, dtype=v.dtype)) for k, v in self.plan.items()}
                           for _ in range(2)]
        if self.mlm_cap is not None:
            self.plan["mlm_cap_t"] = self.mlm_cap[0]
        # AdamW hyper-parameters live on the device (graph nodes stay valid); ring of pinned copies
        self.hyper = views["hyper"]
        self._hyper_host = [pin(torch.zeros(2, 8, dtype=f32)) for _ in range(2)]
        self._slot_events = [None, None]
        self._stage, self._copy_stream = None, None
        self.losses = torch.zeros(8, dtype=f32, device=dev)
        self.losses_host = pin(torch.zeros(8, dtype=f32))
        # ring of `read_losses_lagged`: pinned here rather than on its first call (a page-locked allocation is a
        # device-wide synchronisation point and does not belong inside a running training loop; measured: it was NOT
        # what separates the end-to-end from the device-resident step time, profiles/r03end2_bench.json)
        self._loss_ring = [pin(torch.zeros(8, dtype=f32)) for _ in range(2)]
        self._loss_events = [None, None]
        self._loss_reads = 0
        self.h2d_bytes = sum(int(np.prod(s)) * (8 if d == torch.int64 else 4) for s, d in self.spec.values())
        self.h2d_bytes += sum(v.numel() * v.element_size() for v in self.plan.values() if torch.is_tensor(v)) + self.hyper.numel() * 4
        self.d2h_bytes = self.losses.numel() * 4
        # data-parallel runs reduce in 256 MB pieces on a high-priority stream: measured at 8 GPUs (NVLS
        # all-reduce of the 1.85 GB fp32 gradient arena overlapping backward) 64 MB -> 25.3 ms / step,
        # 256 MB -> 24.5, 256 MB + priority -> 24.3, 512 MB -> 24.9 (profiles/r01e_n8_sweep.txt); one GPU keeps
        # 64 MB pieces, which only pace the AdamW launches that overlap backward
        if bucket_mb is None:
            bucket_mb = 256 if self.world > 1 else 64
        self._buckets = self._make_buckets(float(os.environ.get("K3M_BUCKET_MB", bucket_mb)))
        # side stream of the optimizer pipeline: all-reduce of a gradient bucket (world > 1) followed by
        # AdamW on the same range, while backward is still producing the earlier layers' gradients
        self.overlap_opt = overlap and os.environ.get("K3M_OPT_OVERLAP", "1") != "0"
        prio = -1 if os.environ.get("K3M_COMM_PRIO", "1" if self.world > 1 else "0") != "0" else 0
        self._comm_stream = (torch.cuda.Stream(device=dev, priority=prio)
                             if (self._cuda and (self.world > 1 or self.overlap_opt)) else None)
        # data parallel: gradients cross NVLink in bf16 (0.92 GB per step instead of 1.85 GB) — each
        # finished bucket is cast into a persistent bf16 send buffer on the optimizer stream, all-reduced, and AdamW
        # reads the reduced bf16 gradient (k3m_adamw_g16).  K3M_COMM_BF16=0 reduces the fp32 arena itself.
        if comm_bf16 is None:
            comm_bf16 = os.environ.get("K3M_COMM_BF16", "1") != "0"
        self.comm_bf16 = self.world > 1 and (bool(comm_bf16) or self.dp_fused)
        self._send16 = (torch.zeros(L.trained_end, dtype=torch.bfloat16, device=dev)
                        if (self.comm_bf16 and not self.dp_fused) else None)
        # ... and the persistent GEMM grids leave a few SMs to the collective's CTAs (a GEMM CTA owns its SM: registers
        # and shared memory), which otherwise wait for a GEMM to retire and then make the next one's CTAs wait
        self.reserve_sms = int(os.environ.get("K3M_COMM_SMS", "0")) if self.world > 1 else 0
        if self.reserve_sms > 0 and hasattr(self.ops, "gemm_set_sm_limit"):
            self.ops.gemm_set_sm_limit(148 - self.reserve_sms)
        self.comm_events = None       # comm_timing(): [(start event, end event, bytes)] of one instrumented step
        self.use_graph = use_graph and self._cuda
        self.graph = None
        self.launches_per_step = 0

    def _input_spec(self):
        return input_spec(self.cfg, self.B, self.Lt, self.Lv1, self.Lpv, self.max_num_pv)

    # ---- data parallel over NVLink peer memory ---------------------------------------------------------
    def _setup_peer_memory(self, L, dev):
        """Symmetric (peer-mapped) allocations for the fused data-parallel optimizer (`k3m_adamw_dp`): the bf16 weight
        arena every rank's kernels read, and a bf16 gradient send arena.  torch's symmetric-memory allocator does the
        CUDA IPC / fabric handle exchange; what runs over the mapping is this repo's kernel.  Any failure (no peer
        access, old driver) falls back to the NCCL all-reduce path, loudly."""
        try:
            import torch.distributed._symmetric_memory as symm
            w16 = symm.empty(L.total, dtype=torch.bfloat16, device=dev)
            g16 = symm.empty(L.trained_end, dtype=torch.bfloat16, device=dev)
            w16.zero_()
            g16.zero_()
            self._h_w = symm.rendezvous(w16, self.group)
            self._h_g = symm.rendezvous(g16, self.group)
            self._w_ptrs = [int(x) for x in self._h_w.buffer_ptrs]
            self._g_ptrs = [int(x) for x in self._h_g.buffer_ptrs]
            assert len(self._w_ptrs) == self.world and self._w_ptrs[self.rank] == w16.data_ptr()
            # multicast (NVLS) mapping of the same allocations, when the fabric has one: in-switch reduction
            self._g_mc = self._w_mc = 0
            # measured (profiles/r02t_n2_multimem.txt, r02u_n8_multimem.txt): N = 2 19.18 ms with multimem vs 18.58 with
            # peer loads (no bytes saved at two ranks, 16-byte switch requests); N = 8 18.63 vs 19.21 ms -> from 4 ranks on
            if os.environ.get("K3M_DP_MULTIMEM", "1" if self.world >= 4 else "0") != "0":
                try:
                    g_mc, w_mc = int(self._h_g.multicast_ptr or 0), int(self._h_w.multicast_ptr or 0)
                    if g_mc and w_mc:
                        self._g_mc, self._w_mc = g_mc, w_mc
                except Exception:
                    pass
            self.w16, self._g16 = w16, g16
            self.dp_fused = True
        except Exception as e:      # pragma: no cover - depends on the box
            import sys
            sys.stderr.write("[k3m_b200] peer-memory data-parallel path unavailable (%s: %s); using NCCL all-reduce\n"
                             % (type(e).__name__, e))
            self.dp_fused = False

    def _shard(self, lo, hi):
        """this rank's share [s0, s1) of the bucket [lo, hi): equal 1024-element-aligned pieces"""
        n = hi - lo
        piece = ((n + self.world - 1) // self.world + 1023) // 1024 * 1024
        s0 = min(hi, lo + self.rank * piece)
        return s0, min(hi, s0 + piece)

    def _peer_barrier(self, handle, channel):
        handle.barrier(channel=channel, timeout_ms=int(os.environ.get("K3M_DP_BARRIER_TIMEOUT_MS", "30000")))

    def sync_shards(self, moments=False):
        """Fused data-parallel mode keeps the fp32 master weights (and AdamW moments) of a shard only on the rank that
        owns it.  Collective: broadcast every shard from its owner so that `state_dict()` / a checkpoint written by
        any rank is complete.  No-op otherwise."""
        if not self.dp_fused or self.world == 1:
            return
        rank = self.rank
        try:
            for lo, hi in self._buckets:
                for r in range(self.world):
                    self.rank = r
                    s0, s1 = self._shard(lo, hi)
                    if s1 > s0:
                        src = torch.distributed.get_global_rank(self.group, r)
                        for t in (self.p32,) + ((self.m, self.v) if moments else ()):
                            torch.distributed.broadcast(t[s0:s1], src=src, group=self.group)
        finally:
            self.rank = rank

    def _forward_backward(self):
        """-> the eight fp32 values of `self.losses` as a list of [1] tensors / None."""
        out = self.engine.forward(self.inputs, self.plan, None, self.loss_weights)
        self.engine.on_ready = self._on_ready
        self.engine.backward()
        w = self.loss_weights
        total = out["loss_t"] * w[0] + out["loss_v"] * w[1] + out["loss_pv"] * w[2] + out["loss_lpm"] * w[3]
        self.c_initial, self.c_final = out["c_initial"], out["c_final"]
        return [out[k] for k in LOSS_NAMES] + [total]

    # ---- parameters ------------------------------------------------------------------------
    def init_weights(self, seed):
        """`init_weights` (`vilbert_k3m.py:1940-1951`): N(0, initializer_range) for Linear and
        Embedding weights, zero biases, LayerNorm = (1, 0); `struc_w1/2/3` are created after
        `self.apply(self.init_weights)` (`:2257-2259`) and keep nn.Linear's default
        U(-1/sqrt(in), 1/sqrt(in)) init for weight and bias."""
        L = self.layout
        g = torch.Generator(device=self.device).manual_seed(seed)
        for n, (off, num, shape) in L.offsets.items():
            t = self.p32[off:off + num]
            if n.startswith("struc_w"):
                bound = 1.0 / math.sqrt(L.spec[n.rsplit(".", 1)[0] + ".weight"][1])
                t.uniform_(-bound, bound, generator=g)
            elif "LayerNorm" in n:
                t.fill_(1.0 if n.endswith("weight") else 0.0)
            elif len(shape) == 1:
                t.zero_()
            else:
                t.normal_(0.0, self.cfg.initializer_range, generator=g)
        self.w16.copy_(self.p32)

    def load_state_dict(self, sd):
        """Reference-layout state_dict (999 keys; the tied `cls.predictions.decoder.weight` alias
        and a leading `module.` are accepted, `pretrain.py:745-762`)."""
        L = self.layout
        missing = []
        for n, (off, num, shape) in L.offsets.items():
            t = sd.get(n, sd.get("module." + n))
            if t is None:
                missing.append(n)
                continue
            if tuple(t.shape) != tuple(shape):
                raise ValueError("%s: shape %s, expected %s" % (n, tuple(t.shape), tuple(shape)))
            self.p32[off:off + num].copy_(t.reshape(-1).to(torch.float32))
        if missing:
            raise KeyError("state_dict is missing %d parameters, e.g. %s" % (len(missing), missing[:3]))
        self.w16.copy_(self.p32)

    def state_dict(self):
        """(collective in fused data-parallel mode: every rank must call it, see `sync_shards`)"""
        from .layout import TIED_ALIAS
        self.sync_shards()
        out = {}
        for n in self.layout.spec:
            off, num, shape = self.layout.offsets[n]
            out[n] = self.p32[off:off + num].view(*shape)
        if self.KIND == "pretrain":
            out[TIED_ALIAS[0]] = out[TIED_ALIAS[1]]
        return out

    def grad(self, name):
        off, num, shape = self.layout.offsets[name]
        return self.g32[off:off + num].view(*shape)

    # ---- gradient buckets (data parallel) ----------------------------------------------------
    def _make_buckets(self, bucket_mb):
        """Contiguous fp32 slices of the trained region, listed in the order they become final in
        backward: the decay region from its END towards the start in `bucket_mb` pieces, then the
        (small) no-decay region as one last bucket."""
        L = self.layout
        step = max(1, int(bucket_mb * (1 << 20) // 4))
        if self.dp_fused:
            step = max(1024, step // 1024 * 1024)      # shards of the fused optimizer are 16-byte-vector aligned
        # the embedding tables open the arena and get their gradients last (their backward closes the step):
        # they are a bucket of their own, so that only THEIR all-reduce + AdamW is left when backward ends
        # instead of whatever piece of up to `bucket_mb` happened to straddle them
        emb_end = 0
        for n, (off, num, _) in L.offsets.items():
            if off < L.decay_end and n.split(".")[0] in ("embeddings", "v_embeddings"):
                emb_end = max(emb_end, off + num)
        if any(off < emb_end and n.split(".")[0] not in ("embeddings", "v_embeddings")
               for n, (off, num, _) in L.offsets.items()):
            emb_end = 0                                    # not a clean prefix: fall back to uniform pieces
        emb_end = (emb_end + 63) // 64 * 64                # the next tensor starts 64-aligned: the gap holds zeros
        out, hi = [], L.decay_end
        while hi > emb_end:
            lo = max(emb_end, hi - step)
            out.append((lo, hi))
            hi = lo
        if emb_end > 0:
            out.append((0, emb_end))
        out.append((L.decay_end, L.trained_end))
        # stage prefix -> lowest decay-region offset of its parameters
        self._stage_lo = {}
        for n, (off, num, _) in L.offsets.items():
            if off >= L.decay_end:
                continue
            parts = n.split(".")
            key = ".".join(parts[:3]) if parts[0] == "encoder" else "<tail>"
            if parts[0] in ("embeddings", "v_embeddings"):
                key = "<embed>"
            self._stage_lo[key] = min(self._stage_lo.get(key, L.decay_end), off)
        return out

    def _bucket_step(self, bi, background=False):
        """All-reduce (world > 1) then AdamW of bucket `bi`, on the side stream when there is one.  A
        bucket's parameters are no longer read by backward once its gradients are final (`_on_ready`), so
        updating them under the rest of backward is safe; AdamW is HBM-bound and the GEMMs it runs
        beside are not."""
        lo, hi = self._buckets[bi]
        gi = 1 if bi == len(self._buckets) - 1 else 0       # the last bucket is the no-decay region
        t = self.g32[lo:hi]

        def work():
            g16 = None
            if self.world > 1 and self.dp_fused:
                ev = None
                if self.comm_events is not None:
                    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    ev[0].record()
                # gradients -> this rank's bf16 send arena (zeroing the fp32 arena); every rank's cast must be visible
                # before anyone reads it; then reduce + AdamW + weight broadcast for the shard this rank owns
                self.ops.cast_zero(t, self._g16[lo:hi])
                self._peer_barrier(self._h_g, 0)
                s0, s1 = self._shard(lo, hi)
                if s1 > s0:
                    self.ops.adamw_dp(self.p32[s0:s1], self.m[s0:s1], self.v[s0:s1], self.hyper[gi], self._g_ptrs,
                                      self._w_ptrs, s0, self._g_mc, self._w_mc)
                if ev is not None:
                    ev[1].record()
                    self.comm_events.append((ev[0], ev[1], (hi - lo) * 2))
                return
            if self.world > 1:
                ev = None
                if self.comm_events is not None:
                    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    ev[0].record()
                if self.comm_bf16:
                    g16 = self._send16[lo:hi]
                    self.ops.cast(t, torch.bfloat16, out=g16)
                    torch.distributed.all_reduce(g16, group=self.group)
                else:
                    torch.distributed.all_reduce(t, group=self.group)
                if ev is not None:
                    ev[1].record()
                    self.comm_events.append((ev[0], ev[1], (hi - lo) * (2 if self.comm_bf16 else 4)))
            # gradients are summed over ranks; hyper[:, 6] = 1/world turns that into DDP's mean
            self.ops.adamw(self.p32[lo:hi], t, self.m[lo:hi], self.v[lo:hi], self.w16[lo:hi], self.hyper[gi],
                           zero_grad=True, background=background, g16=g16)

        if self._comm_stream is None:
            work()
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ev)
            work()

    def _on_ready(self, prefixes):
        """Engine callback: backward has finished every weight gradient of `prefixes` (None = the
        heads, fusion gates and aggregator).  Launch every bucket that is now final."""
        if not self.overlap or (self.world == 1 and not self.overlap_opt):
            return
        keys = ["<tail>"] if prefixes is None else prefixes
        for k in keys:
            self._ready_lo = min(self._ready_lo, self._stage_lo.get(k, self._ready_lo))
        while self._next_bucket < len(self._buckets) - 1 and self._buckets[self._next_bucket][0] >= self._ready_lo:
            # backward is still running: the update shares the SMs with its GEMMs (k3m_adamw_bg)
            self._bucket_step(self._next_bucket, background=True)
            self._next_bucket += 1

    def _finish_buckets(self):
        if self.comm_events is not None:
            self._bwd_done = torch.cuda.Event(enable_timing=True)
            self._bwd_done.record()
        while self._next_bucket < len(self._buckets):
            self._bucket_step(self._next_bucket)
            self._next_bucket += 1
        if self.world > 1 and self.dp_fused:
            # every rank's weight stores have landed everywhere before anybody's next forward reads them
            with torch.cuda.stream(self._comm_stream):
                self._peer_barrier(self._h_w, 1)
        if self._comm_stream is not None:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
        if self.comm_events is not None:
            self._opt_done = torch.cuda.Event(enable_timing=True)
            self._opt_done.record()

    def comm_timing(self, batch):
        """One instrumented, non-captured step (every rank must call it): CUDA events around each bucket's gradient
        all-reduce on the optimizer stream, and around the part of the optimizer pipeline that is left when backward
        has finished.  The device is kept behind the host so that the events bracket back-to-back kernels.
        -> dict(comm_ms_total, tail_ms, buckets, bytes) — tail_ms is the exposed part: all-reduce + AdamW still to
        run after the last backward kernel."""
        snap = (self.p32.clone(), self.m.clone(), self.v.clone(), self.w16.clone(), self.rng.clone())
        use_graph, self.use_graph = self.use_graph, False
        self.comm_events = []
        torch.cuda.synchronize()
        torch.cuda._sleep(int(8.0e8))
        self.step(batch)
        torch.cuda.synchronize()
        ev, self.comm_events = self.comm_events, None
        self.use_graph = use_graph
        self.global_step -= 1
        for dst, src in zip((self.p32, self.m, self.v, self.w16, self.rng), snap):
            dst.copy_(src)
        self.g32.zero_()
        torch.cuda.synchronize()
        return dict(comm_ms_total=round(sum(a.elapsed_time(b) for a, b, _ in ev), 3),
                    tail_ms=round(self._bwd_done.elapsed_time(self._opt_done), 3), buckets=len(ev),
                    bytes=int(sum(n for _, _, n in ev)), dtype="bf16" if self.comm_bf16 else "f32",
                    mode=("fused peer-memory reduce-scatter + AdamW + weight all-gather (k3m_adamw_dp, %s)"
                          % ("multimem / in-switch reduction" if getattr(self, "_g_mc", 0) else "peer loads / stores")) if self.dp_fused
                    else "NCCL all-reduce + replicated AdamW",
                    reserve_sms=self.reserve_sms,
                    note="one non-captured step; comm_ms_total = sum over buckets of (cast +) all-reduce time on the "
                         "optimizer stream (they run under backward); tail_ms = optimizer pipeline left after backward's "
                         "last kernel (exposed)")

    # ---- the device side of one step ---------------------------------------------------------
    def _device_step(self):
        L, ops = self.layout, self.ops
        self._ready_lo, self._next_bucket = L.decay_end, 0
        vals = self._forward_backward()
        self._finish_buckets()
        self.rng[1:].add_(1)
        torch.stack([v.reshape(()) for v in vals], out=self.losses[:len(vals)])

    def _capture(self):
        # warm-up on a side stream (lazy module loading, cudaFuncSetAttribute, NCCL setup), then capture
        s = torch.cuda.Stream(device=self.device, priority=getattr(self.ops, "hi_prio", 0))
        s.wait_stream(torch.cuda.current_stream())
        snap = (self.p32.clone(), self.m.clone(), self.v.clone(), self.w16.clone(), self.rng.clone())
        with torch.cuda.stream(s):
            self._device_step()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        for dst, src in zip((self.p32, self.m, self.v, self.w16, self.rng), snap):
            dst.copy_(src)
        self.g32.zero_()
        del snap
        self.graph = torch.cuda.CUDAGraph()
        n0 = self.ops.launches
        # No garbage collection while the capture is open: a collected object whose finaliser frees CUDA resources (an
        # older step's graph and its memory pool, pinned buffers) makes a prohibited call from inside a global-mode
        # capture and invalidates it (seen as a flaky cudaErrorStreamCaptureInvalidated when several step objects
        # are created in one process).  torch.cuda.graph collects once on entry; nothing may be collected after that.
        import gc
        gc.collect()
        was_enabled = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(self.graph, stream=s):
                self._device_step()
        finally:
            if was_enabled:
                gc.enable()
        self.launches_per_step = self.ops.launches - n0

    # ---- host side -----------------------------------------------------------------------------
    def _fill_hyper(self, slot):
        step = self.global_step + 1
        f = 1.0
        if self.total_steps > 0:
            f = optim.warmup_linear_factor(self.global_step, self.warmup_steps, self.total_steps)
        h = self._hyper_host[slot]
        for gi, wd in enumerate((self.weight_decay, 0.0)):
            h[gi] = torch.tensor([self.lr * f, self.betas[0], self.betas[1], self.eps, wd, float(step),
                                  1.0 / self.world, 0.0])

    def step(self, batch):
        """batch: dict of HOST tensors named as the reference's forward() arguments (pinned memory
        makes the copies asynchronous).  Returns the device tensor `self.losses`
        [loss_t, loss_v, loss_pv, loss_itm, loss_lpm, weighted training loss, -, -]; call
        `read_losses()` for the host copy (one device->host sync, like the reference's per-step
        `.cpu()` at `pretrain.py:1071-1075`)."""
        slot = self.global_step & 1
        if self._slot_events[slot] is not None:
            self._slot_events[slot].synchronize()           # staging of step-2 has been consumed
        for k, dst in self.inputs.items():
            src = batch[k]
            if tuple(src.shape) != tuple(dst.shape) or src.dtype != dst.dtype:
                raise ValueError("%s: got %s %s, this step was built for %s %s" % (
                    k, tuple(src.shape), src.dtype, tuple(dst.shape), dst.dtype))
        if self._cuda:
            # host -> device on a copy stream into per-slot device staging (overlaps the previous step's
            # kernels), then a device-to-device hop into the graph's static input buffers on the main stream.
            # The input copies (34 MB) are queued BEFORE the host builds the index plan, so that on a cold pipeline
            # (the first step of a run) the transfer and the plan (2.7 ms of host work) overlap.
            if self._stage is None:
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._stage = []
                for _ in range(2):
                    flat, v = _flat_views(self._flat_spec, self.device)
                    self._stage.append(dict(flat=flat, inputs={k[3:]: t for k, t in v.items() if k.startswith("in:")},
                                            plan={k[5:]: t for k, t in v.items() if k.startswith("plan:")}, hyper=v["hyper"]))
            st = self._stage[slot]
            with torch.cuda.stream(self._copy_stream):
                for k in self.inputs:
                    st["inputs"][k].copy_(batch[k], non_blocking=True)
        plan = build_struct_plan(batch["index_p"].numpy(), batch["index_v"].numpy(), self.Lpv,
                                 self.cfg.num_negative_pv, halves=self.HALVES)
        plan = pad_struct_plan(plan, self.n_items, self.max_num_pv, self.cfg.num_negative_pv)
        if self.mlm_cap is not None:
            mp = build_mlm_plan(batch["masked_lm_labels"].numpy(), batch["masked_lm_labels_pv"].numpy(), self.mlm_cap)
            plan["mlm_rows"], plan["mlm_labels"] = mp["mlm_rows"], mp["mlm_labels"]
        ph = self._plan_host[slot]
        for k, v in plan.items():
            ph[k].copy_(torch.from_numpy(v).reshape(ph[k].shape))
        self._fill_hyper(slot)
        if not self._cuda:
            for k, v in ph.items():
                self.plan[k].copy_(v)
            for k, dst in self.inputs.items():
                dst.copy_(batch[k])
            self.hyper.copy_(self._hyper_host[slot])
        else:
            main = torch.cuda.current_stream()
            with torch.cuda.stream(self._copy_stream):
                for k, v in ph.items():
                    st["plan"][k].copy_(v, non_blocking=True)
                st["hyper"].copy_(self._hyper_host[slot], non_blocking=True)
                landed = torch.cuda.Event()
                landed.record(self._copy_stream)
            main.wait_event(landed)
            self._static_flat.copy_(st["flat"], non_blocking=True)      # inputs + plan + hyper: one memcpy
            ev = torch.cuda.Event()
            ev.record(main)                                 # slot's pinned + device staging free again
            self._slot_events[slot] = ev
        if self.use_graph:
            if self.graph is None:
                self._capture()
            self.graph.replay()
        else:
            n0 = getattr(self.ops, "launches", 0)
            self._device_step()
            self.launches_per_step = getattr(self.ops, "launches", 0) - n0
        self.global_step += 1
        return self.losses

    def read_losses_lagged(self):
        """Asynchronous loss logging: enqueue the device->host copy of THIS step's losses and return the
        PREVIOUS step's (None after the first step), waiting only for that older copy.  The host then
        prepares step i+1 (index plan, H2D) while the device runs step i — the reference's drivers block
        on five `.cpu()` calls per step instead (`pretrain.py:1071-1075`).  `read_losses()` drains."""
        slot = self._loss_reads & 1
        prev = slot ^ 1
        had_prev = self._loss_reads > 0
        self._loss_reads += 1
        out = None
        if had_prev:                       # the copy enqueued one call ago (it is the older one: wait for it first)
            if self._loss_events[prev] is not None:
                self._loss_events[prev].synchronize()
            h = self._loss_ring[prev]
            out = {k: float(h[i]) for i, k in enumerate(LOSS_NAMES + ("loss",))}
        self._loss_ring[slot].copy_(self.losses, non_blocking=True)
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record()
            self._loss_events[slot] = ev
        return out

    def read_losses(self):
        self.losses_host.copy_(self.losses, non_blocking=True)
        if self._cuda:
            torch.cuda.current_stream().synchronize()
        return {k: float(self.losses_host[i]) for i, k in enumerate(LOSS_NAMES + ("loss",))}


class AlignmentStep(PretrainStep):
    """The item-alignment finetune step (`finetune.py:885-1010` around `K3MForItemAlignment.forward`,
    `vilbert_k3m.py:3379-3454`, "ce" head) as one CUDA graph: both items of the `batch_size` pairs run as ONE
    stacked tower pass of 2 * batch_size items, then ClassificationHead + cross-entropy, backward, gradient
    all-reduce and AdamW exactly as in `PretrainStep`.  `step(batch)`: the eleven per-item tensors stacked
    item-1-first along the batch axis (`[2B, ...]`) plus `labels` int64 [B].
    `losses` = [pair loss, 0, 0, 0, 0, pair loss]; `probs` [B, 2] holds the class probabilities."""
    KIND = "alignment"
    HALVES = 2

    def _input_spec(self):
        full = input_spec(self.cfg, self.n_items, self.Lt, self.Lv1, self.Lpv, self.max_num_pv)
        keep = ("input_ids", "token_type_ids", "attention_mask", "input_ids_pv", "token_type_ids_pv",
                "attention_mask_pv", "image_feat", "image_loc", "image_attention_mask")
        spec = {k: full[k] for k in keep}
        spec["labels"] = ((self.B,), torch.int64)
        return spec

    def _forward_backward(self):
        out = self.engine.forward_alignment(self.inputs, self.plan, self.inputs["labels"], None)
        self.engine.on_ready = self._on_ready
        self.engine.backward_alignment()
        self.c_initial, self.c_final, self.probs = out["c_initial"], out["c_final"], out["probs"]
        zero = torch.zeros(1, dtype=torch.float32, device=self.device)
        return [out["loss"], zero, zero, zero, zero, out["loss"]]

