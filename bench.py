#!/usr/bin/env python
"""Benchmark of the K3M pretraining step (fwd + bwd + gradient all-reduce + AdamW).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One rank per GPU (torchrun sets RANK / LOCAL_RANK / WORLD_SIZE for N > 1).  Workload = BASELINE.json
configs[1]: `bert_base_6layer_6conect`, batch 64 per GPU, 36-token titles, 36 regions (+1 global)
x 2048-d, 36 PV tokens, all losses, dropout on, bf16 — weak scaling (global batch 64 * N; at N = 8
this is configs[2]'s global batch 512).  Synthetic data, random-init weights.

Prints ONE JSON line (rank 0).  `value` = samples/s with the batch already resident in HBM (graph
replay only); `e2e` = samples/s through the public call `PretrainStep.step(host_batch)` +
`read_losses()` with pinned host inputs (H2D of ~35 MB per step and D2H of the losses inside the
timed region).  `--impl reference` times the CPU restatement of the reference step (oracle port,
fp32, torch.optim.AdamW) on the host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pretrain samples/sec (fwd+bwd+opt)"
UNIT = "samples/s"
# algorithmic FLOPs of the reference graph as written, fwd+bwd (FlopCounterMode probe)
WORKLOADS = {
    # BASELINE configs[1] (N = 1) / configs[2] (N = 8, global batch 512): the headline
    "base": dict(kind="pretrain", per_gpu=64, dims=dict(len_t=36, len_v=36, len_pv=36, max_num_pv=8),
                 flops_per_unit=159.6e9, metric=METRIC, unit="samples/s",
                 text="K3M pretrain step, bert_base_6layer_6conect, batch 64/GPU, 36 title / 36+1 regions x 2048 / "
                      "36 PV tokens, MLM-t + MRM-KL + MLM-pv + LPM (+ITM value), dropout on"),
    # BASELINE configs[3]: 128-token titles, 100 regions, 128 PV tokens, batch 128 on 8 GPUs = 16 per GPU
    "long": dict(kind="pretrain", per_gpu=16, dims=dict(len_t=128, len_v=100, len_pv=128, max_num_pv=30),
                 flops_per_unit=546.7e9, metric=METRIC + ", long-context variant", unit="samples/s",
                 text="K3M pretrain step, long-context variant (BASELINE configs[3]): batch 16/GPU (128 on 8 GPUs), 128 "
                      "title / 100+1 regions x 2048 / 128 PV tokens, all losses, dropout on"),
    # BASELINE configs[4]: item-alignment finetune (run_finetune_item_alignment.sh:40-43), 256 pairs on 8 GPUs
    "align": dict(kind="alignment", per_gpu=32, dims=dict(len_t=50, len_v=36, len_pv=256, max_num_pv=30),
                  flops_per_unit=1081e9, metric="item-alignment finetune pairs/sec (fwd+bwd+opt)", unit="pairs/s",
                  text="K3MForItemAlignment finetune step (BASELINE configs[4]): 32 pairs/GPU (256 on 8 GPUs), 50 title / "
                       "36+1 regions / 256 PV tokens per item, ce head, dropout on"),
}
DIMS = WORKLOADS["base"]["dims"]
PER_GPU_BATCH = WORKLOADS["base"]["per_gpu"]
FLOPS_PER_SAMPLE = WORKLOADS["base"]["flops_per_unit"]


def ncu_gemm_traffic():
    """`roofline.traffic`: mean dram__bytes_read.sum + dram__bytes_write.sum per launch of the GEMM kernel, read
    from the committed `ncu --set full` capture of THIS round's build (profiles/r03_ncu_gemm_full.csv, written by
    tools/ncu_full_summary.py from the .ncu-rep; the earlier r02 capture when that is absent).  None (unmeasured) when the file is absent."""
    name = next((n for n in ("r03_ncu_gemm_full.csv", "r02_ncu_gemm_full.csv")
                 if os.path.exists(os.path.join(ROOT, "profiles", n))), None)
    if name is None:
        return None, "unmeasured: profiles/r03_ncu_gemm_full.csv absent"
    p = os.path.join(ROOT, "profiles", name)
    import csv
    tot, n = 0.0, 0
    for row in csv.DictReader(open(p)):
        try:
            tot += float(row["dram_bytes_read"]) + float(row["dram_bytes_write"])
            n += 1
        except (KeyError, ValueError):
            continue
    if n == 0:
        return None, "unmeasured: no launches in profiles/%s" % name
    return tot / n, "mean over the %d gemm_kernel launches in profiles/%s (ncu --set full)" % (n, name)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], tflops=d["bf16_tflops"], tflops_sustained=d["bf16_tflops_sustained"],
                    source="measured")
    return dict(hbm_gbs=6650.0, tflops=1590.0, tflops_sustained=1400.0, source="fallback")


def workload_config(n_gpus, workload="base"):
    w = WORKLOADS[workload]
    return {"workload": w["text"], "name": workload,
            "global_batch": w["per_gpu"] * n_gpus, "per_gpu_batch": w["per_gpu"],
            "parallelism": "dp%d" % n_gpus,
            "l2": "no explicit flush: one step streams > 20 GB (weights, moments, gradients) through a 126 MB L2"}


def make_cfg():
    from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT
    cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT)
    cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 1, 10.0, 4
    return cfg


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- CPU arm: the oracle port of the reference step on the host cores ---------------------------------
CPU_ARM_NOTE = ("oracle port (oracle/k3m_oracle.py: fp32 autograd restatement of the reference step + torch.optim.AdamW with "
                "the reference's two groups), train mode (dropout ON, on-host Gumbel noise), batch 8 = BASELINE configs[0] "
                "of the same 36/37/36 workload; the unmodified reference itself cannot travel to the GPU box and measured "
                "~1.4x SLOWER than this port in the build container (4.0 vs 2.7-3.0 s/step, 8 threads), so the ratio "
                "against this arm is conservative")


def cpu_reference_step_rate(steps, warmup, batch=8, threads=None):
    """samples/s of the CPU restatement of the reference step: fp32 forward through autograd, backward,
    torch.optim.AdamW (two groups), dropout on.  One step = BASELINE configs[0]: batch 8 at 36/37/36."""
    import random
    import torch
    from k3m_b200.layout import is_no_decay, param_spec
    from k3m_b200.synth import synth_batch, synth_state_dict
    from oracle.k3m_oracle import Oracle, total_loss
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = make_cfg()
    sd = synth_state_dict(param_spec(cfg), seed=0)
    params = {k: torch.nn.Parameter(v) for k, v in sd.items()}
    groups = [dict(params=[p for n, p in params.items() if not is_no_decay(n)], weight_decay=0.01),
              dict(params=[p for n, p in params.items() if is_no_decay(n)], weight_decay=0.0)]
    opt = torch.optim.AdamW(groups, lr=1e-4, betas=(0.9, 0.98), eps=1e-8)
    orc = Oracle(cfg, params, training=True)
    batches = [synth_batch(cfg, batch, DIMS["len_t"], DIMS["len_v"], DIMS["len_pv"], DIMS["max_num_pv"], seed=100 + i)
               for i in range(2)]
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        random.seed(it)
        out = orc.forward(batches[it % 2], None)
        total_loss(out).backward()
        opt.step()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return batch / sec, sec, threads, batch


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # 3 warm-up + 10 timed steps (~1.5-3 s each on the GPU box's host cores: well under a minute)
    steps = max(1, min(args.steps, 10))
    warm = max(1, min(args.warmup, 3))
    v, sec, threads, batch = cpu_reference_step_rate(steps, warm)
    sample = "%d step(s) of batch %d (BASELINE configs[0]) after %d warm-up, fp32, dropout on, %d threads" % (steps, batch, warm, threads)
    # `config` is the GPU arm's (the driver pairs the two lines by it); what THIS arm actually ran — a bounded sample
    # of that workload — is spelled out in `sample` / `reference_arm_note`
    cfgd = workload_config(args.gpus)
    note = ("bounded sample of the configured workload: batch %d per step (BASELINE configs[0]) instead of 64/GPU, same "
            "36 title / 36+1 regions x 2048 / 36 PV shapes, all losses, dropout ON, fp32, %d host threads; samples/s is "
            "the comparable unit.  " % (batch, threads)) + CPU_ARM_NOTE
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfgd,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "reference_arm_note": note, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---- GPU arm ------------------------------------------------------------------------------------------
HBM_KERNELS = ("k3m_ln_fwd", "k3m_ln_fwd_grouped", "k3m_ln_bwd", "k3m_ln_bwd_grouped", "k3m_colsum", "k3m_colsum_grouped",
               "k3m_xent_fwd_bwd", "k3m_kl_fwd_bwd", "k3m_adamw", "k3m_adamw_bg",
               "k3m_embed_fwd", "k3m_embed_bwd", "k3m_relu_cat3", "k3m_fuse_select_fwd", "k3m_fuse_select_bwd",
               "k3m_relu_bwd_split3", "k3m_axpy", "k3m_cast")


def kernel_breakdown(step, batch):
    """One eager (non-graph) step with every launch bracketed by CUDA events on the launching stream:
    -> ({entry point: (launches, total ms, total FLOPs)}, serial GEMM pass ms, hbm kernel rows).
    Parameters/optimizer state are restored."""
    import torch
    snap = (step.p32.clone(), step.m.clone(), step.v.clone(), step.w16.clone(), step.rng.clone())
    use_graph, step.use_graph = step.use_graph, False
    # rank 0 profiles alone: no collective may be issued here (the other ranks are already at the final
    # barrier), so the step runs as a single-rank step
    world, step.world = step.world, 1
    # ... and with the optimizer after backward instead of under it: AdamW saturates HBM, and a kernel timed
    # while it runs on the other stream is not that kernel's duration (the 2 us axpy measured 38 us)
    overlap_opt, step.overlap_opt = step.overlap_opt, False
    step.ops.timing = []
    step.ops.record = {n: [] for n in ("k3m_gemm_bf16", "k3m_gemm_grouped_bf16") + HBM_KERNELS}
    # keep the device ~400 ms behind the host for the WHOLE eager step (1215 launches at ~40 us of host time
    # each outrun a 100 ms head start, after which the pairs of the late backward kernels measured the
    # host's launch gaps): the event pairs then bracket kernels that run back to back
    torch.cuda._sleep(int(8.0e8))
    step.step(batch)
    torch.cuda.synchronize()
    rows, step.ops.timing = step.ops.timing, None
    rec, step.ops.record = step.ops.record, None
    step.use_graph = use_graph
    step.world = world
    step.overlap_opt = overlap_opt
    step.global_step -= 1
    out = {}
    for name, work, e0, e1 in rows:
        n, ms, fl = out.get(name, (0, 0.0, 0.0))
        out[name] = (n + 1, ms + e0.elapsed_time(e1), fl + work)
    gemm_calls = [(n, a) for n in ("k3m_gemm_bf16", "k3m_gemm_grouped_bf16") for a, _ in rec[n]]
    # GEMM launches in issue order do not matter for a serial pass; single and grouped entry points are replayed together
    gemm_ms = replay_serial(step, gemm_calls)
    hbm = []
    for n in HBM_KERNELS:
        calls = rec[n]
        if not calls:
            continue
        ms = replay_serial(step, [(n, a) for a, _ in calls])
        nbytes = sum(b for _, b in calls)
        hbm.append({"kernel": n, "launches": len(calls), "bytes_per_step": nbytes, "ms_serial": round(ms, 4),
                    "gbs": round(nbytes / (ms * 1e-3) / 1e9, 1) if ms > 0 else None})
    for dst, src in zip((step.p32, step.m, step.v, step.w16, step.rng), snap):
        dst.copy_(src)
    step.g32.zero_()
    torch.cuda.synchronize()
    return out, gemm_ms, hbm


def replay_serial(step, calls, reps=3):
    """The step's launches of one kernel family (same pointers, shapes, epilogues) re-issued back to back on one
    stream and timed with ONE event pair: an event pair per launch costs several microseconds of its own,
    which is a third of a small GEMM.  Operands are whatever the finished step left at those addresses
    (duration does not depend on values); accumulating launches land in the gradient arena / parameters, which the
    caller restores afterwards.  `calls`: [(entry point, ctypes argument tuple)].  -> ms for one pass."""
    import ctypes
    import torch
    s = torch.cuda.Stream(device=step.device)
    s.wait_stream(torch.cuda.current_stream())
    st = ctypes.c_void_p(s.cuda_stream)
    fns = [(getattr(step.ops.lib, n), a[:-1]) for n, a in calls]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for fn, a in fns:                                # warm-up pass
            fn(*a, st)
        torch.cuda._sleep(int(4.0e7))                    # ~20 ms head start for the host
        e0.record(s)
        for _ in range(reps):
            for fn, a in fns:
                fn(*a, st)
        e1.record(s)
    s.synchronize()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def build_step(workload, dev, group):
    """-> (step object, list of pinned host batches, units per step per GPU)"""
    import torch
    from k3m_b200.synth import synth_batch
    from k3m_b200.trainer import AlignmentStep, PretrainStep
    rank = int(os.environ.get("RANK", "0"))
    w = WORKLOADS[workload]
    cfg = make_cfg()
    B, d = w["per_gpu"], w["dims"]
    host = []
    if w["kind"] == "pretrain":
        step = PretrainStep(cfg, B, device=dev, init_seed=0, lr=1e-4, warmup_steps=100, total_steps=100000,
                            training=True, rng_seed=1234, use_graph=True, process_group=group, **d)
        for i in range(4):
            b = synth_batch(cfg, B, d["len_t"], d["len_v"], d["len_pv"], d["max_num_pv"], seed=1234 + 97 * rank + i)
            host.append({k: v.pin_memory() for k, v in b.items()})
    else:
        cfg.loss_type = "ce"
        keys = ("input_ids", "token_type_ids", "attention_mask", "input_ids_pv", "token_type_ids_pv",
                "attention_mask_pv", "index_p", "index_v", "image_feat", "image_loc", "image_attention_mask")
        step = AlignmentStep(cfg, B, device=dev, init_seed=1, lr=1e-4, warmup_steps=100, total_steps=100000,
                             training=True, rng_seed=1234, use_graph=True, process_group=group, **d)
        for i in range(4):
            b = synth_batch(cfg, 2 * B, d["len_t"], d["len_v"], d["len_pv"], d["max_num_pv"], seed=1234 + 97 * rank + i)
            hb = {k: b[k].pin_memory() for k in keys}
            hb["labels"] = torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(i + 7 * rank)).pin_memory()
            host.append(hb)
    return step, host, B


def class_api_rate(dev, steps, warmup=4):
    """The same workload through the DROP-IN CLASS API a reference driver uses (`pretrain.py:986-1110`): host batch ->
    `model(...)` -> `loss = out[0] + out[1] + out[3] + out[9]` -> `loss.backward()` -> `torch.optim.AdamW.step()` ->
    `zero_grad()`, one loss read per step.  From the second step of a geometry on, forward and backward are two CUDA
    graph replays.  -> (samples/s, ms per step)"""
    import torch
    from k3m_b200.layout import is_no_decay
    from k3m_b200.synth import FORWARD_ORDER, synth_batch
    from k3m_b200.vilbert_k3m import BertForMultiModalPreTraining_tri_stru
    cfg = make_cfg()
    torch.manual_seed(0)
    model = BertForMultiModalPreTraining_tri_stru(cfg).to(dev)
    model.train()
    named = list(model.named_parameters())
    groups = [dict(params=[p for n, p in named if not is_no_decay(n)], weight_decay=0.01),
              dict(params=[p for n, p in named if is_no_decay(n)], weight_decay=0.0)]
    opt = torch.optim.AdamW(groups, lr=1e-4, betas=(0.9, 0.98), eps=1e-8, fused=True)
    B, d = PER_GPU_BATCH, DIMS
    host = []
    for i in range(4):
        b = synth_batch(cfg, B, d["len_t"], d["len_v"], d["len_pv"], d["max_num_pv"], seed=4321 + i)
        host.append({k: (v.pin_memory() if k not in ("index_p", "index_v") else v) for k, v in b.items()})

    def one(b):
        out = model(*[b[k] for k in FORWARD_ORDER], input_ids_pv=b["input_ids_pv"], token_type_ids_pv=b["token_type_ids_pv"],
                    attention_mask_pv=b["attention_mask_pv"], masked_lm_labels_pv=b["masked_lm_labels_pv"],
                    next_sentence_label_pv_v=b["next_sentence_label_pv_v"], next_sentence_label_pv_t=b["next_sentence_label_pv_t"],
                    index_p=b["index_p"], index_v=b["index_v"])
        loss = out[0] + out[1] + out[3] + out[9]
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=False)
        return loss

    for i in range(warmup):
        one(host[i % 4])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for i in range(steps):
        loss = one(host[i % 4])
        if last is not None:
            float(last)                  # the previous step's loss: one device -> host read per step, one step late
        last = loss.detach()
    float(last)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    del model, opt
    torch.cuda.empty_cache()
    return B / (ms * 1e-3), ms


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    W = WORKLOADS[args.workload]
    step, host, B = build_step(args.workload, dev, group)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def note(msg):
        if os.environ.get("K3M_BENCH_VERBOSE"):
            sys.stderr.write("[bench rank %d] %s\n" % (rank, msg))
            sys.stderr.flush()

    # warm-up through the public call (captures the graph on the first step)
    for i in range(max(3, args.warmup)):
        step.step(host[i % 4])
        step.read_losses_lagged()          # the timed end-to-end loop's read path is warmed too (its pinned ring, events)
        note("warm-up step %d queued" % i)
    loss0 = step.read_losses()
    note("warm-up done %s" % loss0)
    # ---- (1) device-resident: graph replays only ----
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step.graph.replay()
    e1.record()
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    note("device-resident loop done: %.3f ms/step" % ms_dev)
    # ---- (2) end to end: host batch -> H2D -> step -> D2H losses, every step ----
    barrier()
    e0.record()
    # every step: H2D of its inputs from pinned memory, the step, a D2H read of a step's losses — the
    # read lags one step (the host prepares step i+1 while the device runs step i); the last read drains
    host_s = 0.0
    for i in range(args.steps):
        t_h = time.perf_counter()
        step.step(host[i % 4])
        step.read_losses_lagged()
        host_s += time.perf_counter() - t_h      # host time of the public call (index plan, copies queued, graph launch, lagged read)
    last = step.read_losses()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    note("end-to-end loop done: %.3f ms/step" % ms_e2e)
    clk = clocks.stop() if clocks is not None else None
    comm = None
    if world > 1 and hasattr(step, "comm_timing"):
        comm = step.comm_timing(host[0])           # one instrumented non-graph step, every rank takes part
        barrier()

    def finish():
        """Leave cleanly at N > 1: the captured graph holds NCCL kernels, and tearing the communicator down
        under it has been seen to block forever (N = 2, exit only by the launcher's timeout).  Release the
        graph first, and never let teardown outlive the result: a watchdog ends the process with status 0."""
        sys.stdout.flush()
        sys.stderr.flush()
        if world == 1:
            return
        t = threading.Timer(20.0, lambda: os._exit(0))
        t.daemon = True
        t.start()
        step.graph = None
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        t.cancel()

    if rank != 0:
        finish()
        return 0
    peaks = load_peaks()
    if args.no_breakdown or os.environ.get("K3M_BENCH_SKIP"):
        print(json.dumps({"metric": W["metric"], "value": B * world / (ms_dev * 1e-3), "unit": W["unit"], "n_gpus": world,
                          "ms_per_step": ms_dev, "e2e_ms_per_step": ms_e2e, "e2e_host_ms_per_step": host_s * 1e3 / args.steps, "skip": os.environ.get("K3M_BENCH_SKIP", ""),
                          "launches_per_step": step.launches_per_step, "note": "timing only (no kernel breakdown)",
                          "comm": comm}))
        finish()
        return 0
    kb, gemm_serial_ms, hbm = kernel_breakdown(step, host[0])
    for h in hbm:
        h["frac_of_hbm_peak"] = round(h["gbs"] / peaks["hbm_gbs"], 3) if h["gbs"] else None
    tot_ms = sum(v[1] for v in kb.values())
    shapes = {k: v for k, v in kb.items() if k.startswith("k3m_gemm")}
    merged = {}
    for k, v in kb.items():
        n = k.split(" ")[0]
        a = merged.get(n, (0, 0.0, 0.0))
        merged[n] = (a[0] + v[0], a[1] + v[1], a[2] + v[2])
    kb = merged
    g_n = sum(v[0] for k, v in kb.items() if k.startswith("k3m_gemm"))
    g_ms = sum(v[1] for k, v in kb.items() if k.startswith("k3m_gemm")) or 1e-9
    g_fl = sum(v[2] for k, v in kb.items() if k.startswith("k3m_gemm"))
    gemm_tflops = g_fl / (gemm_serial_ms * 1e-3) / 1e12
    units = B * world
    step_tflops = W["flops_per_unit"] * B / (ms_dev * 1e-3) / 1e12        # per GPU
    traffic, traffic_note = ncu_gemm_traffic()
    line = {
        "metric": W["metric"], "value": units / (ms_dev * 1e-3), "unit": W["unit"], "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(world, args.workload),
        "clocks": clk,
        "e2e": {"value": units / (ms_e2e * 1e-3), "unit": W["unit"], "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": step.h2d_bytes, "d2h_bytes_per_step": step.d2h_bytes,
                "host_ms_per_step": host_s * 1e3 / args.steps},
        "gpu_launches": step.launches_per_step * args.steps * 2,
        "launches_per_step": step.launches_per_step,
        "roofline": {"bound": "tensor", "kernel": "gemm_kernel (k3m_gemm_bf16 / k3m_gemm_grouped_bf16: tcgen05/TMEM/TMA)",
                     "achieved": gemm_tflops, "peak": peaks["tflops"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["tflops"], "traffic": traffic, "traffic_note": traffic_note,
                     "peak_source": peaks["source"],
                     "launches_per_step": g_n, "ms_per_step_in_kernel": gemm_serial_ms,
                     "ms_per_step_in_kernel_event_pair_per_launch": g_ms,
                     "share_of_step_kernel_time": g_ms / tot_ms,
                     "flops_per_step_in_kernel": g_fl,
                     "how": "the step's %d GEMM launches (same operands, shapes, epilogues) re-issued back to back on one "
                            "stream, one CUDA-event pair around the pass, mean of 3 passes; "
                            "ms_per_step_in_kernel_event_pair_per_launch and kernel_ms bracket EVERY launch of one eager "
                            "step with its own event pair, which adds several microseconds to each" % g_n},
        "step_roofline": {"bound": "tensor", "flops_per_unit": W["flops_per_unit"],
                          "flops_note": "reference-equivalent FLOPs (the reference graph as written, SURVEY 8d); the masked-LM "
                                        "head runs on the labelled rows only, so the EXECUTED GEMM FLOPs per step are "
                                        "roofline.flops_per_step_in_kernel (what roofline.achieved is computed from)",
                          "achieved": step_tflops,
                          "peak": peaks["tflops_sustained"], "unit": "TFLOP/s per GPU",
                          "frac": step_tflops / peaks["tflops_sustained"], "frac_of_burst": step_tflops / peaks["tflops"]},
        "hbm_kernels": {"peak_gbs": peaks["hbm_gbs"], "how": "algorithmic bytes of the step's launches of each row-wise kernel "
                        "(inputs read once + outputs written once) / one serial back-to-back replay of exactly those launches",
                        "rows": hbm},
        "kernel_ms": {k: {"launches": v[0], "ms": round(v[1], 4)} for k, v in sorted(kb.items(), key=lambda kv: -kv[1][1])},
        "gemm_shapes": {k: {"launches": v[0], "ms": round(v[1], 4), "tflops": round(v[2] / max(v[1], 1e-9) / 1e9, 1)}
                        for k, v in sorted(shapes.items(), key=lambda kv: -kv[1][1])},
        "losses_first": loss0, "losses_last": last,
    }
    if comm is not None:
        line["comm"] = comm
    if world == 1 and args.workload == "base" and not args.no_class_api:
        del step
        torch.cuda.empty_cache()
        v, ms = class_api_rate(dev, max(5, min(args.steps, 20)))
        line["e2e_class_api"] = {"value": v, "unit": UNIT, "ms_per_step": ms, "api": "BertForMultiModalPreTraining_tri_stru(...) "
                                 "forward -> loss.backward() -> torch.optim.AdamW(fused).step(), host batch in pinned memory, one "
                                 "loss read per step (the drop-in class path of pretrain.py:986-1110; forward and backward replay "
                                 "CUDA graphs)", "ratio_to_e2e": v / line["e2e"]["value"]}
    if world == 1 and not args.no_cpu_baseline and args.workload == "base":
        v, sec, threads, batch = cpu_reference_step_rate(8, 2)       # ~15-25 s of host work on the GPU box
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": "8 steps of batch %d (BASELINE configs[0]) after 2 warm-ups, dropout on, %.1f s/step: %s"
                                          % (batch, sec, CPU_ARM_NOTE)}
    print(json.dumps(line))
    finish()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="k3m_b200", choices=["k3m_b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-class-api", action="store_true", help="skip the class-API end-to-end measurement")
    ap.add_argument("--no-breakdown", action="store_true", help="timing only: skip the per-kernel breakdown / roofline replays")
    ap.add_argument("--workload", default="base", choices=sorted(WORKLOADS),
                    help="base = BASELINE configs[1]/[2] (the headline); long = configs[3]; align = configs[4]")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
