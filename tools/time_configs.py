"""Per-GPU step time of BASELINE.json configs[3] and configs[4] on one B200 (not bench lines: bench.py
measures configs[1]; these are the per-rank shares of the 8-GPU configurations, timed with CUDA events).

  configs[3]  long-context pretrain: 128-token titles, 100 + 1 regions, 128 PV tokens, batch 128 / 8 GPUs = 16 per GPU
  configs[4]  item-alignment finetune (run_finetune_item_alignment.sh: 50 / 36 + 1 / 256 tokens), 256 pairs / 8 GPUs
              = 32 pairs (64 items) per GPU

Prints one JSON line per configuration.  Usage: python tools/time_configs.py [--steps 10] [--warmup 3]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def time_step(step, batches, steps, warmup):
    import torch
    for i in range(max(3, warmup)):
        step.step(batches[i % len(batches)])
    step.read_losses()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step.step(batches[i % len(batches)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, step.read_losses()


def breakdown(step, batch, top=28):
    from bench import kernel_breakdown
    rows, _ = kernel_breakdown(step, batch)
    tot = sum(v[1] for v in rows.values())
    sys.stderr.write("  eager step, %.2f ms of kernels in %d launches\n" % (tot, sum(v[0] for v in rows.values())))
    for name, (n, ms, fl) in sorted(rows.items(), key=lambda kv: -kv[1][1])[:top]:
        sys.stderr.write("  %8.3f ms %5.1f%% %5d x  %s%s\n" % (ms, 100 * ms / tot, n, name,
                                                               "  %.0f TF/s" % (fl / ms / 1e9) if fl else ""))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--only", default="")
    ap.add_argument("--breakdown", action="store_true", help="per entry point device time of one eager step")
    args = ap.parse_args()
    import torch
    from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT
    from k3m_b200.synth import synth_batch
    from k3m_b200.trainer import AlignmentStep, PretrainStep
    from tests.util import ITEM_KEYS
    torch.cuda.set_device(0)

    if args.only in ("", "long"):
        cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT)
        cfg.if_pre_sampling, cfg.margin, cfg.num_negative_pv = 1, 10.0, 4
        B, dims = 16, dict(len_t=128, len_v=100, len_pv=128, max_num_pv=30)
        step = PretrainStep(cfg, B, device="cuda", init_seed=0, lr=1e-4, warmup_steps=100, total_steps=100000,
                            training=True, rng_seed=1234, use_graph=True, **dims)
        hb = [{k: v.pin_memory() for k, v in synth_batch(cfg, B, 128, 100, 128, 30, seed=7 + i).items()}
              for i in range(3)]
        ms, losses = time_step(step, hb, args.steps, args.warmup)
        print(json.dumps({"config": "configs[3] long-context pretrain, per-GPU share", "per_gpu_batch": B,
                          "dims": dims, "ms_per_step": round(ms, 3), "samples_per_s": round(B / ms * 1e3, 1),
                          "gpu_launches": step.launches_per_step,
                          "losses": {k: round(float(v), 4) for k, v in losses.items()}}))
        if args.breakdown:
            breakdown(step, hb[0])
        del step
        torch.cuda.empty_cache()

    if args.only in ("", "align"):
        cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT)
        cfg.if_pre_sampling, cfg.loss_type = 1, "ce"
        B, dims = 32, dict(len_t=50, len_v=36, len_pv=256, max_num_pv=30)
        step = AlignmentStep(cfg, B, device="cuda", init_seed=1, training=True, lr=1e-4, use_graph=True, **dims)
        hb = []
        for i in range(3):
            b = synth_batch(cfg, 2 * B, 50, 36, 256, max_num_pv=30, seed=11 + i)
            d = {k: b[k].pin_memory() for k in ITEM_KEYS}
            d["labels"] = torch.randint(0, 2, (B,), generator=torch.Generator().manual_seed(i)).pin_memory()
            hb.append(d)
        ms, losses = time_step(step, hb, args.steps, args.warmup)
        print(json.dumps({"config": "configs[4] item-alignment finetune step, per-GPU share", "pairs_per_gpu": B,
                          "dims": dims, "ms_per_step": round(ms, 3), "pairs_per_s": round(B / ms * 1e3, 1),
                          "gpu_launches": step.launches_per_step,
                          "losses": {k: round(float(v), 4) for k, v in losses.items()}}))
        if args.breakdown:
            breakdown(step, hb[0])


if __name__ == "__main__":
    main()
