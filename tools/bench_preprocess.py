"""Times the GPU batch preprocessing (csrc/preprocess.cu) at the bench workload's shapes (B = 64, 36 / 36+1 / 36,
2048-d features) with CUDA events, against its HBM roofline (algorithmic bytes: features read once and written
once with the global row, plus the small id / box tensors) and against the CPU restatement of the reference's
per-sample Python path (oracle/preprocess_ref.py) on the host.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from k3m_b200.preprocess import GpuBatchPreprocessor, pack_raw_batch
    from oracle.preprocess_ref import preprocess_batch
    from tests.test_preprocess import synth_raws
    B, Lt, Lpv, R, F, T, P = 64, 36, 36, 36, 2048, 1601, 8
    rng = np.random.RandomState(0)
    raws, u_tok, rnd_tok, u_reg = synth_raws(rng, B, Lt, Lpv, R, F, T, P)
    raw = pack_raw_batch(raws, Lt, Lpv, R)
    pre = GpuBatchPreprocessor(P)
    draws = dict(u_tok=torch.from_numpy(u_tok).cuda(), rnd_tok=torch.from_numpy(rnd_tok).cuda(),
                 u_reg=torch.from_numpy(u_reg).cuda())
    # rotate over copies of the feature tensor larger than L2 together
    raws_dev = [dict(raw, features=raw["features"].clone()) for _ in range(8)]
    for r in raws_dev[:3]:
        pre(r, draws)
    torch.cuda.synchronize()
    n = 40
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=s):
        for i in range(n):
            out = pre(raws_dev[i % 8], draws)
    g.replay()
    torch.cuda.synchronize()
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    bytes_alg = B * R * F * 4 + B * (R + 1) * F * 4 + B * (Lt + Lpv) * 8 * 5 + B * R * (16 + 8 + 8) + B * (R + 1) * 28
    peak = 6500.0
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        preprocess_batch(raws, u_tok, rnd_tok, u_reg, Lt, Lpv, R, P)
    cpu_s = (time.perf_counter() - t0) / reps
    print(json.dumps({"what": "GPU batch preprocessing, B=64, 36/36+1/36 tokens, 2048-d features (2 launches: k3m_prep_batch)",
                      "us_per_batch": round(us, 2), "samples_per_s": round(B / us * 1e6),
                      "roofline": {"bound": "hbm", "achieved": round(bytes_alg / us / 1e3, 1), "peak": peak,
                                   "unit": "GB/s", "frac": round(bytes_alg / us / 1e3 / peak, 3),
                                   "algorithmic_bytes": bytes_alg},
                      "cpu_restatement": {"samples_per_s": round(B / cpu_s, 1), "cores": 1,
                                          "what": "oracle/preprocess_ref.py (numpy, per sample like the reference)"}}))


if __name__ == "__main__":
    main()
