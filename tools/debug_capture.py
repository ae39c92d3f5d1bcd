"""Which call invalidates the stream capture of AlignmentStep (flaky `cudaErrorStreamCaptureInvalidated`)?  After every
kernel-table call, and around the fork / join of `CudaOps.parallel`, ask the runtime for the capture status of the
current stream and report the first point at which it is no longer `active`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200 import ops as ops_mod  # noqa: E402
from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT  # noqa: E402
from k3m_b200.synth import synth_batch  # noqa: E402
from k3m_b200.trainer import AlignmentStep  # noqa: E402
from tests.util import ITEM_KEYS  # noqa: E402

import ctypes  # noqa: E402
from k3m_b200 import _lib  # noqa: E402
state = {"bad": None, "n": 0}


def status():
    st = _lib.load().k3m_debug_capture_status(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return {0: "none", 1: "active", 2: "Invalidated"}.get(st, "queryerror%d" % st)


orig_call = ops_mod.CudaOps._call


def traced(self, name, *args):
    before = status()
    try:
        orig_call(self, name, *args)
    finally:
        after = status()
        state["n"] += 1
        if state["bad"] is None and ("Invalidated" in after + before or "queryerror" in after + before):
            state["bad"] = (state["n"], name, str(before), str(after))
            print("FIRST invalidated capture at call #%d %s: before=%s after=%s" % state["bad"], flush=True)


ops_mod.CudaOps._call = traced
orig_parallel = ops_mod.CudaOps.parallel


def traced_parallel(self, fns):
    b = status()
    orig_parallel(self, fns)
    a = status()
    if state["bad"] is None and ("Invalidated" in a + b or "queryerror" in a + b):
        state["bad"] = (state["n"], "parallel()", b, a)
        print("FIRST invalidated capture around parallel(): before=%s after=%s at call #%d" % (b, a, state["n"]), flush=True)


ops_mod.CudaOps.parallel = traced_parallel
cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT)
cfg.if_pre_sampling, cfg.loss_type = 1, "ce"
B = 4
for trial in range(4):
    state.update(bad=None, n=0)
    step = AlignmentStep(cfg, B, len_t=50, len_v=36, len_pv=256, max_num_pv=30, device="cuda", init_seed=1,
                         training=True, lr=1e-4, use_graph=True)
    batch = synth_batch(cfg, 2 * B, 50, 36, 256, max_num_pv=30, seed=3)
    hb = {k: batch[k] for k in ITEM_KEYS}
    hb["labels"] = torch.tensor([0, 1, 1, 0])
    try:
        for it in range(3):
            step.step(hb)
        print("trial", trial, "ok", step.read_losses()["loss"], "calls", state["n"])
    except Exception as e:
        print("trial", trial, "FAILED", type(e).__name__, str(e)[:200].replace("\n", " "), "first bad:", state["bad"])
        torch.cuda.synchronize()
    del step
