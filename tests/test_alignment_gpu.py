"""GPU: the item-alignment forward/backward (`K3MForItemAlignment`, BASELINE configs[4]) on the CUDA op
table against fixtures made from the unmodified reference (fp32), and against the bf16-storage
restatement on the same device."""
import pytest
import torch

from tests.util import align_case_setup, compare_alignment_with_golden, load_golden, run_engine_alignment

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from k3m_b200.ops import CudaOps
    return CudaOps("cuda")


# bf16 activations: loss 1e-3 (north_star), probabilities 2e-3 absolute, gradients at the bf16 storage
# floor of the tiny configs (tests/test_step_gpu.py header); hard Gumbel pick: 2e-3 / floor-sized bound
# align_base_hard: the real bert_base_6layer_6conect towers at the finetune script's lengths (50 / 36+1 / 256 tokens,
# run_finetune_item_alignment.sh:40-43) from the unmodified reference: pins the > 128-token attention kernels
@pytest.mark.parametrize("name,loss_rtol,grad_tol", [("align_tiny_mean", 1e-3, 0.1), ("align_tiny_hard", 2e-3, 0.4),
                                                     ("align_base_hard", 2e-3, 0.5)])
def test_alignment_matches_reference(ops, name, loss_rtol, grad_tol):
    fx = load_golden(name)
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    out, L, g = run_engine_alignment(cfg, sd, batch, labels, noise, ops, "cuda", torch.bfloat16)
    torch.cuda.synchronize()
    errs = compare_alignment_with_golden(fx, out, L, g, loss_rtol=loss_rtol, grad_tol=grad_tol, prob_atol=2e-3)
    assert not errs, (len(errs), errs[:12])


def test_alignment_matches_bf16_storage_restatement(ops):
    from oracle.ops_ref import RefOps
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    out, L, g = run_engine_alignment(cfg, sd, batch, labels, noise, ops, "cuda", torch.bfloat16)
    outr, _, gr = run_engine_alignment(cfg, sd, batch, labels, noise, RefOps(torch.bfloat16), "cuda", torch.bfloat16)
    assert float(out["loss"]) == pytest.approx(float(outr["loss"]), rel=1e-3)
    assert float((out["probs"] - outr["probs"]).abs().max()) <= 2e-3
    rel = []
    for n, (off, num, _) in L.offsets.items():
        ref = fx["grads"][n]
        if ref is None or ref["maxabs"] < 1e-9:
            continue
        a, b = gr[off:off + num], g[off:off + num]
        rel.append(float((a - b).norm() / (a.norm() + 1e-12)))
    rel.sort()
    assert rel[len(rel) // 2] <= 3e-2 and rel[-1] <= 0.3, (rel[len(rel) // 2], rel[-1])


def test_alignment_training_mode_runs_and_is_deterministic(ops):
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    o1, L, g1 = run_engine_alignment(cfg, sd, batch, labels, noise, ops, "cuda", torch.bfloat16, training=True)
    o2, _, g2 = run_engine_alignment(cfg, sd, batch, labels, noise, ops, "cuda", torch.bfloat16, training=True)
    o3, _, g3 = run_engine_alignment(cfg, sd, batch, labels, noise, ops, "cuda", torch.bfloat16, training=False)
    assert float((g1 - g2).abs().max()) <= 1e-4 * float(g1.abs().max())
    assert float((g1 - g3).abs().max()) > 1e-3 * float(g1.abs().max())
    assert torch.isfinite(g1).all()


def test_alignment_step_graph_on_cuda():
    """`AlignmentStep` (one CUDA graph per finetune step, fused AdamW): first-step loss against the reference
    fixture, then the pair loss falls over a few steps on a fixed batch."""
    from k3m_b200.trainer import AlignmentStep
    from tests.util import ITEM_KEYS
    fx = load_golden("align_tiny_mean")
    cfg, sd, batch, labels, noise = align_case_setup(fx)
    B, Lt, Lv, Lpv, mpv = fx["dims"]
    step = AlignmentStep(cfg, B, len_t=Lt, len_v=Lv, len_pv=Lpv, max_num_pv=mpv, device="cuda", state_dict=sd,
                         training=False, lr=2e-3, use_graph=True)
    hb = {k: batch[k] for k in ITEM_KEYS}
    hb["labels"] = labels
    losses = []
    for it in range(6):
        step.step(hb)
        losses.append(step.read_losses()["loss"])
    assert abs(losses[0] - fx["loss"]) <= 1e-3 * abs(fx["loss"]), (losses[0], fx["loss"])
    assert float((step.probs.cpu() - fx["probs"]).abs().max()) < 0.5          # probs of the LAST step: finite, sane
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    assert step.launches_per_step > 100


def test_alignment_base_config_runs():
    """bert_base_6layer_6conect towers at the finetune script's lengths (50 title tokens, 36 + 1 regions, 256 PV
    tokens: run_finetune_item_alignment.sh:40-43) — PV self-attention and both PV co-attentions take the blocked
    attention kernels."""
    from k3m_b200.config import BertConfig, K3M_BASE_6LAYER_6CONNECT
    from k3m_b200.synth import synth_batch
    from k3m_b200.trainer import AlignmentStep
    from tests.util import ITEM_KEYS
    cfg = BertConfig.from_dict(K3M_BASE_6LAYER_6CONNECT)
    cfg.if_pre_sampling, cfg.loss_type = 1, "ce"
    B = 4
    step = AlignmentStep(cfg, B, len_t=50, len_v=36, len_pv=256, max_num_pv=30, device="cuda", init_seed=1,
                         training=True, lr=1e-4, use_graph=True)
    batch = synth_batch(cfg, 2 * B, 50, 36, 256, max_num_pv=30, seed=3)
    hb = {k: batch[k] for k in ITEM_KEYS}
    hb["labels"] = torch.tensor([0, 1, 1, 0])
    for it in range(3):
        step.step(hb)
    out = step.read_losses()
    assert out["loss"] == out["loss"] and 0.2 < out["loss"] < 2.0, out
