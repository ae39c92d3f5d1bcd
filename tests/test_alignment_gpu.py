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
@pytest.mark.parametrize("name,loss_rtol,grad_tol", [("align_tiny_mean", 1e-3, 0.1), ("align_tiny_hard", 2e-3, 0.4)])
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
