import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k3m_b200.ops import CudaOps
from oracle.ops_ref import RefOps
ops, ref = CudaOps("cuda"), RefOps(torch.float32)
for (B, H, Lq, Lk, d) in ((1, 1, 16, 16, 16), (1, 1, 36, 36, 64), (2, 2, 5, 7, 16), (1, 1, 36, 48, 64), (1, 1, 36, 49, 64)):
    torch.manual_seed(0)
    qb = torch.randn(B * Lq, 3 * H * d).bfloat16().cuda()
    kv = torch.randn(B * Lk, 3 * H * d).bfloat16().cuda()
    mask = torch.zeros(B, Lk).cuda()
    ctx, lse = ops.attn_fwd(qb, 0, kv, H * d, 2 * H * d, mask, B, H, Lq, Lk, d)
    cr, lr = ref.attn_fwd(qb, 0, kv, H * d, 2 * H * d, mask, B, H, Lq, Lk, d)
    e = (ctx.float() - cr).abs()
    print((B, H, Lq, Lk, d), "lse err", float((lse - lr).abs().max()), "ctx err", float(e.max()),
          "bad rows", (e.max(1)[0] > 0.05).nonzero().flatten().tolist()[:20], "bad cols", (e.max(0)[0] > 0.05).nonzero().flatten().tolist()[:40])
