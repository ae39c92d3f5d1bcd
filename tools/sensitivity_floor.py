"""How far do the gradients of one K3M step move when every GEMM result of the PyTorch bf16-storage restatement is
perturbed by 1e-6 relative (i.e. by what a different fp32 accumulation order does)?  CPU only, no CUDA kernel involved:
the intrinsic bf16 sensitivity floor of each fusion mode, which bounds how tightly two correct implementations can agree.
  python tools/sensitivity_floor.py base_mean   ->  median 1.3e-2  p90 2.1e-2  worst 3.7e-2
  python tools/sensitivity_floor.py base_hard   ->  median 7.5e-2  p90 1.0e-1  worst 2.0e-1   (picks forced equal)"""
import sys, torch, random
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.ops_ref import RefOps
from tests.util import case_setup, load_golden, run_engine
from tests.test_step_gpu import _ForcedPickOps
name=sys.argv[1]
fx=load_golden(name); cfg,sd,batch,noise=case_setup(fx)
torch.manual_seed(0)
class P2(RefOps):
    """RefOps whose GEMM result is perturbed by one part in 1e6 before the bf16 store (stands for a different accumulation order)"""
    def gemm(self, a, b, layout, out=None, out_dtype=None, **kw):
        dt = out.dtype if out is not None else (out_dtype or self.adt)
        tmp = super().gemm(a, b, layout, out=None, out_dtype=torch.float32, **{k:v for k,v in kw.items() if k!='accumulate'})
        tmp = tmp * (1.0 + 1e-6*torch.randn_like(tmp))
        if out is None:
            return tmp.to(dt)
        if kw.get('accumulate'):
            out.add_(tmp.to(out.dtype))
        else:
            out.copy_(tmp)
        return out
hard = cfg.if_pre_sampling==1
rec=[]
o1 = _ForcedPickOps(RefOps(torch.bfloat16), rec, False) if hard else RefOps(torch.bfloat16)
outr,L,gr = run_engine(cfg,sd,batch,noise,o1,'cpu',torch.bfloat16,fx['seeds']['python'])
o2 = _ForcedPickOps(P2(torch.bfloat16), rec, True) if hard else P2(torch.bfloat16)
out,_,g = run_engine(cfg,sd,batch,noise,o2,'cpu',torch.bfloat16,fx['seeds']['python'])
rel=[]
for n,(off,num,_) in L.offsets.items():
    ref=fx['grads'][n]
    if ref is None or ref['maxabs']<1e-7: continue
    a,b=gr[off:off+num],g[off:off+num]
    rel.append((float((a-b).norm()/a.norm()),n))
rel.sort()
print(name,'median %.3e p90 %.3e worst %.3e'%(rel[len(rel)//2][0],rel[int(len(rel)*.9)][0],rel[-1][0]), rel[-1][1])
for k in ('loss_t','loss_v','loss_pv','loss_lpm'): print(k,float(out[k]),float(outr[k]))
