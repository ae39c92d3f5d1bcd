"""TEST INFRASTRUCTURE — plain PyTorch restatement of every kernel contract in `include/k3m.h`.

Two uses, both in `tests/` only:
  * `-m "not gpu"`: injected into `k3m_b200.engine.Engine` in place of the CUDA op table so the
    hand-written forward/backward *schedule* (host logic) is checked on CPU in fp32 against the
    oracle / reference gradients;
  * `-m gpu`: each CUDA kernel is compared against the function of the same name here, run with
    torch ops on the same tensors.

The product never imports this module (k3m_b200.ops raises if the CUDA library is missing).
"""
import math

import torch

ACT_NONE, ACT_GELU, ACT_DGELU, ACT_SIGMOID, ACT_TANH, ACT_DTANH, ACT_GELU_GRAD, ACT_MULAUX = 0, 1, 2, 3, 4, 5, 6, 7

_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
_MASK32 = 0xFFFFFFFF


def _mulhilo(m, c):
    cl, ch = c & 0xFFFF, c >> 16
    t = m * cl
    u = m * ch + (t >> 16)
    lo = ((u & 0xFFFF) << 16) | (t & 0xFFFF)
    return u >> 16, lo


PHILOX_ROUNDS = 7


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Philox4x32-7 on int64 tensors holding 32-bit values (same rounds as the CUDA kernels)."""
    for _ in range(PHILOX_ROUNDS):
        hi0, lo0 = _mulhilo(_M0, c0)
        hi1, lo1 = _mulhilo(_M1, c2)
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK32, lo1, (hi0 ^ c3 ^ k1) & _MASK32, lo0
        k0, k1 = (k0 + _W0) & _MASK32, (k1 + _W1) & _MASK32
    return c0, c1, c2, c3


def _rng_words(rng, site, idx):
    """idx: int64 tensor of counter values -> 4 int64 tensors of 32-bit random words."""
    seed, step = int(rng[0]), int(rng[1])
    dev = idx.device
    z = torch.zeros_like(idx)
    return philox4x32(idx & _MASK32, idx >> 32, z + (site & _MASK32), z + (step & _MASK32),
                      torch.tensor(seed & _MASK32, device=dev), torch.tensor((seed >> 32) & _MASK32, device=dev))


def dropout_keep(rng, site, rows, cols, p, ld=None):
    """Bool keep-mask [rows, cols]; element (r, c) has linear index r*ld + c (ld multiple of 8);
    8 consecutive indices share one Philox call, 16 bits each; keep iff u16 >= round(p*65536)."""
    ld = ld or ((cols + 7) // 8 * 8)
    dev = rng.device
    r = torch.arange(rows, device=dev, dtype=torch.int64)[:, None]
    c = torch.arange(cols, device=dev, dtype=torch.int64)[None, :]
    lin = r * ld + c
    w = _rng_words(rng, site, lin >> 3)
    e = lin & 7
    word = torch.where(e < 2, w[0], torch.where(e < 4, w[1], torch.where(e < 6, w[2], w[3])))
    u16 = (word >> (16 * (e & 1))) & 0xFFFF
    return u16 >= int(p * 65536 + 0.5)


def gumbel_noise(rng, site, n):
    """n fp32 Gumbel(0,1) draws g = -log(-log(u)), 4 per Philox call, u = (r>>9 + 0.5) / 2^23
    (exact in fp32, never 0 or 1 -> finite noise)."""
    dev = rng.device
    lin = torch.arange(n, device=dev, dtype=torch.int64)
    w = _rng_words(rng, site, lin >> 2)
    e = lin & 3
    word = torch.where(e == 0, w[0], torch.where(e == 1, w[1], torch.where(e == 2, w[2], w[3])))
    u = ((word >> 9).to(torch.float32) + 0.5) * (1.0 / 8388608.0)
    return -torch.log(-torch.log(u))


def _apply_drop(v, drop, rng):
    if drop is None:
        return v
    p, site = drop
    keep = dropout_keep(rng, site, v.shape[0], v.shape[1], p)
    return torch.where(keep, v / (1.0 - p), torch.zeros_like(v))


def _gelu(x):
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def _dgelu(x):
    return 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0))) + x * torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)


class RefOps(object):
    """Op table with the same method names / argument meaning as `k3m_b200.ops.CudaOps`."""

    def __init__(self, act_dtype=torch.float32):
        self.adt = act_dtype

    def parallel(self, fns):
        """independent closures (concurrent streams in the CUDA op table): sequential here"""
        for f in fns:
            if f is not None:
                f()

    # ---- GEMM ------------------------------------------------------------------------------
    def gemm(self, a, b, layout, out=None, out_dtype=None, bias=None, act=ACT_NONE, aux_in=None,
             aux_out=None, addend=None, drop=None, rng=None, accumulate=False):
        """layout 'nt': a[M,K] b[N,K];  'nn': a[M,K] b[K,N];  'tn': a[K,M] b[K,N]  -> [M,N].
        v = acc (+bias[n]); GELU: aux_out=v, v=gelu(v); GELU_GRAD: aux_out=gelu'(v), v=gelu(v); DGELU: v*=gelu'(aux_in);
        MULAUX: v*=aux_in; SIGMOID/TANH;
        DTANH: v *= 1-aux_in^2; then dropout; then += addend; then (accumulate ? out+=v : out=v)."""
        af, bf = a.float(), b.float()
        if layout == "nt":
            v = af @ bf.t()
        elif layout == "nn":
            v = af @ bf
        elif layout == "tn":
            v = af.t() @ bf
        else:
            raise ValueError(layout)
        if bias is not None:
            v = v + bias.float()[None, :]
        if act == ACT_GELU:
            if aux_out is not None:
                aux_out.copy_(v)
            v = _gelu(v)
        elif act == ACT_GELU_GRAD:
            if aux_out is not None:
                aux_out.copy_(_dgelu(v))
            v = _gelu(v)
        elif act == ACT_MULAUX:
            v = v * aux_in.float()
        elif act == ACT_DGELU:
            v = v * _dgelu(aux_in.float())
        elif act == ACT_SIGMOID:
            v = torch.sigmoid(v)
        elif act == ACT_TANH:
            v = torch.tanh(v)
        elif act == ACT_DTANH:
            v = v * (1.0 - aux_in.float() ** 2)
        v = _apply_drop(v, drop, rng)
        if addend is not None:
            v = v + addend.float()
        if out is None:
            out = torch.zeros(v.shape, dtype=out_dtype or self.adt, device=v.device)
            accumulate = False
        if accumulate:
            out.add_(v.to(out.dtype))
        else:
            out.copy_(v)
        return out

    def gemm_group(self, calls):
        """independent problems of one layout (one grouped launch in the CUDA table): one after the other here"""
        return [self.gemm(**c) for c in calls]

    def colsum_group(self, items):
        for x, out in items:
            self.colsum(x, out)

    def colsum(self, x, out):
        """out[n] += sum_m x[m, n]  (fp32 accumulate; bias gradients)."""
        out.add_(x.float().sum(0))

    # ---- LayerNorm -------------------------------------------------------------------------
    def ln_fwd(self, x, gamma, beta, out=None, drop=None, rng=None, eps=1e-12):
        """TF-style LN (`vilbert_k3m.py:328-332`), fp32 statistics; optional dropout on y."""
        xf = x.float()
        mean = xf.mean(-1)
        var = (xf - mean[:, None]).pow(2).mean(-1)
        rstd = 1.0 / torch.sqrt(var + eps)
        y = (xf - mean[:, None]) * rstd[:, None] * gamma.float() + beta.float()
        y = _apply_drop(y, drop, rng)
        if out is None:
            out = torch.empty(x.shape, dtype=self.adt, device=x.device)
        out.copy_(y)
        return out, mean, rstd

    def ln_fwd_group(self, items, eps=1e-12):
        return [self.ln_fwd(eps=eps, **it) for it in items]

    def ln_bwd_group(self, items):
        return [self.ln_bwd(**it) for it in items]

    def ln_bwd(self, dy, x, gamma, mean, rstd, dgamma, dbeta, drop_out=None, drop_in=None,
               rng=None, dbias=None, want_masked=False):
        """dx of LN; dgamma/dbeta += ; if drop_out: dy is first masked by the post-LN dropout;
        if drop_in: second output dxm = dx * mask_in / keep (grad of the pre-residual GEMM
        output) and dbias += colsum(dxm); returns (dx, dxm)."""
        dyf = dy.float()
        if drop_out is not None:
            dyf = _apply_drop(dyf, drop_out, rng)
        xhat = (x.float() - mean[:, None]) * rstd[:, None]
        dgamma.add_((dyf * xhat).sum(0))
        dbeta.add_(dyf.sum(0))
        g = dyf * gamma.float()
        dx = (g - g.mean(-1, keepdim=True) - xhat * (g * xhat).mean(-1, keepdim=True)) * rstd[:, None]
        dxm = _apply_drop(dx, drop_in, rng) if drop_in is not None else dx
        if dbias is not None:
            dbias.add_(dxm.sum(0))
        dx_o = dx.to(self.adt)
        dxm_o = dxm.to(self.adt) if (drop_in is not None or want_masked) else dx_o
        return dx_o, dxm_o

    # ---- attention -------------------------------------------------------------------------
    def _qkv(self, qbuf, q_off, kvbuf, k_off, v_off, B, H, Lq, Lk, d):
        q = qbuf[:, q_off:q_off + H * d].float().reshape(B, Lq, H, d).permute(0, 2, 1, 3)
        k = kvbuf[:, k_off:k_off + H * d].float().reshape(B, Lk, H, d).permute(0, 2, 1, 3)
        v = kvbuf[:, v_off:v_off + H * d].float().reshape(B, Lk, H, d).permute(0, 2, 1, 3)
        return q, k, v

    def _probs(self, q, k, mask_add, d):
        s = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d) + mask_add.float()[:, None, None, :]
        return torch.softmax(s, -1), torch.logsumexp(s, -1)

    def attn_fwd(self, qbuf, q_off, kvbuf, k_off, v_off, mask_add, B, H, Lq, Lk, d, drop=None,
                 rng=None, out=None):
        """ctx[b*Lq+i, h*d:+d] = dropout(softmax(q k^T / sqrt(d) + mask_add[b])) v ; returns
        (ctx [B*Lq, H*d], lse [B*H*Lq] fp32).  Dropout element (b,h,i,k): row (b*H+h)*Lq+i."""
        q, k, v = self._qkv(qbuf, q_off, kvbuf, k_off, v_off, B, H, Lq, Lk, d)
        p, lse = self._probs(q, k, mask_add, d)
        if drop is not None:
            p = _apply_drop(p.reshape(B * H * Lq, Lk), drop, rng).reshape(B, H, Lq, Lk)
        ctx = torch.matmul(p, v).permute(0, 2, 1, 3).reshape(B * Lq, H * d)
        if out is None:
            out = torch.empty(B * Lq, H * d, dtype=self.adt, device=qbuf.device)
        out.copy_(ctx)
        return out, lse.reshape(-1)

    def attn_bwd(self, dctx, ctx, qbuf, q_off, kvbuf, k_off, v_off, mask_add, lse, dqbuf, dkvbuf,
                 B, H, Lq, Lk, d, drop=None, rng=None):
        """Writes dq into dqbuf[:, q_off + h*d], dk / dv into dkvbuf[:, k_off / v_off + h*d]."""
        q, k, v = self._qkv(qbuf, q_off, kvbuf, k_off, v_off, B, H, Lq, Lk, d)
        p, _ = self._probs(q, k, mask_add, d)
        do = dctx.float().reshape(B, Lq, H, d).permute(0, 2, 1, 3)
        pd = p
        if drop is not None:
            pd = _apply_drop(p.reshape(B * H * Lq, Lk), drop, rng).reshape(B, H, Lq, Lk)
        dv = torch.matmul(pd.transpose(-1, -2), do)
        dpd = torch.matmul(do, v.transpose(-1, -2))
        if drop is not None:
            dpd = _apply_drop(dpd.reshape(B * H * Lq, Lk), drop, rng).reshape(B, H, Lq, Lk)
        ds = p * (dpd - (dpd * p).sum(-1, keepdim=True))
        dq = torch.matmul(ds, k) / math.sqrt(d)
        dk = torch.matmul(ds.transpose(-1, -2), q) / math.sqrt(d)
        dqbuf[:, q_off:q_off + H * d] = dq.permute(0, 2, 1, 3).reshape(B * Lq, H * d).to(dqbuf.dtype)
        dkvbuf[:, k_off:k_off + H * d] = dk.permute(0, 2, 1, 3).reshape(B * Lk, H * d).to(dkvbuf.dtype)
        dkvbuf[:, v_off:v_off + H * d] = dv.permute(0, 2, 1, 3).reshape(B * Lk, H * d).to(dkvbuf.dtype)

    # ---- embeddings ------------------------------------------------------------------------
    def embed_fwd(self, ids, type_ids, word, pos, typ, L, out=None):
        """out[r] = word[ids[r]] + pos[r % L] + typ[type_ids[r]]   (ids flattened [B*L])."""
        r = torch.arange(ids.numel(), device=ids.device)
        e = word[ids.reshape(-1)].float() + pos[r % L].float() + typ[type_ids.reshape(-1)].float()
        if out is None:
            out = torch.empty(e.shape, dtype=self.adt, device=ids.device)
        out.copy_(e)
        return out

    def embed_bwd(self, dx, ids, type_ids, L, dword, dpos, dtyp):
        """Scatter-add of dx rows into the three fp32 embedding gradients; word id 0 is the
        padding row and gets nothing from this path (`vilbert_k3m.py:343-345`)."""
        idf = ids.reshape(-1)
        dxf = dx.float()
        keep = (idf != 0).to(dxf.dtype)[:, None]
        dword.index_add_(0, idf, dxf * keep)
        r = torch.arange(idf.numel(), device=idf.device)
        dpos.index_add_(0, r % L, dxf)
        dtyp.index_add_(0, type_ids.reshape(-1), dxf)

    def loc_embed_fwd(self, loc, w, b, out=None):
        """out[M, Hv] = loc[M,5] @ w[Hv,5]^T + b  (K=5 is below any tensor-core tile)."""
        v = loc.float() @ w.float().t() + b.float()
        if out is None:
            out = torch.empty(v.shape, dtype=self.adt, device=loc.device)
        out.copy_(v)
        return out

    def loc_embed_bwd(self, dy, loc, dw, db):
        dw.add_(dy.float().t() @ loc.float())
        db.add_(dy.float().sum(0))

    # ---- fusion (hard Gumbel selection) ----------------------------------------------------
    def relu_cat3(self, a, b, c, out=None):
        v = torch.cat([torch.relu(a.float()), torch.relu(b.float()), torch.relu(c.float())], 1)
        if out is None:
            out = torch.empty(v.shape, dtype=self.adt, device=a.device)
        out.copy_(v)
        return out

    def _ysoft(self, gates, noise, site, rng):
        M, H3 = gates.shape
        H = H3 // 3
        ak = torch.sigmoid(gates.float()).reshape(M, 3, H)
        if noise is None:
            # element (m, k, h) -> linear index (m*3 + k)*H + h
            noise = gumbel_noise(rng, site, M * 3 * H).reshape(M, 3, H)
        else:
            noise = noise.float().reshape(M, 3, H)
        y = torch.softmax(ak + noise, dim=1)
        return ak, y

    def fuse_select_fwd(self, cat, gates, noise=None, site=0, rng=None, out=None):
        """cat [M,3H] = relu'd sources, gates [M,3H] pre-sigmoid.  y = softmax_k(sigmoid(gate_k)
        + g_k); out = cat[:, argmax_k] (first max wins), per (row, channel)."""
        M, H3 = cat.shape
        H = H3 // 3
        _, y = self._ysoft(gates, noise, site, rng)
        idx = y.max(1, keepdim=True)[1]
        v = torch.gather(cat.float().reshape(M, 3, H), 1, idx).reshape(M, H)
        if out is None:
            out = torch.empty(M, H, dtype=self.adt, device=cat.device)
        out.copy_(v)
        return out

    def fuse_select_bwd(self, dout, cat, gates, noise=None, site=0, rng=None):
        """Straight-through backward: returns (dgates [M,3H], dcat_direct [M,3H])."""
        M, H3 = cat.shape
        H = H3 // 3
        ak, y = self._ysoft(gates, noise, site, rng)
        idx = y.max(1, keepdim=True)[1]
        hard = torch.zeros_like(y).scatter_(1, idx, 1.0)
        do = dout.float()[:, None, :]
        dy = do * cat.float().reshape(M, 3, H)
        dlog = y * (dy - (y * dy).sum(1, keepdim=True))
        dg = dlog * ak * (1.0 - ak)
        return dg.reshape(M, H3).to(self.adt), (do * hard).reshape(M, H3).to(self.adt)

    def gather_rows(self, src, idx):
        """out[i] = src[idx[i]], zero rows where idx < 0 (labelled-row gather of the MLM head)"""
        i = idx.long()
        out = src[i.clamp_min(0)].clone()
        out[i < 0] = 0
        return out

    def scatter_add_rows(self, dsrc, idx, dst):
        i = idx.long()
        ok = i >= 0
        dst[i[ok]] = (dst[i[ok]].float() + dsrc[ok].float()).to(dst.dtype)

    def gate_mul(self, cat, alpha):
        """soft fusion (`:2300-2329`): cat * alpha, stored in the activation dtype."""
        return (cat.float() * alpha.float()).to(self.adt)

    def gate_bwd(self, dgated, cat, alpha):
        d, c, a = dgated.float(), cat.float(), alpha.float()
        return (d * a).to(self.adt), (d * c * a * (1.0 - a)).to(self.adt)

    def relu_bwd_split3(self, dcat, cat, da, db, dc):
        """d{a,b,c} += dcat[:, k*H:(k+1)*H] * (cat[:, k*H:(k+1)*H] > 0)."""
        H = dcat.shape[1] // 3
        for k, t in enumerate((da, db, dc)):
            t.add_((dcat[:, k * H:(k + 1) * H].float() * (cat[:, k * H:(k + 1) * H] > 0)).to(t.dtype))

    def lincomb3(self, a, b, c, wa, wb, wc, out=None, out_dtype=None):
        v = wa * a.float() + wb * b.float() + (wc * c.float() if c is not None else 0.0)
        if out is None:
            out = torch.empty(v.shape, dtype=out_dtype or self.adt, device=a.device)
        out.copy_(v)
        return out

    def axpy(self, x, y, alpha=1.0):
        """y += alpha * x."""
        y.add_((alpha * x.float()).to(y.dtype))

    def cast(self, x, dtype, out=None):
        if out is not None:
            out.copy_(x)
            return out
        return x.to(dtype)

    def scale_dev_(self, x, alpha):
        x.mul_(alpha.to(x.dtype))

    def scale_(self, x, alpha):
        x.mul_(alpha)

    def act_bwd(self, dy, pre, act):
        """dy * act'(pre) (ACT_DGELU)."""
        assert act == ACT_DGELU
        return (dy.float() * _dgelu(pre.float())).to(self.adt)

    def mask_additive(self, mask):
        """(1 - mask) * -10000 as fp32 [B, L] (`vilbert_k3m.py:2561`)."""
        return (1.0 - mask.float()) * -10000.0

    # ---- pooling ---------------------------------------------------------------------------
    def mean_pool_fwd(self, seq, B, L, out_dtype=torch.float32):
        """[B*L, H] -> [B, H]: mean over tokens 1..L-1 (padding included, `:2405-2409`)."""
        H = seq.shape[1]
        return seq.float().reshape(B, L, H)[:, 1:].mean(1).to(out_dtype)

    def mean_pool_bwd(self, dpool, dseq, B, L):
        """dseq[b*L + i] += dpool[b] / (L-1) for i >= 1."""
        H = dseq.shape[1]
        v = dseq.reshape(B, L, H)
        v[:, 1:] += (dpool.float() / (L - 1))[:, None, :].to(dseq.dtype)

    # ---- structure aggregation + LPM (`vilbert_k3m.py:2413-2505`) --------------------------
    def pair_gather_fwd(self, c_init, seq, pair_item, pair_rows):
        """X[p] = cat(c_init[item_p], (seq[r0]+seq[r1])/2, (seq[r2]+seq[r3])/2)  -> (X [P,3H]
        act dtype for the W1 GEMM, PV [P,2H] fp32 for the LPM norms)."""
        s = seq.float()
        pr = pair_rows.long()
        pv = torch.cat([(s[pr[:, 0]] + s[pr[:, 1]]) * 0.5, (s[pr[:, 2]] + s[pr[:, 3]]) * 0.5], 1)
        x = torch.cat([c_init.float()[pair_item.long()], pv], 1)
        return x.to(self.adt), pv

    def pair_gather_bwd(self, dx, dpv, pair_item, pair_rows, dc_init, dseq):
        """dc_init[item] += dx[:, :H]; dseq[r0], dseq[r1] += (dx[:, H:2H] + dpv[:, :H]) / 2; ..."""
        H = dc_init.shape[1]
        d = dx.float()
        dc_init.index_add_(0, pair_item.long(), d[:, :H].to(dc_init.dtype))
        g = (d[:, H:] + dpv.float()) * 0.5
        pr = pair_rows.long()
        for col, sl in ((0, slice(0, H)), (1, slice(0, H)), (2, slice(H, 2 * H)), (3, slice(H, 2 * H))):
            dseq.index_add_(0, pr[:, col], g[:, sl].to(dseq.dtype))

    def struct_fallback_fwd(self, t, c_init, fb):
        """fb [n,2] int32 (row of t, item) or (-1,-1): t[row] = c_init[item]  (`:2452-2456`)."""
        for row, item in fb.tolist():
            if row >= 0:
                t[row] = c_init[item].to(t.dtype)

    def struct_fallback_bwd(self, dt, dc_init, fb):
        for row, item in fb.tolist():
            if row >= 0:
                dc_init[item] += dt[row].to(dc_init.dtype)
                dt[row] = 0

    def struct_attn_fwd(self, t, w2, b2, seg):
        """seg [B,2] int32 (start,end) pair range per item.  b = w2.leaky_relu(t)+b2 ; att =
        softmax over the item's range ; s[i] = sum_j att_j t_j.  -> (s [B,H] fp32, att [B,maxn])."""
        B = seg.shape[0]
        H = t.shape[1]
        segl = seg.tolist()
        maxn = max(e - s for s, e in segl)
        sc = torch.nn.functional.leaky_relu(t.float()) @ w2.float().reshape(-1) + b2.float().reshape(())
        s_out = torch.zeros(B, H, dtype=torch.float32, device=t.device)
        att = torch.zeros(B, maxn, dtype=torch.float32, device=t.device)
        for i, (s, e) in enumerate(segl):
            a = torch.softmax(sc[s:e], 0)
            att[i, :e - s] = a
            s_out[i] = (a[:, None] * t[s:e].float()).sum(0)
        return s_out, att

    def struct_attn_bwd(self, ds, t, att, w2, seg, dt, dw2, db2):
        """dt [P,H] fp32 += ; dw2 [1,H] += ; db2 [1] +=  (items may share a pair range)."""
        segl = seg.tolist()
        w = w2.float().reshape(-1)
        for i, (s, e) in enumerate(segl):
            tt = t[s:e].float()
            a = att[i, :e - s]
            da = tt @ ds[i].float()
            dsc = a * (da - (a * da).sum())
            dlr = torch.where(tt > 0, torch.ones_like(tt), torch.full_like(tt, 0.01))
            dt[s:e] += a[:, None] * ds[i].float()[None, :] + dsc[:, None] * w[None, :] * dlr
            dw2.add_((dsc[:, None] * torch.nn.functional.leaky_relu(tt)).sum(0).reshape(dw2.shape))
            db2.add_(dsc.sum().reshape(db2.shape))

    def lpm_fwd_bwd(self, c_final, pv, terms, margin, dc_final, dpv, n_valid=None):
        """terms [T,4] int32 = (item i, pair j, negative item, negative value-pair).
        pos = |c[i] + p[j] - v[j]|, neg = |c[ni] + p[j] - v[nv]|;
        loss = mean(max(0, neg - pos + margin)) (MarginRankingLoss(pos, neg, +1), `:2502`).
        Gradients of the loss are accumulated into dc_final [B,H] / dpv [P,2H] (fp32)."""
        H = c_final.shape[1]
        tm = terms.long()
        if n_valid is not None:
            tm = tm[:int(n_valid)]
        i, j, ni, nv = tm[:, 0], tm[:, 1], tm[:, 2], tm[:, 3]
        p, v = pv[:, :H], pv[:, H:]
        up = c_final[i] + p[j] - v[j]
        un = c_final[ni] + p[j] - v[nv]
        npos, nneg = up.norm(dim=1), un.norm(dim=1)
        h = nneg - npos + margin
        T = tm.shape[0]
        loss = torch.clamp(h, min=0).sum() / T
        act = (h > 0).float() / T
        gp = -(act / npos.clamp_min(1e-30))[:, None] * up
        gn = (act / nneg.clamp_min(1e-30))[:, None] * un
        dc_final.index_add_(0, i, gp)
        dc_final.index_add_(0, ni, gn)
        dpv[:, :H].index_add_(0, j, gp + gn)
        dpv[:, H:].index_add_(0, j, -gp)
        dpv[:, H:].index_add_(0, nv, -gn)
        return loss.reshape(1)

    # ---- losses ----------------------------------------------------------------------------
    def count_valid(self, labels, ignore):
        """-> fp32 [1] = number of labels != ignore  (== 1 for the MRM image_label==1 count uses
        `count_eq`)."""
        return (labels.reshape(-1) != ignore).sum().float().reshape(1)

    def count_eq(self, labels, value):
        return (labels.reshape(-1) == value).sum().float().reshape(1)

    def xent_fwd_bwd(self, logits, labels, count, n_cols, dlogits=None, grad_scale=1.0):
        """Mean cross-entropy with ignore_index=-1 over rows of logits[:, :n_cols];
        dlogits = (softmax - onehot) / count on labelled rows, 0 elsewhere.  -> loss [1]."""
        lab = labels.reshape(-1)
        lg = logits[:, :n_cols].float()
        lse = torch.logsumexp(lg, 1)
        valid = lab != -1
        safe = torch.where(valid, lab, torch.zeros_like(lab))
        row = (lse - lg.gather(1, safe[:, None])[:, 0]) * valid
        loss = row.sum() / count
        if dlogits is not None:
            g = torch.softmax(lg, 1)
            g.scatter_add_(1, safe[:, None], -torch.ones_like(g[:, :1]))
            g = g * (valid[:, None] * (grad_scale / count))
            if dlogits.shape[1] > n_cols:
                dlogits[:, n_cols:] = 0
            dlogits[:, :n_cols] = g.to(dlogits.dtype)
        return loss.reshape(1)

    def kl_fwd_bwd(self, logits, target, image_label, count, n_cols, Lv1, dlogits=None, grad_scale=1.0):
        """MRM (`:2754-2760`): logits [B*Lv1, ld] fp32 (row 0 of each item = global region, no
        target); target [B*(Lv1-1), n_cols]; loss = sum_{label==1} sum_c xlogy(t,t) - t*logp
        / count."""
        B = logits.shape[0] // Lv1
        lg = logits[:, :n_cols].float().reshape(B, Lv1, n_cols)[:, 1:]
        logp = torch.log_softmax(lg, 2)
        tg = target.float().reshape(B, Lv1 - 1, n_cols)
        w = (image_label.reshape(B, Lv1 - 1) == 1).float()[:, :, None]
        kl = torch.xlogy(tg, tg) - tg * logp
        loss = (kl * w).sum() / count
        if dlogits is not None:
            g = (torch.softmax(lg, 2) * tg.sum(2, keepdim=True) - tg) * w * (grad_scale / count)
            dlogits.zero_()
            dlogits.reshape(B, Lv1, -1)[:, 1:, :n_cols] = g.to(dlogits.dtype)
        return loss.reshape(1)

    def itm_loss(self, pool_t, pool_pv, pool_v, w, b, l1, l2, l3, drop=None, rng=None):
        """ITM head value (`:1881-1887`, `:2828-2832`): CE(Linear(dropout(sum of pooled))) with
        label 1 - ((l1 + l2 + l3) == 0)."""
        x = _apply_drop(pool_t.float() + pool_pv.float() + pool_v.float(), drop, rng)
        lg = x @ w.float().t() + b.float()
        labels = 1 - ((l1 + l2 + l3) == 0).long()
        return torch.nn.functional.cross_entropy(lg, labels, ignore_index=-1).reshape(1)

    def pair_cls_loss(self, logits, labels, dlogits=None, grad_scale=1.0):
        """"ce" branch of K3MForItemAlignment.forward (`:3443-3450`)."""
        lg = logits.float()
        probs = torch.softmax(lg, dim=1)
        loss = torch.nn.functional.cross_entropy(lg, labels.long()).reshape(1)
        if dlogits is not None:
            oh = torch.nn.functional.one_hot((labels != 0).long(), 2).float()
            dlogits[:, :2] = (grad_scale * (probs - oh) / lg.shape[0]).to(dlogits.dtype)
        return loss, probs

    def dropout(self, x, drop, rng, out=None, out_dtype=None):
        y = _apply_drop(x.float(), drop, rng).to(out_dtype or (out.dtype if out is not None else x.dtype))
        if out is not None:
            out.copy_(y)
            return out
        return y

    # ---- optimizer -------------------------------------------------------------------------
    def plan_to_device(self, plan, device):
        out = {k: torch.from_numpy(v).to(device) for k, v in plan.items() if k != "cap_t"}
        if "cap_t" in plan:
            out["mlm_cap_t"] = int(plan["cap_t"])
        return out

    def adamw(self, p32, g32, m, v, p_act, hyper, zero_grad=True, background=False, g16=None):
        """torch.optim.AdamW step on a flat range.  hyper fp32 [8] = (lr, beta1, beta2, eps, wd,
        step, grad_scale, -).  Writes the act-dtype shadow copy and (optionally) zeroes the grad."""
        lr, b1, b2, eps, wd, step, grad_scale = [float(x) for x in hyper[:7]]
        g = (g16.float() if g16 is not None else g32) * grad_scale
        p32.mul_(1.0 - lr * wd)
        m.mul_(b1).add_(g, alpha=1.0 - b1)
        v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
        bc1 = 1.0 - b1 ** step
        bc2 = 1.0 - b2 ** step
        denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
        p32.addcdiv_(m, denom, value=-lr / bc1)
        p_act.copy_(p32)
        if zero_grad:
            g32.zero_()
