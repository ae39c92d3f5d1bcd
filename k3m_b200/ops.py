"""Op table over the C ABI: every method takes torch CUDA tensors, passes raw device pointers,
strides and the current CUDA stream to libk3m_sm100a.so, and returns torch tensors that own the
outputs.  torch is used for memory and streams only — no torch arithmetic runs here."""
import ctypes
import os

import torch

from . import _lib

ACT_NONE, ACT_GELU, ACT_DGELU, ACT_SIGMOID, ACT_TANH, ACT_DTANH, ACT_GELU_GRAD, ACT_MULAUX = 0, 1, 2, 3, 4, 5, 6, 7
_LAYOUT = {"nt": 0, "nn": 1, "tn": 2}
_c = ctypes
_vp, _i, _f, _u, _ll = _c.c_void_p, _c.c_int, _c.c_float, _c.c_uint, _c.c_longlong


def _p(t):
    return _vp(t.data_ptr()) if t is not None else _vp(0)


def _ld(t):
    if t is None:
        return 0
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensor expected"
    return t.stride(0)


def _isf32(t):
    if t.dtype == torch.float32:
        return 1
    assert t.dtype == torch.bfloat16, t.dtype       # (fp32 check build: every tensor is fp32 and takes the branch above)
    return 0


class _GemmDesc(_c.Structure):
    """K3MGemmDesc of include/k3m.h"""
    _fields_ = [("a", _vp), ("lda", _i), ("b", _vp), ("ldb", _i), ("layout", _i), ("M", _i), ("N", _i), ("K", _i),
                ("out", _vp), ("ldc", _i), ("out_f32", _i), ("accumulate", _i), ("bias", _vp), ("act", _i),
                ("aux_in", _vp), ("ld_aux_in", _i), ("aux_out", _vp), ("ld_aux_out", _i), ("addend", _vp),
                ("ld_add", _i), ("addend_f32", _i), ("rng", _vp), ("drop_site", _u), ("drop_p", _f)]


class _ColsumDesc(_c.Structure):
    """K3MColsumDesc of include/k3m.h"""
    _fields_ = [("x", _vp), ("ld", _i), ("out", _vp), ("M", _i), ("N", _i)]


class _LnFwdDesc(_c.Structure):
    """K3MLnFwdDesc of include/k3m.h"""
    _fields_ = [("x", _vp), ("gamma", _vp), ("beta", _vp), ("y", _vp), ("mean", _vp), ("rstd", _vp), ("M", _i), ("H", _i),
                ("rng", _vp), ("site", _u), ("drop_p", _f)]


class _LnBwdDesc(_c.Structure):
    """K3MLnBwdDesc of include/k3m.h"""
    _fields_ = [("dy", _vp), ("x", _vp), ("gamma", _vp), ("mean", _vp), ("rstd", _vp), ("dx", _vp), ("dxm", _vp),
                ("dgamma", _vp), ("dbeta", _vp), ("dbias", _vp), ("M", _i), ("H", _i), ("rng", _vp), ("site_out", _u),
                ("p_out", _f), ("site_in", _u), ("p_in", _f)]


class CudaOps(object):
    def __init__(self, device=None, act_dtype=torch.bfloat16):
        """`act_dtype=torch.float32` selects the fp32 CHECK build of the library (tests only: every activation and
        weight in fp32, CUDA-core GEMM / attention); the product is bf16."""
        if not torch.cuda.is_available():
            raise _lib.K3MError("k3m_b200 needs a CUDA device (sm_100a); there is no CPU path")
        assert act_dtype in (torch.bfloat16, torch.float32)
        self.adt = act_dtype
        self.lib = _lib.load() if act_dtype == torch.bfloat16 else _lib.load_f32()
        self.device = torch.device(device if device is not None else "cuda")
        self.launches = 0
        # profiling aid (bench.py): when a list, every launch is bracketed by CUDA events on the
        # launching stream and (entry point, work, event pair) is appended; never on in the product path
        self.timing = None
        self.record = None          # {entry point: [argument tuples]} while a measurement collects launches
        self._work = 0.0
        self._bytes = 0.0           # algorithmic HBM bytes of the next launch (set by HBM-bound ops while profiling)
        self._tag = ""
        self._side = []
        # auxiliary stream per launching stream for off-critical-path work (weight gradients):
        # stream handle -> [aux stream, tensors kept alive until the join, pending flag]
        self._aux = {}
        import os
        self.use_aux = os.environ.get("K3M_WGRAD_STREAM", "1") != "0"
        # the dgrad chain is the critical path of backward: its streams outrank the auxiliary (weight-gradient)
        # and optimizer streams, whose kernels then fill the SMs the chain leaves idle instead of delaying it
        self.hi_prio = -1 if os.environ.get("K3M_STREAM_PRIO", "0") != "0" else 0
        # grouped GEMM launches (k3m_gemm_grouped_bf16); K3M_GEMM_GROUP=0 issues every problem as its own launch
        self.use_groups = os.environ.get("K3M_GEMM_GROUP", "1") != "0"
        self.use_parallel = os.environ.get("K3M_PARALLEL", "1") != "0"      # concurrent streams for independent launches
        # k3m_adamw_bg (a variant sized to co-reside with a GEMM CTA) measured SLOWER: GEMM + AdamW on two streams took
        # 3.6 ms together against 2.6 + 0.8 ms alone (tools/micro_overlap.py, profiles/r02_overlap_micro.txt) — a
        # co-resident CTA starves next to 18 spinning GEMM warps and its HBM traffic slows the GEMM's L2 fills.  Off.
        self.use_adamw_bg = os.environ.get("K3M_ADAMW_BG", "0") != "0"
        self.skip = frozenset(x for x in os.environ.get("K3M_BENCH_SKIP", "").split(",") if x)

    def parallel(self, fns):
        """Run independent closures on concurrent CUDA streams (fork after the current stream, join back
        into it).  The K3M step has three independent connection layers per stage and independent text /
        image stacks between them; each alone leaves SMs idle (M = 2304 rows -> 54..108 tiles).  Works
        under CUDA-graph capture (the side streams join the capture).  Closures run in list order on the
        host, so dropout-site numbering is the same as in a sequential run."""
        fns = [f for f in fns if f is not None]
        if len(fns) <= 1 or self.timing is not None or not self.use_parallel:
            for f in fns:
                f()
            return
        main = torch.cuda.current_stream()
        while len(self._side) < len(fns) - 1:
            self._side.append(torch.cuda.Stream(device=self.device, priority=self.hi_prio))
        fork = torch.cuda.Event()
        fork.record(main)
        joins = []
        for i, f in enumerate(fns):
            if i == 0:
                f()
                continue
            st = self._side[i - 1]
            st.wait_event(fork)
            with torch.cuda.stream(st):
                f()
                self.join_aux()
                ev = torch.cuda.Event()
                ev.record(st)
            joins.append(ev)
        for ev in joins:
            main.wait_event(ev)

    def aux(self, fn, keep=()):
        """Run `fn` (kernels whose results nothing on the critical path reads: weight and bias gradients)
        on the auxiliary stream of the current stream, ordered after everything queued on it so far.
        `keep`: tensors `fn` reads; they stay referenced until `join_aux`, because the caching allocator
        would otherwise hand their memory to the next allocation on the launching stream while the
        auxiliary stream has not consumed them yet."""
        if self.timing is not None or not self.use_aux:
            fn()
            return
        cur = torch.cuda.current_stream()
        rec = self._aux.get(cur.cuda_stream)
        if rec is None:
            rec = self._aux[cur.cuda_stream] = [torch.cuda.Stream(device=self.device), [], False]
        ev = torch.cuda.Event()
        ev.record(cur)
        rec[0].wait_event(ev)
        with torch.cuda.stream(rec[0]):
            fn()
        rec[1].extend(keep)
        rec[2] = True

    def join_aux(self):
        """The current stream waits for its auxiliary stream (no-op if nothing is pending)."""
        cur = torch.cuda.current_stream()
        rec = self._aux.get(cur.cuda_stream)
        if rec is None or not rec[2]:
            return
        ev = torch.cuda.Event()
        ev.record(rec[0])
        cur.wait_event(ev)
        del rec[1][:]
        rec[2] = False

    def _st(self):
        return _vp(torch.cuda.current_stream().cuda_stream)

    def _call(self, name, *args):
        if self.skip and name in self.skip:      # measurement aid only (bench.py --skip): exposed time of a kernel family
            return
        self.launches += 1
        if self.record is not None and name in self.record:
            self.record[name].append((args, self._bytes))
        if self.timing is None:
            _lib.check(getattr(self.lib, name)(*args), name, self.lib)
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(getattr(self.lib, name)(*args), name, self.lib)
        e1.record()
        self.timing.append((name + self._tag, self._work, e0, e1))
        self._work = 0.0
        self._bytes = 0.0
        self._tag = ""

    @staticmethod
    def _nb(*tensors):
        """bytes of the given tensors (None skipped): the algorithmic HBM traffic of a row-wise kernel"""
        return float(sum(t.numel() * t.element_size() for t in tensors if t is not None))

    def _drop(self, drop, rng):
        if drop is None:
            return _vp(0), _u(0), _f(0.0)
        return _p(rng), _u(drop[1]), _f(drop[0])

    # ---- GEMM ------------------------------------------------------------------------------
    def gemm(self, a, b, layout, out=None, out_dtype=None, bias=None, act=ACT_NONE, aux_in=None,
             aux_out=None, addend=None, drop=None, rng=None, accumulate=False):
        assert a.dtype == self.adt and b.dtype == self.adt
        if layout == "nt":
            M, K = a.shape
            N = b.shape[0]
            assert b.shape[1] == K
        elif layout == "nn":
            M, K = a.shape
            N = b.shape[1]
            assert b.shape[0] == K
        else:
            K, M = a.shape
            N = b.shape[1]
            assert b.shape[0] == K
        if out is None:
            out = torch.empty(M, N, dtype=out_dtype or self.adt, device=a.device)
            accumulate = False
        assert tuple(out.shape) == (M, N)
        r, s, p = self._drop(drop, rng)
        self._work = 2.0 * M * N * K
        if self.timing is not None:
            self._tag = " %s %dx%dx%d%s%s%s%s%s" % (layout, M, N, K, " bias" if bias is not None else "", " act%d" % act if act else "",
                                                   " drop" if drop is not None else "", " add" if addend is not None else "",
                                                   " acc" if accumulate else (" f32" if out.dtype == torch.float32 else ""))
        self._call("k3m_gemm_bf16", _p(a), _i(_ld(a)), _p(b), _i(_ld(b)), _i(_LAYOUT[layout]),
                   _i(M), _i(N), _i(K), _p(out), _i(_ld(out)), _i(_isf32(out)), _i(1 if accumulate else 0),
                   _p(bias), _i(act), _p(aux_in), _i(_ld(aux_in)), _p(aux_out), _i(_ld(aux_out)),
                   _p(addend), _i(_ld(addend)), _i(_isf32(addend) if addend is not None else 0),
                   r, s, p, self._st())
        return out

    def gemm_group(self, calls, group_f32_store=False):
        """Several independent GEMMs of ONE layout as one persistent launch (`k3m_gemm_grouped_bf16`): `calls` is a
        list of dicts with the keyword arguments of `gemm` (a, b, layout, out, out_dtype, bias, act, aux_in, aux_out,
        addend, drop, rng, accumulate).  Returns the list of outputs.  Small problems (the three connection layers
        of a stage, a text and an image layer, a stage's weight gradients) then fill whole waves of CTA pairs
        together instead of each paying its own launch latency with half the SMs idle."""
        n = len(calls)

        def f32_store(c):        # fp32 outputs that do not accumulate: issued singly unless asked (`group_f32_store`) —
            # the step's few such problems (fusion gates) measured no faster grouped (profiles/r03o_fuse_group_ab.txt)
            if self.adt == torch.float32 or group_f32_store:
                return False
            o = c.get("out")
            dt = o.dtype if o is not None else (c.get("out_dtype") or self.adt)
            return dt == torch.float32 and not (o is not None and c.get("accumulate", False))
        if n == 1 or not self.use_groups or any(f32_store(c) for c in calls):
            return [self.gemm(**c) for c in calls]
        outs = []
        for i0 in range(0, n, 8):
            chunk = calls[i0:i0 + 8]
            if len(chunk) == 1:
                outs.append(self.gemm(**chunk[0]))
                continue
            arr = (_GemmDesc * len(chunk))()
            work, tags = 0.0, []
            for d, c in zip(arr, chunk):
                a, b, layout = c["a"], c["b"], c["layout"]
                assert a.dtype == self.adt and b.dtype == self.adt and layout == chunk[0]["layout"]
                if layout == "nt":
                    (M, K), N = a.shape, b.shape[0]
                    assert b.shape[1] == K
                elif layout == "nn":
                    (M, K), N = a.shape, b.shape[1]
                    assert b.shape[0] == K
                else:
                    (K, M), N = a.shape, b.shape[1]
                    assert b.shape[0] == K
                out, accumulate = c.get("out"), bool(c.get("accumulate", False))
                if out is None:
                    out = torch.empty(M, N, dtype=c.get("out_dtype") or self.adt, device=a.device)
                    accumulate = False
                assert tuple(out.shape) == (M, N)
                bias, aux_in, aux_out, addend = c.get("bias"), c.get("aux_in"), c.get("aux_out"), c.get("addend")
                drop, rng = c.get("drop"), c.get("rng")
                d.a, d.lda, d.b, d.ldb, d.layout = a.data_ptr(), _ld(a), b.data_ptr(), _ld(b), _LAYOUT[layout]
                d.M, d.N, d.K = M, N, K
                d.out, d.ldc, d.out_f32, d.accumulate = out.data_ptr(), _ld(out), _isf32(out), 1 if accumulate else 0
                d.bias = bias.data_ptr() if bias is not None else None
                d.act = c.get("act", ACT_NONE)
                d.aux_in, d.ld_aux_in = (aux_in.data_ptr(), _ld(aux_in)) if aux_in is not None else (None, 0)
                d.aux_out, d.ld_aux_out = (aux_out.data_ptr(), _ld(aux_out)) if aux_out is not None else (None, 0)
                d.addend, d.ld_add = (addend.data_ptr(), _ld(addend)) if addend is not None else (None, 0)
                d.addend_f32 = _isf32(addend) if addend is not None else 0
                if drop is not None:
                    d.rng, d.drop_site, d.drop_p = rng.data_ptr(), drop[1], drop[0]
                else:
                    d.rng, d.drop_site, d.drop_p = None, 0, 0.0
                work += 2.0 * M * N * K
                tags.append("%dx%dx%d" % (M, N, K))
                outs.append(out)
            self._work = work
            if self.timing is not None:
                self._tag = " %s group[%s]" % (chunk[0]["layout"], " ".join(tags))
            self._call("k3m_gemm_grouped_bf16", arr, _i(len(chunk)), self._st())
        return outs

    def gemm_last_variant(self):
        return self.lib.k3m_gemm_last_variant().decode()

    def colsum_group(self, items):
        """items: [(x, out)] -> out += column sums of x, one launch (`k3m_colsum_grouped`)"""
        ok = all(x.dtype == self.adt and x.dtype != torch.float32 and x.shape[1] % 8 == 0 and _ld(x) % 8 == 0
                 and x.data_ptr() % 16 == 0 for x, _ in items)
        if len(items) == 1 or not self.use_groups or not ok:
            for x, out in items:
                self.colsum(x, out)
            return
        for i0 in range(0, len(items), 8):
            chunk = items[i0:i0 + 8]
            arr = (_ColsumDesc * len(chunk))()
            for d, (x, out) in zip(arr, chunk):
                d.x, d.ld, d.out, d.M, d.N = x.data_ptr(), _ld(x), out.data_ptr(), x.shape[0], x.shape[1]
            if self.timing is not None:
                self._bytes = sum(self._nb(x, out) for x, out in chunk)
            self._call("k3m_colsum_grouped", arr, _i(len(chunk)), self._st())

    def colsum(self, x, out):
        if self.timing is not None:
            self._bytes = self._nb(x, out)
        self._call("k3m_colsum", _p(x), _i(_ld(x)), _i(_isf32(x)), _p(out), _i(x.shape[0]), _i(x.shape[1]),
                   self._st())

    # ---- LayerNorm -------------------------------------------------------------------------
    def ln_fwd(self, x, gamma, beta, out=None, drop=None, rng=None, eps=1e-12):
        M, H = x.shape
        assert x.is_contiguous()
        if out is None:
            out = torch.empty(M, H, dtype=self.adt, device=x.device)
        assert out.is_contiguous()
        mean = torch.empty(M, dtype=torch.float32, device=x.device)
        rstd = torch.empty(M, dtype=torch.float32, device=x.device)
        r, s, p = self._drop(drop, rng)
        if self.timing is not None:
            self._bytes = self._nb(x, out, mean, rstd, gamma, beta)
        self._call("k3m_ln_fwd", _p(x), _p(gamma), _p(beta), _p(out), _p(mean), _p(rstd), _i(M), _i(H),
                   _f(eps), r, s, p, self._st())
        return out, mean, rstd

    def ln_fwd_group(self, items, eps=1e-12):
        """Independent LayerNorms as one launch (`k3m_ln_fwd_grouped`): items = dicts with the arguments of `ln_fwd`
        (x, gamma, beta, out, drop, rng).  -> [(y, mean, rstd)]."""
        if len(items) == 1 or not self.use_groups:
            return [self.ln_fwd(eps=eps, **it) for it in items]
        res, nbytes = [], 0.0
        arr = (_LnFwdDesc * len(items))()
        for d, it in zip(arr, items):
            x, out, drop = it["x"], it.get("out"), it.get("drop")
            M, H = x.shape
            assert x.is_contiguous()
            if out is None:
                out = torch.empty(M, H, dtype=self.adt, device=x.device)
            assert out.is_contiguous()
            mean = torch.empty(M, dtype=torch.float32, device=x.device)
            rstd = torch.empty(M, dtype=torch.float32, device=x.device)
            d.x, d.gamma, d.beta, d.y = x.data_ptr(), it["gamma"].data_ptr(), it["beta"].data_ptr(), out.data_ptr()
            d.mean, d.rstd, d.M, d.H = mean.data_ptr(), rstd.data_ptr(), M, H
            d.rng, d.site, d.drop_p = (it["rng"].data_ptr(), drop[1], drop[0]) if drop is not None else (None, 0, 0.0)
            nbytes += self._nb(x, out, mean, rstd, it["gamma"], it["beta"])
            res.append((out, mean, rstd))
        if self.timing is not None:
            self._bytes = nbytes
        self._call("k3m_ln_fwd_grouped", arr, _i(len(items)), _f(eps), self._st())
        return res

    def ln_bwd_group(self, items):
        """Independent LayerNorm backwards as one launch: items = dicts with the arguments of `ln_bwd`.
        -> [(dx, dxm or dx)]."""
        if len(items) == 1 or not self.use_groups:
            return [self.ln_bwd(**it) for it in items]
        # one launch per register class of the kernel: hidden <= 768 runs two CTAs per SM (3 vectors per lane), hidden
        # 1024 only one — a mixed launch would run every problem at the wide kernel's occupancy (55 vs 22 us)
        wide = [i for i, it in enumerate(items) if it["x"].shape[1] > 768]
        if 0 < len(wide) < len(items):
            res = [None] * len(items)
            for idx in (wide, [i for i in range(len(items)) if i not in wide]):
                for i, r in zip(idx, self.ln_bwd_group([items[i] for i in idx])):
                    res[i] = r
            return res
        res, nbytes = [], 0.0
        arr = (_LnBwdDesc * len(items))()
        for d, it in zip(arr, items):
            dy, x = it["dy"], it["x"]
            M, H = x.shape
            assert dy.is_contiguous() and x.is_contiguous()
            drop_out, drop_in, dbias = it.get("drop_out"), it.get("drop_in"), it.get("dbias")
            dx = torch.empty(M, H, dtype=self.adt, device=x.device)
            dxm = torch.empty(M, H, dtype=self.adt, device=x.device) if drop_in is not None else None
            po = drop_out if drop_out is not None else (0.0, 0)
            pi = drop_in if drop_in is not None else (0.0, 0)
            d.dy, d.x, d.gamma, d.mean, d.rstd = dy.data_ptr(), x.data_ptr(), it["gamma"].data_ptr(), it["mean"].data_ptr(), it["rstd"].data_ptr()
            d.dx, d.dxm = dx.data_ptr(), (dxm.data_ptr() if dxm is not None else None)
            d.dgamma, d.dbeta = it["dgamma"].data_ptr(), it["dbeta"].data_ptr()
            d.dbias = dbias.data_ptr() if dbias is not None else None
            d.M, d.H = M, H
            d.rng = it["rng"].data_ptr() if it.get("rng") is not None else None
            d.site_out, d.p_out, d.site_in, d.p_in = po[1], po[0], pi[1], pi[0]
            nbytes += self._nb(dy, x, dx, dxm, it["mean"], it["rstd"], it["gamma"])
            res.append((dx, dxm if dxm is not None else dx))
        if self.timing is not None:
            self._bytes = nbytes
        self._call("k3m_ln_bwd_grouped", arr, _i(len(items)), self._st())
        return res

    def ln_bwd(self, dy, x, gamma, mean, rstd, dgamma, dbeta, drop_out=None, drop_in=None, rng=None,
               dbias=None, want_masked=False):
        M, H = x.shape
        assert dy.is_contiguous() and x.is_contiguous()
        dx = torch.empty(M, H, dtype=self.adt, device=x.device)
        dxm = torch.empty(M, H, dtype=self.adt, device=x.device) if drop_in is not None else None
        po = drop_out if drop_out is not None else (0.0, 0)
        pi = drop_in if drop_in is not None else (0.0, 0)
        if self.timing is not None:
            self._bytes = self._nb(dy, x, dx, dxm, mean, rstd, gamma)
        self._call("k3m_ln_bwd", _p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dx), _p(dxm), _p(dgamma),
                   _p(dbeta), _p(dbias), _i(M), _i(H), _p(rng), _u(po[1]), _f(po[0]), _u(pi[1]), _f(pi[0]),
                   self._st())
        return dx, (dxm if dxm is not None else dx)

    # ---- attention -------------------------------------------------------------------------
    def attn_fwd(self, qbuf, q_off, kvbuf, k_off, v_off, mask_add, B, H, Lq, Lk, d, drop=None, rng=None,
                 out=None):
        if out is None:
            out = torch.empty(B * Lq, H * d, dtype=self.adt, device=qbuf.device)
        lse = torch.empty(2, B * H * Lq, dtype=torch.float32, device=qbuf.device)   # row 1: backward scratch
        r, s, p = self._drop(drop, rng)
        if self.timing is not None:
            self._tag, self._work = " %dx%dx%d h%d" % (Lq, Lk, d, H), 4.0 * B * H * Lq * Lk * d
        self._call("k3m_attn_fwd", _p(qbuf), _i(_ld(qbuf)), _i(q_off), _p(kvbuf), _i(_ld(kvbuf)), _i(k_off),
                   _i(v_off), _p(mask_add), _i(B), _i(H), _i(Lq), _i(Lk), _i(d), _p(out), _i(_ld(out)),
                   _p(lse), r, s, p, self._st())
        return out, lse

    def attn_bwd(self, dctx, ctx, qbuf, q_off, kvbuf, k_off, v_off, mask_add, lse, dqbuf, dkvbuf, B, H, Lq,
                 Lk, d, drop=None, rng=None):
        r, s, p = self._drop(drop, rng)
        if self.timing is not None:
            self._tag, self._work = " %dx%dx%d h%d" % (Lq, Lk, d, H), 10.0 * B * H * Lq * Lk * d
        self._call("k3m_attn_bwd", _p(dctx), _i(_ld(dctx)), _p(qbuf), _i(_ld(qbuf)), _i(q_off), _p(kvbuf),
                   _i(_ld(kvbuf)), _i(k_off), _i(v_off), _p(mask_add), _p(lse), _p(dqbuf), _i(_ld(dqbuf)),
                   _p(dkvbuf), _i(_ld(dkvbuf)), _i(B), _i(H), _i(Lq), _i(Lk), _i(d), r, s, p, self._st())

    # ---- embeddings ------------------------------------------------------------------------
    def embed_fwd(self, ids, type_ids, word, pos, typ, L, out=None):
        M, H = ids.numel(), word.shape[1]
        if out is None:
            out = torch.empty(M, H, dtype=self.adt, device=word.device)
        if self.timing is not None:
            self._bytes = 2.0 * self._nb(out) + ids.numel() * 16.0
        self._call("k3m_embed_fwd", _p(ids), _p(type_ids), _p(word), _p(pos), _p(typ), _p(out), _i(M), _i(L),
                   _i(H), self._st())
        return out

    def embed_bwd(self, dx, ids, type_ids, L, dword, dpos, dtyp):
        M, H = dx.shape
        if self.timing is not None:      # read dx, fp32 read-modify-write of the touched word / position / type rows
            self._bytes = self._nb(dx) + 3 * 8.0 * M * H
        self._call("k3m_embed_bwd", _p(dx), _p(ids), _p(type_ids), _p(dword), _p(dpos), _p(dtyp), _i(M), _i(L),
                   _i(H), self._st())

    def loc_embed_fwd(self, loc, w, b, out=None):
        M, N = loc.shape[0], w.shape[0]
        assert loc.dtype == torch.float32 and loc.shape[1] == 5 and loc.is_contiguous()
        if out is None:
            out = torch.empty(M, N, dtype=self.adt, device=loc.device)
        self._call("k3m_loc_embed_fwd", _p(loc), _p(w), _p(b), _p(out), _i(M), _i(N), self._st())
        return out

    def loc_embed_bwd(self, dy, loc, dw, db):
        M, N = dy.shape
        assert dy.is_contiguous()
        self._call("k3m_loc_embed_bwd", _p(dy), _p(loc), _p(dw), _p(db), _i(M), _i(N), self._st())

    # ---- fusion ----------------------------------------------------------------------------
    def relu_cat3(self, a, b, c, out=None):
        M, H = a.shape
        if out is None:
            out = torch.empty(M, 3 * H, dtype=self.adt, device=a.device)
        if self.timing is not None:
            self._bytes = 2.0 * self._nb(out)
        self._call("k3m_relu_cat3", _p(a), _i(_ld(a)), _p(b), _i(_ld(b)), _p(c), _i(_ld(c)), _p(out), _i(M),
                   _i(H), self._st())
        return out

    def fuse_select_fwd(self, cat, gates, noise=None, site=0, rng=None, out=None):
        M, H = cat.shape[0], cat.shape[1] // 3
        assert gates.dtype == torch.float32 and gates.is_contiguous() and cat.is_contiguous()
        if out is None:
            out = torch.empty(M, H, dtype=self.adt, device=cat.device)
        if self.timing is not None:
            self._bytes = self._nb(cat, gates, out)
        self._call("k3m_fuse_select_fwd", _p(cat), _p(gates), _p(noise), _p(rng), _u(site), _p(out),
                   _i(_ld(out)), _i(M), _i(H), self._st())
        return out

    def fuse_select_bwd(self, dout, cat, gates, noise=None, site=0, rng=None):
        M, H = cat.shape[0], cat.shape[1] // 3
        dg = torch.empty(M, 3 * H, dtype=self.adt, device=cat.device)
        dc = torch.empty(M, 3 * H, dtype=self.adt, device=cat.device)
        if self.timing is not None:
            self._bytes = self._nb(cat, gates, dg, dc) + 2.0 * M * H
        self._call("k3m_fuse_select_bwd", _p(dout), _i(_ld(dout)), _p(cat), _p(gates), _p(noise), _p(rng),
                   _u(site), _p(dg), _p(dc), _i(M), _i(H), self._st())
        return dg, dc

    def gather_rows(self, src, idx):
        n, H = idx.numel(), src.shape[1]
        out = torch.empty(n, H, dtype=src.dtype, device=src.device)
        self._call("k3m_gather_rows", _p(src), _i(_ld(src)), _p(idx), _p(out), _i(n), _i(H), self._st())
        return out

    def scatter_add_rows(self, dsrc, idx, dst):
        assert dsrc.is_contiguous()
        self._call("k3m_scatter_add_rows", _p(dsrc), _p(idx), _p(dst), _i(_ld(dst)), _i(idx.numel()), _i(dsrc.shape[1]),
                   self._st())

    def gate_mul(self, cat, alpha):
        assert cat.is_contiguous() and alpha.is_contiguous() and cat.shape == alpha.shape
        out = torch.empty_like(cat)
        self._call("k3m_gate_mul", _p(cat), _p(alpha), _p(out), _ll(cat.numel()), self._st())
        return out

    def gate_bwd(self, dgated, cat, alpha):
        assert dgated.is_contiguous() and cat.is_contiguous() and alpha.is_contiguous()
        dcat, dscore = torch.empty_like(cat), torch.empty_like(cat)
        self._call("k3m_gate_bwd", _p(dgated), _p(cat), _p(alpha), _p(dcat), _p(dscore), _ll(cat.numel()), self._st())
        return dcat, dscore

    def relu_bwd_split3(self, dcat, cat, da, db, dc):
        M, H = da.shape
        if self.timing is not None:      # read dcat, cat; read-modify-write the three gradient slices
            self._bytes = self._nb(dcat, cat) + 2 * 3 * 2.0 * M * H
        self._call("k3m_relu_bwd_split3", _p(dcat), _p(cat), _p(da), _i(_ld(da)), _p(db), _i(_ld(db)), _p(dc),
                   _i(_ld(dc)), _i(M), _i(H), self._st())

    def lincomb3(self, a, b, c, wa, wb, wc, out=None, out_dtype=None):
        rows, cols = a.shape
        if out is None:
            out = torch.empty(rows, cols, dtype=out_dtype or a.dtype, device=a.device)
        self._call("k3m_lincomb3", _p(a), _i(_ld(a)), _p(b), _i(_ld(b)), _p(c), _i(_ld(c)), _f(wa), _f(wb),
                   _f(wc), _i(_isf32(a)), _p(out), _i(_ld(out)), _i(_isf32(out)), _i(rows), _i(cols), self._st())
        return out

    def axpy(self, x, y, alpha=1.0):
        rows, cols = x.shape
        if self.timing is not None:
            self._bytes = self._nb(x) + 2.0 * rows * cols * y.element_size()
        self._call("k3m_axpy", _p(x), _i(_ld(x)), _i(_isf32(x)), _p(y), _i(_ld(y)), _i(_isf32(y)), _f(alpha),
                   _i(rows), _i(cols), self._st())

    def cast(self, x, dtype, out=None):
        assert x.is_contiguous()
        y = out if out is not None else torch.empty(x.shape, dtype=dtype, device=x.device)
        assert y.is_contiguous() and y.dtype == dtype and y.numel() == x.numel()
        if self.timing is not None:
            self._bytes = self._nb(x, y)
        self._call("k3m_cast", _p(x), _i(_isf32(x)), _p(y), _i(_isf32(y)), _ll(x.numel()), self._st())
        return y

    def scale_(self, x, alpha):
        assert x.dtype == torch.float32 and x.is_contiguous()
        self._call("k3m_scale", _p(x), _f(alpha), _ll(x.numel()), self._st())

    def scale_dev_(self, x, alpha):
        """x *= alpha[0] with alpha an fp32 device tensor."""
        assert alpha.dtype == torch.float32 and alpha.numel() == 1
        x2 = x if x.dim() == 2 else x.reshape(1, -1)
        self._call("k3m_scale_dev", _p(x2), _i(_ld(x2)), _i(_isf32(x2)), _p(alpha), _i(x2.shape[0]), _i(x2.shape[1]),
                   self._st())

    def act_bwd(self, dy, pre, act):
        assert act == ACT_DGELU and dy.is_contiguous() and pre.is_contiguous()
        dx = torch.empty_like(dy)
        self._call("k3m_dgelu_mul", _p(dy), _p(pre), _p(dx), _ll(dy.numel()), self._st())
        return dx

    def mask_additive(self, mask):
        assert mask.dtype == torch.int64 and mask.is_contiguous()
        out = torch.empty(mask.shape, dtype=torch.float32, device=mask.device)
        self._call("k3m_mask_additive", _p(mask), _p(out), _i(mask.numel()), self._st())
        return out

    # ---- pooling ---------------------------------------------------------------------------
    def mean_pool_fwd(self, seq, B, L, out_dtype=torch.float32):
        H = seq.shape[1]
        out = torch.empty(B, H, dtype=out_dtype, device=seq.device)
        self._call("k3m_mean_pool_fwd", _p(seq), _i(_ld(seq)), _p(out), _i(_isf32(out)), _i(B), _i(L), _i(H),
                   self._st())
        return out

    def mean_pool_bwd(self, dpool, dseq, B, L):
        assert dpool.dtype == torch.float32 and dpool.is_contiguous()
        self._call("k3m_mean_pool_bwd", _p(dpool), _p(dseq), _i(_ld(dseq)), _i(B), _i(L), _i(dseq.shape[1]),
                   self._st())

    # ---- structure aggregator + LPM ----------------------------------------------------------
    def pair_gather_fwd(self, c_init, seq, pair_item, pair_rows):
        P, H = pair_item.numel(), c_init.shape[1]
        X = torch.empty(P, 3 * H, dtype=self.adt, device=seq.device)
        PV = torch.empty(P, 2 * H, dtype=torch.float32, device=seq.device)
        self._call("k3m_pair_gather_fwd", _p(c_init), _p(seq), _i(_ld(seq)), _p(pair_item), _p(pair_rows), _p(X),
                   _p(PV), _i(P), _i(H), self._st())
        return X, PV

    def pair_gather_bwd(self, dx, dpv, pair_item, pair_rows, dc_init, dseq):
        P, H = pair_item.numel(), dc_init.shape[1]
        scratch = torch.zeros(dseq.shape[0], H, dtype=torch.float32, device=dseq.device)
        self._call("k3m_pair_gather_bwd", _p(dx), _p(dpv), _p(pair_item), _p(pair_rows), _p(dc_init), _p(scratch),
                   _i(P), _i(H), self._st())
        self.axpy(scratch, dseq, 1.0)

    def struct_fallback_fwd(self, t, c_init, fb):
        self._call("k3m_struct_fallback_fwd", _p(t), _p(c_init), _p(fb), _i(fb.shape[0]), _i(t.shape[1]), self._st())

    def struct_fallback_bwd(self, dt, dc_init, fb):
        self._call("k3m_struct_fallback_bwd", _p(dt), _p(dc_init), _p(fb), _i(fb.shape[0]), _i(dt.shape[1]), self._st())

    def struct_attn_fwd(self, t, w2, b2, seg):
        B, H = seg.shape[0], t.shape[1]
        maxn = self._maxn(seg)
        s = torch.empty(B, H, dtype=torch.float32, device=t.device)
        att = torch.zeros(B, maxn, dtype=torch.float32, device=t.device)
        self._call("k3m_struct_attn_fwd", _p(t), _p(w2), _p(b2), _p(seg), _p(s), _p(att), _i(B), _i(maxn), _i(H),
                   self._st())
        return s, att

    def _maxn(self, seg):
        n = getattr(seg, "_k3m_maxn", None)
        if n is None:
            raise _lib.K3MError("struct plan tensors must come from CudaOps.plan_to_device")
        return n

    def plan_to_device(self, plan, device):
        """numpy plan (struct_plan.build_struct_plan) -> device int32 tensors; the longest pair range
        is recorded on the host so no device->host sync is needed later."""
        out = {}
        for k, v in plan.items():
            if k == "cap_t":
                out["mlm_cap_t"] = int(v)
                continue
            out[k] = torch.from_numpy(v).to(device, non_blocking=True)
        seg = plan["seg"]
        out["seg"]._k3m_maxn = int((seg[:, 1] - seg[:, 0]).max())
        return out

    def struct_attn_bwd(self, ds, t, att, w2, seg, dt, dw2, db2):
        B, H = seg.shape[0], t.shape[1]
        self._call("k3m_struct_attn_bwd", _p(ds), _p(t), _p(att), _p(w2), _p(seg), _p(dt), _p(dw2), _p(db2),
                   _i(B), _i(att.shape[1]), _i(H), self._st())

    def lpm_fwd_bwd(self, c_final, pv, terms, margin, dc_final, dpv, n_valid=None):
        loss = torch.zeros(1, dtype=torch.float32, device=c_final.device)
        self._call("k3m_lpm_fwd_bwd", _p(c_final), _p(pv), _p(terms), _i(terms.shape[0]), _p(n_valid),
                   _i(c_final.shape[1]),
                   _f(margin), _p(loss), _p(dc_final), _p(dpv), self._st())
        return loss

    # ---- losses ----------------------------------------------------------------------------
    def count_valid(self, labels, ignore):
        out = torch.empty(1, dtype=torch.float32, device=labels.device)
        self._call("k3m_count_labels", _p(labels), _i(labels.numel()), _ll(ignore), _i(0), _p(out), self._st())
        return out

    def count_eq(self, labels, value):
        out = torch.empty(1, dtype=torch.float32, device=labels.device)
        self._call("k3m_count_labels", _p(labels), _i(labels.numel()), _ll(value), _i(1), _p(out), self._st())
        return out

    def xent_fwd_bwd(self, logits, labels, count, n_cols, dlogits=None, grad_scale=1.0):
        assert dlogits is None or dlogits.data_ptr() == logits.data_ptr(), "gradient is written in place"
        loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
        if self.timing is not None:      # logits read once, gradient written once (the row stays in L1/L2 between passes)
            self._bytes = 2.0 * self._nb(logits) if dlogits is not None else self._nb(logits)
        self._call("k3m_xent_fwd_bwd", _p(logits), _i(_ld(logits)), _p(labels), _p(count), _i(logits.shape[0]),
                   _i(n_cols), _f(grad_scale), _p(loss), _i(1 if dlogits is not None else 0), self._st())
        return loss

    def kl_fwd_bwd(self, logits, target, image_label, count, n_cols, Lv1, dlogits=None, grad_scale=1.0):
        assert logits.dtype == torch.float32 and target.dtype == torch.float32 and target.is_contiguous()
        assert dlogits is None or _ld(dlogits) == _ld(logits)
        loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
        if self.timing is not None:
            self._bytes = self._nb(logits, target, dlogits)
        self._call("k3m_kl_fwd_bwd", _p(logits), _i(_ld(logits)), _p(target), _p(image_label), _p(count),
                   _i(logits.shape[0] // Lv1), _i(Lv1), _i(n_cols), _f(grad_scale), _p(loss), _p(dlogits),
                   self._st())
        return loss

    def itm_loss(self, pool_t, pool_pv, pool_v, w, b, l1, l2, l3, drop=None, rng=None):
        B, H = pool_t.shape
        loss = torch.zeros(1, dtype=torch.float32, device=pool_t.device)
        r, s, p = self._drop(drop, rng)
        self._call("k3m_itm_loss", _p(pool_t), _p(pool_pv), _p(pool_v), _p(w), _p(b), _p(l1), _p(l2), _p(l3),
                   _i(B), _i(H), r, s, p, _p(loss), self._st())
        return loss

    def pair_cls_loss(self, logits, labels, dlogits=None, grad_scale=1.0):
        """[B, 2] fp32 logits -> (loss [1], probs [B, 2]); dlogits: bf16 [B, >=2] gets the gradient."""
        assert logits.dtype == torch.float32 and logits.shape[1] == 2
        B = logits.shape[0]
        loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
        probs = torch.empty(B, 2, dtype=torch.float32, device=logits.device)
        self._call("k3m_pair_cls_loss", _p(logits), _i(_ld(logits)), _p(labels), _i(B), _f(grad_scale), _p(loss),
                   _p(probs), _p(dlogits), _i(_ld(dlogits)), self._st())
        return loss, probs

    def dropout(self, x, drop, rng, out=None, out_dtype=None):
        """nn.Dropout outside a GEMM epilogue; `drop` = (p, site) or None (copy / cast only)."""
        rows, cols = x.shape
        if out is None:
            out = torch.empty(rows, cols, dtype=out_dtype or x.dtype, device=x.device)
        r, s, p = self._drop(drop, rng)
        self._call("k3m_dropout", _p(x), _i(_ld(x)), _i(_isf32(x)), _p(out), _i(_ld(out)), _i(_isf32(out)),
                   _i(rows), _i(cols), r, s, p, self._st())
        return out

    # ---- optimizer -------------------------------------------------------------------------
    def cast_zero(self, g32, out16):
        """fp32 gradient range -> bf16 send buffer, zeroing the fp32 range (data-parallel step)"""
        assert g32.dtype == torch.float32 and out16.dtype == torch.bfloat16 and g32.numel() == out16.numel()
        self._call("k3m_cast_zero", _p(g32), _p(out16), _ll(g32.numel()), self._st())

    def adamw_dp(self, p32, m, v, hyper, g_ptrs, w_ptrs, offset, g_mc=0, w_mc=0):
        """Gradient reduce-scatter + AdamW + weight all-gather over NVLink peer memory for the shard
        [offset, offset + p32.numel()) of the arena; g_ptrs / w_ptrs: every rank's bf16 gradient / weight arena base."""
        world = len(g_ptrs)
        ga = (_vp * world)(*[int(x) for x in g_ptrs])
        wa = (_vp * world)(*[int(x) for x in w_ptrs])
        if self.timing is not None:
            self._bytes = p32.numel() * (24.0 + 2.0 * world * 2)
        self._call("k3m_adamw_dp", _p(p32), _p(m), _p(v), _p(hyper), ga, wa, _i(world), _ll(int(offset)),
                   _ll(p32.numel()), _vp(int(g_mc) or None), _vp(int(w_mc) or None), self._st())

    def gemm_set_sm_limit(self, n):
        self.lib.k3m_gemm_set_sm_limit(_i(int(n)))

    def adamw(self, p32, g32, m, v, p_act, hyper, zero_grad=True, background=False, g16=None):
        """`background`: the variant sized to co-reside with a running GEMM (buckets stepped under backward).
        `g16`: read the gradient from this bf16 buffer (all-reduced in bf16) instead of `g32`, which is zeroed."""
        if self.timing is not None:      # read p, g, m, v; write p, m, v, bf16 shadow, zeroed g: 34 B per parameter
            self._bytes = p32.numel() * (34.0 if zero_grad else 30.0)
        if g16 is not None:
            self._call("k3m_adamw_g16", _p(p32), _p(g16), _p(g32), _p(m), _p(v), _p(p_act), _p(hyper), _ll(p32.numel()),
                       _i(1 if zero_grad else 0), self._st())
            return
        self._call("k3m_adamw_bg" if (background and self.use_adamw_bg) else "k3m_adamw", _p(p32), _p(g32), _p(m), _p(v), _p(p_act), _p(hyper), _ll(p32.numel()),
                   _i(1 if zero_grad else 0), self._st())
