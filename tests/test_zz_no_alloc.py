"""k3m_prep_regions_noalloc (include/k3m.h): the region preprocessing with its scratch (zero_flag / count0) in a
static device array of the library instead of caller memory.  It has no counterpart in the reference; it exists for
callers that cannot allocate.  Checked here through the C ABI against k3m_prep_regions on the same inputs — bit for
bit, the bar of every preprocessing test (tests/test_preprocess.py) — on two calls in a row with different batches
(the second call finds the first call's scratch contents), plus the size limit.

Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

_vp, _i = ctypes.c_void_p, ctypes.c_int


def _p(t):
    return _vp(t.data_ptr())


def _inputs(seed, B, R, F):
    rng = np.random.RandomState(seed)
    nb = rng.randint(1, R + 1, B).astype(np.int32)
    nb[0] = R
    w, h = rng.randint(200, 900, B).astype(np.float64), rng.randint(200, 900, B).astype(np.float64)
    x1, y1 = rng.rand(B, R) * w[:, None] * 0.6, rng.rand(B, R) * h[:, None] * 0.6
    boxes = np.stack([x1, y1, x1 + rng.rand(B, R) * w[:, None] * 0.4, y1 + rng.rand(B, R) * h[:, None] * 0.4], 2)
    valid = np.arange(R)[None, :] < nb[:, None]
    boxes = (boxes * valid[:, :, None]).astype(np.float32)
    whp = np.stack([w, h, w * h], 1).astype(np.float32)
    # 40% of the regions fall under the 0.15 selection threshold so that zeroed / kept / unselected all occur
    u = np.where(rng.rand(B, R) < 0.4, rng.rand(B, R) * 0.15, 0.15 + rng.rand(B, R) * 0.85)
    feat = (rng.randn(B, R, F) * valid[:, :, None]).astype(np.float32)
    mk = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()      # noqa: E731
    return dict(boxes=mk(boxes), num_boxes=mk(nb), whp=mk(whp), u=mk(u.astype(np.float64)), feat=mk(feat))


def _outputs(B, R, F, fill):
    # `fill`: the outputs start from different garbage in the two arms, so that a skipped write shows
    e = lambda *s, dtype: torch.full(s, fill, dtype=dtype, device="cuda")        # noqa: E731
    return dict(image_feat=e(B, R + 1, F, dtype=torch.float32), image_loc=e(B, R + 1, 5, dtype=torch.float32),
                image_mask=e(B, R + 1, dtype=torch.int64), image_label=e(B, R, dtype=torch.int64))


def _run(lib, x, B, R, F, noalloc):
    from k3m_b200 import _lib
    o = _outputs(B, R, F, 7 if noalloc else 3)
    st = _vp(torch.cuda.current_stream().cuda_stream)
    head = (_p(x["boxes"]), _p(x["num_boxes"]), _p(x["whp"]), _p(x["u"]), _p(x["feat"]), _i(B), _i(R), _i(F),
            _p(o["image_feat"]), _p(o["image_loc"]), _p(o["image_mask"]), _p(o["image_label"]))
    if noalloc:
        _lib.check(lib.k3m_prep_regions_noalloc(*head, st), "k3m_prep_regions_noalloc")
    else:
        zero_flag = torch.empty(B, R, dtype=torch.uint8, device="cuda")
        count0 = torch.empty(B, dtype=torch.int32, device="cuda")
        _lib.check(lib.k3m_prep_regions(*head, _p(zero_flag), _p(count0), st), "k3m_prep_regions")
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in o.items()}


@pytest.mark.gpu
def test_noalloc_entry_point_matches_prep_regions_on_two_calls_in_a_row():
    from k3m_b200 import _lib
    lib = _lib.load()
    # two different geometries back to back: the second call reuses scratch the first one filled (other strides,
    # a smaller batch), which must not leak into its result
    for seed, (B, R, F) in enumerate([(64, 36, 2048), (16, 100, 512)]):
        x = _inputs(seed, B, R, F)
        want = _run(lib, x, B, R, F, noalloc=False)
        before = torch.cuda.memory_allocated()
        got = _run(lib, x, B, R, F, noalloc=True)
        assert torch.cuda.memory_allocated() == before
        for k, w in want.items():
            assert got[k].dtype == w.dtype and np.array_equal(got[k], w), (seed, k)
        assert (want["image_label"] == 1).any() and (want["image_label"] == -1).any()
        assert np.isfinite(want["image_feat"]).all() and (want["image_feat"][:, 1:] == 0).all(2).any()


@pytest.mark.gpu
def test_noalloc_entry_point_rejects_sizes_beyond_the_static_scratch():
    from k3m_b200 import _lib
    lib = _lib.load()
    x = _inputs(5, 2, 4, 8)
    o = _outputs(2, 4, 8, 0)
    st = _vp(torch.cuda.current_stream().cuda_stream)
    for B, R in ((4097, 4), (4096, 257)):       # B > 4096; B * R > 1 MiB.  Rejected before anything is launched.
        rc = lib.k3m_prep_regions_noalloc(_p(x["boxes"]), _p(x["num_boxes"]), _p(x["whp"]), _p(x["u"]), _p(x["feat"]),
                                          _i(B), _i(R), _i(8), _p(o["image_feat"]), _p(o["image_loc"]),
                                          _p(o["image_mask"]), _p(o["image_label"]), st)
        assert rc == -1
        msg = lib.k3m_last_error_string().decode()
        assert "exceeds the static scratch" in msg and "k3m_prep_regions" in msg
    torch.cuda.synchronize()
