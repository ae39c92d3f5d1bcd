"""On-disk records -> GPU batch preprocessing (SURVEY §8 f N3 + N2): an LMDBSerializer-layout file is read back,
batched on the device and preprocessed by the CUDA kernels; every output equals the numpy restatement of the
reference's `BertPreprocessBatch` run on the same records and draws."""
import numpy as np
import pytest
import torch

from k3m_b200 import records as R
from oracle.preprocess_ref import preprocess_batch
from tests.test_preprocess import synth_raws
from tests.test_records_cpu import dumps, write_lmdb


class IdTokenizer(object):
    """Titles / PV strings of this test are space-separated token ids."""

    def encode(self, text):
        return [int(t) for t in text.split()]


@pytest.mark.gpu
def test_lmdb_records_through_gpu_preprocessing(tmp_path):
    from k3m_b200.preprocess import GpuBatchPreprocessor
    B, Lt, Lpv, Rg, F, T, P = 8, 36, 36, 36, 2048, 1601, 8
    rng = np.random.RandomState(11)
    raws, u_tok, rnd_tok, u_reg = synth_raws(rng, 2 * B, Lt, Lpv, Rg, F, T, P)
    items = []
    for i, r in enumerate(raws):
        n = r["boxes"].shape[0] if i != 3 else 0                      # one record without boxes
        items.append(["id%d" % i, " ".join(map(str, r["tokens"])), " ".join(map(str, r["tokens_pv"])), "c",
                      r["image_h"], r["image_w"], n, r["boxes"][:n], r["features"][:n], r["targets"][:n]])
    keys = ["{:08}".format(i).encode("ascii") for i in range(len(items))]
    kv = {k: dumps(it) for k, it in zip(keys, items)}
    kv[b"__keys__"] = dumps(keys)
    path = str(tmp_path / "train_feat.lmdb")
    write_lmdb(path, kv, leaf_cap=4)

    rf, tok = R.open_records(path), IdTokenizer()
    pre = GpuBatchPreprocessor(P)
    nb = 0
    for bi, raw in enumerate(R.iter_raw_batches(rf, tok, B, Lt, Lpv, Rg, device="cuda")):
        sl = slice(bi * B, (bi + 1) * B)
        draws = dict(u_tok=torch.from_numpy(np.ascontiguousarray(u_tok[sl], np.float64)).cuda(),
                     rnd_tok=torch.from_numpy(np.ascontiguousarray(rnd_tok[sl], np.int64)).cuda(),
                     u_reg=torch.from_numpy(np.ascontiguousarray(u_reg[sl], np.float64)).cuda())
        got = {k: v.cpu().numpy() for k, v in pre(raw, draws).items()}
        samples = [R.record_to_sample(it, tok) for it in items[sl]]
        want = preprocess_batch(samples, u_tok[sl], rnd_tok[sl], u_reg[sl], Lt, Lpv, Rg, P)
        for k, w in want.items():
            assert got[k].dtype == w.dtype and np.array_equal(got[k], w), (bi, k)
        nb += 1
    assert nb == 2
