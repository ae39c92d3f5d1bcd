"""Record reader (SURVEY §8 f N3, `k3m_b200/records.py`): msgpack_numpy decoding against `msgpack` round trips of
the published encoding, the LMDB reader against a test-side writer of the same page layout (lmdb itself is not
installed: file-format parity is unpinned, see the module header), records -> raw samples against the reference
rules (`concept_cap_dataset_struc.py:564-606,963-968`)."""
import struct

import msgpack
import numpy as np
import pytest

from k3m_b200 import records as R

PS = 4096


# ---- msgpack_numpy `encode` as published (0.4.x): what tensorpack's `dumps` applies as `default=`
def np_encode(obj):
    if isinstance(obj, np.ndarray):
        if obj.dtype.kind == "V":
            kind, descr = b"V", obj.dtype.descr
        else:
            kind, descr = b"", obj.dtype.str
        return {b"nd": True, b"type": descr, b"kind": kind, b"shape": obj.shape, b"data": obj.tobytes()}
    if isinstance(obj, (np.bool_, np.number)):
        return {b"nd": False, b"type": obj.dtype.str, b"data": obj.tobytes()}
    if isinstance(obj, complex):
        return {b"complex": True, b"data": repr(obj)}
    return obj


def dumps(obj):
    return msgpack.packb(obj, default=np_encode, use_bin_type=True)


# ---- LMDB data-file writer (format version 1, 64-bit): meta x2, leaf / branch / overflow pages
def _page_hdr(pgno, flags, lower, upper):
    return struct.pack("<QHHHH", pgno, 0, flags, lower, upper)


def _fill_page(pgno, flags, nodes):
    """nodes: list of bytes (already even-sized); placed from the top of the page downwards like mdb_node_add."""
    page = bytearray(PS)
    upper, ptrs = PS, []
    for nd in nodes:
        upper -= len(nd)
        page[upper:upper + len(nd)] = nd
        ptrs.append(upper)
    lower = 16 + 2 * len(nodes)
    assert lower <= upper, "page overflow in the test writer"
    page[:16] = _page_hdr(pgno, flags, lower, upper)
    page[16:lower] = struct.pack("<%dH" % len(ptrs), *ptrs)
    return bytes(page)


def _even(b):
    return b + b"\0" * (len(b) & 1)


def write_lmdb(path, kv, leaf_cap=5, branch_cap=4, inline_max=1000, stale_meta=True):
    kv = sorted(kv.items())
    pages = {}                                       # pgno -> bytes
    nxt = [2]
    n_ovf = n_leaf = n_branch = 0

    def alloc(n=1):
        p = nxt[0]
        nxt[0] += n
        return p

    level = []                                       # (first key, pgno)
    for s in range(0, len(kv), leaf_cap):
        nodes = []
        for k, v in kv[s:s + leaf_cap]:
            if len(v) > inline_max:
                npg = (16 + len(v) + PS - 1) // PS
                pg = alloc(npg)
                blob = bytearray(npg * PS)
                blob[:16] = struct.pack("<QHHI", pg, 0, 0x04, npg)
                blob[16:16 + len(v)] = v
                for i in range(npg):
                    pages[pg + i] = bytes(blob[i * PS:(i + 1) * PS])
                n_ovf += npg
                nodes.append(_even(struct.pack("<HHHH", len(v) & 0xFFFF, len(v) >> 16, 0x01, len(k)) + k
                                   + struct.pack("<Q", pg)))
            else:
                nodes.append(_even(struct.pack("<HHHH", len(v) & 0xFFFF, len(v) >> 16, 0, len(k)) + k + v))
        pg = alloc()
        pages[pg] = _fill_page(pg, 0x02, nodes)
        n_leaf += 1
        level.append((kv[s][0], pg))
    depth = 1 if level else 0
    while len(level) > 1:
        up = []
        for s in range(0, len(level), branch_cap):
            nodes = []
            for i, (k, child) in enumerate(level[s:s + branch_cap]):
                k = b"" if i == 0 else k             # first branch key is implicit
                nodes.append(_even(struct.pack("<HHHH", child & 0xFFFF, (child >> 16) & 0xFFFF, child >> 32, len(k))
                                   + k))
            pg = alloc()
            pages[pg] = _fill_page(pg, 0x01, nodes)
            n_branch += 1
            up.append((level[s][0], pg))
        level = up
        depth += 1
    root = level[0][1] if level else (1 << 64) - 1

    def meta(pgno, txnid, root_, entries, depth_):
        m = bytearray(PS)
        m[:16] = _page_hdr(pgno, 0x08, 0, 0)
        free_db = struct.pack("<IHHQQQQQ", PS, 0, 0, 0, 0, 0, 0, (1 << 64) - 1)
        main_db = struct.pack("<IHHQQQQQ", 0, 0, depth_, n_branch, n_leaf, n_ovf, entries, root_)
        body = struct.pack("<IIQQ", 0xBEEFC0DE, 1, 0, 1 << 30) + free_db + main_db + struct.pack("<QQ", nxt[0] - 1, txnid)
        m[16:16 + len(body)] = body
        return bytes(m)

    # meta 0 carries an older, empty transaction: the reader must pick meta 1 (higher txnid)
    pages[0] = meta(0, 0, (1 << 64) - 1, 0, 0) if stale_meta else meta(0, 2, root, len(kv), depth)
    pages[1] = meta(1, 1, root, len(kv), depth)
    with open(path, "wb") as f:
        for p in range(nxt[0]):
            f.write(pages[p])
    return depth


class CharTokenizer(object):
    """Stands in for bert-base-chinese (one id per character, no specials) like pytorch_transformers 1.1.0's
    `encode`."""

    def encode(self, text):
        return [1000 + (ord(c) % 20000) for c in text]


def make_item(rng, i, n):
    return ["item%d" % i, "标题 %d" % i, "颜色:红;尺码:%d;" % i, "cat", 600 + i, 800 + i, n,
            (rng.random((n, 4)) * 500).astype(np.float32), rng.standard_normal((n, 64)).astype(np.float32),
            rng.random((n, 17)).astype(np.float32)]


def assert_item_equal(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if isinstance(y, np.ndarray):
            assert x.dtype == y.dtype and x.shape == y.shape and np.array_equal(x, y)
        else:
            assert x == y and type(x) is type(y)


def test_msgpack_numpy_roundtrip():
    rng = np.random.default_rng(0)
    for dt in (np.float32, np.float64, np.int64, np.int32, np.uint8, np.bool_, np.float16):
        a = (rng.random((3, 5)) * 100).astype(dt)
        b = R.decode_record(dumps(a))
        assert b.dtype == a.dtype and b.shape == a.shape and np.array_equal(a, b)
    assert R.decode_record(dumps(np.zeros((0, 4), np.float32))).shape == (0, 4)
    s = R.decode_record(dumps(np.float32(1.5)))
    assert s == np.float32(1.5) and s.dtype == np.float32
    assert R.decode_record(dumps(3 + 4j)) == 3 + 4j
    big = np.dtype(">f4")
    assert R.decode_record(dumps(np.arange(4, dtype=big))).dtype == big
    st = np.zeros(3, dtype=[("a", "<i4"), ("b", "<f8")])
    st["a"] = [1, 2, 3]
    got = R.decode_record(dumps(st))
    assert got.dtype == st.dtype and np.array_equal(got, st)
    # str keys (writers without use_bin_type) decode the same
    a = np.arange(6, dtype=np.int64).reshape(2, 3)
    old = msgpack.packb({"nd": True, "type": a.dtype.str, "kind": "", "shape": a.shape, "data": a.tobytes()},
                        use_bin_type=True)
    assert np.array_equal(R.decode_record(old), a)
    item = make_item(rng, 7, 5)
    assert_item_equal(R.decode_record(dumps(item)), item)
    assert_item_equal(R.decode_record(memoryview(dumps(item))), item)


@pytest.mark.parametrize("n,leaf_cap,branch_cap,depth", [(0, 5, 4, 0), (1, 5, 4, 1), (5, 5, 4, 1), (17, 5, 4, 2),
                                                         (97, 3, 3, 5)])
def test_lmdb_reader_tree_shapes(tmp_path, n, leaf_cap, branch_cap, depth):
    rng = np.random.default_rng(n)
    kv = {}
    for i in range(n):
        size = int(rng.integers(0, 3000)) if i % 3 else int(rng.integers(5000, 20000))   # inline and overflow
        kv[("%08d" % i).encode()] = rng.integers(0, 256, size, dtype=np.uint8).tobytes()
    path = str(tmp_path / "t.lmdb")
    assert write_lmdb(path, kv, leaf_cap, branch_cap) == depth
    with R.LMDBReader(path) as db:
        assert len(db) == n and db.page_size == PS
        assert [k for k, _ in db.items()] == sorted(kv)
        for k, v in db.items():
            assert bytes(v) == kv[k]
        for k in kv:
            assert bytes(db.get(k)) == kv[k]
        assert db.get(b"00000000x") is None and db.get(b"") is None and db.get(b"zzzz", 7) == 7


def test_lmdb_reader_meta_choice_and_errors(tmp_path):
    kv = {b"a": b"1", b"b": b"2" * 5000}
    p = str(tmp_path / "m.lmdb")
    write_lmdb(p, kv, stale_meta=False)              # meta 0 newer (txnid 2)
    with R.LMDBReader(p) as db:
        assert bytes(db.get(b"b")) == kv[b"b"]
    bad = tmp_path / "bad.lmdb"
    bad.write_bytes(b"\0" * 8192)
    with pytest.raises(R.LMDBFormatError):
        R.LMDBReader(str(bad))
    empty = tmp_path / "empty.lmdb"
    empty.write_bytes(b"")
    with pytest.raises(R.LMDBFormatError):
        R.LMDBReader(str(empty))


def test_record_file_and_samples(tmp_path):
    """tensorpack LMDBSerializer layout: '{:08}' keys, `__keys__`, msgpack_numpy values (feature rows make every
    record an overflow value, as in the real files)."""
    rng = np.random.default_rng(3)
    items = [make_item(rng, i, n) for i, n in enumerate([5, 0, 36, 9, 1, 12, 3])]
    keys = ["{:08}".format(i).encode("ascii") for i in range(len(items))]
    kv = {k: dumps(it) for k, it in zip(keys, items)}
    kv[b"__keys__"] = dumps(keys)
    path = str(tmp_path / "train_feat.lmdb")
    write_lmdb(path, kv, leaf_cap=3)
    rf = R.open_records(path)
    assert len(rf) == len(items)
    for got, want in zip(rf, items):
        assert_item_equal(got, want)
    assert len(R.RecordFile(path).shard(1, 2)) == 3
    assert R.RecordFile(path).shard(1, 2).record(0)[0] == "item1"

    tok = CharTokenizer()
    s = R.record_to_sample(rf.record(0), tok)
    assert s["tokens"] == tok.encode(items[0][1]) and s["tokens_pv"] == tok.encode(items[0][2])
    assert np.array_equal(s["boxes"], items[0][7]) and s["image_w"] == 800 and s["image_h"] == 600
    assert np.array_equal(s["features"], items[0][8]) and np.array_equal(s["targets"], items[0][9])
    # num_boxes <= 0: one dummy box on an 800 x 800 canvas, zero feature / target rows (`:571-576`)
    z = R.record_to_sample(rf.record(1), tok, v_feature_size=64, v_target_size=17)
    assert z["image_w"] == 800 and z["image_h"] == 800 and z["boxes"].dtype == np.float32
    assert np.array_equal(z["boxes"], np.array([[0.1, 0.1, 799.9, 799.9]], np.float32))
    assert z["features"].shape == (1, 64) and not z["features"].any() and z["targets"].shape == (1, 17)

    # a walk without `__keys__` gives the same records
    kv.pop(b"__keys__")
    path2 = str(tmp_path / "nokeys.lmdb")
    write_lmdb(path2, kv, leaf_cap=3)
    assert [r[0] for r in R.RecordFile(path2)] == [it[0] for it in items]

    # pair records: label + two items (`:963-968`)
    pair = [1] + items[0] + items[2]
    label, s1, s2 = R.pair_record_to_samples(R.decode_record(dumps(pair)), tok)
    assert label == 1 and s1["item_id"] == "item0" and s2["item_id"] == "item2" and s2["features"].shape == (36, 64)
    with pytest.raises(ValueError):
        R.pair_record_to_samples(pair[:-1], tok)

    # NumpySerializer flavour
    buf = np.empty(len(items), dtype=object)
    for i, it in enumerate(items):
        buf[i] = it
    npz = str(tmp_path / "train_feat.npz")
    np.savez(npz, buffer=buf)
    nf = R.open_records(npz)
    assert len(nf) == len(items)
    assert_item_equal(nf.record(2), items[2])


def test_iter_raw_batches_cpu(tmp_path):
    """records -> padded raw batch (host tensors here; the GPU half is `tests/test_preprocess.py`)."""
    rng = np.random.default_rng(5)
    items = [make_item(rng, i, n) for i, n in enumerate([5, 0, 40, 9, 1])]
    batches = list(R.iter_raw_batches(items, CharTokenizer(), 2, Lt=8, Lpv=12, R=36, device="cpu", drop_last=False,
                                      v_feature_size=64, v_target_size=17))
    assert [b["tokens"].shape[0] for b in batches] == [2, 2, 1]
    b0, b1 = batches[0], batches[1]
    assert b0["num_boxes"].tolist() == [5, 1] and b1["num_boxes"].tolist() == [36, 9]      # truncated to R
    assert b0["n_tok"].tolist() == [min(len(items[0][1]), 6), min(len(items[1][1]), 6)]     # L - 2
    assert b0["features"].shape == (2, 36, 64) and b0["whp"][1].tolist() == [800.0, 800.0, 640000.0]
    assert np.array_equal(b1["features"][0].numpy(), items[2][8][:36])


# ---- LMDB layout: an independent walker and hostile files (the writer above is still this repo's own: parity with a
# real `lmdb` writer stays unpinned) ------------------------------------------------------------------
def _walk_independent(path):
    """A second, deliberately different reader of the LMDB 0.9 page format (recursive descent, plain slicing, no shared
    code with k3m_b200.records): meta page with the larger txnid -> root -> branch children left to right -> leaves."""
    buf = open(path, "rb").read()
    ps = struct.unpack_from("<I", buf, 16 + 24)[0]
    metas = []
    for pg in (0, 1):
        o = pg * ps
        root = struct.unpack_from("<Q", buf, o + 16 + 24 + 48 + 40)[0]
        txn = struct.unpack_from("<Q", buf, o + 16 + 24 + 48 + 48 + 8)[0]
        metas.append((txn, root))
    root = max(metas)[1]
    out = {}

    def visit(pg):
        o = pg * ps
        flags, lower = struct.unpack_from("<HH", buf, o + 10)
        for i in range((lower - 16) // 2):
            p = struct.unpack_from("<H", buf, o + 16 + 2 * i)[0]
            lo, hi, nf, ks = struct.unpack_from("<HHHH", buf, o + p)
            key = buf[o + p + 8:o + p + 8 + ks]
            if flags & 0x01:
                visit(lo | hi << 16 | nf << 32)
            else:
                size = lo | hi << 16
                d = o + p + 8 + ks
                if nf & 0x01:
                    d = struct.unpack_from("<Q", buf, d)[0] * ps + 16
                out[key] = buf[d:d + size]
    if root != (1 << 64) - 1:
        visit(root)
    return out


@pytest.mark.parametrize("n,leaf_cap,branch_cap", [(1, 5, 4), (7, 3, 2), (60, 4, 3), (400, 5, 4)])
def test_lmdb_reader_agrees_with_an_independent_walker(tmp_path, n, leaf_cap, branch_cap):
    rng = np.random.default_rng(n)
    kv = {("%08d" % i).encode(): bytes(rng.integers(0, 256, size=int(rng.integers(1, 3000)), dtype=np.uint8)) for i in range(n)}
    path = str(tmp_path / "data.mdb")
    write_lmdb(path, kv, leaf_cap=leaf_cap, branch_cap=branch_cap)
    want = _walk_independent(path)
    assert want == kv
    with R.LMDBReader(path) as r:
        got = {k: bytes(v) for k, v in r.items()}
        assert got == want and len(r) == n
        for k in list(kv)[::7]:
            assert bytes(r.get(k)) == kv[k]
        assert r.get(b"zzzz") is None


@pytest.mark.timeout(120)
def test_lmdb_reader_survives_truncated_and_bit_flipped_files(tmp_path):
    """Hostile input: every prefix length class and 400 random bit flips of a depth-3 file.  The reader may return data
    or raise a clean error; it must never hang, recurse without bound or read outside the mapping."""
    rng = np.random.default_rng(5)
    kv = {("%08d" % i).encode(): bytes(rng.integers(0, 256, size=int(rng.integers(1, 2500)), dtype=np.uint8)) for i in range(80)}
    good = str(tmp_path / "good.mdb")
    write_lmdb(good, kv, leaf_cap=4, branch_cap=3)
    blob = open(good, "rb").read()
    ok_err = (R.LMDBFormatError, struct.error, ValueError, IndexError, OverflowError)

    def exercise(data, name):
        p = str(tmp_path / name)
        with open(p, "wb") as f:
            f.write(data)
        try:
            r = R.LMDBReader(p)
        except ok_err:
            return "open"
        try:
            n = 0
            for k, v in r.items():
                assert len(v) <= len(data)
                n += 1
                assert n <= 10 * len(kv)
            for k in list(kv)[::9]:
                r.get(k)
            return "read"
        except ok_err:
            return "err"
        finally:
            r.close()

    outcomes = []
    for cut in (0, 1, 100, 151, 152, PS - 1, PS, PS + 152, 2 * PS, 2 * PS + 17, len(blob) // 2, len(blob) - PS, len(blob) - 1):
        outcomes.append(exercise(blob[:cut], "cut%d.mdb" % cut))
    for t in range(400):
        b = bytearray(blob)
        # flips concentrated on structure: page headers, node pointers, node headers and the meta pages
        pg = int(rng.integers(0, len(blob) // PS))
        off = pg * PS + int(rng.integers(0, 64 if t % 2 else PS))
        b[off] ^= 1 << int(rng.integers(0, 8))
        if t % 5 == 0:
            b[pg * PS + 12:pg * PS + 16] = bytes(rng.integers(0, 256, size=4, dtype=np.uint8))      # lower / upper
        outcomes.append(exercise(bytes(b), "flip.mdb"))
    assert "err" in outcomes or "open" in outcomes           # at least some corruptions are detected
    assert outcomes.count("read") > 50                       # and harmless flips (value bytes) still read
