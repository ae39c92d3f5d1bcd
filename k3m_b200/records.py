"""On-disk record reader (SURVEY §8 f row N3): the files the reference trains from, read without tensorpack /
lmdb / msgpack_numpy (none of them is installed here; `msgpack` itself is, and is used for the container
format only).

What the reference writes (`data_prepare.py:496-544`): `tensorpack.dataflow.LMDBSerializer.save(ds, file)` — one
LMDB data file (`subdir=False`), key `'{:08}'.format(i)` -> `dumps(datapoint)`, plus a `__keys__` entry holding
`dumps(list of keys)`; `dumps` is msgpack (`use_bin_type=True`) with msgpack_numpy's ndarray encoding.  A
pretraining datapoint is the 10-list `[item_id, title, pv, category, image_h, image_w, num_boxes, boxes [n, 4],
features [n, 2048], cls_prob [n, 1601]]` (`data_prepare.py:365`), a pair datapoint is `[label] + item 1 + item 2`
(21 entries, `data_prepare.py:490`, unpacked at `concept_cap_dataset_struc.py:963-968`).  On Windows the reference
falls back to `NumpySerializer` (`.npz` with an object array `buffer`); `open_records` reads that too.

What the reference does with a record before masking (`concept_cap_dataset_struc.py:564-606`): substitute one
dummy 800 x 800 box when `num_boxes <= 0`, then `tokenizer.encode` the title and the PV string.  `record_to_sample`
does exactly that and returns the dict `preprocess.pack_raw_batch` takes, so the chain is
    LMDBReader -> decode_record -> record_to_sample -> pack_raw_batch -> GpuBatchPreprocessor (CUDA).

Parity: the msgpack_numpy encoding is restated from the package's published format (dict `nd` / `type` / `kind` /
`shape` / `data`) and checked against `msgpack` round trips in `tests/test_records_cpu.py`; the LMDB page layout
(0.9.x data format version 1) is restated from lmdb.h / mdb.c's published structures and checked against a
test-side writer of the same layout only — `lmdb` is absent, so the file format is **parity unpinned**."""
import mmap
import struct

import msgpack
import numpy as np

MAX_MSGPACK_LEN = 1000000000                      # `concept_cap_dataset_struc.py:37`

# --------------------------------------------------------------------------------------------- msgpack_numpy


def _get(d, k):
    """Keys arrive as bytes (packed with use_bin_type=True) or str (older writers / raw=False)."""
    kb = k.encode()
    return d[kb] if kb in d else d[k]


def _s(x):
    return x.decode() if isinstance(x, (bytes, bytearray)) else x


def _ndarray_hook(obj):
    """msgpack_numpy `decode`: {nd: True, type, kind, shape, data} -> ndarray; {nd: False, type, data} -> numpy
    scalar; {complex: True, data: repr} -> complex.  Anything else passes through."""
    if b"nd" in obj or "nd" in obj:
        if _get(obj, "nd") is True:
            descr = _get(obj, "type")
            kind = obj.get(b"kind", obj.get("kind", b""))
            if _s(kind) == "V":                   # structured dtype: list of (name, format[, shape]) tuples
                descr = [tuple(_s(t) if isinstance(t, (bytes, str)) else tuple(t) if isinstance(t, list) else t
                               for t in d) for d in descr]
            else:
                descr = _s(descr)
            a = np.frombuffer(_get(obj, "data"), dtype=np.dtype(descr))
            return a.reshape(_get(obj, "shape"))
        return np.frombuffer(_get(obj, "data"), dtype=np.dtype(_s(_get(obj, "type"))))[0]
    if b"complex" in obj or "complex" in obj:
        return complex(_s(_get(obj, "data")))
    return obj


def decode_record(buf):
    """tensorpack `loads` (`tensorpack/utils/serialize.py`, msgpack flavour) / the reference's `deserialize_lmdb`
    (`concept_cap_dataset_struc.py:80-88`): msgpack with the ndarray hook, str for str, bytes for bin."""
    return msgpack.unpackb(buf, raw=False, object_hook=_ndarray_hook, strict_map_key=False,
                           max_bin_len=MAX_MSGPACK_LEN, max_array_len=MAX_MSGPACK_LEN,
                           max_map_len=MAX_MSGPACK_LEN, max_str_len=MAX_MSGPACK_LEN)


# ------------------------------------------------------------------------------------------------------ LMDB

_MAGIC, _VERSION = 0xBEEFC0DE, 1
_HDR = 16                                          # PAGEHDRSZ: pgno u64, pad u16, flags u16, lower u16, upper u16
_P_BRANCH, _P_LEAF, _P_OVERFLOW, _P_META, _P_LEAF2 = 0x01, 0x02, 0x04, 0x08, 0x20
_F_BIGDATA, _F_SUBDATA, _F_DUPDATA = 0x01, 0x02, 0x04
_INVALID = (1 << 64) - 1


class LMDBFormatError(ValueError):
    pass


class LMDBReader(object):
    """Read-only view of one LMDB data file's main database: `get(key)`, `keys()`, `items()`, `len()`.
    Values are returned as `memoryview`s into the mapping (zero copy; `decode_record` accepts them).
    Supports what tensorpack's writer produces: plain keys, inline and overflow values, any tree depth.
    Not supported (raises): named sub-databases, DUPSORT data, LEAF2 pages."""

    def __init__(self, path):
        self._f = open(path, "rb")
        try:
            self._m = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        except ValueError:
            self._f.close()
            raise LMDBFormatError("%s: empty file" % path)
        self._v = memoryview(self._m)
        metas = []
        psize = None
        for off in (0, None):
            if off is None:
                if psize is None:
                    break
                off = psize
            if len(self._m) < off + 152:
                continue
            flags, = struct.unpack_from("<H", self._m, off + 10)
            magic, version = struct.unpack_from("<II", self._m, off + 16)
            if magic != _MAGIC or not flags & _P_META:
                if off == 0:
                    raise LMDBFormatError("%s: not an LMDB data file (magic %#x)" % (path, magic))
                continue
            if version != _VERSION:
                raise LMDBFormatError("%s: LMDB data version %d, reader knows %d" % (path, version, _VERSION))
            ps, = struct.unpack_from("<I", self._m, off + 40)          # mm_dbs[FREE_DBI].md_pad = page size
            psize = psize or ps
            dbflags, depth = struct.unpack_from("<HH", self._m, off + 92)
            entries, root = struct.unpack_from("<QQ", self._m, off + 120)
            last_pg, txnid = struct.unpack_from("<QQ", self._m, off + 136)
            metas.append((txnid, dict(flags=dbflags, depth=depth, entries=entries, root=root, last_pg=last_pg)))
        if not metas:
            raise LMDBFormatError("%s: no valid meta page" % path)
        self.page_size = psize
        self.meta = max(metas, key=lambda t: t[0])[1]                  # the newer of the two meta pages
        if self.meta["flags"] & 0x04:                                  # MDB_DUPSORT
            raise LMDBFormatError("DUPSORT databases are not supported")

    def close(self):
        self._f.close()
        try:
            self._v.release()
            self._m.close()
        except BufferError:                       # value views still alive: the mapping goes when they do
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __len__(self):
        return self.meta["entries"]

    # -- pages and nodes
    def _page(self, pgno):
        off = pgno * self.page_size
        if pgno < 0 or off + self.page_size > len(self._m):
            raise LMDBFormatError("page %d beyond the end of the file" % pgno)
        flags, lower = struct.unpack_from("<HH", self._m, off + 10)
        n = (lower - _HDR) >> 1
        if lower < _HDR or lower > self.page_size:
            if not flags & _P_OVERFLOW:
                raise LMDBFormatError("page %d: corrupt header (lower = %d)" % (pgno, lower))
            n = 0
        return off, flags, n

    def _node(self, off, i):
        p, = struct.unpack_from("<H", self._m, off + _HDR + 2 * i)
        if p < _HDR or p + 8 > self.page_size:
            raise LMDBFormatError("node pointer %d outside its page" % p)
        lo, hi, nflags, ksize = struct.unpack_from("<HHHH", self._m, off + p)
        if p + 8 + ksize > self.page_size:
            raise LMDBFormatError("node key runs past its page")
        return off + p + 8, lo, hi, nflags, ksize

    def _leaf_value(self, kpos, lo, hi, nflags, ksize):
        if nflags & (_F_SUBDATA | _F_DUPDATA):
            raise LMDBFormatError("sub-database / duplicate nodes are not supported")
        dsize = lo | hi << 16
        d = kpos + ksize
        if nflags & _F_BIGDATA:
            pg, = struct.unpack_from("<Q", self._m, d)
            ooff, oflags, _ = self._page(pg)
            if not oflags & _P_OVERFLOW:
                raise LMDBFormatError("page %d is not an overflow page" % pg)
            d = ooff + _HDR
        elif (d % self.page_size) + dsize > self.page_size:
            raise LMDBFormatError("inline value runs past its page")
        if d + dsize > len(self._m):
            raise LMDBFormatError("value runs past the end of the file")
        return self._v[d:d + dsize]

    def get(self, key, default=None):
        key = bytes(key)
        pg = self.meta["root"]
        if pg == _INVALID:
            return default
        for _level in range(64):                       # a B+tree of 2^64 pages is < 64 levels deep: a longer walk is a cycle
            off, flags, n = self._page(pg)
            if flags & _P_LEAF2:
                raise LMDBFormatError("LEAF2 pages are not supported")
            if flags & _P_BRANCH:
                lo_i, hi_i = 0, n - 1                  # last node whose key <= `key`; node 0 = minus infinity
                while lo_i < hi_i:
                    mid = (lo_i + hi_i + 1) >> 1
                    kpos, _, _, _, ks = self._node(off, mid)
                    if bytes(self._v[kpos:kpos + ks]) <= key:
                        lo_i = mid
                    else:
                        hi_i = mid - 1
                _, lo, hi, nflags, _ = self._node(off, lo_i)
                pg = lo | hi << 16 | nflags << 32
            elif flags & _P_LEAF:
                a, b = 0, n - 1
                while a <= b:
                    mid = (a + b) >> 1
                    kpos, lo, hi, nflags, ks = self._node(off, mid)
                    k = bytes(self._v[kpos:kpos + ks])
                    if k == key:
                        return self._leaf_value(kpos, lo, hi, nflags, ks)
                    if k < key:
                        a = mid + 1
                    else:
                        b = mid - 1
                return default
            else:
                raise LMDBFormatError("page %d: unexpected flags %#x" % (pg, flags))
        raise LMDBFormatError("tree deeper than 64 levels (page cycle)")

    def items(self):
        """(key, value) in key order: depth-first walk of the tree."""
        if self.meta["root"] == _INVALID:
            return
        stack = [(self.meta["root"], 0)]
        budget = 4 * (len(self._m) // self.page_size + 1)        # every page is entered a bounded number of times
        while stack:
            pg, i = stack.pop()
            budget -= 1
            if budget < 0 or len(stack) > 64:
                raise LMDBFormatError("page cycle in the tree")
            off, flags, n = self._page(pg)
            if flags & _P_BRANCH:
                if i < n:
                    stack.append((pg, i + 1))
                    _, lo, hi, nflags, _ = self._node(off, i)
                    stack.append((lo | hi << 16 | nflags << 32, 0))
            elif flags & _P_LEAF and not flags & _P_LEAF2:
                for j in range(n):
                    kpos, lo, hi, nflags, ks = self._node(off, j)
                    yield bytes(self._v[kpos:kpos + ks]), self._leaf_value(kpos, lo, hi, nflags, ks)
            else:
                raise LMDBFormatError("page %d: unexpected flags %#x" % (pg, flags))

    def keys(self):
        return [k for k, _ in self.items()]


# ------------------------------------------------------------------------------------------- record datasets


class RecordFile(object):
    """tensorpack `LMDBData` + `LMDBDataPoint` for one serialized file (`LMDBSerializer.load`): the key list comes
    from `__keys__` when present (else from a walk that skips it), `len()` excludes it, `record(i)` decodes the
    i-th datapoint.  `shard(rank, world)` keeps every world-th key — one disjoint slice per GPU process."""

    def __init__(self, path):
        self.db = LMDBReader(path)
        blob = self.db.get(b"__keys__")
        if blob is not None:
            self._keys = [k if isinstance(k, bytes) else k.encode("ascii") for k in decode_record(blob)]
        else:
            self._keys = [k for k in self.db.keys() if k != b"__keys__"]

    def shard(self, rank, world):
        self._keys = self._keys[rank::world]
        return self

    def __len__(self):
        return len(self._keys)

    def record(self, i):
        v = self.db.get(self._keys[i])
        if v is None:
            raise KeyError(self._keys[i])
        return decode_record(v)

    def __iter__(self):
        for i in range(len(self)):
            yield self.record(i)

    def close(self):
        self.db.close()


class NpzRecordFile(object):
    """`NumpySerializer` files (`data_prepare.py:503-505`): `np.load(file, allow_pickle=True)['buffer']`
    (`concept_cap_dataset_struc.py:225-227`)."""

    def __init__(self, path):
        self._buf = list(np.load(path, allow_pickle=True)["buffer"])

    def shard(self, rank, world):
        self._buf = self._buf[rank::world]
        return self

    def __len__(self):
        return len(self._buf)

    def record(self, i):
        return self._buf[i]

    def __iter__(self):
        return iter(self._buf)

    def close(self):
        pass


def open_records(path):
    return NpzRecordFile(path) if str(path).endswith(".npz") else RecordFile(path)


def _encode(tokenizer, text):
    """`tokenizer.encode(text)` of pytorch_transformers 1.1.0 adds no special tokens (`:605-606`); newer
    tokenizers need to be told."""
    try:
        return list(tokenizer.encode(text, add_special_tokens=False))
    except TypeError:
        return list(tokenizer.encode(text))


def record_to_sample(item, tokenizer, v_feature_size=2048, v_target_size=1601):
    """One 10-field item record -> the per-sample dict of `preprocess.pack_raw_batch`
    (`concept_cap_dataset_struc.py:564-606`: dummy box for `num_boxes <= 0`, then tokenisation)."""
    item_id, title, pv, _category, image_h, image_w, num_boxes, boxes, features, cls_prob = item
    num_boxes = int(num_boxes)
    if num_boxes <= 0:
        image_h, image_w, num_boxes = 800, 800, 1
        boxes = np.array([[0.1, 0.1, image_w - 0.1, image_h - 0.1]], dtype=np.float32)
        features = np.zeros((1, v_feature_size), np.float32)
        cls_prob = np.zeros((1, v_target_size), np.float32)
    return dict(item_id=item_id, tokens=_encode(tokenizer, title), tokens_pv=_encode(tokenizer, pv),
                boxes=np.asarray(boxes, np.float32)[:num_boxes], image_w=image_w, image_h=image_h,
                features=np.asarray(features, np.float32)[:num_boxes],
                targets=np.asarray(cls_prob, np.float32)[:num_boxes])


def pair_record_to_samples(pair, tokenizer, **kw):
    """One 21-field pair record -> (label, sample 1, sample 2) (`concept_cap_dataset_struc.py:963-968`)."""
    if len(pair) != 21:
        raise ValueError("pair record has %d fields, expected 21" % len(pair))
    return int(pair[0]), record_to_sample(pair[1:11], tokenizer, **kw), record_to_sample(pair[11:21], tokenizer, **kw)


def iter_raw_batches(records, tokenizer, batch_size, Lt, Lpv, R, device="cuda", drop_last=True, **sizes):
    """Records -> padded raw batches on `device` (input of `GpuBatchPreprocessor`); `sizes` = v_feature_size /
    v_target_size of the dummy rows when a record has no boxes."""
    from .preprocess import pack_raw_batch
    pending = []
    for rec in records:
        pending.append(record_to_sample(rec, tokenizer, **sizes))
        if len(pending) == batch_size:
            yield pack_raw_batch(pending, Lt, Lpv, R, device=device)
            pending = []
    if pending and not drop_last:
        yield pack_raw_batch(pending, Lt, Lpv, R, device=device)
