"""ctypes front-end for the reader-side checkers.  TEST INFRASTRUCTURE ONLY.

* ``PortIO()`` -- oracle/liboracle_port.so: wh_oracle_io.c, the plain-C restatement of the Criteo and
  libsvm text parsers and of CityHash64 (<= 16 bytes).
* ``RefIter`` -- oracle/_ref/libwh_ref_io.so: the reference's own MinibatchIter + CriteoParser compiled
  unchanged (ref_io.cc + shims_io/); only where it was built from /root/reference or travelled along.
* ``minibatches()`` -- the assembly rule of MinibatchIter::Next (learn/base/minibatch_iter.h:74-125) in numpy,
  for the Port arm: exact mb_size rows across chunk borders, shuffle buffer + negative sampling on glibc
  rand(), values dropped when all are 1.
"""
import ctypes as C
import os

import numpy as np

from oracle import oracle as O
from wormhole_b200.rowblock import RowBlock

_HERE = os.path.dirname(os.path.abspath(__file__))
RAND_MAX = 2147483647


class PortIO:
    def __init__(self):
        path = os.path.join(_HERE, "liboracle_port.so")
        if not os.path.exists(path):
            O.build()
        self.L = L = C.CDLL(path)
        L.wh_o_cityhash64.restype = C.c_uint64
        L.wh_o_cityhash64.argtypes = [C.c_char_p, C.c_size_t]

    def cityhash64(self, s):
        return int(self.L.wh_o_cityhash64(s, len(s)))

    @staticmethod
    def _lines(text):
        return text.count(b"\n") + text.count(b"\r") + 1

    def parse_criteo(self, text, is_train=True):
        """-> RowBlock (binary features). Raises ValueError(code, line) on a malformed line."""
        m = self._lines(text)
        off, lab, idx = np.zeros(m + 1, np.uint64), np.zeros(m, np.float32), np.zeros(39 * m, np.uint64)
        rows, nnz = C.c_size_t(0), C.c_size_t(0)
        rc = self.L.wh_o_parse_criteo(text, C.c_size_t(len(text)), int(is_train), C.byref(rows), C.byref(nnz),
                                      O._p(off), O._p(lab), O._p(idx))
        if rc:
            raise ValueError(rc, rows.value)
        return RowBlock(off[:rows.value + 1], idx[:nnz.value], None, lab[:rows.value])

    def parse_adfea(self, text):
        t = len(text.split()) + 1
        off, lab, idx = np.zeros(t + 1, np.uint64), np.zeros(t, np.float32), np.zeros(t, np.uint64)
        rows, nnz = C.c_size_t(0), C.c_size_t(0)
        rc = self.L.wh_o_parse_adfea(text, C.c_size_t(len(text)), C.byref(rows), C.byref(nnz), O._p(off), O._p(lab), O._p(idx))
        if rc:
            raise ValueError(rc, rows.value)
        return RowBlock(off[:rows.value + 1], idx[:nnz.value], None, lab[:rows.value])

    def parse_libsvm(self, text):
        m = self._lines(text)
        t = len(text.split()) + 1
        off, lab = np.zeros(m + 1, np.uint64), np.zeros(m, np.float32)
        idx, val = np.zeros(t, np.uint64), np.zeros(t, np.float32)
        rows, nnz, nv = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
        rc = self.L.wh_o_parse_libsvm(text, C.c_size_t(len(text)), C.byref(rows), C.byref(nnz), C.byref(nv),
                                      O._p(off), O._p(lab), O._p(idx), O._p(val))
        if rc:
            raise ValueError(rc, rows.value)
        return RowBlock(off[:rows.value + 1], idx[:nnz.value], val[:nnz.value] if nv.value else None, lab[:rows.value])


class RefIter:
    """dmlc::data::MinibatchIter<uint64_t> of the reference (learn/base/minibatch_iter.h), as an iterator
    over RowBlocks."""

    @staticmethod
    def available():
        return os.path.exists(os.path.join(_HERE, "_ref", "libwh_ref_io.so"))

    def __init__(self, path, fmt, mb_size, part=0, nparts=1, shuf_buf=0, neg_sampling=1.0):
        self.L = L = C.CDLL(os.path.join(_HERE, "_ref", "libwh_ref_io.so"))
        L.ref_iter_open.restype = C.c_void_p
        L.ref_cityhash64.restype = C.c_uint64
        self.h = C.c_void_p(L.ref_iter_open(path.encode(), part, nparts, fmt.encode(), mb_size, shuf_buf,
                                            C.c_float(neg_sampling)))

    def cityhash64(self, s):
        return int(self.L.ref_cityhash64(s, C.c_size_t(len(s))))

    def __iter__(self):
        return self

    def __next__(self):
        rows, nnz, hv = C.c_size_t(0), C.c_size_t(0), C.c_int(0)
        if not self.L.ref_iter_next(self.h, C.byref(rows), C.byref(nnz), C.byref(hv)):
            raise StopIteration
        off, lab = np.zeros(rows.value + 1, np.uint64), np.zeros(rows.value, np.float32)
        idx = np.zeros(nnz.value, np.uint64)
        val = np.zeros(nnz.value, np.float32) if hv.value else None
        self.L.ref_iter_copy(self.h, O._p(off), O._p(lab), O._p(idx), O._p(val))
        return RowBlock(off, idx, val, lab)

    def close(self):
        if self.h:
            self.L.ref_iter_close(self.h)
            self.h = None


def _concat(blocks):
    off = [np.zeros(1, np.uint64)]
    base = 0
    for b in blocks:
        off.append(b.offset[1:] + np.uint64(base))
        base += b.nnz
    has_val = any(b.value is not None for b in blocks)
    val = None
    if has_val:
        val = np.concatenate([b.value if b.value is not None else np.ones(b.nnz, np.float32) for b in blocks])
    return RowBlock(np.concatenate(off), np.concatenate([b.index for b in blocks]), val,
                    np.concatenate([b.label for b in blocks]))


def _take(b, rows):
    rows = np.asarray(rows, np.int64)
    lens = (b.offset[rows + 1] - b.offset[rows]).astype(np.int64)
    off = np.zeros(len(rows) + 1, np.uint64)
    off[1:] = np.cumsum(lens)
    pos = np.concatenate([np.arange(int(b.offset[r]), int(b.offset[r + 1])) for r in rows]) if len(rows) else \
        np.zeros(0, np.int64)
    pos = pos.astype(np.int64)
    return RowBlock(off, b.index[pos], None if b.value is None else b.value[pos], b.label[rows])


def minibatches(chunks, mb_size, shuf_buf=0, neg_sampling=1.0, rand=None):
    """MinibatchIter::Next over an iterable of parsed chunks (RowBlocks). `rand` = callable returning the
    next glibc rand() value (needed when shuf_buf > 0)."""
    if shuf_buf:
        assert shuf_buf > mb_size                                   # :37
        src = minibatches(chunks, shuf_buf)                         # :38-39 the buffer reader
    else:
        src = iter(chunks)
    cur, start, rdp = None, 0, None
    while True:
        parts, have = [], 0
        while have < mb_size:
            if cur is None or start == cur.rows:
                cur = next(src, None)
                if cur is None:
                    break
                start = 0
                if shuf_buf:
                    if rdp is None or len(rdp) != cur.rows:         # :88-91
                        rdp = list(range(cur.rows))
                    for i in range(1, len(rdp)):                    # std::random_shuffle of libstdc++ :92
                        j = rand() % (i + 1)
                        if i != j:
                            rdp[i], rdp[j] = rdp[j], rdp[i]
            n = min(cur.rows - start, mb_size - have)               # :98
            if not shuf_buf:
                parts.append(cur.slice(start, start + n))
                have += n
            else:
                keep = []
                for i in range(start, start + n):                   # :102-110
                    j = rdp[i]
                    if neg_sampling < 1.0 and cur.label[j] <= 0 and \
                            np.float32(rand()) / np.float32(RAND_MAX) > np.float32(1 - np.float32(neg_sampling)):
                        continue
                    keep.append(j)
                parts.append(_take(cur, keep))
                have += len(keep)
            start += n
        if not parts:
            return
        mb = _concat(parts)
        if mb.value is not None and np.all(mb.value == 1):          # :116-118
            mb = RowBlock(mb.offset, mb.index, None, mb.label)
        if mb.rows == 0:
            return                                                  # :122 out_blk_.size > 0
        yield mb


# ---- crb: CompressedRowBlock::Compress (learn/base/compressed_row_block.h:18-49,100-112) + dmlc-core's RecordIOWriter ----
# Test infrastructure like the rest of oracle/. LZ4 itself is not restated: the blocks are compressed by liblz4 through
# pyarrow's "lz4_raw" codec (LZ4's block format), which is what pins the library's own decoder.
CRB_MAGIC, RECORDIO_MAGIC = 1196140743, 0xced7230a


def lz4_block(data):
    import pyarrow as pa
    return pa.compress(data, codec="lz4_raw", asbytes=True)


def crb_record(blk):
    """One RowBlock -> the bytes CompressedRowBlock::Compress<uint64_t> appends to its string."""
    import struct
    nrows = blk.rows
    off = np.ascontiguousarray(blk.offset, np.uint64)
    nnz = int(off[-1] - off[0])
    value = blk.value
    if value is not None and np.all(value[:nnz] == 1):
        value = None                                    # :27-35
    out = [struct.pack("<iii", CRB_MAGIC, 8, nrows)]
    for arr in (np.ascontiguousarray(blk.label, np.float32), off, np.ascontiguousarray(blk.index, np.uint64),
                None if value is None else np.ascontiguousarray(value, np.float32), None):   # weight: NULL
        if arr is None:
            out.append(struct.pack("<i", 0))
        else:
            c = lz4_block(arr.tobytes())
            out.append(struct.pack("<i", len(c)) + c)
    return b"".join(out)


def recordio_write(records):
    """dmlc::RecordIOWriter::WriteRecord (dmlc-core, un-vendored; restated from its published format): a payload that holds
    the magic at a 4-byte boundary is split there into parts flagged 1 / 2 / 3, otherwise one part flagged 0."""
    import struct
    magic = struct.pack("<I", RECORDIO_MAGIC)
    out = []
    for rec in records:
        cuts = [i for i in range(0, len(rec) - 3, 4) if rec[i:i + 4] == magic]
        parts, last = [], 0
        for c in cuts:
            parts.append(rec[last:c]); last = c + 4
        parts.append(rec[last:])
        for k, part in enumerate(parts):
            flag = 0 if len(parts) == 1 else (1 if k == 0 else (3 if k == len(parts) - 1 else 2))
            out.append(magic + struct.pack("<I", (flag << 29) | len(part)) + part + b"\0" * (-len(part) % 4))
    return b"".join(out)
