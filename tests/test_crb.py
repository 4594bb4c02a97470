"""The crb reader (learn/base/crb_parser.h, learn/base/compressed_row_block.h) against a writer that restates the
reference's Compress and dmlc-core's RecordIO framing, with the LZ4 blocks produced by liblz4 itself (pyarrow's lz4_raw
codec): that pins the library's LZ4 block decoder and record handling by round trip through an independent encoder."""
import struct

import numpy as np
import pytest

from oracle.oracle_io import RECORDIO_MAGIC, crb_record, recordio_write
from tests.util import random_block, wide_ids
from wormhole_b200 import capi
from wormhole_b200.reader import MinibatchIter
from wormhole_b200.rowblock import RowBlock

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


def blocks(rng, n):
    out = []
    for i in range(n):
        b = random_block(rng, int(rng.integers(1, 4000)), 20, 10 ** 6, with_value=(i % 3 == 0))
        ids = wide_ids(rng, b.nnz) if i % 2 else b.index           # incompressible and compressible id arrays
        val = b.value
        if i % 3 == 0 and i % 2:
            val = np.ones(b.nnz, np.float32)                       # all ones: the writer drops the array
        out.append(RowBlock(b.offset, ids, val, b.label))
    return out


def concat(blks):
    off = [np.zeros(1, np.uint64)]
    for b in blks:
        off.append(b.offset[1:] - b.offset[0] + off[-1][-1])
    has_val = any(b.value is not None and not np.all(b.value == 1) for b in blks)
    val = np.concatenate([np.ones(b.nnz, np.float32) if (b.value is None or np.all(b.value == 1)) else b.value for b in blks]) if has_val else None
    return RowBlock(np.concatenate(off), np.concatenate([b.index for b in blks]), val, np.concatenate([b.label for b in blks]))


def same(a, b):
    assert a.rows == b.rows and a.nnz == b.nnz
    assert np.array_equal(a.offset, b.offset) and np.array_equal(a.index, b.index) and np.array_equal(a.label, b.label)
    assert (a.value is None) == (b.value is None)
    if a.value is not None:
        assert np.array_equal(a.value.view(np.uint32), b.value.view(np.uint32))


def test_crb_records_round_trip(gpu):
    rng = np.random.default_rng(31)
    blks = blocks(rng, 12)
    data = recordio_write([crb_record(b) for b in blks])
    same(capi.parse_crb(gpu, data).to_host(), concat(blks))


def test_crb_record_containing_the_magic_is_rejoined(gpu):
    """A payload with the RecordIO magic at a word boundary is written as parts (flags 1, 2, 3). Incompressible id arrays
    (LZ4 stores them as literals) that hold the magic bytes once, at each of the four byte alignments in turn: in at
    least one of the four records the magic lands on a word boundary and the writer has to split."""
    rng = np.random.default_rng(32)
    blks = []
    for shift in range(4):
        raw = bytearray(wide_ids(rng, 3000).tobytes())
        raw[1000 + shift:1004 + shift] = struct.pack("<I", RECORDIO_MAGIC)
        off = np.arange(0, 3001, 3, dtype=np.uint64)
        blks.append(RowBlock(off, np.frombuffer(bytes(raw), np.uint64).copy(), None, (rng.random(1000) < .5).astype(np.float32)))
    data = recordio_write([crb_record(b) for b in blks])
    nparts = sum(1 for i in range(0, len(data) - 7, 4) if data[i:i + 4] == struct.pack("<I", RECORDIO_MAGIC))
    assert nparts > len(blks)                                      # some record really was split
    same(capi.parse_crb(gpu, data).to_host(), concat(blks))


def test_crb_errors(gpu):
    blk = random_block(np.random.default_rng(33), 50, 5, 1000)
    rec = crb_record(blk)
    with pytest.raises(capi.WormholeError):
        capi.parse_crb(gpu, recordio_write([b"\x01\x02\x03\x04" + rec[4:]]))          # wrong data format
    with pytest.raises(capi.WormholeError):
        capi.parse_crb(gpu, recordio_write([rec[:4] + struct.pack("<i", 4) + rec[8:]]))   # wrong indextype
    with pytest.raises(capi.WormholeError):
        capi.parse_crb(gpu, recordio_write([rec[:-7]]))                                # truncated LZ4 block
    assert capi.parse_crb(gpu, b"").rows == 0


def test_minibatch_iter_over_crb_parts(gpu, tmp_path):
    """MinibatchIter(type = "crb") over two parts of one file: exactly minibatch_size rows per Next(), every row once."""
    rng = np.random.default_rng(34)
    blks = blocks(rng, 30)
    path = tmp_path / "data.crb"
    path.write_bytes(recordio_write([crb_record(b) for b in blks]))
    want = concat(blks)
    got_rows, got_idx = 0, []
    for part in range(2):
        it = MinibatchIter(gpu, str(path), part, 2, "crb", 1000, chunk_bytes=64 << 10)
        it.BeforeFirst()
        sizes = []
        while it.Next():
            h = it.Value().to_host()
            sizes.append(h.rows); got_idx.append(h.index)
        assert all(s == 1000 for s in sizes[:-1]) and 0 < sizes[-1] <= 1000
        got_rows += sum(sizes)
    assert got_rows == want.rows
    assert np.array_equal(np.concatenate(got_idx), want.index)
