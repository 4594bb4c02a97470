"""dmlc::data::MinibatchIter (learn/base/minibatch_iter.h:29-148) with the parsing and the minibatch assembly on
the GPU: text chunks go to the device as they are, wh_parse_text turns them into device-resident row blocks,
and Next() assembles exactly `minibatch_size` rows across chunk borders in a device container that
Worker.submit(..., mem=WH_DEVICE) consumes directly. The host only reads the file, cuts it at line ends and
runs the reference's rand()-driven bookkeeping (shuffle order, negative sampling) on 4 bytes per row.

Same constructor arguments and methods as the reference class; `Value()` returns a capi.Rows.
"""
import ctypes
import os

import numpy as np

from . import capi

_libc = None


def _rand():
    """glibc rand(): the generator std::random_shuffle and the negative sampling of the reference draw from."""
    global _libc
    if _libc is None:
        _libc = ctypes.CDLL("libc.so.6")
    return _libc.rand()


RAND_MAX = 2147483647


def _is_brk(c):
    return c in (10, 13)


def part_range(path, part_index, num_parts):
    """Byte range of one part: the even split of the file, both ends moved forward to the next line start
    (InputSplit's rule: a part owns the lines that START inside its byte range)."""
    total = os.path.getsize(path)

    def align(f, x):
        if x == 0 or x >= total:
            return min(x, total)
        f.seek(x)
        while True:                      # rest of the current line
            buf = f.read(1 << 16)
            if not buf:
                return total
            for i, c in enumerate(buf):
                if _is_brk(c):
                    x += i
                    break
            else:
                x += len(buf)
                continue
            break
        f.seek(x)
        while True:                      # its line break(s)
            buf = f.read(1 << 12)
            if not buf:
                return total
            for i, c in enumerate(buf):
                if not _is_brk(c):
                    return x + i
            x += len(buf)

    with open(path, "rb") as f:
        return align(f, total * part_index // num_parts), align(f, total * (part_index + 1) // num_parts)


RECORDIO_MAGIC = 0xced7230a


def _record_head(buf, i):
    """Does a RecordIO record START at byte i of buf (magic, then a header word whose flag says whole record or first
    part)? dmlc-core's InputSplit for recordio finds part boundaries the same way."""
    if i + 8 > len(buf) or int.from_bytes(buf[i:i + 4], "little") != RECORDIO_MAGIC:
        return False
    return (int.from_bytes(buf[i + 4:i + 8], "little") >> 29) in (0, 1)


def crb_part_range(path, part_index, num_parts):
    """Byte range of one part of a RecordIO file: the even split, both ends moved forward to the next record head
    (4-byte aligned: records are padded to words)."""
    total = os.path.getsize(path)

    def align(f, x):
        x = min(total, (x + 3) & ~3)
        while x < total:
            f.seek(x)
            buf = f.read(1 << 16)
            for i in range(0, len(buf) - 7, 4):
                if _record_head(buf, i):
                    return x + i
            x += max(4, (len(buf) - 8) & ~3)
        return total

    with open(path, "rb") as f:
        return align(f, total * part_index // num_parts), align(f, total * (part_index + 1) // num_parts)


class _CrbParser:
    """CRBParser (learn/base/crb_parser.h) over one part of one file: next() -> the rows of the next chunk of records."""

    def __init__(self, ctx, uri, part_index, num_parts, chunk_bytes):
        self.ctx, self.uri, self.chunk_bytes = ctx, uri, chunk_bytes
        self.begin, self.end = crb_part_range(uri, part_index, num_parts)
        self.f = open(uri, "rb")
        self.before_first()

    def before_first(self):
        self.pos = self.begin
        self.bytes_read = 0

    def next(self):
        while self.pos < self.end:
            # whole records up to chunk_bytes (at least one; a split record's parts stay together)
            start, p, open_parts = self.pos, self.pos, False
            while p < self.end:
                self.f.seek(p)
                hdr = self.f.read(8)
                if len(hdr) < 8 or int.from_bytes(hdr[:4], "little") != RECORDIO_MAGIC:
                    raise capi.WormholeError("crb: no RecordIO magic at byte %d of %s" % (p, self.uri))
                lrec = int.from_bytes(hdr[4:], "little")
                flag, n = lrec >> 29, lrec & ((1 << 29) - 1)
                nxt = p + 8 + ((n + 3) & ~3)
                if p > start and not open_parts and nxt - start > self.chunk_bytes:
                    break
                open_parts = flag in (1, 2)
                p = nxt
            self.f.seek(start)
            buf = self.f.read(p - start)
            self.pos = p
            self.bytes_read += len(buf)
            rows = capi.parse_crb(self.ctx, buf)
            if rows.rows:
                return rows
            rows.close()
        return None


class _ChunkParser:
    """ParserImpl over one part of one text file: Next() -> the next parsed chunk (capi.Rows) or None."""

    def __init__(self, ctx, uri, part_index, num_parts, fmt, chunk_bytes):
        self.ctx, self.uri, self.fmt, self.chunk_bytes = ctx, uri, fmt, chunk_bytes
        self.begin, self.end = part_range(uri, part_index, num_parts)
        self.f = open(uri, "rb")
        self.before_first()

    def before_first(self):
        self.pos = self.begin
        self.bytes_read = 0
        self.f.seek(self.begin)

    def next(self):
        while self.pos < self.end:
            want = min(self.chunk_bytes, self.end - self.pos)
            buf = bytearray(self.f.read(want))
            if self.pos + len(buf) < self.end:          # finish the current line and take its break(s)
                in_break = _is_brk(buf[-1])
                while True:
                    more = self.f.read(4096)
                    if not more:
                        break
                    cut = None
                    for i, c in enumerate(more):
                        if _is_brk(c):
                            in_break = True
                        elif in_break:
                            cut = i
                            break
                    if cut is None:
                        buf += more
                    else:
                        buf += more[:cut]
                        break
                self.f.seek(self.pos + len(buf))
            buf = bytes(buf)
            self.pos += len(buf)
            self.bytes_read += len(buf)
            rows = capi.parse_text(self.ctx, self.fmt, buf)
            if rows.rows:
                return rows
            rows.close()
        return None


class MinibatchIter:
    def __init__(self, ctx, uri, part_index, num_parts, type, minibatch_size, shuf_buf=0, negative_sampling=1.0,
                 chunk_bytes=8 << 20, ring=4, rand=_rand):
        self.ctx, self.mb_size, self.shuf_buf, self.neg = ctx, int(minibatch_size), int(shuf_buf), negative_sampling
        self.rand = rand
        if type not in capi.FMT and type != "crb":
            raise capi.WormholeError("unknown datatype %s" % type)  # :58
        if shuf_buf:
            if not shuf_buf > minibatch_size:                                                      # :37
                raise capi.WormholeError("shuf_buf must exceed minibatch_size")
            self.buf_reader = MinibatchIter(ctx, uri, part_index, num_parts, type, shuf_buf, 0, 1.0, chunk_bytes, 2)
            self.parser = None
        else:
            self.parser = (_CrbParser(ctx, uri, part_index, num_parts, chunk_bytes) if type == "crb" else   # :50-52
                           _ChunkParser(ctx, uri, part_index, num_parts, type, chunk_bytes))
            self.buf_reader = None
        # the worker reads a submitted minibatch asynchronously: hand out `ring` containers in turn
        self.ring = [capi.Rows(ctx) for _ in range(ring)]
        self.turn = 0
        self.cur, self.start, self.end = None, 0, 0
        self.rdp, self.cur_label, self.out = None, None, None

    def BeforeFirst(self):
        if self.parser:
            self.parser.before_first()
        if self.buf_reader:
            self.buf_reader.BeforeFirst()
        self.cur, self.start, self.end = None, 0, 0

    def Next(self):
        out = self.ring[self.turn]
        self.turn = (self.turn + 1) % len(self.ring)
        out.clear()
        have = 0
        while have < self.mb_size:
            if self.start == self.end:
                if self.shuf_buf == 0:
                    if self.cur is not None:
                        self.cur.close()
                    self.cur = self.parser.next()
                    if self.cur is None:
                        self.start = self.end = 0
                        break
                else:
                    if not self.buf_reader.Next():
                        self.cur = None
                        self.start = self.end = 0
                        break
                    self.cur = self.buf_reader.Value()
                    n = self.cur.rows
                    if self.rdp is None or len(self.rdp) != n:                                      # :88-91
                        self.rdp = list(range(n))
                    rdp, rand = self.rdp, self.rand
                    for i in range(1, n):                                   # std::random_shuffle (libstdc++) :92
                        j = rand() % (i + 1)
                        if i != j:
                            rdp[i], rdp[j] = rdp[j], rdp[i]
                    self.cur_label = self.cur.labels() if self.neg < 1.0 else None
                self.start, self.end = 0, self.cur.rows
            n = min(self.end - self.start, self.mb_size - have)                                     # :98
            if self.shuf_buf == 0:
                out.append(self.cur, self.start, n)
                have += n
            else:
                keep = []
                thr = np.float32(1) - np.float32(self.neg)
                for i in range(self.start, self.start + n):                                         # :102-110
                    j = self.rdp[i]
                    if self.neg < 1.0 and self.cur_label[j] <= 0 and \
                            np.float32(self.rand()) / np.float32(RAND_MAX) > thr:
                        continue
                    keep.append(j)
                if keep:
                    out.append(self.cur, row_ids=keep)
                have += len(keep)
            self.start += n
        out.drop_unit_values()                                                                      # :116-118
        self.ctx.sync()
        self.out = out
        return out.rows > 0

    def Value(self):
        return self.out

    def BytesRead(self):
        return self.parser.bytes_read if self.parser else self.buf_reader.BytesRead()

    def __iter__(self):
        while self.Next():
            yield self.Value()
