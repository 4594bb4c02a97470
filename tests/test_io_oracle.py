"""The reader-side checkers against each other: the plain-C parsers of oracle/wh_oracle_io.c
versus the reference's own CriteoParser / MinibatchIter (oracle/_ref/libwh_ref_io.so), and the numpy assembly
rule versus MinibatchIter::Next including its rand()-driven shuffle and negative sampling. CPU only."""
import ctypes
import os

import numpy as np
import pytest

from oracle.oracle_io import PortIO, RefIter, minibatches
from tests.util import adfea_text, agaricus_text, criteo_text, libsvm_text

libc = ctypes.CDLL("libc.so.6")
needs_ref = pytest.mark.skipif(not RefIter.available(), reason="oracle/_ref/libwh_ref_io.so not built")


def same(a, b):
    assert a.rows == b.rows and a.nnz == b.nnz
    assert np.array_equal(a.offset, b.offset) and np.array_equal(a.index, b.index)
    assert np.array_equal(a.label, b.label)
    assert (a.value is None) == (b.value is None)
    if a.value is not None:
        assert np.array_equal(a.value, b.value)


def test_cityhash64_known_answers():
    p = PortIO()
    assert p.cityhash64(b"") == 0x9ae16a3b2f90404f                 # k2, city.cc HashLen0to16
    assert p.cityhash64(b"abc") == 2640714258260161385             # published value (python-cityhash README)


def test_criteo_parser_rejects_what_the_reference_misparses():
    p = PortIO()
    good = b"1\t" + b"\t".join([b"5"] * 13) + b"\t" + b"\t".join([b"0a1b2c3d"] * 26) + b"\n"
    assert p.parse_criteo(good).rows == 1 and p.parse_criteo(good).nnz == 39
    for bad, code in [(b"\t" + good, 1), (b"1\t5\t5\n", 2), (good.replace(b"0a1b2c3d\n", b"0a1b\n"), 6),
                      (good[:-1] + b"\textra\n", 7), (b"1\t" + b"\t" * 13 + b"\t\t\n", 4)]:
        with pytest.raises(ValueError) as e:
            p.parse_criteo(bad)
        assert e.value.args[0] == code


@needs_ref
@pytest.mark.parametrize("crlf", [False, True])
def test_criteo_port_vs_reference(tmp_path, crlf, monkeypatch):
    rng = np.random.default_rng(5)
    text = criteo_text(rng, 3000, crlf=crlf)
    f = tmp_path / "criteo.txt"
    f.write_bytes(text)
    port = PortIO().parse_criteo(text)
    assert port.rows == 3000
    monkeypatch.setenv("WH_ORACLE_CHUNK", "20000")                  # many chunks: assembly crosses their borders
    it = RefIter(str(f), "criteo", 10 ** 6)
    assert it.cityhash64(b"abc") == 2640714258260161385
    same(next(it), port)
    it.close()
    got = list(RefIter(str(f), "criteo", 700))
    want = list(minibatches([port], 700))
    assert [b.rows for b in got] == [700, 700, 700, 700, 200]
    for a, b in zip(got, want):
        same(a, b)
    test_text = b"\n".join(line.split(b"\t", 1)[1] for line in text.replace(b"\r", b"").split(b"\n") if line) + b"\n"
    f.write_bytes(test_text)
    same(next(RefIter(str(f), "criteo_test", 10 ** 6)), PortIO().parse_criteo(test_text, is_train=False))


@needs_ref
@pytest.mark.parametrize("with_value", ["values", "ones", "bare"])
def test_libsvm_port_vs_reference_reader(tmp_path, with_value):
    text = libsvm_text(np.random.default_rng(6), 2000, with_value=with_value, key_space=2 ** 63)
    f = tmp_path / "a.libsvm"
    f.write_bytes(text)
    port = PortIO().parse_libsvm(text)
    got = next(RefIter(str(f), "libsvm", 10 ** 6))
    want = next(minibatches([port], 10 ** 6))
    same(got, want)
    assert (got.value is not None) == (with_value == "values")                    # ":1" everywhere -> values dropped (:116-118)


def test_agaricus_text_parses_to_the_committed_arrays():
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "agaricus_train.npz"))
    b = PortIO().parse_libsvm(agaricus_text())
    assert b.rows == 6513 and b.nnz == 143286
    assert np.array_equal(b.offset, z["offset"].astype(np.uint64))
    assert np.array_equal(b.index, z["index"].astype(np.uint64))
    assert np.array_equal(b.label, z["label"].astype(np.float32))
    assert np.all(b.value == 1)


@needs_ref
@pytest.mark.parametrize("neg", [1.0, 0.3])
def test_shuffle_and_negative_sampling_follow_rand(tmp_path, neg, monkeypatch):
    text = criteo_text(np.random.default_rng(7), 2500)
    f = tmp_path / "c.txt"
    f.write_bytes(text)
    monkeypatch.setenv("WH_ORACLE_CHUNK", "50000")
    libc.srand(11)
    got = list(RefIter(str(f), "criteo", 300, shuf_buf=1000, neg_sampling=neg))
    libc.srand(11)
    want = list(minibatches([PortIO().parse_criteo(text)], 300, shuf_buf=1000, neg_sampling=neg, rand=libc.rand))
    assert len(got) == len(want) and len(got) >= 3
    for a, b in zip(got, want):
        same(a, b)


@needs_ref
def test_parts_cover_the_file_once(tmp_path):
    text = criteo_text(np.random.default_rng(8), 1000, short_lines=False)
    f = tmp_path / "c.txt"
    f.write_bytes(text)
    whole = PortIO().parse_criteo(text)
    rows = 0
    idx = []
    for part in range(3):
        for b in RefIter(str(f), "criteo", 10 ** 6, part=part, nparts=3):
            rows += b.rows
            idx.append(b.index)
    assert rows == whole.rows and np.array_equal(np.concatenate(idx), whole.index)


@needs_ref
def test_adfea_port_vs_reference_reader(tmp_path, monkeypatch):
    text = adfea_text(np.random.default_rng(9), 2000)
    f = tmp_path / "a.adfea"
    f.write_bytes(text)
    port = PortIO().parse_adfea(text)
    assert port.rows == 2000 and port.value is None
    monkeypatch.setenv("WH_ORACLE_CHUNK", "30000")
    got = list(RefIter(str(f), "adfea", 600))
    want = list(minibatches([port], 600))
    assert [b.rows for b in got] == [600, 600, 600, 200]
    for a, b in zip(got, want):
        same(a, b)
    with pytest.raises(ValueError):
        PortIO().parse_adfea(b"1 2 x 5:3\n")


@needs_ref
@pytest.mark.parametrize("crlf", [False, True])
def test_part_ranges_of_the_gpu_reader_equal_the_reference_split(tmp_path, crlf):
    """wormhole_b200.reader.part_range (host logic of the GPU reader: which bytes part k of n owns) against the rows the
    reference's MinibatchIter reads for the same part: every part holds exactly the same rows."""
    from wormhole_b200.reader import part_range
    text = criteo_text(np.random.default_rng(10), 700, crlf=crlf, short_lines=False)
    f = tmp_path / "c.txt"
    f.write_bytes(text)
    covered = 0
    for nparts in (1, 3, 7):
        for k in range(nparts):
            b, e = part_range(str(f), k, nparts)
            mine = PortIO().parse_criteo(text[b:e]) if e > b else None
            ref = list(RefIter(str(f), "criteo", 10 ** 6, part=k, nparts=nparts))
            if mine is None or mine.rows == 0:
                assert not ref
            else:
                assert len(ref) == 1
                same(ref[0], mine)
            covered += (e - b) if nparts == 7 else 0
    assert covered == len(text)
