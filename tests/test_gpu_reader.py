"""The GPU reader against the CPU checkers: wh_parse_text versus the C restatement of the
Criteo / libsvm parsers (bit-exact: ids, offsets, labels, values), the device-side MinibatchIter versus the
reference's own MinibatchIter (oracle/_ref/libwh_ref_io.so, when it travelled) or the numpy assembly rule, and
text file -> reader -> worker versus the checker fed with the checker's parse."""
import ctypes

import numpy as np
import pytest

from oracle import oracle as O
from oracle.oracle_io import PortIO, RefIter, minibatches
from tests.util import adfea_text, agaricus_text, criteo_text, libsvm_text
from wormhole_b200 import capi
from wormhole_b200.reader import MinibatchIter, part_range

pytestmark = pytest.mark.gpu
libc = ctypes.CDLL("libc.so.6")


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


def same(a, b):
    assert a.rows == b.rows and a.nnz == b.nnz
    assert np.array_equal(a.offset, b.offset) and np.array_equal(a.index, b.index)
    assert np.array_equal(a.label, b.label)
    assert (a.value is None) == (b.value is None)
    if a.value is not None:
        assert np.array_equal(a.value.view(np.uint32), b.value.view(np.uint32))


def test_long_literals_like_atof(gpu):
    """Values written with repr(float) / "%.17g" (17 significant digits), with exponents beyond +-22, with 40 digits, and
    around the float range: bit-identical to the checker's atof (learn/base/criteo_parser.h / dmlc-core's libsvm parser)."""
    rng = np.random.default_rng(5)
    vals = [repr(float(x)) for x in rng.standard_normal(400)]
    vals += ["%.17g" % x for x in rng.standard_normal(400) * 10.0 ** rng.integers(-30, 30, 400)]
    vals += ["0.30000000000000004", "1e-30", "1E+30", "2.5e-40", "3e38", "4e38", "1e-46", "7e-46", "123456789012345678901234567890.5e-10",
             "0.00000000000000000000000000000000000012345678901234567", "9007199254740993", "9007199254740992.5", "1.7976931348623157e308",
             "4.9406564584124654e-324", "1e400", "1e-400", "00012.5000e+01", ".5e1", "5.e-1"]
    lines = []
    for i in range(0, len(vals), 4):
        lines.append("%s " % vals[i] + " ".join("%d:%s" % (7 * i + j, v) for j, v in enumerate(vals[i:i + 4])))
    text = ("\n".join(lines) + "\n").encode()
    same(capi.parse_text(gpu, "libsvm", text).to_host(), PortIO().parse_libsvm(text))


@pytest.mark.parametrize("crlf,trail", [(False, True), (True, True), (False, False)])
def test_criteo_parse_bit_exact(gpu, crlf, trail):
    text = criteo_text(np.random.default_rng(21), 5000, crlf=crlf, trailing_newline=trail)
    same(capi.parse_text(gpu, "criteo", text).to_host(), PortIO().parse_criteo(text))
    test_text = b"\n".join(l.split(b"\t", 1)[1] for l in text.replace(b"\r", b"").split(b"\n") if l) + b"\n"
    same(capi.parse_text(gpu, "criteo_test", test_text).to_host(), PortIO().parse_criteo(test_text, is_train=False))


def test_criteo_known_row(gpu):
    # one hand-checkable row: field 0 = "5", categorical 0 = "0a1b2c3d", everything else empty
    line = b"1\t5" + b"\t" * 13 + b"0a1b2c3d" + b"\t" * 25 + b"\n"
    b = capi.parse_text(gpu, "criteo", line).to_host()
    p = PortIO()
    assert b.rows == 1 and list(b.label) == [1.0]
    assert list(b.index) == [p.cityhash64(b"5") >> 10, (p.cityhash64(b"0a1b2c3d") >> 10) | (13 << 54)]


def test_malformed_criteo_lines_are_named(gpu):
    good = b"1\t" + b"\t".join([b"5"] * 13) + b"\t" + b"\t".join([b"0a1b2c3d"] * 26) + b"\n"
    for bad in (b"1\t5\t5\n", good.replace(b"0a1b2c3d\n", b"0a1b\n"), good[:-1] + b"\textra\n"):
        with pytest.raises(capi.WormholeError, match="line 3"):
            capi.parse_text(gpu, "criteo", good + good + bad + good)
    with pytest.raises(capi.WormholeError, match="inf / nan / hex"):
        capi.parse_text(gpu, "libsvm", b"1 3:0x1p3\n")
    v = capi.parse_text(gpu, "libsvm", b"1 3:0.12345678901234567890\n").to_host().value   # long literals convert like atof
    assert v[0] == np.float32(0.12345678901234567890)
    with pytest.raises(capi.WormholeError, match="some features"):
        capi.parse_text(gpu, "libsvm", b"1 3:1 4\n")


def test_empty_and_blank_text(gpu):
    for t in (b"", b"\n\n\r\n", b"   \n"):
        r = capi.parse_text(gpu, "libsvm", t)
        assert r.rows == 0 and r.nnz == 0
    assert capi.parse_text(gpu, "criteo", b"\n\n").rows == 0


@pytest.mark.parametrize("mode", ["values", "ones", "bare"])
def test_libsvm_parse_bit_exact(gpu, mode):
    text = libsvm_text(np.random.default_rng(22), 4000, with_value=mode, key_space=2 ** 63)
    same(capi.parse_text(gpu, "libsvm", text).to_host(), PortIO().parse_libsvm(text))


def test_adfea_parse_bit_exact(gpu, tmp_path):
    text = adfea_text(np.random.default_rng(26), 4000)
    same(capi.parse_text(gpu, "adfea", text).to_host(), PortIO().parse_adfea(text))
    for bad, msg in ((b"7 2 1 5:3 9:1\n8 1 x 4:2\n", "line 2"), (b"7 2 1 5:3\n8 1\n", "fewer than three"), (b"7 2 1 5:3 9\n", "out of order")):
        with pytest.raises(capi.WormholeError, match=msg):
            capi.parse_text(gpu, "adfea", bad)
    f = tmp_path / "a.adfea"
    f.write_bytes(text)
    got = [b.to_host() for b in MinibatchIter(gpu, str(f), 0, 1, "adfea", 700, chunk_bytes=40000)]
    want = list(minibatches([PortIO().parse_adfea(text)], 700))
    assert len(got) == len(want) == 6
    for a, b in zip(got, want):
        same(a, b)


def test_agaricus_text(gpu):
    b = capi.parse_text(gpu, "libsvm", agaricus_text()).to_host()
    same(b, PortIO().parse_libsvm(agaricus_text()))
    assert b.rows == 6513 and b.nnz == 143286


def test_device_resident_text(gpu):
    import torch
    text = criteo_text(np.random.default_rng(23), 1000)
    d = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
    same(capi.parse_text(gpu, "criteo", d.data_ptr(), mem=capi.WH_DEVICE, length=len(text)).to_host(),
         PortIO().parse_criteo(text))


@pytest.mark.parametrize("fmt,shuf,neg", [("criteo", 0, 1.0), ("criteo", 1000, 1.0), ("criteo", 1000, 0.3),
                                          ("libsvm", 0, 1.0)])
def test_minibatch_iter_matches_the_reference_reader(gpu, tmp_path, fmt, shuf, neg, monkeypatch):
    rng = np.random.default_rng(24)
    text = criteo_text(rng, 2500) if fmt == "criteo" else libsvm_text(rng, 2500, with_value="values")
    f = tmp_path / "data.txt"
    f.write_bytes(text)
    libc.srand(5)
    got = [b.to_host() for b in MinibatchIter(gpu, str(f), 0, 1, fmt, 300, shuf, neg, chunk_bytes=50000)]
    libc.srand(5)
    if RefIter.available():
        monkeypatch.setenv("WH_ORACLE_CHUNK", "70000")          # different chunking: the result must not depend on it
        want = list(RefIter(str(f), fmt, 300, shuf_buf=shuf, neg_sampling=neg))
    else:
        whole = PortIO().parse_criteo(text) if fmt == "criteo" else PortIO().parse_libsvm(text)
        want = list(minibatches([whole], 300, shuf_buf=shuf, neg_sampling=neg, rand=libc.rand))
    assert len(got) == len(want) and len(got) >= 3
    for a, b in zip(got, want):
        same(a, b)


def test_parts_cover_the_file_once(gpu, tmp_path):
    text = criteo_text(np.random.default_rng(25), 1200, crlf=True)
    f = tmp_path / "c.txt"
    f.write_bytes(text)
    whole = PortIO().parse_criteo(text)
    idx, rows = [], 0
    for part in range(3):
        for b in MinibatchIter(gpu, str(f), part, 3, "criteo", 250, chunk_bytes=30000):
            h = b.to_host()
            rows += h.rows
            idx.append(h.index)
    assert rows == whole.rows and np.array_equal(np.concatenate(idx), whole.index)
    assert part_range(str(f), 0, 3)[0] == 0 and part_range(str(f), 2, 3)[1] == len(text)


def test_file_to_model_equals_the_checker(gpu, tmp_path, checker):
    """agaricus as TEXT through the GPU reader into the worker, against the checker trained on the checker's
    parse of the same text: the end-to-end path of learn/linear/guide/demo.conf minus the scheduler."""
    f = tmp_path / "agaricus.txt"
    f.write_bytes(agaricus_text())
    max_key = 2 ** 20
    srv = capi.Server(gpu, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key, depth=2)
    job = checker.linear_create(O.ALGO_FTRL, O.LOSS_LOGIT, 1.0, 0.1, 0.1, 1.0, max_key)
    whole = PortIO().parse_libsvm(agaricus_text())
    pend = []
    for got, want in zip(MinibatchIter(gpu, str(f), 0, 1, "libsvm", 1000, chunk_bytes=100000),
                         minibatches([whole], 1000)):
        assert not got.has_value and want.value is None              # every value is 1: dropped (:116-118)
        if len(pend) == 2:
            wk.wait(pend.pop(0))
        pend.append(wk.submit(got, True, capi.WH_DEVICE, got.nnz))
        checker.linear_minibatch(job, want)
    for t in pend:
        wk.wait(t)
    keys, w, z, cg = checker.linear_model(job)
    st = srv.get_state(keys)
    assert np.allclose(st[0], w, rtol=1e-4, atol=1e-6)
    assert srv.nnz_w() == int((w != 0).sum())
