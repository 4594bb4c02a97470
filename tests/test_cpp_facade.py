"""The C++ host facade (wormhole_b200/host: dmlc::Localizer, SpMV, SpMM, BinClassEval, CreateLoss,
ps::KVWorker, AsgdServer, AsgdWorker) compiled with g++ against the C-ABI library and run on the GPU;
its outputs are compared with the CPU checker exactly as the ctypes path's are. The CPU part checks
that the headers compile and link (no GPU needed for that)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from tests.util import FLIP_ABS, STRAGGLERS, assert_state_close, random_block

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_KEY = 2 ** 20
CONF = 'algo = FTRL\nloss = LOGIT\nlambda_l1 = .5   # comment\nlambda_l2 = .1\nlr_eta = .1\nlr_beta = 1\nmax_concurrency = 1\n'


def build_binary(tmp):
    from wormhole_b200 import build
    build.build()
    exe = os.path.join(str(tmp), "facade_test")
    cmd = ["/usr/bin/g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "wormhole_b200", "host"), os.path.join(ROOT, "tests", "cpp", "facade_test.cc"),
           "-o", exe, "-L" + os.path.join(ROOT, "wormhole_b200"), "-lwormhole_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_facade_headers_compile_and_link(tmp_path):
    assert os.path.exists(build_binary(tmp_path))


def put(f, a):
    f.write(struct.pack("<Q", len(a)))
    f.write(np.ascontiguousarray(a).tobytes())


def get(f, dtype):
    n, = struct.unpack("<Q", f.read(8))
    return np.frombuffer(f.read(n * np.dtype(dtype).itemsize), dtype=dtype).copy()


@pytest.mark.gpu
def test_facade_matches_the_checker(tmp_path, checker):
    from oracle import oracle as O
    exe = build_binary(tmp_path)
    rng = np.random.default_rng(77)
    blks = [random_block(rng, 3000, 14, 30000, with_value=(m % 2 == 1), empty_rows=False) for m in range(4)]
    fin, fout, fconf = (str(tmp_path / n) for n in ("in.bin", "out.bin", "conf.txt"))
    with open(fin, "wb") as f:
        put(f, np.array([MAX_KEY, len(blks)], np.uint64))
        for b in blks:
            put(f, b.offset); put(f, b.index)
            put(f, b.value if b.value is not None else np.zeros(0, np.float32)); put(f, b.label)
    open(fconf, "w").write(CONF)
    r = subprocess.run([exe, fin, fout, fconf], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr

    D = checker.localize(blks[0], MAX_KEY)
    U = len(D.uniq)
    with open(fout, "rb") as f:
        assert np.array_equal(get(f, np.uint64), D.uniq)
        assert np.array_equal(get(f, np.uint32), D.cnt)
        assert np.array_equal(get(f, np.uint64), D.offset)
        assert np.array_equal(get(f, np.uint32), D.index)
        val = (np.float32(0.01) * ((np.arange(U, dtype=np.uint64) * np.uint64(2654435761)) % np.uint64(97)).astype(np.float32)
               - np.float32(0.4)).astype(np.float32)
        d = (np.float32(0.001) * (np.arange(D.rows) % 1000).astype(np.float32) - np.float32(0.5)).astype(np.float32)
        xw, g = get(f, np.float32), get(f, np.float32)
        assert np.array_equal(xw, checker.spmv_times(D, val))
        gw = checker.spmv_trans_times(D, d, U)
        assert np.linalg.norm(g.astype(np.float64) - gw) <= 1e-5 * np.linalg.norm(gw.astype(np.float64))
        prog, grad = get(f, np.float64), get(f, np.float32)
        pc, gc = checker.linear_loss(O.LOSS_LOGIT, D, val, 2)
        assert prog[1] == pc[1] and prog[4] == pc[4] and abs(prog[0] - pc[0]) <= 1e-4 * abs(pc[0])
        assert np.linalg.norm(grad.astype(np.float64) - gc) <= 1e-5 * np.linalg.norm(gc.astype(np.float64))
        ev = get(f, np.float32)
        ec = checker.bin_eval(D.label, xw, 2)
        assert abs(ev[0] - ec["auc"]) <= 1e-3 and ev[1] == ec["acc"] and abs(ev[2] - ec["logit_objv"]) <= 1e-4 * ec["logit_objv"]
        dim = 4
        V = (np.float32(0.001) * (np.arange(U * dim) % 613).astype(np.float32) - np.float32(0.3)).astype(np.float32)
        xv, gv = get(f, np.float32), get(f, np.float32)
        xvc = checker.spmm_times(D, V, dim)
        assert np.array_equal(xv, xvc)
        gvc = checker.spmm_trans_times(D, xvc, U * dim, dim)
        assert np.linalg.norm(gv.astype(np.float64) - gvc) <= 1e-5 * np.linalg.norm(gvc.astype(np.float64))
        # the whole path: AsgdWorker::ProcessMinibatch over 4 minibatches
        progs = get(f, np.float64).reshape(-1, 6)
        keys, w, z, cg, nnz = get(f, np.uint64), get(f, np.float32), get(f, np.float32), get(f, np.float32), get(f, np.int64)
    job = checker.linear_create(O.ALGO_FTRL, O.LOSS_LOGIT, 0.5, 0.1, 0.1, 1.0, MAX_KEY)
    pcs = np.array([checker.linear_minibatch(job, b) for b in blks])
    ck, cw, cz, ccg = checker.linear_model(job)
    assert np.array_equal(keys, ck)
    assert np.array_equal(progs[:, 1], pcs[:, 1]) and np.array_equal(progs[:, 4], pcs[:, 4])
    assert np.all(np.abs(progs[:, 0] - pcs[:, 0]) <= 1e-4 * np.abs(pcs[:, 0]))
    for a, b in ((w, cw), (z, cz), (cg, ccg)):
        assert_state_close(a, b, what="facade", stragglers=STRAGGLERS, flip_abs=FLIP_ABS)
    assert abs(int(nnz[0]) - int((cw != 0).sum())) <= 2 and progs[:, 5].sum() == nnz[0]


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,shuf,neg", [("criteo", 0, 1.0), ("criteo", 900, 0.4), ("libsvm", 0, 1.0)])
def test_cpp_minibatch_iter_and_device_minibatches(tmp_path, checker, fmt, shuf, neg):
    """dmlc::data::MinibatchIter of the C++ facade (GPU reader) against the CPU assembly rule on the CPU parse of
    the same text, same rand() seed; and AsgdWorker::ProcessMinibatch(DeviceMinibatch) against the checker."""
    import ctypes
    from oracle import oracle as O
    from oracle.oracle_io import PortIO, minibatches
    from tests.util import criteo_text, libsvm_text
    exe = build_binary(tmp_path)
    rng = np.random.default_rng(78)
    text = criteo_text(rng, 2000) if fmt == "criteo" else libsvm_text(rng, 2000, with_value="values", key_space=2 ** 40)
    f, fout = tmp_path / "data.txt", str(tmp_path / "reader.bin")
    f.write_bytes(text)
    r = subprocess.run([exe, "--reader", str(f), fmt, "300", str(shuf), str(neg), "9", str(MAX_KEY), fout],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(9)
    whole = PortIO().parse_criteo(text) if fmt == "criteo" else PortIO().parse_libsvm(text)
    want = list(minibatches([whole], 300, shuf_buf=shuf, neg_sampling=neg, rand=libc.rand))
    job = checker.linear_create(O.ALGO_FTRL, O.LOSS_LOGIT, 0.5, 0.1, 0.1, 1.0, MAX_KEY)
    with open(fout, "rb") as fh:
        for b in want:
            assert np.array_equal(get(fh, np.uint64), b.offset)
            assert np.array_equal(get(fh, np.uint64), b.index)
            v = get(fh, np.float32)
            assert (len(v) == 0) == (b.value is None) and (b.value is None or np.array_equal(v, b.value))
            assert np.array_equal(get(fh, np.float32), b.label)
        assert len(get(fh, np.uint64)) == 0                       # no further minibatch
        progs = get(fh, np.float64).reshape(-1, 6)
        nnz_w, bytes_read = get(fh, np.int64)
    pcs = np.array([checker.linear_minibatch(job, b) for b in want])
    assert np.array_equal(progs[:, 1], pcs[:, 1]) and np.array_equal(progs[:, 4], pcs[:, 4])
    assert np.all(np.abs(progs[:, 0] - pcs[:, 0]) <= 1e-4 * np.abs(pcs[:, 0]))
    assert bytes_read == len(text)


def build_difacto_binary(tmp):
    from wormhole_b200 import build
    build.build()
    exe = os.path.join(str(tmp), "difacto_facade_test")
    cmd = ["/usr/bin/g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "wormhole_b200", "host"), os.path.join(ROOT, "tests", "cpp", "difacto_facade_test.cc"),
           "-o", exe, "-L" + os.path.join(ROOT, "wormhole_b200"), "-lwormhole_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_difacto_facade_compiles_and_links(tmp_path):
    assert os.path.exists(build_difacto_binary(tmp_path))


@pytest.mark.gpu
def test_difacto_facade_equals_the_ctypes_path(tmp_path):
    """difacto::AsyncSGD (C++) and capi.Difacto (ctypes) are two bindings of one C ABI: same conf, same minibatches,
    same hash-seeded embedding init -> identical progress vectors; the table rows follow learn/difacto/progress.h."""
    from wormhole_b200 import capi
    from wormhole_b200.solver import DifactoProgress
    exe = build_difacto_binary(tmp_path)
    rng = np.random.default_rng(79)
    blks = [random_block(rng, 2000, 14, 3000, with_value=(m % 2 == 1), empty_rows=False) for m in range(3)]
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        put(f, np.array([MAX_KEY, len(blks)], np.uint64))
        for b in blks:
            put(f, b.offset); put(f, b.index)
            put(f, b.value if b.value is not None else np.zeros(0, np.float32)); put(f, b.label)
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    with open(fout, "rb") as f:
        one_call = [get(f, t) for t in (np.uint32, np.int32, np.float32, np.float32)]
        by_messages = [get(f, t) for t in (np.uint32, np.int32, np.float32, np.float32)]
        progs_msg = get(f, np.float64).reshape(-1, 8)
        progs = get(f, np.float64).reshape(-1, 8)
        rows = get(f, np.uint8).tobytes().decode().splitlines()
    # the worker driven through ps::KVWorker (count push with cmd, ZVPull with deps, difacto::Loss, ZVPush) leaves the
    # same model as the one-call path, and sees the same objective / AUC / counts on the way
    assert len(one_call[0]) > 1000 and (one_call[1] > 1).any()
    for a, b in zip(one_call, by_messages):
        assert np.array_equal(a, b)
    assert np.array_equal(progs_msg[:, :6], progs[:, :6])
    ctx = capi.Context(0)
    d = capi.Difacto(ctx, l1=1.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=5, threshold=5, v_l2=0.0, init_scale=.01,
                     max_key=MAX_KEY, seed=7)
    want = np.array([d.process_minibatch(b, True, dp) for dp in range(2) for b in blks])
    d.close(); ctx.close()
    assert np.array_equal(progs, want)
    assert progs[:, 7].sum() > 0                      # embeddings were created
    table = DifactoProgress()
    for p, row in zip(want, rows):
        table.data = list(p)
        assert table.PrintStr() == row
