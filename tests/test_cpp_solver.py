"""tools/linear_b200.cc -- the reference's `linear.dmlc <conf> key=value ...` entry on top of the C++ facade
(host/solver/minibatch_solver.h: MinibatchScheduler::Run, host/base/minibatch_iter.h, AsgdServer / AsgdWorker): compiled
with g++ against the C-ABI library. CPU: it compiles and links. GPU: the demo.conf job on agaricus text prints the
same console output, progress rows included, as the Python mirror (wormhole_b200/solver.py, itself checked against the
CPU checker in tests/test_gpu_solver.py) run with the same rand() seed, and writes the same model file."""
import ctypes
import io
import os
import re
import subprocess

import pytest

from tests.util import agaricus_text

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_binary(tmp):
    from wormhole_b200 import build
    build.build()
    exe = os.path.join(str(tmp), "linear_b200")
    cmd = ["/usr/bin/g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "wormhole_b200", "host"), os.path.join(ROOT, "tools", "linear_b200.cc"),
           "-o", exe, "-L" + os.path.join(ROOT, "wormhole_b200"), "-lwormhole_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_linear_b200_compiles_and_links(tmp_path):
    assert os.path.exists(build_binary(tmp_path))


def test_rank_hub_without_a_gpu(tmp_path):
    """RunRanks / Peer::Allgather (the few host messages of `linear_b200 -gpus S`: IPC handles, step flags, progress sums):
    every rank gets every rank's bytes in rank order; a rank that leaves early fails the job instead of hanging it."""
    from wormhole_b200 import build
    build.build()
    exe = os.path.join(str(tmp_path), "hub_test")
    cmd = ["/usr/bin/g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "wormhole_b200", "host"), os.path.join(ROOT, "tests", "cpp", "hub_test.cc"),
           "-o", exe, "-L" + os.path.join(ROOT, "wormhole_b200"), "-lwormhole_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for ranks in ("2", "3", "8"):
        r = subprocess.run([exe, ranks], capture_output=True, text=True, timeout=60)
        assert r.returncode == 0 and "hub_test ok" in r.stdout, r.stdout + r.stderr
    r = subprocess.run([exe, "3", "quit"], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "hub_test failed" in r.stdout, r.stdout + r.stderr


def strip_seconds(log):
    """progress rows start with the elapsed seconds: drop that column"""
    return [re.sub(r"^\s*\d+  ", "", l) if re.match(r"^\s*\d+  \s*[\d.e+]+", l) else l for l in log.splitlines()]


@pytest.mark.gpu
def test_linear_b200_prints_what_the_python_mirror_prints(tmp_path):
    from wormhole_b200.solver import Config, MinibatchScheduler
    exe = build_binary(tmp_path)
    text = agaricus_text()
    lines = text.split(b"\n")
    train, val = tmp_path / "agaricus.txt.train", tmp_path / "agaricus.txt.test"
    train.write_bytes(text)
    val.write_bytes(b"\n".join(lines[:1500]) + b"\n")
    conf = tmp_path / "demo.conf"
    conf.write_text('# demo conf\n\ntrain_data = "%s"\nval_data = "%s"\nmodel_out = "%s"\nmax_data_pass = 2\n'
                    % (train, val, tmp_path / "cpp_model"))
    args = ["print_sec=1000", "num_parts_per_file=3", "lr_eta=.1"]
    r = subprocess.run([exe, str(conf)] + args + ["-seed", "17"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    # the Python mirror, same conf and overrides, same rand() seed
    pconf = Config.from_file(str(conf), args)
    pconf.model_out = str(tmp_path / "py_model")
    out = io.StringIO()
    ctypes.CDLL("libc.so.6").srand(17)
    assert MinibatchScheduler(pconf, out=out).Run()
    assert strip_seconds(r.stdout) == strip_seconds(out.getvalue())
    a = open(str(tmp_path / "cpp_model_part-0"), "rb").read()
    b = open(str(tmp_path / "py_model_part-0"), "rb").read()
    assert len(a) > 0 and a == b


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["FTRL", "SGD"])
def test_linear_b200_on_two_gpus(tmp_path, algo):
    """linear_b200 -gpus 2: one process per GPU, rank g = worker g + server of key range g, the parts of two files of
    different length (the shorter rank idles through the last steps). The model parts must equal the CPU checker's
    simulation of the same message order: per step every rank with a minibatch pulls the model of the step before, then
    each owner applies the ranks' slices in ascending rank order, one push message each."""
    import numpy as np
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    from oracle import oracle as O
    from tests.util import FLIP_ABS, STRAGGLERS, assert_state_close
    from wormhole_b200.rowblock import RowBlock
    exe = build_binary(tmp_path)
    rng = np.random.default_rng(91)
    MAX_KEY, MB, PASSES = 2 ** 16, 500, 2
    L1, L2, ETA, BETA = 0.5, 0.1, 0.1, 1.0

    def make_file(path, rows):
        nnz = rng.integers(1, 11, rows)
        off = np.zeros(rows + 1, np.uint64); off[1:] = np.cumsum(nnz)
        n = int(off[-1])
        ids = np.where(rng.random(n) < 0.5, rng.integers(0, 300, n) * 211, rng.integers(0, 10 ** 9, n)).astype(np.uint64)
        lab = (rng.random(rows) < 0.4).astype(np.float32)
        with open(path, "w") as f:
            for r in range(rows):
                f.write(" ".join([str(int(lab[r]))] + ["%d:1" % int(i) for i in ids[int(off[r]):int(off[r + 1])]]) + "\n")
        blks = []
        for a in range(0, rows, MB):
            b = min(rows, a + MB)
            o = off[a:b + 1] - off[a]
            blks.append(RowBlock(o.astype(np.uint64), ids[int(off[a]):int(off[b])].copy(), None, lab[a:b].copy()))
        return blks
    streams = [make_file(str(tmp_path / "part-0.txt"), 3300), make_file(str(tmp_path / "part-1.txt"), 1700)]   # 7 and 4 minibatches
    conf = tmp_path / "two.conf"
    conf.write_text('train_data = "%s"\nval_data = "%s"\nmodel_out = "%s"\ndata_format = "libsvm"\nalgo = %s\n'
                    'minibatch = %d\nmax_data_pass = %d\nrand_shuffle = 0\nneg_sampling = 1\nmax_concurrency = 2\n'
                    'lambda_l1 = %g\nlambda_l2 = %g\nlr_eta = %g\nlr_beta = %g\nprint_sec = 1000\nnum_parts_per_file = 1\n'
                    % (tmp_path / "part-.*", tmp_path / "part-0.txt", tmp_path / "m", algo, MB, PASSES, L1, L2, ETA, BETA))
    r = subprocess.run([exe, str(conf), "-gpus", "2", "-max_key", str(MAX_KEY), "-seed", "5"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    # the same schedule on the CPU checker
    chk = O.best()
    a = {"FTRL": O.ALGO_FTRL, "SGD": O.ALGO_SGD}[algo]
    world = 2
    bounds = [(g * MAX_KEY) // world for g in range(world + 1)]
    W, Z, CG = (np.zeros(MAX_KEY, np.float32) for _ in range(3))
    t = [1] * world
    objv = ex = 0.0
    for _ in range(PASSES):
        for step in range(max(len(s) for s in streams)):
            msgs = []
            for rk in range(world):
                if step >= len(streams[rk]):
                    continue
                D = chk.localize(streams[rk][step], MAX_KEY)
                keys = D.uniq.astype(np.int64)
                prog, grad = chk.linear_loss(O.LOSS_LOGIT, D, W[keys], 2)
                msgs.append((keys, grad, np.searchsorted(keys, bounds)))
            for g in range(world):
                for keys, grad, cut in msgs:
                    sl = slice(cut[g], cut[g + 1])
                    w, z, cg = W[keys[sl]].copy(), Z[keys[sl]].copy(), CG[keys[sl]].copy()
                    _, t[g] = chk.handle_push(a, L1, L2, ETA, BETA, grad[sl].copy(), w, z, cg, t[g])
                    W[keys[sl]], Z[keys[sl]], CG[keys[sl]] = w, z, cg
    rec = np.dtype([("k", "<u8"), ("w", "<f4")])
    got = np.zeros(MAX_KEY, np.float32)
    for g in range(world):
        raw = open(str(tmp_path / ("m_part-%d" % g)), "rb").read()
        assert len(raw) % 12 == 0 and len(raw) > 0
        part = np.frombuffer(raw, dtype=rec)
        assert np.all((part["k"] >= bounds[g]) & (part["k"] < bounds[g + 1]))
        got[part["k"].astype(np.int64)] = part["w"]
    bad = np.nonzero(np.abs(got - W) > 1e-4)[0]
    if len(bad):     # say where the differing keys occur before failing
        print(r.stdout)
        where = []
        for k in bad[:20]:
            occ = [(rk, st, int((streams[rk][st].index % np.uint64(MAX_KEY) == np.uint64(k)).sum()))
                   for rk in range(world) for st in range(len(streams[rk]))]
            where.append((int(k), float(got[k]), float(W[k]), [o for o in occ if o[2]]))
        print("differing keys (key, got, want, [(rank, step, entries)]):", where)
    assert_state_close(got, W, what="two-GPU C++ host w", stragglers=STRAGGLERS, flip_abs=FLIP_ABS)
    assert (W != 0).sum() > 100
    # console: rank 0 prints the merged rows; every training pass saw all 5 000 rows, validation the 3 300 of part-0
    rows = [l.split() for l in strip_seconds(r.stdout) if re.match(r"^\s*[\d.e+]+\s", l)]
    assert r.stdout.startswith("Connected 2 servers and 2 workers"), r.stdout
    seen = [float(x[1]) for x in rows]
    assert seen == [5000.0, 3300.0] * PASSES, r.stdout
