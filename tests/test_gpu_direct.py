"""The shared-GPU training step (csrc/direct.cu: keys address the shard directly, two launches per minibatch) against
the chain of launches it replaces (WH_DIRECT=0: localize -> pull -> forward/backward -> push) and against the CPU
checker. Gradients are sums of integers in both, so for binary features the two device paths must leave bit-identical
models; the checker is held to the parity tolerances."""
import os

import numpy as np
import pytest

from tests.util import FLIP_ABS, STRAGGLERS, assert_state_close, random_block
from wormhole_b200 import capi, synth
from wormhole_b200.rowblock import RowBlock

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


class env:
    """Environment switches the library reads when a worker is created / a step is launched."""
    def __init__(self, **kv):
        self.kv = kv

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        for k, v in self.kv.items():
            os.environ[k] = str(v)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def run(gpu, blks, algo, loss, max_key, depth=1, train_mask=None, **switches):
    switches.setdefault("WH_DX_MIN_COVERAGE", 0)      # these tests want the direct step whatever the table covers
    with env(**switches):
        srv = capi.Server(gpu, algo, 0.5, 0.1, 0.1, 1.0, 0, max_key)
        wk = capi.Worker(gpu, srv, loss, max_key, depth=depth)
        progs, pending = [], []
        for i, b in enumerate(blks):
            if len(pending) == depth:
                progs.append(wk.wait(pending.pop(0)))
            pending.append(wk.submit(b, True if train_mask is None else train_mask[i]))
        progs += [wk.wait(t) for t in pending]
        keys = np.unique(np.concatenate([b.index % np.uint64(max_key) for b in blks]))
        state = srv.get_state(keys)
        nnz, t = srv.nnz_w(), srv.sgd_t()
        wk.close(); srv.close()
    return np.array(progs), keys, state, nnz, t


def same_model(a, b):
    pa, ka, sa, na, ta = a
    pb, kb, sb, nb, tb = b
    assert np.array_equal(ka, kb) and na == nb and ta == tb
    for x, y in zip(sa, sb):
        assert np.array_equal(x, y)
    assert np.array_equal(pa[:, 1], pb[:, 1]) and np.array_equal(pa[:, 3:], pb[:, 3:])   # accuracy, count, rows, new_w
    assert np.allclose(pa[:, 0], pb[:, 0], rtol=1e-12, atol=0)                          # fp64 sums, block order differs
    assert np.allclose(pa[:, 2], pb[:, 2], rtol=0, atol=1e-12)                          # AUC: integer pair counts


@pytest.mark.parametrize("algo", [capi.ALGO_SGD, capi.ALGO_ADAGRAD, capi.ALGO_FTRL])
@pytest.mark.parametrize("loss", [capi.LOSS_LOGIT, capi.LOSS_SQUARE_HINGE])
@pytest.mark.parametrize("max_key", [2 ** 20, 1000003])
def test_direct_equals_chain_binary(gpu, algo, loss, max_key):
    rng = np.random.default_rng(7)
    blks = [random_block(rng, 5000, 30, 40000) for _ in range(5)]     # ragged, empty rows, repeated ids in a row
    blks += [synth.criteo_minibatch(3, rows=20000)]
    a = run(gpu, blks, algo, loss, max_key)
    b = run(gpu, blks, algo, loss, max_key, WH_DIRECT=0)
    same_model(a, b)


def test_direct_small_tiles_and_long_rows(gpu):
    """Tiles of 4 160 entries (the smallest the kernel takes) with rows of up to 9 000: whole rows that do not fit a tile go
    through the chunked path, the others through ragged tiles."""
    rng = np.random.default_rng(11)
    blks = []
    for _ in range(3):
        b = random_block(rng, 300, 40, 5000)
        nnz = np.diff(b.offset.astype(np.int64))
        nnz[rng.integers(0, 300, 4)] = rng.integers(5000, 9000, 4)
        off = np.zeros(301, np.uint64); off[1:] = np.cumsum(nnz)
        ids = rng.integers(0, 5000, int(off[-1]), dtype=np.uint64)
        blks.append(RowBlock(off, ids, None, b.label))
    a = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 16, WH_DX_ECAP=64)
    b = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 16, WH_DIRECT=0)
    same_model(a, b)


def test_direct_lanes_and_evaluation_passes(gpu):
    """Four minibatches in flight, training and evaluation interleaved, equal strictly sequential processing; the
    three-group variant of the kernel leaves the same bits."""
    blks = [synth.criteo_minibatch(m, rows=30000) for m in range(7)]
    mask = [True, True, False, True, False, True, True]
    a = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 24, depth=1, train_mask=mask)
    b = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 24, depth=4, train_mask=mask)
    c = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 24, depth=1, train_mask=mask, WH_DIRECT=0)
    same_model(a, b)
    same_model(a, c)


def test_direct_staged_codes_change_nothing(gpu):
    """WH_DX_STAGED=1: the codes of a minibatch are made by a launch of their own, ahead of the step."""
    blks = [synth.criteo_minibatch(m, rows=20000) for m in range(6)] + [random_block(np.random.default_rng(3), 4000, 30, 40000)]
    a = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 22, depth=3)
    b = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 22, depth=3, WH_DX_STAGED=1)
    same_model(a, b)


def test_direct_table_refresh_changes_nothing(gpu):
    blks = [synth.criteo_minibatch(m, rows=20000) for m in range(8)]
    a = run(gpu, blks, capi.ALGO_ADAGRAD, capi.LOSS_LOGIT, 2 ** 22, depth=3)
    b = run(gpu, blks, capi.ALGO_ADAGRAD, capi.LOSS_LOGIT, 2 ** 22, depth=3, WH_HOT_REFRESH=4)
    same_model(a, b)


def test_direct_sees_changes_made_behind_its_back(gpu):
    """wh_server_set_state / wh_server_push between minibatches change weights the worker mirrors for its table keys."""
    max_key = 2 ** 20
    blks = [synth.criteo_minibatch(m, rows=10000) for m in range(4)]
    out = []
    for direct in (1, 0):
        with env(WH_DIRECT=direct, WH_DX_MIN_COVERAGE=0):
            srv = capi.Server(gpu, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, max_key)
            wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key)
            progs = [wk.process_minibatch(blks[0]), wk.process_minibatch(blks[1])]
            keys = np.unique(blks[1].index % np.uint64(max_key))
            w, z, cg = srv.get_state(keys)
            srv.set_state(keys, (w + 0.25).astype(np.float32), z, cg)
            progs.append(wk.process_minibatch(blks[2]))
            srv.push(keys[:5000], np.full(5000, 0.5, np.float32))
            progs.append(wk.process_minibatch(blks[3]))
            allk = np.unique(np.concatenate([b.index % np.uint64(max_key) for b in blks]))
            out.append((np.array(progs), allk, srv.get_state(allk), srv.nnz_w(), srv.sgd_t()))
            wk.close(); srv.close()
    same_model(out[0], out[1])


@pytest.mark.parametrize("algo", [capi.ALGO_ADAGRAD, capi.ALGO_FTRL])
def test_direct_valued_features_vs_checker(gpu, checker, algo):
    """Non-binary features: the fixed-point scale comes from max|value| on the device (|dual| <= 1 is known).
    Two comparisons. (a) Against the checker's own pieces with ONE change -- X^T dual summed exactly (fp64 over the fp32
    terms) instead of in fp32 row order: the device must agree with that to 2e-6, it adds integers. (b) Against the
    reference as it is. With values of size ~1 and keys that occur 600 times per minibatch the reference's fp32
    running sums carry their own rounding into z: the exact-sum emulation itself ends 1.43e-5 away from the reference on
    this data (CPU only, no GPU involved), so (b) allows twice the usual size for the counted stragglers."""
    from oracle import oracle as O
    rng = np.random.default_rng(13)
    max_key = 2 ** 18
    job = checker.linear_create(algo, O.LOSS_LOGIT, 0.5, 0.1, 0.1, 1.0, max_key)
    srv = capi.Server(gpu, algo, 0.5, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key)
    os.environ["WH_DX_MIN_COVERAGE"] = "0"
    W, Z, CG = (np.zeros(max_key, np.float32) for _ in range(3))
    for m in range(6):
        blk = random_block(rng, 4000, 16, 30000, with_value=True, empty_rows=(m % 2 == 0))
        pg, pc = wk.process_minibatch(blk), checker.linear_minibatch(job, blk)
        assert pg[4] == pc[4] and abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0]) + 1e-6 and abs(pg[1] - pc[1]) <= 1e-3
        # the emulation: checker's Localize / SpMV::Times / handle, exact sum in between
        D = checker.localize(blk, max_key)
        keys = D.uniq.astype(np.int64)
        xw = checker.spmv_times(D, W[keys])
        y = np.where(D.label > 0, 1.0, -1.0).astype(np.float32)
        dual = (-y / (np.float32(1) + np.exp((y * xw).astype(np.float64)).astype(np.float32))).astype(np.float32)
        rows = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
        g = np.zeros(len(keys))
        np.add.at(g, D.index, (dual[rows] * D.value).astype(np.float32).astype(np.float64))
        w, z, cg = W[keys].copy(), Z[keys].copy(), CG[keys].copy()
        checker.handle_push(algo, 0.5, 0.1, 0.1, 1.0, g.astype(np.float32), w, z, cg, 1)
        W[keys], Z[keys], CG[keys] = w, z, cg
        keys, w, z, cg = checker.linear_model(job)
        got = srv.get_state(keys)
        k = keys.astype(np.int64)
        for name, x, e, r in zip(("w", "z", "cg"), got, (W[k], Z[k], CG[k]), (w, z, cg)):
            assert_state_close(x, e, rtol=2e-6, what="direct valued vs exact-sum emulation " + name, flip_abs=FLIP_ABS)
            assert_state_close(x, r, what="direct valued " + name, stragglers=STRAGGLERS, flip_abs=FLIP_ABS, loose_factor=2.0)
    os.environ.pop("WH_DX_MIN_COVERAGE", None)
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_direct_criteo_shape_vs_checker(gpu, checker):
    """Two full-size minibatches (100 000 x 39, 2^24 keys, FTRL): the benchmark's own configuration."""
    from oracle import oracle as O
    max_key = 2 ** 24
    job = checker.linear_create(O.ALGO_FTRL, O.LOSS_LOGIT, 1.0, 0.1, 0.1, 1.0, max_key)
    srv = capi.Server(gpu, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key)
    for m in range(2):
        blk = synth.criteo_minibatch(m)
        pg, pc = wk.process_minibatch(blk), checker.linear_minibatch(job, blk)
        assert pg[1] == pc[1] and abs(pg[0] - pc[0]) / 100000 <= 1e-4 and abs(pg[5] - pc[5]) <= 2
        if m > 0:
            assert abs(pg[2] - pc[2]) <= 1e-4     # AUC (minibatch 0 scores are all equal: any tie order is valid)
    keys, w, z, cg = checker.linear_model(job)
    for name, x, y in zip(("w", "z", "cg"), srv.get_state(keys), (w, z, cg)):
        assert_state_close(x, y, what="direct criteo " + name, stragglers=STRAGGLERS, flip_abs=FLIP_ABS)
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_direct_kdd_shape_equals_chain(gpu):
    """The KDD2012 shape (ids over 54.7 M values, flat: the table covers under half of the entries, key space 2^26) takes the
    shared-GPU step by default; three minibatches in flight, bit for bit the chain of launches."""
    blks = [synth.kdd_minibatch(m, rows=150000) for m in range(4)]
    a = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, synth.KDD_VOCAB, depth=3)
    b = run(gpu, blks, capi.ALGO_FTRL, capi.LOSS_LOGIT, synth.KDD_VOCAB, depth=3, WH_DIRECT=0)
    same_model(a, b)
    assert a[3] > 1000


def test_bench_line_parity_leg_one_gpu():
    """bench.py's own parity_check leg at N = 1 (the shared-GPU step against the checker, outside the timed regions)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--steps", "4", "--warmup", "3", "--rows", "20000",
                          "--pool", "4", "--no-others", "--no-cpu"], capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["parity_check"]["ok"] and line["gpu_launches"] > 0 and line["hot"]["keys_placed"] > 0, line
