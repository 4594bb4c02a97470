"""The BASELINE.json configurations at their full sizes, one or two minibatches each, against the CPU checker
(which finishes a 100k x 39 minibatch in under a second): config 2/3 (Criteo shape, FTRL / AdaGrad / SGD x logit /
squared hinge), config 4 (KDD2012 shape: 1M ragged rows, 2^26 keys) and config 5 (DiFacto dim 16 on Criteo shape).
Plus size-independent properties of the localizer at those sizes."""
import ctypes

import numpy as np
import pytest

from tests.util import FLIP_ABS, STRAGGLERS, assert_state_close
from wormhole_b200 import capi, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def criteo():
    return [synth.criteo_minibatch(m) for m in range(2)]


def state_close(got, want, rtol=1e-5):
    for name, a, b in zip(("w", "z", "cg"), got, want):
        assert_state_close(a, b, rtol=rtol, what="fullsize " + name, stragglers=STRAGGLERS, flip_abs=FLIP_ABS)


@pytest.mark.parametrize("algo,loss", [(capi.ALGO_FTRL, capi.LOSS_LOGIT), (capi.ALGO_ADAGRAD, capi.LOSS_SQUARE_HINGE),
                                       (capi.ALGO_SGD, capi.LOSS_LOGIT)])
def test_criteo_shape_minibatches(gpu, checker, criteo, algo, loss):
    max_key = 2 ** 24
    job = checker.linear_create(algo, loss, 1.0, 0.1, 0.1, 1.0, max_key)
    srv = capi.Server(gpu, algo, 1.0, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, loss, max_key)
    for blk in criteo:
        pg, pc = wk.process_minibatch(blk), checker.linear_minibatch(job, blk)
        assert pg[1] == pc[1] and pg[4] == pc[4] == 100000
        # logloss = objv / rows within the north star's 1e-4. The gap is the reference's own fp32 accumulation over
        # 50k terms per thread: on the first minibatch (w = 0) the exact objective is 100000 ln 2 = 69314.718, which
        # the device returns, while the reference's sum reads 69322.49.
        assert (np.isinf(pg[0]) and np.isinf(pc[0])) or abs(pg[0] - pc[0]) / 100000 <= 1e-4   # SGD overshoots: inf on both
        if not np.any(srv.get_state(blk.index[:1] % np.uint64(max_key))[0]) and loss == capi.LOSS_LOGIT and blk is criteo[0]:
            assert abs(pg[0] - 100000 * np.log(2.0)) <= 1e-2
        assert abs(pg[5] - pc[5]) <= 2
    keys, w, z, cg = checker.linear_model(job)
    state_close(srv.get_state(keys), (w, z, cg))
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_criteo_shape_pass_of_50_minibatches(gpu, checker):
    """One pass of 50 minibatches (5 M examples) in the benchmark's configuration -- FTRL L1+L2, logit, 2^24 keys, four
    minibatches in flight -- beside the reference's own worker + server: the numbers the scheduler prints (learn/linear/
    progress.h:21-23: logloss = sum objv / sum examples, accuracy and AUC = means of the per-minibatch values) agree
    within the north star's 1e-4, the AUC of every 100 000-row minibatch within 1e-4, |w|_0 within a few keys on the
    L1 threshold."""
    max_key, N, depth = 2 ** 24, 50, 4
    job = checker.linear_create(capi.ALGO_FTRL, capi.LOSS_LOGIT, 1.0, 0.1, 0.1, 1.0, max_key)
    srv = capi.Server(gpu, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key, depth=depth)
    pg, pc, pending = [], [], []
    for m in range(N):
        blk = synth.criteo_minibatch(500 + m)
        if len(pending) == depth:
            pg.append(wk.wait(pending.pop(0)[0]))
        pending.append((wk.submit(blk), blk))          # (the block stays alive until its ticket is collected)
        pc.append(checker.linear_minibatch(job, blk))
    pg += [wk.wait(t) for t, _ in pending]
    pg, pc = np.array(pg), np.array(pc)
    assert np.array_equal(pg[:, 3:5], pc[:, 3:5])                              # count, examples
    assert np.all(np.abs(pg[1:, 2] - pc[1:, 2]) <= 1e-4)                       # AUC per minibatch (0: all scores tie)
    assert np.all(np.abs(pg[:, 1] - pc[:, 1]) <= 1e-4)                         # accuracy per minibatch (a fraction of its rows)
    logloss = lambda p: p[:, 0].sum() / p[:, 4].sum()
    assert abs(logloss(pg) - logloss(pc)) <= 1e-4
    assert abs(pg[1:, 2].mean() - pc[1:, 2].mean()) <= 1e-4
    assert abs(pg[:, 1].mean() - pc[:, 1].mean()) <= 1e-4
    nnz_ref = int((checker.linear_model(job)[1] != 0).sum())
    assert abs(srv.nnz_w() - nnz_ref) <= max(5, 2e-4 * nnz_ref)
    assert pg[-1, 2] > 0.6 and logloss(pg[-10:]) < logloss(pg[:10])            # it did learn something
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_criteo_localize_properties(gpu, criteo):
    blk = criteo[0]
    a = gpu.localize(blk, 2 ** 24, method=capi.LOC_DIRECT)
    b = gpu.localize(blk, 2 ** 24, method=capi.LOC_SORT)
    for x, y in ((a.uniq, b.uniq), (a.cnt, b.cnt), (a.index, b.index)):      # two device strategies, one answer
        assert np.array_equal(x, y)
    assert np.all(np.diff(a.uniq.astype(np.int64)) > 0) and int(a.cnt.sum()) == blk.nnz
    assert np.array_equal(a.uniq[a.index], blk.index % np.uint64(2 ** 24))
    again = gpu.localize(blk, 2 ** 24)                                          # idempotent: scratch is left clean
    assert np.array_equal(again.index, a.index)


def test_kdd_shape_minibatch(gpu, checker):
    """Config 4: 1M rows, ~11 entries per row, one power-law vocabulary, max_key = 2^26, FTRL with L1."""
    max_key = 2 ** 26
    blk = synth.kdd_minibatch(0)
    assert blk.rows == 1000000
    got, want = gpu.localize(blk, max_key), checker.localize(blk, max_key)
    for x, y in ((got.uniq, want.uniq), (got.cnt, want.cnt), (got.offset, want.offset), (got.index, want.index)):
        assert np.array_equal(x, y)
    job = checker.linear_create(capi.ALGO_FTRL, capi.LOSS_LOGIT, 1.0, 0.0, 0.1, 1.0, max_key)
    srv = capi.Server(gpu, capi.ALGO_FTRL, 1.0, 0.0, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, max_key)
    pg, pc = wk.process_minibatch(blk), checker.linear_minibatch(job, blk)
    # w = 0: the exact objective is 1e6 ln 2 = 693147.18; the reference's fp32 sum over 500k terms per thread
    # returns 689894 (0.5 % low). The device must hit the exact value; the reference only has to be near it.
    assert pg[1] == pc[1] and abs(pg[0] - 1e6 * np.log(2.0)) <= 0.1 and abs(pg[0] - pc[0]) <= 1e-2 * abs(pc[0])
    keys, w, z, cg = checker.linear_model(job)
    state_close(srv.get_state(keys), (w, z, cg))
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_difacto_dim16_criteo_shape(gpu, checker, criteo):
    """Config 5: embedding dim 16, threshold 16, lambda_l1 4, lr_eta .01 (doc/tutorial/criteo_kaggle.rst:112-127)."""
    max_key, seed = 2 ** 24, 5
    kw = dict(l1=4.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=False, dim=16, threshold=16, v_l2=1e-4, init_scale=.01)
    # rand() is process-global: draw the stream for the device first, then let the checker re-seed for itself
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(ctypes.c_uint(seed))
    stream = np.array([libc.rand() for _ in range(1 << 20)], np.int32)
    d = capi.Difacto(gpu, max_key=max_key, **kw)
    d.set_init_stream(stream)
    job = checker.difacto_create(max_key=max_key, seed=seed, **kw)
    for blk in criteo:
        pg, pc = d.process_minibatch(blk, True, 0), checker.difacto_minibatch(job, blk, True, 0)
        # count, new_ex, new_V exact; new_w up to a couple of keys whose z sits within rounding of the L1 threshold
        assert np.array_equal(pg[[4, 5, 7]], pc[[4, 5, 7]]) and abs(pg[6] - pc[6]) <= 2
        assert abs(pg[0] - pc[0]) / 100000 <= 1e-4 and abs(pg[2] - pc[2]) / 100000 <= 1e-4
    keys, cnt, size, w, cg = checker.difacto_model(job, 16)
    gcnt, gsize, gw, gcg = d.get_state(keys)
    assert np.array_equal(gcnt, cnt) and np.array_equal(gsize, size) and (size == 17).sum() > 1000
    state_close((gw, gcg), (w, cg), rtol=2e-5)
    d.close(); checker.difacto_destroy(job)
