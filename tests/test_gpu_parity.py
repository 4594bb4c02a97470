"""Parity of the CUDA path (through the C-ABI, host-pointer convention) against the CPU checker:
oracle/_ref (the reference's own headers) when it travelled with the snapshot, else the C port.

Bars (BASELINE.json north_star): localized keys / index remaps / counts bit-exact; forward
products bit-exact; gradients and weights within 1e-5 relative in fp32, where "relative" for a
sum of signed terms means relative to the sum of |terms| (the reference's own fp32 accumulation
error is of that order and depends on its summation order); AUC within 1e-4."""
import numpy as np
import pytest

from tests.util import FLIP_ABS, STRAGGLERS, assert_state_close, random_block, wide_ids
from wormhole_b200 import capi
from wormhole_b200.rowblock import RowBlock

pytestmark = pytest.mark.gpu

RTOL = 1e-5   # north_star: "within 1e-5 relative in fp32"
NO = capi.NO_MAX_KEY


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


def same_loc(a, b, counts=True):
    assert np.array_equal(a.uniq, b.uniq)
    if counts:
        assert np.array_equal(a.cnt, b.cnt)
    assert np.array_equal(a.offset, b.offset)
    assert np.array_equal(a.index, b.index)
    assert (a.value is None) == (b.value is None)
    if a.value is not None:
        assert np.array_equal(a.value, b.value)
    assert np.array_equal(a.label, b.label)


def abs_terms(D, x, U):
    """sum_j |x_i * val_ij| per output column: the scale of a transposed product's rounding."""
    rows = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
    t = np.abs(x[rows].astype(np.float64) * (1.0 if D.value is None else D.value.astype(np.float64)))
    out = np.zeros(U)
    np.add.at(out, D.index, t)
    return out


def close_sum(got, want, scale, rtol=RTOL):
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    bound = rtol * scale + 1e-30
    bad = err > bound
    assert not bad.any(), "max err/bound = %g at %d" % ((err / bound).max(), int(np.argmax(err / bound)))


# ---------------------------------------------------------------- (1) localizer ------------

def test_known_answers(gpu):
    assert capi.reverse_bytes(5) == 0x5000000000000000
    blk = RowBlock([0, 4, 6], [5, 1000000007, 5, 42, 42, 7], None, [1, 0])
    r = gpu.localize(blk)
    assert [hex(int(x)) for x in r.uniq] == ["0x5000000000000000", "0x7000000000000000",
                                             "0x70aca9b300000000", "0xa200000000000000"]
    assert r.index.tolist() == [0, 2, 0, 3, 3, 1]
    assert r.cnt.tolist() == [2, 1, 1, 2]


@pytest.mark.parametrize("max_key", [NO, 2 ** 24, 1000003, 7, 2 ** 28])
@pytest.mark.parametrize("with_value", [False, True])
def test_localize_bit_exact(gpu, checker, max_key, with_value):
    rng = np.random.default_rng(11)
    methods = [capi.LOC_SORT] + ([capi.LOC_DIRECT] if max_key <= 2 ** 28 else [])
    for rows, mx, space in [(1, 3, 10), (50, 8, 100), (3000, 40, 2 ** 40), (20000, 39, 2 ** 62)]:
        blk = random_block(rng, rows, mx, space, with_value, empty_rows=(rows > 1))
        if blk.nnz == 0:
            continue
        want = checker.localize(blk, max_key)
        for m in methods:
            same_loc(gpu.localize(blk, max_key, method=m), want)
        same_loc(gpu.localize(blk, max_key, float_cnt=True), checker.localize(blk, max_key, float_cnt=True))
        same_loc(gpu.localize(blk, max_key, counts=False), want, counts=False)


def test_localize_wide_keys_all_digits(gpu, checker):
    rng = np.random.default_rng(12)
    n = 100000
    blk = RowBlock(np.arange(0, n + 1, 4), wide_ids(rng, n), None, np.zeros(n // 4))
    same_loc(gpu.localize(blk, NO), checker.localize(blk, NO))
    same_loc(gpu.localize(blk, 2 ** 24, method=capi.LOC_SORT), checker.localize(blk, 2 ** 24))


def test_localize_criteo_shape(gpu, checker):
    from wormhole_b200 import synth
    blk = synth.criteo_minibatch(3, rows=100000)
    want = checker.localize(blk, 2 ** 24)
    same_loc(gpu.localize(blk, 2 ** 24, method=capi.LOC_DIRECT), want)
    same_loc(gpu.localize(blk, 2 ** 24, method=capi.LOC_SORT), want)
    # properties that hold at any size: ascending unique keys, ranks consistent, counts sum to nnz
    got = gpu.localize(blk, 2 ** 24)
    assert np.all(np.diff(got.uniq.astype(np.int64)) > 0)
    assert np.array_equal(got.uniq[got.index], blk.index % np.uint64(2 ** 24))
    assert int(got.cnt.sum()) == blk.nnz


def test_localize_device_input(gpu, checker):
    import torch
    rng = np.random.default_rng(13)
    blk = random_block(rng, 5000, 30, 2 ** 30, True, empty_rows=True)
    dev = RowBlock.__new__(RowBlock)
    dev.offset = torch.from_numpy(blk.offset.astype(np.int64)).cuda()
    dev.index = torch.from_numpy(blk.index.astype(np.int64)).cuda()
    dev.value = torch.from_numpy(blk.value).cuda()
    dev.label = torch.from_numpy(blk.label).cuda()
    torch.cuda.synchronize()
    same_loc(gpu.localize(dev, 2 ** 24, mem=capi.WH_DEVICE), checker.localize(blk, 2 ** 24))


def test_remap_with_foreign_dict(gpu, checker):
    rng = np.random.default_rng(14)
    blk = random_block(rng, 2000, 10, 300, True)
    full = checker.localize(blk, 2 ** 24).uniq
    for keep in (full[::2], full[:1], full[-3:], np.array([2 ** 30], np.uint64), full):
        a, b = gpu.remap(blk, keep, 2 ** 24), checker.remap(blk, keep, 2 ** 24)
        assert np.array_equal(a.offset, b.offset) and np.array_equal(a.index, b.index)
        if b.value is not None:
            assert np.array_equal(a.value, b.value)
        assert np.array_equal(a.label, b.label)


def test_empty_block(gpu):
    blk = RowBlock([0], np.zeros(0, np.uint64), None, np.zeros(0, np.float32))
    r = gpu.localize(blk, 2 ** 24)
    assert len(r.uniq) == 0 and len(r.index) == 0


def test_errors_are_reported(gpu):
    blk = RowBlock([0, 1], [3], None, [1])
    with pytest.raises(capi.WormholeError):
        gpu.localize(blk, 2 ** 40, method=capi.LOC_DIRECT)   # direct addressing needs a hashed space
    with pytest.raises(capi.WormholeError):
        gpu.linear_loss(3, gpu.localize(blk), np.zeros(1, np.float32))  # not a Config::Loss value (loss.h:178-180)


def test_square_loss_closed_form(gpu, checker):
    """Config::SQUARE = 1/2 (p - y)^2 (learn/linear/config.proto:30-34). The reference declares it and never implemented
    it (learn/linear/loss.h:163-166), so there is NO ORACLE: the device is held to the closed form in fp64 -- dual = Xw - y,
    grad = X^T dual, objv = sum 1/2 dual^2 -- standalone and through the worker (AdaGrad update via the checker's handle)."""
    from oracle import oracle as O
    rng = np.random.default_rng(77)
    max_key = 2 ** 20
    srv = capi.Server(gpu, capi.ALGO_ADAGRAD, 0.01, 0.1, 0.1, 1.0, 0, max_key)
    wk = capi.Worker(gpu, srv, capi.LOSS_SQUARE, max_key)
    W, CG = np.zeros(max_key, np.float32), np.zeros(max_key, np.float32)
    for m in range(4):
        blk = random_block(rng, 3000, 12, 20000, with_value=(m % 2 == 1), empty_rows=False)
        D = checker.localize(blk, max_key)
        keys = D.uniq.astype(np.int64)
        w = W[keys].copy()
        prog, grad = gpu.linear_loss(capi.LOSS_SQUARE, D, w)[:2]
        rows = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
        val = np.ones(len(D.index)) if D.value is None else D.value.astype(np.float64)
        xw = np.zeros(D.rows); np.add.at(xw, rows, w[D.index].astype(np.float64) * val)
        y = np.where(D.label > 0, 1.0, -1.0)
        dual = xw - y
        g = np.zeros(len(keys)); np.add.at(g, D.index, dual[rows] * val)
        absg = np.zeros(len(keys)); np.add.at(absg, D.index, np.abs(dual[rows] * val))
        assert np.all(np.abs(grad - g) <= 1e-5 * absg + 1e-6)
        assert abs(prog[0] - 0.5 * (dual ** 2).sum()) <= 1e-5 * 0.5 * (dual ** 2).sum()
        pg = wk.process_minibatch(blk)
        assert abs(pg[0] - prog[0]) <= 1e-9 * abs(prog[0]) and pg[4] == D.rows
        cg = CG[keys].copy()
        checker.handle_push(O.ALGO_ADAGRAD, 0.01, 0.1, 0.1, 1.0, grad.astype(np.float32), w, None, cg, 1)
        W[keys], CG[keys] = w, cg
        gw, _, gcg = srv.get_state(D.uniq)
        assert_state_close(gw, w, what="square w", stragglers=STRAGGLERS, flip_abs=FLIP_ABS)
        assert_state_close(gcg, cg, what="square cg", stragglers=STRAGGLERS, flip_abs=FLIP_ABS)
    wk.close(); srv.close()


# ---------------------------------------------------------------- (2)/(5) SpMV, SpMM -------

@pytest.mark.parametrize("with_value", [False, True])
def test_spmv(gpu, checker, with_value):
    rng = np.random.default_rng(21)
    for rows, mx, space in [(1, 5, 10), (800, 30, 5000), (30000, 39, 2 ** 30), (300, 300, 100)]:
        blk = random_block(rng, rows, mx, space, with_value)
        if blk.nnz == 0:
            continue
        D = checker.localize(blk, 2 ** 24)
        U = len(D.uniq)
        x = rng.standard_normal(U).astype(np.float32)
        d = rng.standard_normal(D.rows).astype(np.float32)
        y0 = rng.standard_normal(D.rows).astype(np.float32)   # rows without entries keep y (spmv.h:79)
        assert np.array_equal(gpu.spmv_times(D, x, y0.copy()), checker.spmv_times(D, x, y0.copy()))
        close_sum(gpu.spmv_trans_times(D, d, U), checker.spmv_trans_times(D, d, U), abs_terms(D, d, U))


def test_spmv_trans_is_order_independent(gpu, checker):
    """Fixed-point accumulation: bit-identical run to run, and closer to the fp64 truth than 1e-5."""
    rng = np.random.default_rng(22)
    blk = random_block(rng, 50000, 39, 2000, False)      # hot columns: ~1000 entries per key
    D = checker.localize(blk, 2 ** 24)
    U = len(D.uniq)
    d = rng.standard_normal(D.rows).astype(np.float32)
    a, b = gpu.spmv_trans_times(D, d, U), gpu.spmv_trans_times(D, d, U)
    assert np.array_equal(a, b)
    rows = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
    truth = np.zeros(U)
    np.add.at(truth, D.index, d[rows].astype(np.float64))
    close_sum(a, truth.astype(np.float32), abs_terms(D, d, U), rtol=1e-6)
    close_sum(checker.spmv_trans_times(D, d, U), truth.astype(np.float32), abs_terms(D, d, U))


@pytest.mark.parametrize("dim", [1, 5, 16])
@pytest.mark.parametrize("with_value", [False, True])
def test_spmm(gpu, checker, dim, with_value):
    rng = np.random.default_rng(23)
    blk = random_block(rng, 2000, 30, 5000, with_value)
    D = checker.localize(blk, 2 ** 24)
    U = len(D.uniq)
    X = rng.standard_normal(U * dim).astype(np.float32)
    Y = rng.standard_normal(D.rows * dim).astype(np.float32)
    Z = rng.standard_normal(U * dim).astype(np.float32)
    assert np.array_equal(gpu.spmm_times(D, X, dim), checker.spmm_times(D, X, dim))
    scale = np.stack([abs_terms(D, np.ascontiguousarray(Y.reshape(-1, dim)[:, k]), U) for k in range(dim)], 1).reshape(-1)
    close_sum(gpu.spmm_trans_times(D, Y, U * dim, dim), checker.spmm_trans_times(D, Y, U * dim, dim), scale)
    close_sum(gpu.spmm_trans_times(D, Y, U * dim, dim, 0.5, Z),
              checker.spmm_trans_times(D, Y, U * dim, dim, 0.5, Z), scale + np.abs(Z) * 0.5)


# ---------------------------------------------------------------- (3) loss ------------------

def test_bin_eval(gpu, checker):
    rng = np.random.default_rng(31)
    for n in (1, 2, 31, 33, 777, 20000, 300000, 1200000):   # > 2^20 takes the multi-launch sort
        label = (rng.random(n) < 0.3).astype(np.float32)
        pred = (rng.standard_normal(n) * 3).astype(np.float32)
        a, b = gpu.bin_eval(label, pred), checker.bin_eval(label, pred, 2)
        assert a["acc"] == b["acc"]
        # the objective is a sum over examples: the device sums the fp32 terms in fp64, the
        # reference in fp32 (binary_class_evaluation.h:66-74), whose rounding grows with n.
        y = np.where(label > 0, 1, -1).astype(np.float32)
        terms = np.log((1 + np.exp(-y * pred, dtype=np.float32)).astype(np.float32), dtype=np.float32)
        exact = float(terms.astype(np.float64).sum())
        assert abs(a["logit_objv"] - exact) <= 2e-6 * exact + 1e-6
        ref_tol = (2e-6 + 4e-10 * n) * exact + 1e-6
        assert abs(a["logit_objv"] - b["logit_objv"]) <= ref_tol
        assert abs(a["logloss"] - b["logloss"]) <= ref_tol
        assert abs(a["auc"] - b["auc"]) <= 1e-4       # reference accumulates the area in fp32
        assert a["auc_lo"] - 1e-4 <= b["auc"] <= a["auc_hi"] + 1e-4
    # equal scores: any tie order of the reference's unstable sort lies inside [auc_lo, auc_hi]
    label = (rng.random(5000) < 0.5).astype(np.float32)
    pred = rng.integers(0, 7, 5000).astype(np.float32)
    a, b = gpu.bin_eval(label, pred), checker.bin_eval(label, pred, 2)
    assert a["auc_lo"] - 1e-5 <= b["auc"] <= a["auc_hi"] + 1e-5
    assert a["auc_lo"] <= a["auc"] <= a["auc_hi"]
    one = np.ones(10, np.float32)
    assert gpu.bin_eval(one, one)["auc"] == 1.0      # single class (binary_class_evaluation.h:36)


@pytest.mark.parametrize("loss", [capi.LOSS_LOGIT, capi.LOSS_SQUARE_HINGE])
@pytest.mark.parametrize("with_value", [False, True])
def test_linear_loss(gpu, checker, loss, with_value):
    rng = np.random.default_rng(32)
    blk = random_block(rng, 20000, 25, 30000, with_value)
    D = checker.localize(blk, 2 ** 24)
    U = len(D.uniq)
    w = (rng.standard_normal(U) * 0.3).astype(np.float32)
    pg, gg, xw, dual = gpu.linear_loss(loss, D, w, want_xw=True)
    pc, gc = checker.linear_loss(loss, D, w, 2)
    assert np.array_equal(xw, checker.spmv_times(D, w))                      # Xw bit-exact
    assert pg[1] == pc[1] and pg[3] == pc[3] and pg[4] == pc[4]             # acc, count, new_ex
    # objv: the reference adds the per-example terms in fp32 (rounding grows with the row count,
    # ~2e-5 relative at a few thousand rows); the device adds the same fp32 terms in fp64.
    assert abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0])                           # objv (north_star: logloss 1e-4)
    assert abs(pg[2] - pc[2]) <= 1e-4                                        # auc
    mul = 2.0 if loss == capi.LOSS_SQUARE_HINGE else 1.0
    close_sum(gg, gc, mul * abs_terms(D, dual, U))
    # normwise: ||g_gpu - g_ref|| <= 1e-5 ||g_ref||
    assert np.linalg.norm(gg.astype(np.float64) - gc) <= RTOL * np.linalg.norm(gc.astype(np.float64))


# ---------------------------------------------------------------- (4) server handles -------

@pytest.mark.parametrize("algo", [capi.ALGO_SGD, capi.ALGO_ADAGRAD, capi.ALGO_FTRL])
def test_handle_push_bit_exact(gpu, checker, algo):
    """Same gradients in -> bit-identical {w, z, sq_cum_grad}, |w|_0 and SGD t out."""
    rng = np.random.default_rng(41)
    n = 20000
    for eta, beta, l1, l2 in [(None, None, 1.0, 0.0), (0.1, 1.0, 1.0, 0.1), (0.05, 2.0, 0.0, 0.0)]:
        srv = capi.Server(gpu, algo, l1, l2, eta, beta, 1000, 1000 + 4 * n)
        keys = (1000 + np.sort(rng.choice(4 * n, n, replace=False))).astype(np.uint64)
        w = (rng.standard_normal(n) * (rng.random(n) < 0.7)).astype(np.float32)
        z = (rng.standard_normal(n)).astype(np.float32)
        cg = np.abs(rng.standard_normal(n)).astype(np.float32)
        srv.set_state(keys, w, z, cg)
        nnz0 = srv.nnz_w()
        assert nnz0 == int((w != 0).sum())
        t = 1
        for step in range(3):
            g = (rng.standard_normal(n) * 2).astype(np.float32)
            srv.push(keys, g)
            dn, t = checker.handle_push(algo, l1, l2, eta, beta, g, w, z, cg, t)
            nnz0 += dn
            gw, gz, gc = srv.get_state(keys)
            assert np.array_equal(gw, w)
            if algo == capi.ALGO_FTRL:
                assert np.array_equal(gz, z)
            if algo != capi.ALGO_SGD:
                assert np.array_equal(gc, cg)
            assert srv.nnz_w() == nnz0
            assert np.array_equal(srv.pull(keys), w)
        if algo == capi.ALGO_SGD:
            assert srv.sgd_t() == t
        srv.close()


def test_server_rejects_foreign_keys(gpu):
    srv = capi.Server(gpu, capi.ALGO_FTRL, key_begin=100, key_end=200)
    with pytest.raises(capi.WormholeError):
        srv.pull(np.array([5], np.uint64))
    with pytest.raises(capi.WormholeError):
        srv.push(np.array([200], np.uint64), np.ones(1, np.float32))
    srv.close()


# ---------------------------------------------------------------- the whole path ------------

def weights_close(gpu_state, ref_state, rtol=1e-5, what=""):
    """Per-entry fields within rtol of the field's magnitude scale; reports the worst offender."""
    for name, a, b in zip(("w", "z", "cg"), gpu_state, ref_state):
        assert_state_close(a, b, rtol=rtol, what=what + name, stragglers=STRAGGLERS, flip_abs=FLIP_ABS)


@pytest.mark.parametrize("algo", [capi.ALGO_SGD, capi.ALGO_ADAGRAD, capi.ALGO_FTRL])
@pytest.mark.parametrize("loss", [capi.LOSS_LOGIT, capi.LOSS_SQUARE_HINGE])
@pytest.mark.parametrize("max_key", [2 ** 20, 1000003])
def test_process_minibatch_vs_reference(gpu, checker, algo, loss, max_key):
    """AsgdWorker::ProcessMinibatch, 1 worker + 1 server, max_concurrency = 1."""
    rng = np.random.default_rng(51)
    l1, l2, eta, beta = 0.5, 0.1, 0.1, 1.0
    job = checker.linear_create(algo, loss, l1, l2, eta, beta, max_key)
    srv = capi.Server(gpu, algo, l1, l2, eta, beta, 0, max_key)
    wk = capi.Worker(gpu, srv, loss, max_key)
    nnz_w = 0
    for m in range(8):
        blk = random_block(rng, 3000, 12, 20000, with_value=(m % 2 == 1), empty_rows=False)
        pg = wk.process_minibatch(blk)
        pc = checker.linear_minibatch(job, blk)
        assert pg[4] == pc[4] and pg[3] == pc[3]
        assert abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0]) + 1e-6, (m, pg, pc)   # fp32 vs fp64 accumulation
        assert abs(pg[1] - pc[1]) <= 1e-3
        nnz_w += pg[5]
        keys, w, z, cg = checker.linear_model(job)
        weights_close(srv.get_state(keys), (w, z, cg))
    assert abs(nnz_w - srv.nnz_w()) == 0
    keys, w, z, cg = checker.linear_model(job)
    assert abs(srv.nnz_w() - int((w != 0).sum())) <= max(2, 1e-3 * len(w))
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_default_key_mode_uses_sort_path(gpu, checker):
    """ps::FLAGS_max_key at its default: nibble-reversed 64-bit keys (learn/base/localizer.h:116-121).
    The dense shard cannot hold them; the localizer + loss still must agree."""
    rng = np.random.default_rng(52)
    blk = random_block(rng, 4000, 22, 127, False, empty_rows=False)   # agaricus-like ids
    D = gpu.localize(blk)
    same_loc(D, checker.localize(blk))
    w = (rng.standard_normal(len(D.uniq)) * 0.2).astype(np.float32)
    pg, gg = gpu.linear_loss(capi.LOSS_LOGIT, D, w)
    pc, gc = checker.linear_loss(capi.LOSS_LOGIT, D, w, 2)
    assert abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0])
    assert np.linalg.norm(gg.astype(np.float64) - gc) <= RTOL * np.linalg.norm(gc.astype(np.float64))


def test_minibatches_in_flight_equal_sequential(gpu):
    """submit/wait with 3 minibatches in flight gives the very numbers of one-at-a-time processing:
    every pull is ordered after all earlier pushes (max_concurrency changes overlap, not results)."""
    rng = np.random.default_rng(53)
    blks = [random_block(rng, 4000, 20, 30000, with_value=(m % 3 == 0), empty_rows=False) for m in range(7)]
    states, progs = [], []
    for depth in (1, 3):
        srv = capi.Server(gpu, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, 2 ** 20)
        wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, 2 ** 20, depth=depth)
        out, pending = [], []
        for b in blks:
            if len(pending) == depth:
                out.append(wk.wait(pending.pop(0)))
            pending.append(wk.submit(b))
        out += [wk.wait(t) for t in pending]
        keys = np.unique(np.concatenate([b.index % np.uint64(2 ** 20) for b in blks]))
        states.append(srv.get_state(keys)); progs.append(np.array(out))
        with pytest.raises(capi.WormholeError):
            wk.wait(12345)                      # not in flight
        wk.close(); srv.close()
    assert np.array_equal(progs[0], progs[1])
    for a, b in zip(states[0], states[1]):
        assert np.array_equal(a, b)


def test_unhashed_64bit_keys_end_to_end(gpu, checker):
    """ps::FLAGS_max_key at its default (learn/base/localizer.h:116-121): nibble-reversed ids over the whole
    64-bit range, radix-sort localizer, open-addressing shard; including a table that has to probe a lot."""
    from tests.util import wide_ids
    rng = np.random.default_rng(54)
    pool = wide_ids(rng, 3000)                       # 3000 distinct 64-bit feature ids
    job = checker.linear_create(capi.ALGO_FTRL, capi.LOSS_LOGIT, 0.5, 0.1, 0.1, 1.0, capi.NO_MAX_KEY)
    srv = capi.Server(gpu, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, capi.NO_MAX_KEY, hash_slots=4096)   # 73 % full
    wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, capi.NO_MAX_KEY)
    for m in range(5):
        nnz = rng.integers(1, 15, 2000)
        off = np.zeros(2001, np.uint64); off[1:] = np.cumsum(nnz)
        blk = RowBlock(off, pool[rng.integers(0, len(pool), int(off[-1]))], None, (rng.random(2000) < 0.4).astype(np.float32))
        pg, pc = wk.process_minibatch(blk), checker.linear_minibatch(job, blk)
        assert pg[1] == pc[1] and abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0]) and abs(pg[5] - pc[5]) <= 1
    keys, w, z, cg = checker.linear_model(job)
    weights_close(srv.get_state(keys), (w, z, cg))
    assert np.array_equal(srv.pull(np.array([12345], np.uint64)), [0.0])      # a key never seen: default entry
    with pytest.raises(capi.WormholeError):                                     # more keys than slots: reported
        small = capi.Server(gpu, capi.ALGO_SGD, 0, 0, 0.1, 1.0, 0, capi.NO_MAX_KEY, hash_slots=16)
        small.push(pool[:100], np.ones(100, np.float32))
    wk.close(); srv.close(); checker.linear_destroy(job)


def test_hot_key_table_changes_nothing(gpu, monkeypatch):
    """The worker's hot-key table (csrc/hot.cuh) is a cache: with it (codes, fused forward/backward, shared-memory
    accumulators) and without it (WH_HOT=0: ranks, separate kernels) the progress numbers and the model are
    bit-identical -- the fixed-point gradient sums do not depend on where they were added up. Skewed ids, odd entry
    counts, train and evaluation minibatches, every updater."""
    rng = np.random.default_rng(55)
    blks = []
    for m in range(6):
        rows = 3001 + m
        nnz = rng.integers(1, 25, rows)
        off = np.zeros(rows + 1, np.uint64); off[1:] = np.cumsum(nnz)
        ids = (rng.zipf(1.3, int(off[-1])) % 50000).astype(np.uint64) * np.uint64(2654435761)   # power law, hashed
        blks.append(RowBlock(off, ids, None, (rng.random(rows) < 0.3).astype(np.float32)))
    # (the last case has a directory too large for the cooperative scan: look-back scan + separate table-key launches)
    for algo, loss, mk in [(capi.ALGO_FTRL, capi.LOSS_LOGIT, 2 ** 22), (capi.ALGO_ADAGRAD, capi.LOSS_SQUARE_HINGE, 2 ** 22),
                           (capi.ALGO_SGD, capi.LOSS_LOGIT, 2 ** 22), (capi.ALGO_SGD, capi.LOSS_LOGIT, 2 ** 28)]:
        res = []
        # third: Handle::Pull folded into the scan's emit; fourth: the table rebuilt every 3 minibatches
        for hot, emit, refresh in (("1", "0", "0"), ("0", "0", "0"), ("1", "1", "0"), ("1", "0", "3")):
            monkeypatch.setenv("WH_HOT", hot)
            monkeypatch.setenv("WH_PULL_AT_EMIT", emit)
            monkeypatch.setenv("WH_HOT_REFRESH", refresh)
            srv = capi.Server(gpu, algo, 0.5, 0.1, 0.1, 1.0, 0, mk)
            wk = capi.Worker(gpu, srv, loss, mk, depth=2)
            pend, out = [], []
            for m, b in enumerate(blks):
                if len(pend) == 2:
                    out.append(wk.wait(pend.pop(0)))
                pend.append(wk.submit(b, train=(m != 3)))
            out += [wk.wait(t) for t in pend]
            keys = np.unique(np.concatenate([b.index % np.uint64(mk) for b in blks]))
            res.append((np.array(out), srv.get_state(keys), srv.nnz_w()))
            wk.close(); srv.close()
        for other in res[1:]:
            assert np.array_equal(res[0][0], other[0])
            for a, b in zip(res[0][1], other[1]):
                assert np.array_equal(a, b)
            assert res[0][2] == other[2]


def test_bounded_staleness_option(gpu):
    """wh_worker_set_max_delay: d = 0 is the default (sequential numbers); d > 0 lets a pull overtake the worker's
    own last d pushes, as the reference does with max_concurrency minibatches in flight -- a different but close
    trajectory; the arguments are validated."""
    rng = np.random.default_rng(56)
    blks = [random_block(rng, 3000, 20, 20000, with_value=False, empty_rows=False) for _ in range(8)]
    objv = {}
    for d in (0, 2):
        srv = capi.Server(gpu, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, 2 ** 20)
        wk = capi.Worker(gpu, srv, capi.LOSS_LOGIT, 2 ** 20, depth=3)
        with pytest.raises(capi.WormholeError):
            wk.set_max_delay(3)                          # needs d < depth
        wk.set_max_delay(d)
        pend, out = [], []
        for b in blks:
            if len(pend) == 3:
                out.append(wk.wait(pend.pop(0)))
            pend.append(wk.submit(b))
            if len(pend) == 1 and d:
                with pytest.raises(capi.WormholeError):
                    wk.set_max_delay(0)                  # not while minibatches are in flight
        out += [wk.wait(t) for t in pend]
        objv[d] = np.array(out)[:, 0]
        wk.close(); srv.close()
    assert objv[0][0] == objv[2][0]                      # the first minibatch sees the same (empty) model
    # stale by at most two minibatches: close (and identical whenever the GPU was done before the next submit --
    # max_delay is a permission, not an obligation)
    assert np.all(np.abs(objv[2] - objv[0]) <= 0.05 * objv[0])
