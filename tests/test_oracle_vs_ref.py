"""Pins the C port (oracle/wh_oracle*.c) against oracle/_ref -- the reference's own headers
compiled unchanged -- on seeded inputs, bit for bit.  Runs where oracle/_ref exists (the build
container, or a box the .so travelled to); elsewhere the golden vectors do this job."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.util import random_block, wide_ids
from wormhole_b200.rowblock import RowBlock

MAXKEYS = [O.NO_MAX_KEY, 2 ** 24, 1000003, 7]


def same_loc(a, b):
    assert np.array_equal(a.uniq, b.uniq)
    assert np.array_equal(a.cnt, b.cnt)
    assert np.array_equal(a.offset, b.offset)
    assert np.array_equal(a.index, b.index)
    assert (a.value is None) == (b.value is None)
    if a.value is not None:
        assert np.array_equal(a.value, b.value)
    assert np.array_equal(a.label, b.label)


def test_known_answers(port, ref):
    # SURVEY.md section 8c spot checks
    for o in (port, ref):
        assert o.reverse_bytes(5) == 0x5000000000000000
        assert o.l1l2_solve(1, 0, 3, 2) == 1.0
        blk = RowBlock([0, 4, 6], [5, 1000000007, 5, 42, 42, 7], None, [1, 0])
        r = o.localize(blk)
        assert [hex(int(x)) for x in r.uniq] == ["0x5000000000000000", "0x7000000000000000",
                                                 "0x70aca9b300000000", "0xa200000000000000"]
        assert r.index.tolist() == [0, 2, 0, 3, 3, 1]
        assert r.cnt.tolist() == [2, 1, 1, 2]


@pytest.mark.parametrize("max_key", MAXKEYS)
@pytest.mark.parametrize("with_value", [False, True])
def test_localize(port, ref, max_key, with_value):
    rng = np.random.default_rng(1)
    for rows, mx, space in [(1, 3, 10), (50, 8, 100), (3000, 40, 2 ** 40)]:
        blk = random_block(rng, rows, mx, space, with_value, empty_rows=(rows > 1))
        if blk.nnz == 0:
            continue
        same_loc(port.localize(blk, max_key), ref.localize(blk, max_key))
        same_loc(port.localize(blk, max_key, float_cnt=True), ref.localize(blk, max_key, float_cnt=True))
    blk = RowBlock(np.arange(0, 4001, 4), wide_ids(rng, 4000), None, np.zeros(1000))
    same_loc(port.localize(blk, max_key), ref.localize(blk, max_key))


def test_remap_with_foreign_dict(port, ref):
    rng = np.random.default_rng(2)
    blk = random_block(rng, 500, 10, 300, True)
    full = ref.localize(blk, 2 ** 24).uniq
    for keep in (full[::2], full[:1], full[-3:], np.array([2 ** 30], np.uint64)):
        same_loc_nouniq = (port.remap(blk, keep, 2 ** 24), ref.remap(blk, keep, 2 ** 24))
        a, b = same_loc_nouniq
        assert np.array_equal(a.offset, b.offset) and np.array_equal(a.index, b.index)
        assert np.array_equal(a.value, b.value) and np.array_equal(a.label, b.label)


@pytest.mark.parametrize("with_value", [False, True])
def test_spmv_spmm(port, ref, with_value):
    rng = np.random.default_rng(3)
    blk = random_block(rng, 800, 30, 5000, with_value)
    D = ref.localize(blk, 2 ** 24)
    U = len(D.uniq)
    x = rng.standard_normal(U).astype(np.float32)
    d = rng.standard_normal(D.rows).astype(np.float32)
    for nt in (1, 2, 3):
        assert np.array_equal(port.spmv_times(D, x), ref.spmv_times(D, x, nt=nt))
        assert np.array_equal(port.spmv_trans_times(D, d, U), ref.spmv_trans_times(D, d, U, nt=nt))
    for dim in (1, 5, 16):
        X = rng.standard_normal(U * dim).astype(np.float32)
        Y = rng.standard_normal(D.rows * dim).astype(np.float32)
        Z = rng.standard_normal(U * dim).astype(np.float32)
        assert np.array_equal(port.spmm_times(D, X, dim), ref.spmm_times(D, X, dim))
        assert np.array_equal(port.spmm_trans_times(D, Y, U * dim, dim),
                              ref.spmm_trans_times(D, Y, U * dim, dim))
        assert np.array_equal(port.spmm_trans_times(D, Y, U * dim, dim, 0.5, Z),
                              ref.spmm_trans_times(D, Y, U * dim, dim, 0.5, Z))


@pytest.mark.parametrize("nt", [1, 2])
def test_bin_eval(port, ref, nt):
    rng = np.random.default_rng(4)
    for n in (1, 2, 777, 20000):
        label = (rng.random(n) < 0.3).astype(np.float32)
        pred = rng.standard_normal(n).astype(np.float32) * 3  # distinct scores w.p. ~1
        a, b = port.bin_eval(label, pred, nt), ref.bin_eval(label, pred, nt)
        for k in ("auc", "acc", "logloss", "logit_objv"):
            assert np.float32(a[k]) == np.float32(b[k]), (k, n, a[k], b[k])
    # single-class and all-equal-score corner cases of AUC (binary_class_evaluation.h:36)
    one = np.ones(10, np.float32)
    assert port.bin_eval(one, one)["auc"] == ref.bin_eval(one, one)["auc"] == 1.0


@pytest.mark.parametrize("loss", [O.LOSS_LOGIT, O.LOSS_SQUARE_HINGE])
@pytest.mark.parametrize("with_value", [False, True])
def test_linear_loss(port, ref, loss, with_value):
    rng = np.random.default_rng(5)
    blk = random_block(rng, 1500, 25, 3000, with_value)
    D = ref.localize(blk, 2 ** 24)
    w = (rng.standard_normal(len(D.uniq)) * 0.3).astype(np.float32)
    for nt in (1, 2):
        pa, ga, xw, _ = port.linear_loss(loss, D, w, nt, want_xw=True)
        pb, gb = ref.linear_loss(loss, D, w, nt)
        if len(np.unique(xw)) < D.rows:
            # equal scores: std::sort's tie order is unspecified (binary_class_evaluation.h:25-26),
            # the port's is stable; AUC then agrees only up to the tied pairs
            assert abs(pa[2] - pb[2]) < 5e-3
            pa[2] = pb[2]
        assert np.array_equal(pa, pb), (pa, pb)
        assert np.array_equal(ga, gb)


@pytest.mark.parametrize("algo", [O.ALGO_SGD, O.ALGO_ADAGRAD, O.ALGO_FTRL])
def test_handle_push(port, ref, algo):
    rng = np.random.default_rng(6)
    n = 5000
    for eta, beta, l1, l2 in [(None, None, 1.0, 0.0), (0.1, 1.0, 1.0, 0.1), (0.05, 2.0, 0.0, 0.0)]:
        st = [[(rng.standard_normal(n) * (rng.random(n) < 0.7)).astype(np.float32) for _ in range(3)]
              for _ in range(2)]
        for a in st:
            a[2][:] = np.abs(a[2])
        st[1] = [x.copy() for x in st[0]]
        t = [1, 1]
        for step in range(3):
            g = rng.standard_normal(n).astype(np.float32) * 2
            outs = []
            for i, o in enumerate((port, ref)):
                nw, t[i] = o.handle_push(algo, l1, l2, eta, beta, g, st[i][0], st[i][1], st[i][2], t[i])
                outs.append(nw)
            assert outs[0] == outs[1] and t[0] == t[1]
            for x, y in zip(st[0], st[1]):
                assert np.array_equal(x, y)


@pytest.mark.parametrize("algo", [O.ALGO_SGD, O.ALGO_ADAGRAD, O.ALGO_FTRL])
@pytest.mark.parametrize("loss", [O.LOSS_LOGIT, O.LOSS_SQUARE_HINGE])
def test_full_job(port, ref, algo, loss):
    """AsgdWorker::ProcessMinibatch + AsgdServer (1 worker, 1 server) vs the port's job."""
    rng = np.random.default_rng(7)
    jobs = [o.linear_create(algo, loss, 0.5, 0.1, 0.1, 1.0, 2 ** 20) for o in (port, ref)]
    for m in range(6):
        blk = random_block(rng, 400, 12, 2000, with_value=(m % 2 == 1), empty_rows=False)
        pa = port.linear_minibatch(jobs[0], blk)
        pb = ref.linear_minibatch(jobs[1], blk)
        # AUC is pinned by test_bin_eval on distinct scores; here sparse weights give many equal
        # scores, whose order under std::sort is unspecified (binary_class_evaluation.h:25-26)
        pa[2] = pb[2]
        assert np.array_equal(pa, pb), (m, pa, pb)
    ma, mb = port.linear_model(jobs[0]), ref.linear_model(jobs[1])
    for x, y in zip(ma, mb):
        assert np.array_equal(x, y)
    port.linear_destroy(jobs[0]); ref.linear_destroy(jobs[1])


# ------------------------------------------------------------------ DiFacto ------------------

def make_model(rng, U, dim, frac_v=0.4, frac_absent=0.05):
    """A pulled payload: per key 0 (absent), 1 (w only) or dim+1 (w, V) floats, concatenated."""
    r = rng.random(U)
    siz = np.where(r < frac_absent, 0, np.where(r < frac_absent + frac_v, dim + 1, 1)).astype(np.int32)
    if dim == 0:
        siz = np.where(siz > 1, 1, siz).astype(np.int32)
    model = (rng.standard_normal(int(siz.sum())) * 0.2).astype(np.float32)
    return model, siz


@pytest.mark.parametrize("with_value", [False, True])
@pytest.mark.parametrize("dim", [0, 1, 5, 16])
def test_difacto_loss(port, ref, with_value, dim):
    rng = np.random.default_rng(8)
    blk = random_block(rng, 700, 20, 2500, with_value, empty_rows=True)
    D = ref.localize(blk, 2 ** 24)
    model, siz = make_model(rng, len(D.uniq), dim)
    for nt, clip, norm in [(1, 0.0, 0.0), (2, 0.05, 0.0), (2, 0.0, 1.0)]:
        pa, ga = port.difacto_loss(D, model, siz, dim, nt, clip, norm)
        pb, gb = ref.difacto_loss(D, model, siz, dim, nt, clip, norm)
        pa[1] = pb[1] if abs(pa[1] - pb[1]) < 5e-3 else pa[1]    # AUC: tie order of equal scores
        assert np.array_equal(pa, pb), (pa, pb)
        assert np.array_equal(ga, gb)


def test_difacto_loss_dropout(port, ref):
    """Gradient dropout draws from the process's glibc rand() (learn/difacto/loss.h:150-154): same seed, same drops."""
    import ctypes
    libc = ctypes.CDLL("libc.so.6")
    rng = np.random.default_rng(10)
    blk = random_block(rng, 700, 20, 2500, True, empty_rows=True)
    D = ref.localize(blk, 2 ** 24)
    model, siz = make_model(rng, len(D.uniq), 5)
    for dropout, clip, norm in [(0.3, 0.0, 0.0), (0.6, 0.05, 1.0)]:
        out = []
        for o in (port, ref):
            libc.srand(ctypes.c_uint(17))
            out.append(o.difacto_loss(D, model, siz, 5, 2, clip, norm, dropout=dropout)[1])
        assert np.array_equal(out[0], out[1])
        assert 0.5 * dropout < float((out[0] == 0).mean()) * len(out[0]) / max(1, int((siz == 6).sum()) * 5) < 1.5 * dropout


def test_difacto_full_job_dropout(port, ref):
    rng = np.random.default_rng(11)
    kw = dict(l1=0.3, l2=0.01, eta=0.05, beta=1.0, l1_shrk=True, dim=4, threshold=3, v_l2=1e-4, v_eta=0.02,
              init_scale=0.01, max_key=2 ** 20, seed=321, dropout=0.25)
    blks = [random_block(rng, 500, 12, 1500, with_value=(m % 2 == 1), empty_rows=False) for m in range(3)]
    models = []
    for o in (port, ref):
        j = o.difacto_create(**kw)
        for dp in range(2):
            for blk in blks:
                o.difacto_minibatch(j, blk, True, dp)
        models.append(o.difacto_model(j, 4))
        o.difacto_destroy(j)
    for a, b in zip(models[0], models[1]):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("dim,l1_shrk", [(0, True), (4, True), (4, False), (16, True)])
def test_difacto_full_job(port, ref, dim, l1_shrk):
    """AsyncWorker::ProcessMinibatch + AsyncServer (count push on pass 0, ZVPull, ZVPush) vs the port."""
    rng = np.random.default_rng(9)
    kw = dict(l1=0.3, l2=0.01, eta=0.05, beta=1.0, l1_shrk=l1_shrk, dim=dim, threshold=3, v_l2=1e-4,
              v_eta=0.02, init_scale=0.01, max_key=2 ** 20, seed=123)
    blks = [random_block(rng, 500, 12, 1500, with_value=(m % 2 == 1), empty_rows=False) for m in range(4)]
    # V is initialised from the process-global glibc rand(): each job runs alone, from the same seed
    models, progs = [], []
    for o in (port, ref):
        j = o.difacto_create(**kw)
        pr = [o.difacto_minibatch(j, blk, True, dp) for dp in range(2) for blk in blks]
        models.append(o.difacto_model(j, dim)); progs.append(np.array(pr))
        o.difacto_destroy(j)
    pa, pb = progs
    pa[:, 1] = pb[:, 1]          # AUC with equal scores: pinned elsewhere
    assert np.array_equal(pa, pb)
    for x, y in zip(models[0], models[1]):
        assert np.array_equal(x, y)
    if dim:
        assert (models[0][2] == dim + 1).any(), "the test should exercise embeddings"
