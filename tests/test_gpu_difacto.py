"""DiFacto on the GPU against the CPU checker: the factorization-machine loss (learn/difacto/loss.h) on a
ZVPull-shaped payload, and AsyncWorker::ProcessMinibatch + AsyncServer (learn/difacto/async_sgd.h) end to
end, including the count push, lazily created embeddings and their rand()-ordered initial values."""
import ctypes

import numpy as np
import pytest

from tests.test_oracle_vs_ref import make_model
from tests.util import random_block
from wormhole_b200 import capi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    c = capi.Context(0)
    yield c
    c.close()


def glibc_rand_stream(seed, n):
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(ctypes.c_uint(seed))
    return np.array([libc.rand() for _ in range(n)], np.int32)


@pytest.mark.parametrize("with_value", [False, True])
@pytest.mark.parametrize("dim", [0, 1, 5, 16])
def test_difacto_loss(gpu, checker, with_value, dim):
    rng = np.random.default_rng(61)
    blk = random_block(rng, 4000, 20, 6000, with_value, empty_rows=True)
    D = checker.localize(blk, 2 ** 24)
    model, siz = make_model(rng, len(D.uniq), dim)
    for clip, norm in [(0.0, 0.0), (0.05, 0.0), (0.0, 1.0)]:
        pg, gg = capi.difacto_loss(gpu, D, model, siz, dim, clip, norm)
        pc, gc = checker.difacto_loss(D, model, siz, dim, 2, clip, norm)
        assert pg[4] == pc[4] and pg[5] == pc[5]
        assert abs(pg[0] - pc[0]) <= 1e-4 * abs(pc[0]) and abs(pg[2] - pc[2]) <= 1e-4 * abs(pc[2])   # objv, objv_w
        assert abs(pg[1] - pc[1]) <= 1e-3                                                          # auc (ties)
        assert np.linalg.norm(gg.astype(np.float64) - gc) <= 2e-5 * np.linalg.norm(gc.astype(np.float64)), (clip, norm)


@pytest.mark.parametrize("dim,l1_shrk", [(0, True), (4, True), (4, False), (16, True)])
def test_difacto_process_minibatch(gpu, checker, dim, l1_shrk):
    rng = np.random.default_rng(62)
    max_key, seed = 2 ** 20, 321
    kw = dict(l1=0.3, l2=0.01, eta=0.05, beta=1.0, l1_shrk=l1_shrk, dim=dim, threshold=3, v_l2=1e-4, v_eta=0.02,
              init_scale=0.01)
    blks = [random_block(rng, 2000, 12, 4000, with_value=(m % 2 == 1), empty_rows=False) for m in range(4)]
    job = checker.difacto_create(max_key=max_key, seed=seed, **kw)
    pcs = np.array([checker.difacto_minibatch(job, b, True, dp) for dp in range(2) for b in blks])
    keys, ccnt, csize, cw, ccg = checker.difacto_model(job, dim)
    checker.difacto_destroy(job)

    d = capi.Difacto(gpu, max_key=max_key, **kw)
    d.set_init_stream(glibc_rand_stream(seed, 1 << 20))     # the reference's V init, in arrival order
    pgs = np.array([d.process_minibatch(b, True, dp) for dp in range(2) for b in blks])
    gcnt, gsize, gw, gcg = d.get_state(keys)
    d.close()
    assert np.array_equal(gcnt, ccnt) and np.array_equal(gsize, csize)
    if dim:
        assert (csize == dim + 1).any()
    assert np.array_equal(pgs[:, 4:], pcs[:, 4:])                      # count, new_ex, new_w, new_V
    assert np.all(np.abs(pgs[:, 0] - pcs[:, 0]) <= 1e-4 * np.abs(pcs[:, 0]))
    assert np.all(np.abs(pgs[:, 2] - pcs[:, 2]) <= 1e-4 * np.abs(pcs[:, 2]))
    for a, b in ((gw, cw), (gcg, ccg)):
        err = np.abs(a - b)
        assert np.all(err <= 2e-5 * np.maximum(np.abs(b), 1.0)), float(err.max())


def test_difacto_loss_dropout(gpu, checker):
    """learn/difacto/loss.h:150-154: `(T)rand() / RAND_MAX > 1 - dropout` zeroes a gradient float, one rand() per float of
    the keys that carry an embedding, after clipping and before normalisation. The checker draws from glibc rand() after
    srand(seed); the GPU takes the same values as an explicit stream."""
    rng = np.random.default_rng(63)
    blk = random_block(rng, 3000, 20, 5000, True, empty_rows=True)
    D = checker.localize(blk, 2 ** 24)
    dim = 6
    model, siz = make_model(rng, len(D.uniq), dim)
    libc = ctypes.CDLL("libc.so.6")
    for dropout, clip, norm in [(0.3, 0.0, 0.0), (0.5, 0.05, 0.0), (0.2, 0.0, 1.0), (1.0, 0.0, 0.0)]:
        libc.srand(ctypes.c_uint(99))
        pc, gc = checker.difacto_loss(D, model, siz, dim, 2, clip, norm, dropout=dropout)
        stream = glibc_rand_stream(99, int((np.asarray(siz) == dim + 1).sum()) * dim)
        pg, gg = capi.difacto_loss(gpu, D, model, siz, dim, clip, norm, dropout=dropout, rand_values=stream)
        assert np.array_equal(gg == 0, gc == 0), (dropout, clip, norm)          # the same floats are dropped
        nz = int((gc != 0).sum())
        assert np.linalg.norm(gg.astype(np.float64) - gc) <= 2e-5 * np.linalg.norm(gc.astype(np.float64)), (dropout, clip, norm)
        emb = np.concatenate([np.arange(1, s) + o for o, s in zip(np.cumsum(siz) - siz, siz) if s > 1])
        dropped = float((gc[emb] == 0).mean())
        assert abs(dropped - dropout) < 0.03, (dropout, dropped, nz)
    with pytest.raises(capi.WormholeError, match="rand"):
        capi.difacto_loss(gpu, D, model, siz, dim, dropout=0.3, rand_values=stream[:10])


def test_difacto_minibatches_with_dropout(gpu, checker):
    """End to end with embedding dropout. The checker runs the reference's worker and server in one process, so V
    initialisation and dropout draw from ONE rand() sequence: the GPU job takes that sequence as its init stream and lets
    dropout share it."""
    rng = np.random.default_rng(64)
    max_key, seed, dim = 2 ** 20, 4321, 4
    kw = dict(l1=0.3, l2=0.01, eta=0.05, beta=1.0, l1_shrk=True, dim=dim, threshold=3, v_l2=1e-4, v_eta=0.02,
              init_scale=0.01, dropout=0.25)
    blks = [random_block(rng, 2000, 12, 4000, with_value=(m % 2 == 1), empty_rows=False) for m in range(4)]
    job = checker.difacto_create(max_key=max_key, seed=seed, **kw)
    pcs = np.array([checker.difacto_minibatch(job, b, True, dp) for dp in range(2) for b in blks])
    keys, ccnt, csize, cw, ccg = checker.difacto_model(job, dim)
    checker.difacto_destroy(job)

    d = capi.Difacto(gpu, max_key=max_key, **kw)
    with pytest.raises(capi.WormholeError, match="wh_difacto_set_dropout_stream"):
        d.process_minibatch(blks[0], True, 0)
    d.close()
    d = capi.Difacto(gpu, max_key=max_key, **kw)
    d.set_init_stream(glibc_rand_stream(seed, 1 << 21))
    d.set_dropout_stream(shared_with_init=True)
    pgs = np.array([d.process_minibatch(b, True, dp) for dp in range(2) for b in blks])
    gcnt, gsize, gw, gcg = d.get_state(keys)
    d.close()
    assert np.array_equal(gcnt, ccnt) and np.array_equal(gsize, csize) and (csize == dim + 1).any()
    assert np.array_equal(pgs[:, 4:], pcs[:, 4:])
    assert np.all(np.abs(pgs[:, 0] - pcs[:, 0]) <= 1e-4 * np.abs(pcs[:, 0]))
    for a, b in ((gw, cw), (gcg, ccg)):
        err = np.abs(a - b)
        assert np.all(err <= 2e-5 * np.maximum(np.abs(b), 1.0)), float(err.max())

    # a worker with a rand() sequence of its own (separate processes, as ps-lite runs them): the phases of a sharded job
    # and the one-call path consume it identically
    import torch
    from wormhole_b200.dist import ShardedDifacto
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)
    own = np.random.default_rng(5).integers(0, 2 ** 31 - 1, 1 << 20).astype(np.int32)
    out = []
    for phased in (False, True):
        d = capi.Difacto(gpu, max_key=max_key, seed=3, **kw)
        d.set_dropout_stream(own)
        run = ShardedDifacto(gpu, d, max_key, device=torch.device("cuda", 0)).process_minibatch if phased else d.process_minibatch
        progs = np.array([run(b, True, dp) for dp in range(2) for b in blks])
        torch.cuda.synchronize()
        out.append((progs, d.get_state(keys)))
        d.close()
    assert np.array_equal(out[0][0], out[1][0])
    for a, b in zip(out[0][1], out[1][1]):
        assert np.array_equal(a, b)
