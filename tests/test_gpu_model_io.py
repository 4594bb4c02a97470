"""The train -> save -> load -> predict loop: model files hold {key, w} for the
non-zero weights (Entry::Save / Empty, learn/linear/async_sgd.h:59-75); loading restores w and |w|_0 but not
the optimizer state (TLoad, :59-62); the PRED workload writes sigmoid(Xw) or Xw per row (loss.h:55-63)."""
import numpy as np
import pytest

from tests.util import random_block
from wormhole_b200 import capi

pytestmark = pytest.mark.gpu
MAX_KEY = 2 ** 20


def test_save_load_predict(tmp_path, checker):
    from oracle import oracle as O
    ctx = capi.Context(0)
    rng = np.random.default_rng(81)
    blks = [random_block(rng, 3000, 12, 20000, with_value=(m % 2 == 0), empty_rows=False) for m in range(5)]
    srv = capi.Server(ctx, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, MAX_KEY)
    wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, MAX_KEY)
    for b in blks[:4]:
        wk.process_minibatch(b)
    path = str(tmp_path / "model_part-0")        # IterServer::ModelName: <model_out>_part-<rank>
    srv.save(path)
    rec = np.fromfile(path, dtype=np.dtype([("k", "<u8"), ("w", "<f4")]))
    assert len(rec) == srv.nnz_w() and np.all(rec["w"] != 0) and np.all(np.diff(rec["k"].astype(np.int64)) > 0)
    w_all, _, _ = srv.get_state(rec["k"])
    assert np.array_equal(w_all, rec["w"])

    # a fresh server loads it: same weights, same |w|_0, zeroed optimizer state
    srv2 = capi.Server(ctx, capi.ALGO_FTRL, 0.5, 0.1, 0.1, 1.0, 0, MAX_KEY)
    srv2.load(path)
    assert srv2.nnz_w() == len(rec)
    w2, z2, cg2 = srv2.get_state(rec["k"])
    assert np.array_equal(w2, rec["w"]) and not z2.any() and not cg2.any()

    # predictions of the loaded model == the checker's Xw / sigmoid on the held-out block
    wk2 = capi.Worker(ctx, srv2, capi.LOSS_LOGIT, MAX_KEY)
    D = checker.localize(blks[4], MAX_KEY)
    dense = np.zeros(MAX_KEY, np.float32); dense[rec["k"].astype(np.int64)] = rec["w"]
    xw = checker.spmv_times(D, dense[D.uniq.astype(np.int64)])
    assert np.array_equal(wk2.predict(blks[4], prob_predict=False), xw)
    want = (1.0 / (1.0 + np.exp(-xw, dtype=np.float32).astype(np.float64))).astype(np.float32)
    got = wk2.predict(blks[4], prob_predict=True)
    assert np.max(np.abs(got - want)) <= 2e-7
    # an evaluation pass (Workload::VAL) leaves the model alone
    before = srv2.get_state(rec["k"])
    wk2.process_minibatch(blks[4], train=False)
    for a, b in zip(before, srv2.get_state(rec["k"])):
        assert np.array_equal(a, b)
    for o in (wk, wk2, srv, srv2):
        o.close()
    ctx.close()
