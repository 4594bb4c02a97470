"""The allocation-free transposed product (wh_spmv_trans_times_static) against the regular entry
point (wh_spmv_trans_times) and against the CPU checker, through the C-ABI.

Why the variant exists: some callers cannot allocate device memory -- a worker that shares its GPU
with other minibatch streams (an allocation synchronises the whole device and would stall them),
or a process whose device memory is already fully claimed (an allocation would fail). The variant
keeps its scratch in a fixed array inside the library instead, shared by all calls on a GPU, so
this test makes two calls in a row with different data and sizes: the second must not see anything
the first left behind. It runs last (file name) so that a failure here cannot hide other tests,
and it must pass both alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

from tests.util import random_block
from wormhole_b200 import capi

pytestmark = pytest.mark.gpu

MAX_OUTPUTS = 262144   # WH_STATIC_SCRATCH_MAX_OUTPUTS in include/wormhole_b200.h


def _abs_terms(D, x, U):
    rows = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
    t = np.abs(x[rows].astype(np.float64) * (1.0 if D.value is None else D.value.astype(np.float64)))
    out = np.zeros(U)
    np.add.at(out, D.index, t)
    return out


def test_static_scratch_variant_matches_allocating_entry_point(checker):
    import torch
    L = capi.lib()
    ctx = capi.Context(0)
    rng = np.random.default_rng(99)
    stream = torch.cuda.current_stream().cuda_stream
    # call 1: values present, larger; call 2: binary features, smaller output (stale scratch beyond
    # and inside its range must not leak into the result)
    for rows, mx, space, with_value in [(20000, 30, 50000, True), (3000, 12, 900, False)]:
        blk = random_block(rng, rows, mx, space, with_value)
        D = ctx.localize(blk, 2 ** 24)
        U = len(D.uniq)
        assert U <= MAX_OUTPUTS
        x = rng.standard_normal(D.rows).astype(np.float32)
        want = ctx.spmv_trans_times(D, x, U)              # the existing, allocating entry point

        d_off = torch.from_numpy(D.offset.view(np.int64)).cuda()
        d_idx = torch.from_numpy(D.index.view(np.int32)).cuda()
        d_val = torch.from_numpy(D.value).cuda() if D.value is not None else None
        d_x = torch.from_numpy(x).cuda()
        d_y = torch.full((U,), float("nan"), dtype=torch.float32, device="cuda")
        lb = capi._LocalBlock()
        lb.size = D.rows
        lb.offset, lb.label = C.c_void_p(d_off.data_ptr()), None
        lb.index = C.c_void_p(d_idx.data_ptr())
        lb.value = C.c_void_p(d_val.data_ptr()) if d_val is not None else None
        torch.cuda.synchronize()
        rc = L.wh_spmv_trans_times_static(C.c_void_p(stream), C.byref(lb), C.c_size_t(len(D.index)),
                                          C.c_void_p(d_x.data_ptr()), C.c_void_p(d_y.data_ptr()), C.c_size_t(U))
        assert rc == 0, L.wh_last_error().decode()
        torch.cuda.synchronize()
        got = d_y.cpu().numpy()
        # same tolerance as tests/test_gpu_parity.py uses for transposed products
        err = np.abs(got.astype(np.float64) - want.astype(np.float64))
        assert np.all(err <= 1e-5 * _abs_terms(D, x, U) + 1e-30), float(err.max())
        assert not np.isnan(got).any()
        # the two entry points share their kernels, so also hold the result against the CPU checker's
        # SpMV::TransTimes (learn/base/spmv.h:95-119), with the bound the parity tests use for that product
        ref = checker.spmv_trans_times(D, x, U).astype(np.float64)
        assert np.all(np.abs(got.astype(np.float64) - ref) <= 1e-5 * _abs_terms(D, x, U) + 1e-30)

    # sizes beyond the fixed scratch are refused with a message that names the limit
    rc = L.wh_spmv_trans_times_static(C.c_void_p(stream), C.byref(lb), C.c_size_t(len(D.index)),
                                      C.c_void_p(d_x.data_ptr()), C.c_void_p(d_y.data_ptr()),
                                      C.c_size_t(MAX_OUTPUTS + 1))
    assert rc != 0
    msg = L.wh_last_error().decode()
    assert str(MAX_OUTPUTS) in msg and "limit" in msg
    ctx.close()
