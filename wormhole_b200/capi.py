"""ctypes binding of include/wormhole_b200.h plus a thin Python mirror of the reference's
operator interface for this path (same names and argument meaning as the C++ headers under
learn/base and learn/linear), so parity tests read like the reference's call sites.

There is no CPU fallback: loading fails loudly when libwormhole_b200.so is missing, and every
compute call needs a CUDA device (wh_create fails otherwise).
"""
import ctypes as C
import os

import numpy as np

from .rowblock import LocalBlock, RowBlock

HERE = os.path.dirname(os.path.abspath(__file__))
# (WH_LIB: another build of the same library -- kernel variants compiled with different constants, timing experiments)
LIB_PATH = os.environ.get("WH_LIB") or os.path.join(HERE, "libwormhole_b200.so")

WH_HOST, WH_DEVICE = 0, 1
LOSS_SQUARE, LOSS_LOGIT, LOSS_SQUARE_HINGE = 1, 2, 4
ALGO_SGD, ALGO_ADAGRAD, ALGO_FTRL = 1, 2, 3
LOC_AUTO, LOC_DIRECT, LOC_SORT, LOC_WITH_COUNTS = 0, 1, 2, 16
NO_MAX_KEY = 2 ** 64 - 1


class WormholeError(RuntimeError):
    pass


class _RowBlock(C.Structure):
    _fields_ = [("size", C.c_size_t), ("offset", C.c_void_p), ("label", C.c_void_p),
                ("weight", C.c_void_p), ("index", C.c_void_p), ("value", C.c_void_p)]


class _LocalBlock(C.Structure):
    _fields_ = [("size", C.c_size_t), ("offset", C.c_void_p), ("label", C.c_void_p),
                ("index", C.c_void_p), ("value", C.c_void_p)]


class _ServerConfig(C.Structure):
    _fields_ = [("algo", C.c_int), ("lambda_l1", C.c_float), ("lambda_l2", C.c_float),
                ("has_lr_eta", C.c_int), ("lr_eta", C.c_float), ("has_lr_beta", C.c_int),
                ("lr_beta", C.c_float), ("key_begin", C.c_uint64), ("key_end", C.c_uint64),
                ("hash_slots", C.c_uint64)]


_lib = None


def lib():
    """The loaded extension. Raises if it was not built: the product has no other path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WormholeError(
                "wormhole_b200: %s is missing. Build it with `python -m wormhole_b200.build` "
                "(or __graft_entry__.build()); there is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.wh_last_error.restype = C.c_char_p
        L.wh_version.restype = C.c_char_p
        L.wh_kernel_launches.restype = C.c_uint64
        L.wh_reverse_bytes.restype = C.c_uint64
        L.wh_reverse_bytes.argtypes = [C.c_uint64]
        for n in ("wh_batch_rows", "wh_batch_nnz", "wh_batch_num_uniq"):
            getattr(L, n).restype = C.c_size_t
            getattr(L, n).argtypes = [C.c_void_p]
        for n in ("wh_batch_has_value", "wh_batch_has_label", "wh_batch_destroy", "wh_destroy",
                  "wh_sync", "wh_server_destroy", "wh_worker_destroy", "wh_server_sgd_t", "wh_difacto_destroy"):
            getattr(L, n).argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise WormholeError("wormhole_b200 error %d: %s" % (rc, lib().wh_last_error().decode()))


def _ptr(a):
    """Pointer of a numpy array, a torch tensor (host or device) or None."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    return C.c_void_p(a.data_ptr())  # torch.Tensor


def _rowblock(blk):
    r = _RowBlock()
    r.size = blk.rows
    r.offset, r.label, r.weight = _ptr(blk.offset), _ptr(blk.label), None
    r.index, r.value = _ptr(blk.index), _ptr(blk.value)
    return r


def _localblock(D):
    r = _LocalBlock()
    r.size = D.rows
    r.offset, r.label, r.index, r.value = _ptr(D.offset), _ptr(D.label), _ptr(D.index), _ptr(D.value)
    return r


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def reverse_bytes(x):
    """dmlc::ReverseBytes (learn/base/localizer.h:16-26)."""
    return int(lib().wh_reverse_bytes(C.c_uint64(int(x))))


def kernel_launches():
    return int(lib().wh_kernel_launches())


class Context:
    """One GPU (whContext)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib().wh_create(C.c_int(device), C.byref(self.h)))
        self.device = device

    def set_stream(self, cuda_stream):
        _check(lib().wh_set_stream(self.h, C.c_void_p(cuda_stream)))

    def sync(self):
        _check(lib().wh_sync(self.h))

    def close(self):
        if self.h:
            lib().wh_destroy(self.h)
            self.h = None

    # ---- Localizer<uint64_t> (learn/base/localizer.h) --------------------------------------
    def localize(self, blk, max_key=NO_MAX_KEY, method=LOC_AUTO, counts=True, float_cnt=False,
                 mem=WH_HOST):
        """Localizer::Localize(blk, &localized, &uniq_idx, &idx_frq) -> LocalBlock (host arrays)."""
        b = C.c_void_p()
        rb = _rowblock(blk)
        m = method | (LOC_WITH_COUNTS if counts else 0)
        _check(lib().wh_localize(self.h, C.byref(rb), C.c_int(mem), C.c_uint64(max_key), C.c_int(m),
                                 C.byref(b)))
        return self._collect(b, counts, float_cnt)

    def remap(self, blk, dict_keys, max_key=NO_MAX_KEY):
        """CountUniqIndex(blk) then RemapIndex(blk, dict_keys, &localized)."""
        d = np.ascontiguousarray(dict_keys, np.uint64)
        b = C.c_void_p()
        rb = _rowblock(blk)
        _check(lib().wh_localize_with_dict(self.h, C.byref(rb), C.c_int(WH_HOST), C.c_uint64(max_key),
                                           _ptr(d), C.c_size_t(len(d)), C.byref(b)))
        return self._collect(b, True, False)

    def _collect(self, b, counts, float_cnt):
        L = lib()
        try:
            U, n, rows = L.wh_batch_num_uniq(b), L.wh_batch_nnz(b), L.wh_batch_rows(b)
            out = LocalBlock()
            out.uniq = np.zeros(U, np.uint64)
            out.cnt = np.zeros(U, np.float32 if float_cnt else np.uint32) if counts else None
            out.offset = np.zeros(rows + 1, np.uint64)
            out.index = np.zeros(n, np.uint32)
            out.value = np.zeros(n, np.float32) if L.wh_batch_has_value(b) else None
            out.label = np.zeros(rows, np.float32) if L.wh_batch_has_label(b) else None
            _check(L.wh_batch_copy(self.h, b, _ptr(out.uniq),
                                   _ptr(out.cnt) if counts and not float_cnt else None,
                                   _ptr(out.cnt) if counts and float_cnt else None,
                                   _ptr(out.offset), _ptr(out.index), _ptr(out.value), _ptr(out.label)))
            return out
        finally:
            L.wh_batch_destroy(b)

    # ---- SpMV / SpMM (learn/base/spmv.h, spmm.h) ---------------------------------------------
    def spmv_times(self, D, x, y=None):
        x = _f32(x)
        y = np.zeros(D.rows, np.float32) if y is None else y
        lb = _localblock(D)
        _check(lib().wh_spmv_times(self.h, C.byref(lb), C.c_int(WH_HOST), _ptr(x), C.c_size_t(len(x)), _ptr(y)))
        return y

    def spmv_trans_times(self, D, x, ysize):
        x = _f32(x)
        y = np.empty(ysize, np.float32)
        lb = _localblock(D)
        _check(lib().wh_spmv_trans_times(self.h, C.byref(lb), C.c_int(WH_HOST), _ptr(x), _ptr(y), C.c_size_t(ysize)))
        return y

    def spmm_times(self, D, x, dim):
        x = _f32(x)
        y = np.zeros(D.rows * dim, np.float32)
        lb = _localblock(D)
        _check(lib().wh_spmm_times(self.h, C.byref(lb), C.c_int(WH_HOST), _ptr(x), C.c_size_t(len(x)), _ptr(y), C.c_int(dim)))
        return y

    def spmm_trans_times(self, D, x, ysize, dim, p=0.0, z=None):
        x, z = _f32(x), _f32(z)
        y = np.zeros(ysize, np.float32)
        lb = _localblock(D)
        _check(lib().wh_spmm_trans_times(self.h, C.byref(lb), C.c_int(WH_HOST), _ptr(x), C.c_float(p), _ptr(z),
                                         _ptr(y), C.c_size_t(ysize), C.c_int(dim)))
        return y

    # ---- BinClassEval / ScalarLoss (learn/base/binary_class_evaluation.h, learn/linear/loss.h) --
    def bin_eval(self, label, predict):
        label, predict = _f32(label), _f32(predict)
        out = np.zeros(6, np.float32)
        _check(lib().wh_bin_eval(self.h, _ptr(label), _ptr(predict), C.c_size_t(len(label)), C.c_int(WH_HOST), _ptr(out)))
        return dict(auc=out[0], acc=out[1], logloss=out[2], logit_objv=out[3], auc_lo=out[4], auc_hi=out[5])

    def linear_loss(self, loss, D, w, want_grad=True, want_xw=False):
        """CreateLoss(loss)->Init(D, w); Evaluate(&prog); CalcGrad(&grad)."""
        w = _f32(w)
        prog = np.zeros(6, np.float64)
        grad = np.empty(len(w), np.float32) if want_grad else None
        xw = np.empty(D.rows, np.float32) if want_xw else None
        dual = np.empty(D.rows, np.float32) if want_xw else None
        lb = _localblock(D)
        _check(lib().wh_linear_loss(self.h, C.c_int(loss), C.byref(lb), C.c_int(WH_HOST), _ptr(w),
                                    C.c_size_t(len(w)), _ptr(prog), _ptr(grad), _ptr(xw), _ptr(dual)))
        return (prog, grad, xw, dual) if want_xw else (prog, grad)


class Server:
    """One model shard with its update handle (AsgdServer::CreateServer + SGD/AdaGrad/FTRL handle,
    learn/linear/async_sgd.h:83-213). eta / beta None = key absent from the conf."""

    def __init__(self, ctx, algo=ALGO_FTRL, l1=1.0, l2=0.0, eta=None, beta=None, key_begin=0,
                 key_end=2 ** 24, hash_slots=0):
        self.ctx = ctx
        cfg = _ServerConfig(algo, l1, l2, int(eta is not None), eta or 0.0, int(beta is not None),
                            beta or 0.0, key_begin, key_end, hash_slots)
        self.h = C.c_void_p()
        _check(lib().wh_server_create(ctx.h, C.byref(cfg), C.byref(self.h)))
        self.algo = algo

    def pull(self, keys):
        keys = np.ascontiguousarray(keys, np.uint64)
        vals = np.empty(len(keys), np.float32)
        _check(lib().wh_server_pull(self.h, _ptr(keys), C.c_size_t(len(keys)), C.c_int(WH_HOST), _ptr(vals)))
        return vals

    def push(self, keys, grads):
        keys, grads = np.ascontiguousarray(keys, np.uint64), _f32(grads)
        _check(lib().wh_server_push(self.h, _ptr(keys), C.c_size_t(len(keys)), C.c_int(WH_HOST), _ptr(grads)))

    def pull_dev(self, keys_ptr, n, vals_ptr):
        """One pull message on device arrays (addresses); enqueued on the context's stream."""
        _check(lib().wh_server_pull(self.h, C.c_void_p(keys_ptr), C.c_size_t(n), C.c_int(WH_DEVICE), C.c_void_p(vals_ptr)))

    def push_dev(self, keys_ptr, n, grads_ptr):
        """One push message on device arrays (addresses); enqueued on the context's stream."""
        _check(lib().wh_server_push(self.h, C.c_void_p(keys_ptr), C.c_size_t(n), C.c_int(WH_DEVICE), C.c_void_p(grads_ptr)))

    def save(self, path):
        """AsgdServer::SaveModel: {key, w} records of the non-zero weights."""
        _check(lib().wh_server_save(self.h, path.encode()))

    def load(self, path):
        _check(lib().wh_server_load(self.h, path.encode()))

    def set_active_sources(self, mask):
        """Bit r: rank r sends a push message in the next step (SGD's per-message t advances for those only)."""
        _check(lib().wh_server_set_active_sources(self.h, C.c_uint64(mask)))

    def ipc_export(self):
        """The bytes other ranks need to map this shard (wh_server_ipc_export)."""
        buf = C.create_string_buffer(192)
        _check(lib().wh_server_ipc_export(self.h, buf))
        return buf.raw

    def ipc_attach(self, rank, world, max_key, handles):
        """handles: the ipc_export() bytes of ranks 0..world-1, in rank order."""
        blob = b"".join(handles)
        assert len(blob) == 192 * world
        _check(lib().wh_server_ipc_attach(self.h, C.c_int(rank), C.c_int(world), C.c_uint64(max_key), blob))

    def nnz_w(self):
        out = C.c_int64()
        _check(lib().wh_server_nnz_w(self.h, C.byref(out)))
        return out.value

    def sgd_t(self):
        return lib().wh_server_sgd_t(self.h)

    def get_state(self, keys):
        keys = np.ascontiguousarray(keys, np.uint64)
        w, z, cg = (np.zeros(len(keys), np.float32) for _ in range(3))
        _check(lib().wh_server_get_state(self.h, _ptr(keys), C.c_size_t(len(keys)), C.c_int(WH_HOST),
                                         _ptr(w), _ptr(z), _ptr(cg)))
        return w, z, cg

    def set_state(self, keys, w=None, z=None, cg=None):
        keys = np.ascontiguousarray(keys, np.uint64)
        _check(lib().wh_server_set_state(self.h, _ptr(keys), C.c_size_t(len(keys)), C.c_int(WH_HOST),
                                         _ptr(_f32(w)), _ptr(_f32(z)), _ptr(_f32(cg))))

    def close(self):
        if self.h:
            lib().wh_server_destroy(self.h)
            self.h = None


class _DifactoConfig(C.Structure):
    _fields_ = [("lambda_l1", C.c_float), ("lambda_l2", C.c_float), ("lr_eta", C.c_float), ("lr_beta", C.c_float),
                ("l1_shrk", C.c_int), ("dim", C.c_int), ("threshold", C.c_int), ("v_lambda_l2", C.c_float),
                ("v_has_lr_eta", C.c_int), ("v_lr_eta", C.c_float), ("v_has_lr_beta", C.c_int), ("v_lr_beta", C.c_float),
                ("init_scale", C.c_float), ("grad_clipping", C.c_float), ("grad_normalization", C.c_float),
                ("dropout", C.c_float), ("key_begin", C.c_uint64), ("key_end", C.c_uint64), ("max_key", C.c_uint64),
                ("seed", C.c_uint64)]


def difacto_loss(ctx, D, model, model_siz, dim, clip=0.0, norm=0.0, want_grad=True, dropout=0.0, rand_values=None):
    """difacto::Loss(data, model, model_siz, conf); Evaluate; CalcGrad -> (prog[8], grad or None). dropout > 0 takes the
    caller's next glibc rand() values (one per embedding-gradient float)."""
    model = _f32(model)
    siz = np.ascontiguousarray(model_siz, np.int32)
    prog = np.zeros(8, np.float64)
    grad = np.zeros(len(model), np.float32) if want_grad else None
    lb = _localblock(D)
    if dropout > 0:
        rv = np.ascontiguousarray(rand_values if rand_values is not None else [], np.int32)
        _check(lib().wh_difacto_loss_dropout(ctx.h, C.byref(lb), _ptr(model), C.c_size_t(len(model)), _ptr(siz), C.c_size_t(len(siz)),
                                             C.c_int(dim), C.c_float(clip), C.c_float(norm), C.c_float(dropout), _ptr(rv),
                                             C.c_size_t(len(rv)), _ptr(prog), _ptr(grad)))
        return prog, grad
    _check(lib().wh_difacto_loss(ctx.h, C.byref(lb), _ptr(model), C.c_size_t(len(model)), _ptr(siz), C.c_size_t(len(siz)),
                                 C.c_int(dim), C.c_float(clip), C.c_float(norm), _ptr(prog), _ptr(grad)))
    return prog, grad


class Difacto:
    """AsyncServer + AsyncWorker of learn/difacto/async_sgd.h on one GPU."""

    def __init__(self, ctx, l1=1.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=0, threshold=0, v_l2=0.0, v_eta=None,
                 v_beta=None, init_scale=.01, clip=0.0, norm=0.0, max_key=2 ** 24, key_begin=0, key_end=None, seed=0,
                 dropout=0.0):
        self.ctx, self.dim = ctx, dim
        cfg = _DifactoConfig(l1, l2, eta, beta, int(l1_shrk), dim, threshold, v_l2, int(v_eta is not None), v_eta or 0.0,
                             int(v_beta is not None), v_beta or 0.0, init_scale, clip, norm, dropout, key_begin,
                             key_end if key_end is not None else max_key, max_key, seed)
        self.h = C.c_void_p()
        _check(lib().wh_difacto_create(ctx.h, C.byref(cfg), C.byref(self.h)))

    def set_init_stream(self, rand_values):
        v = np.ascontiguousarray(rand_values, np.int32)
        _check(lib().wh_difacto_set_init_stream(self.h, _ptr(v), C.c_size_t(len(v))))

    def set_dropout_stream(self, rand_values=None, shared_with_init=False):
        """The worker's next rand() values for gradient dropout; shared_with_init: draw from the init stream instead
        (worker and server in one process)."""
        v = np.ascontiguousarray(rand_values if rand_values is not None else [], np.int32)
        _check(lib().wh_difacto_set_dropout_stream(self.h, _ptr(v), C.c_size_t(len(v)), C.c_int(shared_with_init)))

    def process_minibatch(self, blk, train=True, data_pass=0, mem=WH_HOST):
        prog = np.zeros(8, np.float64)
        rb = _rowblock(blk)
        _check(lib().wh_difacto_process_minibatch(self.h, C.byref(rb), C.c_int(mem), C.c_int(train), C.c_int(data_pass), _ptr(prog)))
        return prog

    # ---- the minibatch in phases (key-range shards on several GPUs; device addresses) ----
    def begin(self, blk, mem=WH_HOST):
        pu, pc, n = C.c_void_p(), C.c_void_p(), C.c_size_t()
        rb = _rowblock(blk)
        self._begun = blk
        _check(lib().wh_difacto_begin(self.h, C.byref(rb), C.c_int(mem), C.byref(pu), C.byref(pc), C.byref(n)))
        return (pu.value or 0), (pc.value or 0), n.value

    def finish(self, w_ptr, has_ptr, v_ptr, train, gw_ptr, g_ptr):
        prog = np.zeros(8, np.float64)
        _check(lib().wh_difacto_finish(self.h, C.c_void_p(w_ptr), C.c_void_p(has_ptr), C.c_void_p(v_ptr), C.c_int(train),
                                       C.c_void_p(gw_ptr), C.c_void_p(g_ptr), _ptr(prog)))
        return prog

    def shard_count_push(self, keys_ptr, cnt_ptr, n):
        _check(lib().wh_difacto_shard_count_push(self.h, C.c_void_p(keys_ptr), C.c_void_p(cnt_ptr), C.c_size_t(n)))

    def shard_pull(self, keys_ptr, n, w_ptr, has_ptr, v_ptr):
        _check(lib().wh_difacto_shard_pull(self.h, C.c_void_p(keys_ptr), C.c_size_t(n), C.c_void_p(w_ptr), C.c_void_p(has_ptr), C.c_void_p(v_ptr)))

    def shard_push(self, keys_ptr, n, gw_ptr, g_ptr, has_ptr):
        _check(lib().wh_difacto_shard_push(self.h, C.c_void_p(keys_ptr), C.c_size_t(n), C.c_void_p(gw_ptr), C.c_void_p(g_ptr), C.c_void_p(has_ptr)))

    def shard_progress(self):
        a, b = C.c_double(), C.c_double()
        _check(lib().wh_difacto_shard_progress(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def get_state(self, keys):
        keys = np.ascontiguousarray(keys, np.uint64)
        n, dim = len(keys), self.dim
        cnt = np.zeros(n, np.uint32); size = np.zeros(n, np.int32)
        w = np.zeros((n, dim + 1), np.float32); cg = np.zeros((n, dim + 2), np.float32)
        _check(lib().wh_difacto_get_state(self.h, _ptr(keys), C.c_size_t(n), _ptr(cnt), _ptr(size), _ptr(w), _ptr(cg)))
        return cnt, size, w, cg

    def close(self):
        if self.h:
            lib().wh_difacto_destroy(self.h)
            self.h = None


class Worker:
    """AsgdWorker (learn/linear/async_sgd.h:229-305) bound to one server on the same GPU."""

    def __init__(self, ctx, server, loss=LOSS_LOGIT, max_key=NO_MAX_KEY, depth=1):
        self.ctx, self.server, self.depth = ctx, server, depth
        self.h = C.c_void_p()
        _check(lib().wh_worker_create_ex(ctx.h, server.h if server is not None else None, C.c_int(loss),
                                         C.c_uint64(max_key), C.c_int(depth), C.byref(self.h)))
        self._keep = {}

    def set_max_delay(self, d):
        _check(lib().wh_worker_set_max_delay(self.h, C.c_int(d)))

    def table_keys(self):
        """-> (keys held by the worker's table of frequent keys, ms the last census took, censuses taken)."""
        keys = np.zeros(16384, np.uint32)
        n, ms, cnt = C.c_size_t(), C.c_double(), C.c_int()
        _check(lib().wh_worker_table_keys(self.h, _ptr(keys), C.c_size_t(len(keys)), C.byref(n), C.byref(ms), C.byref(cnt)))
        return keys[:min(n.value, len(keys))], ms.value, cnt.value

    def submit(self, blk, train=True, mem=WH_HOST, nnz=0):
        """Enqueue one minibatch (up to `depth` in flight); returns a ticket for wait()."""
        t = C.c_uint64()
        rb = _rowblock(blk)
        _check(lib().wh_worker_submit(self.h, C.byref(rb), C.c_int(mem), C.c_size_t(nnz), C.c_int(train), C.byref(t)))
        self._keep[t.value] = blk   # the arrays must outlive the call
        return t.value

    def submit_idle(self, train=True):
        """A step of a sharded job in which this rank has no minibatch (ragged streams): joins the collective barriers,
        serves the others' pulls, applies their pushes. Returns the ticket; its progress is all zero."""
        t = C.c_uint64()
        _check(lib().wh_worker_submit_idle(self.h, C.c_int(train), C.byref(t)))
        return t.value

    def predict(self, blk, prob_predict=True, mem=WH_HOST):
        """The PRED workload: one prediction per row (sigmoid of Xw if prob_predict)."""
        out = np.zeros(blk.rows, np.float32)
        rb = _rowblock(blk)
        _check(lib().wh_worker_predict(self.h, C.byref(rb), C.c_int(mem), C.c_int(prob_predict), _ptr(out)))
        return out

    def ipc_export(self, max_uniq):
        buf = C.create_string_buffer(64 * self.depth)
        _check(lib().wh_worker_ipc_export(self.h, C.c_size_t(max_uniq), buf))
        return buf.raw

    def ipc_attach(self, handles):
        _check(lib().wh_worker_ipc_attach(self.h, b"".join(handles)))

    def begin(self, blk, mem=WH_HOST, nnz=0):
        """First half of ProcessMinibatch (Localize). -> (device address of the ascending unique keys, U)."""
        p, n = C.c_void_p(), C.c_size_t()
        rb = _rowblock(blk)
        self._begun = blk
        _check(lib().wh_worker_begin(self.h, C.byref(rb), C.c_int(mem), C.c_size_t(nnz), C.byref(p), C.byref(n)))
        return (p.value or 0), n.value

    def finish(self, weights_ptr, grads_ptr, train=True):
        """The pull callback: Init / Evaluate / CalcGrad with the pulled weights (device addresses)."""
        prog = np.zeros(6, np.float64)
        _check(lib().wh_worker_finish(self.h, C.c_void_p(weights_ptr), C.c_void_p(grads_ptr), C.c_int(train), _ptr(prog)))
        return prog

    def wait(self, ticket):
        prog = np.zeros(6, np.float64)
        _check(lib().wh_worker_wait(self.h, C.c_uint64(ticket), _ptr(prog)))
        self._keep.pop(ticket, None)
        return prog

    def process_minibatch(self, blk, train=True, mem=WH_HOST):
        """ProcessMinibatch(mb, wl): -> progress {objv, acc, auc, count, new_ex, new_w}."""
        prog = np.zeros(6, np.float64)
        rb = _rowblock(blk)
        _check(lib().wh_worker_process_minibatch(self.h, C.byref(rb), C.c_int(mem), C.c_int(train), _ptr(prog)))
        return prog

    def close(self):
        if self.h:
            lib().wh_worker_destroy(self.h)
            self.h = None


FMT = {"libsvm": 0, "criteo": 1, "criteo_test": 2, "adfea": 3}


class Rows:
    """A device-resident RowBlockContainer<uint64_t> (whRows): the product of wh_parse_text and the
    container MinibatchIter assembles minibatches in. Duck-types the block argument of Worker.submit
    (rows / offset / label / index / value are DEVICE addresses; pass mem=WH_DEVICE, nnz=self.nnz)."""

    class _DevPtr:
        def __init__(self, p):
            self.p = p

        def data_ptr(self):
            return self.p

    def __init__(self, ctx, handle=None):
        self.ctx = ctx
        if handle is None:
            handle = C.c_void_p()
            _check(lib().wh_rows_create(ctx.h, C.byref(handle)))
        self.h = handle

    @property
    def rows(self):
        L = lib()
        L.wh_rows_size.restype = C.c_size_t
        return int(L.wh_rows_size(self.h))

    @property
    def nnz(self):
        L = lib()
        L.wh_rows_nnz.restype = C.c_size_t
        return int(L.wh_rows_nnz(self.h))

    @property
    def has_value(self):
        return bool(lib().wh_rows_has_value(self.h))

    def _view(self):
        v = _RowBlock()
        _check(lib().wh_rows_view(self.h, C.byref(v)))
        return v

    offset = property(lambda self: Rows._DevPtr(self._view().offset))
    label = property(lambda self: Rows._DevPtr(self._view().label))
    index = property(lambda self: Rows._DevPtr(self._view().index))
    value = property(lambda self: Rows._DevPtr(self._view().value) if self.has_value else None)

    def clear(self):
        _check(lib().wh_rows_clear(self.h))

    def append(self, src, begin=0, n=None, row_ids=None):
        if row_ids is not None:
            ids = np.ascontiguousarray(row_ids, dtype=np.uint32)
            _check(lib().wh_rows_append(self.ctx.h, self.h, src.h, C.c_size_t(0), C.c_size_t(len(ids)), _ptr(ids)))
        else:
            n = src.rows - begin if n is None else n
            _check(lib().wh_rows_append(self.ctx.h, self.h, src.h, C.c_size_t(begin), C.c_size_t(n), None))

    def drop_unit_values(self):
        _check(lib().wh_rows_drop_unit_values(self.ctx.h, self.h))

    def labels(self):
        lab = np.empty(self.rows, np.float32)
        _check(lib().wh_rows_copy(self.ctx.h, self.h, None, _ptr(lab), None, None))
        return lab

    def to_host(self):
        """-> RowBlock on the host (tests, debugging)."""
        rows, nnz = self.rows, self.nnz
        off, lab, idx = np.empty(rows + 1, np.uint64), np.empty(rows, np.float32), np.empty(nnz, np.uint64)
        val = np.empty(nnz, np.float32) if self.has_value else None
        _check(lib().wh_rows_copy(self.ctx.h, self.h, _ptr(off), _ptr(lab), _ptr(idx), _ptr(val)))
        return RowBlock(off, idx, val, lab)

    def close(self):
        if self.h:
            try:
                lib().wh_rows_destroy(self.h)
            except TypeError:        # interpreter shutdown: the module globals are already gone
                pass
            self.h = None

    __del__ = close


def parse_crb(ctx, data, into=None):
    """wh_parse_crb: one chunk of whole RecordIO records of CompressedRowBlocks (host bytes) -> Rows."""
    h = C.c_void_p() if into is None else into.h
    buf = (C.c_char * len(data)).from_buffer_copy(data)
    _check(lib().wh_parse_crb(ctx.h, C.cast(buf, C.c_void_p), C.c_size_t(len(data)), C.byref(h)))
    return Rows(ctx, h) if into is None else into


def parse_text(ctx, fmt, text, mem=WH_HOST, length=None, into=None):
    """wh_parse_text: one chunk of whole lines -> Rows. text: bytes / numpy uint8 (host) or a device address with
    `length`. into: a Rows to refill (its buffers are reused) instead of creating one."""
    h = C.c_void_p() if into is None else into.h
    if mem == WH_HOST:
        buf = (C.c_char * len(text)).from_buffer_copy(text) if not isinstance(text, np.ndarray) else None
        p = C.cast(buf, C.c_void_p) if buf is not None else _ptr(text)
        n = len(text)
    else:
        p, n = C.c_void_p(int(text)), int(length)
    _check(lib().wh_parse_text(ctx.h, FMT[fmt], p, C.c_size_t(n), mem, C.byref(h)))
    return Rows(ctx, h) if into is None else into
