"""ctypes front-end for the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

* ``Port()`` -- oracle/liboracle_port.so, the plain-C restatement (wh_oracle*.c).
* ``Ref()``  -- oracle/_ref/libwh_ref_{linear,difacto}.so, the reference's own headers
  compiled unchanged (ref_linear.cc / ref_difacto.cc + shims/).  Exists only where it was
  built from /root/reference (the build container) or travelled with the snapshot.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing under wormhole_b200/ does.

Both classes expose the same methods, so a parity test reads
``check(gpu.localize(blk), oracle.localize(blk))`` whichever checker is available.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

LOSS_SQUARE, LOSS_LOGIT, LOSS_SQUARE_HINGE = 1, 2, 4     # learn/linear/config.proto:30-38
ALGO_SGD, ALGO_ADAGRAD, ALGO_FTRL = 1, 2, 3              # learn/linear/config.proto:45-52
NO_MAX_KEY = 2 ** 64 - 1                                 # ps::FLAGS_max_key default


def build(verbose=False):
    """make -C oracle: always the port; _ref only where /root/reference exists."""
    r = subprocess.run(["make", "-C", _HERE, "-s"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class LocalBlock:
    """A localized minibatch: what Localizer::Localize returns."""

    def __init__(self):
        self.uniq = self.cnt = self.offset = self.index = self.value = self.label = None

    @property
    def rows(self):
        return len(self.offset) - 1


class _Base:
    pre = ""

    def _setup_common(self, L):
        pre = self.pre
        for n in ("loc_nuniq", "loc_nnz", "loc_rows"):
            getattr(L, pre + n).restype = C.c_size_t
            getattr(L, pre + n).argtypes = [C.c_void_p]
        for n in ("loc_has_value", "loc_has_label"):
            getattr(L, pre + n).argtypes = [C.c_void_p]
        getattr(L, pre + "loc_free").argtypes = [C.c_void_p]
        getattr(L, pre + "loc_get").argtypes = [C.c_void_p] * 8
        getattr(L, pre + "reverse_bytes").restype = C.c_uint64
        getattr(L, pre + "reverse_bytes").argtypes = [C.c_uint64]
        getattr(L, pre + "l1l2_solve").restype = C.c_float
        getattr(L, pre + "l1l2_solve").argtypes = [C.c_float] * 4
        getattr(L, pre + "handle_push").restype = C.c_int64
        getattr(L, pre + "linear_create").restype = C.c_void_p
        getattr(L, pre + "linear_model_size").restype = C.c_size_t
        getattr(L, pre + "linear_model_size").argtypes = [C.c_void_p]
        getattr(L, pre + "linear_destroy").argtypes = [C.c_void_p]

    def _f(self, name):
        return getattr(self.L, self.pre + name)

    # -- keys ---------------------------------------------------------------------------
    def reverse_bytes(self, x):
        return int(self._f("reverse_bytes")(C.c_uint64(int(x))))

    def _collect(self, h, want_float_cnt=False):
        h = C.c_void_p(h)
        out = LocalBlock()
        U, n, rows = self._f("loc_nuniq")(h), self._f("loc_nnz")(h), self._f("loc_rows")(h)
        out.uniq = np.zeros(U, np.uint64)
        out.cnt = np.zeros(U, np.float32 if want_float_cnt else np.uint32)
        out.offset = np.zeros(rows + 1, np.uint64)
        out.index = np.zeros(n, np.uint32)
        out.value = np.zeros(n, np.float32) if self._f("loc_has_value")(h) else None
        out.label = np.zeros(rows, np.float32) if self._f("loc_has_label")(h) else None
        self._f("loc_get")(h, _p(out.uniq), None if want_float_cnt else _p(out.cnt),
                           _p(out.cnt) if want_float_cnt else None, _p(out.offset),
                           _p(out.index), _p(out.value), _p(out.label))
        self._f("loc_free")(h)
        return out

    # -- math ---------------------------------------------------------------------------
    def l1l2_solve(self, l1, l2, z, eta):
        return float(self._f("l1l2_solve")(l1, l2, z, eta))

    def handle_push(self, algo, l1, l2, eta, beta, grad, w, z=None, cg=None, t=1):
        """Applies one push message in place; eta/beta None = key absent from the conf.
        Returns (delta |w|_0, t after)."""
        n = len(grad)
        tt = C.c_int(t)
        grad = _f32(grad)
        r = self._f("handle_push")(
            C.c_int(algo), C.c_float(l1), C.c_float(l2), C.c_int(eta is not None),
            C.c_float(eta or 0), C.c_int(beta is not None), C.c_float(beta or 0), C.c_size_t(n),
            _p(grad), _p(w), _p(z), _p(cg), C.byref(tt))
        return int(r), tt.value

    # -- full job -----------------------------------------------------------------------
    def linear_create(self, algo=ALGO_FTRL, loss=LOSS_LOGIT, l1=1.0, l2=0.0, eta=None, beta=None,
                      max_key=NO_MAX_KEY):
        return C.c_void_p(self._f("linear_create")(
            C.c_int(algo), C.c_int(loss), C.c_float(l1), C.c_float(l2), C.c_int(eta is not None),
            C.c_float(eta or 0), C.c_int(beta is not None), C.c_float(beta or 0),
            C.c_uint64(max_key)))

    def linear_minibatch(self, job, blk, train=True):
        prog = np.zeros(6, np.float64)
        self._f("linear_minibatch")(job, _p(blk.index), _p(blk.offset), _p(blk.value),
                                    _p(blk.label), C.c_size_t(blk.rows), C.c_int(train), _p(prog))
        return prog

    def linear_model(self, job):
        n = self._f("linear_model_size")(job)
        keys = np.zeros(n, np.uint64)
        w, z, cg = (np.zeros(n, np.float32) for _ in range(3))
        self._f("linear_model_get")(job, _p(keys), _p(w), _p(z), _p(cg))
        return keys, w, z, cg

    def linear_destroy(self, job):
        self._f("linear_destroy")(job)


class _DifactoMixin:
    """DiFacto: learn/difacto/loss.h and learn/difacto/async_sgd.h (same calls on both checkers)."""

    def _dl(self):
        raise NotImplementedError

    def difacto_loss(self, D, model, model_siz, dim, nt=2, clip=0.0, norm=0.0, want_grad=True, dropout=0.0):
        """Loss(data, model, model_siz, conf); Evaluate; CalcGrad -> (prog[8], grad or None). dropout > 0 draws from
        this process's glibc rand() (learn/difacto/loss.h:150-154): srand() first."""
        L, pre = self._dl()
        model = _f32(model)
        siz = np.ascontiguousarray(model_siz, np.int32)
        prog = np.zeros(8, np.float64)
        grad = np.zeros(len(model), np.float32) if want_grad else None
        getattr(L, pre + "difacto_loss")(
            C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value), _p(D.label), _p(model),
            C.c_size_t(len(model)), _p(siz), C.c_size_t(len(siz)), C.c_int(dim), C.c_int(nt), C.c_float(clip),
            C.c_float(dropout), C.c_float(norm), _p(prog), _p(grad))
        return prog, grad

    def difacto_create(self, l1=1.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=0, threshold=0, v_l2=0.0,
                       v_eta=None, v_beta=None, init_scale=.01, clip=0.0, norm=0.0, nt=2, max_key=NO_MAX_KEY,
                       seed=0, dropout=0.0):
        L, pre = self._dl()
        f = getattr(L, pre + "difacto_create")
        f.restype = C.c_void_p
        return C.c_void_p(f(C.c_float(l1), C.c_float(l2), C.c_float(eta), C.c_float(beta), C.c_int(l1_shrk),
                            C.c_int(dim), C.c_int(threshold), C.c_float(v_l2), C.c_int(v_eta is not None),
                            C.c_float(v_eta or 0), C.c_int(v_beta is not None), C.c_float(v_beta or 0),
                            C.c_float(init_scale), C.c_float(dropout), C.c_float(clip), C.c_float(norm), C.c_int(nt),
                            C.c_uint64(max_key), C.c_uint(seed)))

    def difacto_minibatch(self, job, blk, train=True, data_pass=0):
        L, pre = self._dl()
        prog = np.zeros(8, np.float64)
        getattr(L, pre + "difacto_minibatch")(job, _p(blk.index), _p(blk.offset), _p(blk.value), _p(blk.label),
                                              C.c_size_t(blk.rows), C.c_int(train), C.c_int(data_pass), _p(prog))
        return prog

    def difacto_model(self, job, dim):
        """-> keys, fea_cnt, size, w[n, dim+1] = {w0, V}, cg[n, dim+2] = {sqc_grad_0, z_0, sqc_grad_V}"""
        L, pre = self._dl()
        f = getattr(L, pre + "difacto_model_size")
        f.restype = C.c_size_t
        n = f(job)
        keys = np.zeros(n, np.uint64); cnt = np.zeros(n, np.uint32); size = np.zeros(n, np.int32)
        w = np.zeros((n, dim + 1), np.float32); cg = np.zeros((n, dim + 2), np.float32)
        getattr(L, pre + "difacto_model_get")(job, C.c_int(dim), _p(keys), _p(cnt), _p(size), _p(w), _p(cg))
        return keys, cnt, size, w, cg

    def difacto_destroy(self, job):
        L, pre = self._dl()
        getattr(L, pre + "difacto_destroy")(job)


class Port(_Base, _DifactoMixin):
    """The plain-C restatement (always available; built on demand)."""
    pre = "orc_"
    kind = "port"

    def __init__(self):
        path = os.path.join(_HERE, "liboracle_port.so")
        if not os.path.exists(path):
            build()
        self.L = L = C.CDLL(path)
        self._setup_common(L)
        L.orc_localize.restype = C.c_void_p
        L.orc_remap.restype = C.c_void_p
        for n in ("orc_auc", "orc_accuracy", "orc_logit_objv", "orc_logloss"):
            getattr(L, n).restype = C.c_float

    def _dl(self):
        return self.L, "orc_"

    def localize(self, blk, max_key=NO_MAX_KEY, nt=2, float_cnt=False):
        h = self.L.orc_localize(_p(blk.index), _p(blk.offset), _p(blk.value), _p(blk.label),
                                C.c_size_t(blk.rows), C.c_uint64(max_key))
        return self._collect(h, float_cnt)

    def remap(self, blk, dict_keys, max_key=NO_MAX_KEY, nt=2):
        d = np.ascontiguousarray(dict_keys, np.uint64)
        h = self.L.orc_remap(_p(blk.index), _p(blk.offset), _p(blk.value), _p(blk.label),
                             C.c_size_t(blk.rows), C.c_uint64(max_key), _p(d), C.c_size_t(len(d)))
        return self._collect(h)

    def spmv_times(self, D, x, y=None, nt=2):
        y = np.zeros(D.rows, np.float32) if y is None else y
        self.L.orc_spmv_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                              _p(_f32(x)), _p(y))
        return y

    def spmv_trans_times(self, D, x, ysize, nt=2):
        y = np.empty(ysize, np.float32)
        self.L.orc_spmv_trans_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                                    _p(_f32(x)), _p(y), C.c_size_t(ysize))
        return y

    def spmm_times(self, D, x, dim, nt=2):
        y = np.empty(D.rows * dim, np.float32)
        self.L.orc_spmm_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                              _p(_f32(x)), _p(y), C.c_int(dim))
        return y

    def spmm_trans_times(self, D, x, ysize, dim, p=0.0, z=None, nt=2):
        y = np.empty(ysize, np.float32)
        z = _f32(z)
        self.L.orc_spmm_trans_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                                    _p(_f32(x)), C.c_float(p), _p(z), _p(y), C.c_size_t(ysize),
                                    C.c_int(dim))
        return y

    def bin_eval(self, label, predict, nt=2):
        label, predict = _f32(label), _f32(predict)
        n = C.c_size_t(len(label))
        return dict(auc=self.L.orc_auc(_p(label), _p(predict), n),
                    acc=self.L.orc_accuracy(_p(label), _p(predict), n, C.c_float(0)),
                    logloss=self.L.orc_logloss(_p(label), _p(predict), n, C.c_int(nt)),
                    logit_objv=self.L.orc_logit_objv(_p(label), _p(predict), n, C.c_int(nt)))

    def linear_loss(self, loss, D, w, nt=2, want_grad=True, want_xw=False):
        """-> (prog[6], grad or None[, xw, dual])"""
        w = _f32(w)
        prog = np.zeros(6, np.float64)
        grad = np.empty(len(w), np.float32) if want_grad else None
        xw = np.empty(D.rows, np.float32) if want_xw else None
        dual = np.empty(D.rows, np.float32) if want_xw else None
        self.L.orc_linear_loss(C.c_int(loss), C.c_size_t(D.rows), _p(D.offset), _p(D.index),
                               _p(D.value), _p(D.label), _p(w), C.c_size_t(len(w)), C.c_int(nt),
                               _p(prog), _p(grad), _p(xw), _p(dual))
        return (prog, grad, xw, dual) if want_xw else (prog, grad)


class Ref(_Base, _DifactoMixin):
    """The reference's own headers (oracle/_ref). Raises FileNotFoundError if absent."""
    pre = "ref_"
    kind = "reference"

    @staticmethod
    def available():
        return os.path.exists(os.path.join(_HERE, "_ref", "libwh_ref_linear.so"))

    def __init__(self):
        path = os.path.join(_HERE, "_ref", "libwh_ref_linear.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.L = L = C.CDLL(path)
        self._setup_common(L)
        L.ref_localize.restype = C.c_void_p
        L.ref_remap.restype = C.c_void_p

    def num_procs(self):
        return int(self.L.ref_num_procs())

    def _dl(self):
        if not hasattr(self, "_LD"):
            self._LD = C.CDLL(os.path.join(_HERE, "_ref", "libwh_ref_difacto.so"))
        return self._LD, "ref_"

    def localize(self, blk, max_key=NO_MAX_KEY, nt=2, float_cnt=False):
        h = self.L.ref_localize(_p(blk.index), _p(blk.offset), _p(blk.value), _p(blk.label),
                                C.c_size_t(blk.rows), C.c_uint64(max_key), C.c_int(nt),
                                C.c_int(2 if float_cnt else 1))
        return self._collect(h, float_cnt)

    def remap(self, blk, dict_keys, max_key=NO_MAX_KEY, nt=2):
        d = np.ascontiguousarray(dict_keys, np.uint64)
        h = self.L.ref_remap(_p(blk.index), _p(blk.offset), _p(blk.value), _p(blk.label),
                             C.c_size_t(blk.rows), C.c_uint64(max_key), C.c_int(nt), _p(d),
                             C.c_size_t(len(d)))
        return self._collect(h)

    def spmv_times(self, D, x, y=None, nt=2):
        y = np.zeros(D.rows, np.float32) if y is None else y
        self.L.ref_spmv_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                              _p(_f32(x)), _p(y), C.c_int(nt))
        return y

    def spmv_trans_times(self, D, x, ysize, nt=2):
        y = np.empty(ysize, np.float32)
        self.L.ref_spmv_trans_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                                    _p(_f32(x)), _p(y), C.c_size_t(ysize), C.c_int(nt))
        return y

    def spmm_times(self, D, x, dim, nt=2):
        x = _f32(x)
        y = np.empty(D.rows * dim, np.float32)
        self.L.ref_spmm_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value), _p(x),
                              C.c_size_t(len(x)), _p(y), C.c_size_t(len(y)), C.c_int(nt))
        return y

    def spmm_trans_times(self, D, x, ysize, dim, p=0.0, z=None, nt=2):
        x, z = _f32(x), _f32(z)
        y = np.zeros(ysize, np.float32)
        self.L.ref_spmm_trans_times(C.c_size_t(D.rows), _p(D.offset), _p(D.index), _p(D.value),
                                    _p(x), C.c_size_t(len(x)), C.c_float(p), _p(z), _p(y),
                                    C.c_size_t(ysize), C.c_int(nt))
        return y

    def bin_eval(self, label, predict, nt=2):
        label, predict = _f32(label), _f32(predict)
        out = np.zeros(5, np.float32)
        self.L.ref_bin_eval(_p(label), _p(predict), C.c_size_t(len(label)), C.c_int(nt), _p(out))
        return dict(auc=out[0], acc=out[1], logloss=out[2], logit_objv=out[3], copc=out[4])

    def linear_loss(self, loss, D, w, nt=2, want_grad=True, want_xw=False):
        w = _f32(w)
        prog = np.zeros(6, np.float64)
        grad = np.empty(len(w), np.float32) if want_grad else None
        self.L.ref_linear_loss(C.c_int(loss), C.c_size_t(D.rows), _p(D.offset), _p(D.index),
                               _p(D.value), _p(D.label), _p(w), C.c_size_t(len(w)), C.c_int(nt),
                               _p(prog), _p(grad))
        if want_xw:
            return prog, grad, self.spmv_times(D, w, nt=nt), None
        return prog, grad


def best():
    """The strongest checker available here: the reference itself, else the port."""
    return Ref() if Ref.available() else Port()
