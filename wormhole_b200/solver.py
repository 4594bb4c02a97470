"""The scheduler-side behaviour around the hot path, single process, one GPU:

  Config             learn/linear/config.proto:3-79 read the way learn/base/arg_parser.h:40-63 reads it
                     (conf file and/or `key=value` arguments, '=' meaning ':' outside quotes)
  Progress           learn/linear/progress.h:7-37 (same header and row format)
  MinibatchWorker    learn/solver/minibatch_solver.h:284-315 (Process: reader -> ProcessMinibatch with
                     max_concurrency minibatches in flight; 10 M-row minibatches and concurrency 10 for VAL/PRED)
  MinibatchScheduler learn/solver/minibatch_solver.h:85-192 (Run / Iterate / ShowProgress: load model -> PRED pass,
                     TRAIN + VAL passes per data pass, stop criteria, periodic and final model files)

ps-lite's node roles collapse: the scheduler loop, the workload pool (file parts handed out in order, or in
rand() order like learn/base/workload_pool.h:132-149 when `shuffle_parts`) and the one worker live in this
process; server 0 is the whServer on the same GPU. Text is parsed and minibatches are assembled on the GPU
(wormhole_b200/reader.py)."""
import os
import re
import sys
import time

import numpy as np

from . import capi
from .reader import MinibatchIter, _rand

TRAIN, VAL, PRED = 1, 2, 3      # learn/base/workload.h Workload::Type

_LOSS = {"SQUARE": 1, "LOGIT": 2, "SQUARE_HINGE": 4}
_ALGO = {"SGD": 1, "ADAGRAD": 2, "FTRL": 3}


class Config:
    """linear::Config with the defaults of learn/linear/config.proto."""
    DEFAULTS = dict(train_data="", val_data="", data_format="libsvm", model_out="", model_in="", predict_out="",
                    loss="LOGIT", lambda_l1=1.0, lambda_l2=0.0, algo="FTRL", minibatch=1000, max_data_pass=10,
                    lr_eta=.01, save_iter=-1, load_iter=-1, local_data=False, num_parts_per_file=10,
                    rand_shuffle=10, neg_sampling=1.0, prob_predict=True, dropout=0.0, print_sec=1.0, lr_beta=1.0,
                    num_threads=2, max_concurrency=2, key_cache=True, msg_compression=True, fixed_bytes=0)

    def __init__(self, text="", **kw):
        self.__dict__.update(self.DEFAULTS)
        self.given = set()
        self.merge(text)
        for k, v in kw.items():
            self._set(k, v)

    @staticmethod
    def from_file(path, args=()):
        c = Config(open(path).read())
        c.merge(" ".join(args))
        return c

    def has(self, key):
        """protobuf has_<key>(): set explicitly (async_sgd.h:205-206 only overrides the handle defaults then)."""
        return key in self.given

    def _set(self, k, v):
        if k not in self.DEFAULTS:
            raise capi.WormholeError("conf: unknown field %r" % k)        # TextFormat::Parse fails -> CHECK
        d = self.DEFAULTS[k]
        if isinstance(d, bool):
            v = v if isinstance(v, bool) else str(v).lower() in ("true", "1", "t")
        elif isinstance(d, int):
            v = int(v)
        elif isinstance(d, float):
            v = float(np.float32(float(v)))                                # proto `float` fields
        else:
            v = str(v)
        setattr(self, k, v)
        self.given.add(k)

    def merge(self, text):
        # ArgParser::ReplaceStr: '=' -> ':' outside quotes; then protobuf text format `key: value`
        out, q = [], False
        for ch in text:
            if ch in "'\"":
                q = not q
            out.append(":" if ch == "=" and not q else ch)
        text = re.sub(r"#[^\n]*", "", "".join(out))
        for m in re.finditer(r"([A-Za-z_][A-Za-z_0-9]*)\s*:\s*(\"[^\"]*\"|'[^']*'|[^\s]+)", text):
            v = m.group(2)
            if v[0] in "'\"":
                v = v[1:-1]
            self._set(m.group(1), v)


class Progress:
    """learn/linear/progress.h: data = [objv, acc, auc, count, new_ex, new_w]."""

    def __init__(self):
        self.data = [0.0] * 6
        self.ttl_ex = 0.0
        self.nnz_w = 0.0

    @staticmethod
    def HeadStr():
        return "  ttl #ex   inc #ex    |w|_0       logloss  accuracy     AUC"

    def PrintStr(self):
        objv, acc, auc, count, new_ex, new_w = self.data
        self.ttl_ex += new_ex
        self.nnz_w += new_w
        if new_ex == 0:
            return ""
        return "%8.3g  %8.3g  %11.6g  %8.6f  %8.6f  %8.6f" % (self.ttl_ex, new_ex, self.nnz_w, objv / new_ex,
                                                             acc / count, auc / count)


class DifactoProgress:
    """learn/difacto/progress.h: data = [objv, auc, objv_w, copc, count, new_ex, new_w, new_V]."""

    def __init__(self):
        self.data = [0.0] * 8
        self.ttl_ex = self.nnz_w = self.nnz_V = 0.0

    @staticmethod
    def HeadStr():
        return "  ttl #ex   inc #ex |  |w|_0  logloss_w |   |V|_0    logloss    AUC"

    def PrintStr(self):
        objv, auc, objv_w, _copc, count, new_ex, new_w, new_V = self.data
        self.ttl_ex += new_ex
        self.nnz_w += new_w
        self.nnz_V += new_V
        if new_ex == 0:
            return ""
        return "%9.4g  %7.2g | %9.4g  %6.4f | %9.4g  %7.5f  %7.5f " % (
            self.ttl_ex, new_ex, self.nnz_w, objv_w / new_ex, self.nnz_V, objv / new_ex, auc / count)


class DifactoStop:
    """AsyncScheduler::Stop, learn/difacto/async_sgd.h:31-49: abort a training pass whose objective exceeds
    max_objv; stop after a validation pass that improved the objective by less than min_objv_decr."""

    def __init__(self, early_stop=False, min_objv_decr=1e-5, max_objv=None, out=sys.stdout):
        self.early_stop, self.min_objv_decr, self.max_objv, self.out = early_stop, min_objv_decr, max_objv, out
        self.pre_val_objv = 100.0

    def __call__(self, objv, new_ex, train):
        cur = objv / new_ex
        if train:
            return self.max_objv is not None and cur > self.max_objv
        diff = self.pre_val_objv - cur
        self.pre_val_objv = cur
        if self.early_stop and diff < self.min_objv_decr:
            self.out.write("The decrease of validation objective is smaller than the minimal requirement: "
                           "%g vs %g\n" % (diff, self.min_objv_decr))
            return True
        return False


def match_files(pattern):
    """learn/base/match_file.h: the files of dirname(pattern) whose name matches the regex basename(pattern)."""
    if os.path.isfile(pattern):
        return [pattern]
    d, base = os.path.split(pattern)
    d = d or "."
    rx = re.compile(base)
    return sorted(os.path.join(d, f) for f in os.listdir(d) if rx.fullmatch(f) or rx.match(f))


class MinibatchWorker:
    """AsgdWorker on one GPU: whWorker with `max(max_concurrency, ...)` lanes on a local whServer."""
    val_mb_size = 10000000        # minibatch_solver.h:239
    val_concurrent_mb = 10        # :245

    def __init__(self, ctx, conf, server, max_key=capi.NO_MAX_KEY):
        self.ctx, self.conf, self.server = ctx, conf, server
        self.depth = max(1, min(conf.max_concurrency, 8))
        self.worker = capi.Worker(ctx, server, _LOSS[conf.loss], max_key, depth=self.depth)
        self.max_key = max_key

    def process(self, wl_type, filename, k, n, report, predict_stream=None):
        """MinibatchWorker::Process for one workload (file part k of n). report(prog6) per finished minibatch."""
        conf = self.conf
        train = wl_type == TRAIN
        mb_size = conf.minibatch if train else self.val_mb_size
        shuffle = conf.minibatch * conf.rand_shuffle if train else 0
        neg = conf.neg_sampling if train else 1.0
        max_mb = 1 if wl_type == PRED else min(self.depth, conf.max_concurrency if train else self.val_concurrent_mb)
        reader = MinibatchIter(self.ctx, filename, k, n, conf.data_format, mb_size, shuffle, neg, ring=self.depth + 2)
        reader.BeforeFirst()
        pend = []
        while reader.Next():
            mb = reader.Value()
            while len(pend) >= max_mb:                                     # WaitMinibatch(max_mb)
                report(self.worker.wait(pend.pop(0)))
            if wl_type == PRED:
                # async_sgd.h:262-265: Evaluate + report as in a VAL pass, then Predict (loss.h:55-63)
                report(self.worker.wait(self.worker.submit(mb, False, capi.WH_DEVICE, mb.nnz)))
                out = self.worker.predict(mb, conf.prob_predict, capi.WH_DEVICE)
                if predict_stream is not None:
                    predict_stream.write("".join("%.6g\n" % p for p in out))
            else:
                pend.append(self.worker.submit(mb, train, capi.WH_DEVICE, mb.nnz))
        for t in pend:                                                     # WaitMinibatch(1)
            report(self.worker.wait(t))


class MinibatchScheduler:
    """Run() of learn/solver/minibatch_solver.h:85-147 for learn/linear (AsgdScheduler, async_sgd.h:304-324)."""

    def __init__(self, conf, ctx=None, max_key=capi.NO_MAX_KEY, hash_slots=1 << 22, out=sys.stdout,
                 shuffle_parts=False):
        self.conf, self.out, self.shuffle_parts = conf, out, shuffle_parts
        self.ctx = ctx or capi.Context(0)
        dense = max_key != capi.NO_MAX_KEY
        eta = conf.lr_eta if conf.has("lr_eta") else None
        beta = conf.lr_beta if conf.has("lr_beta") else None
        self.server = capi.Server(self.ctx, _ALGO[conf.algo], conf.lambda_l1, conf.lambda_l2, eta, beta, 0,
                                  max_key, hash_slots=0 if dense else hash_slots)
        self.worker = MinibatchWorker(self.ctx, conf, self.server, max_key)
        self.prog = Progress()
        self.rows = []                 # the progress rows printed, as (seconds, text): what tests inspect

    def _print(self, s):
        self.out.write(s + "\n")
        self.out.flush()

    def model_name(self, base, it):                                          # iter_solver.h:113-117, rank 0
        return base + ("_iter-%d" % it if it >= 0 else "") + "_part-0"

    def stop(self, prog, is_train):                                          # minibatch_solver.h default: never
        return False

    def iterate(self, it, wl_type):
        conf = self.conf
        is_train = wl_type == TRAIN
        if is_train:
            data = conf.train_data
            self._print("Training: iter = %d" % it)
        else:
            data = conf.val_data
            if wl_type == PRED:
                self._print("Predicting")
            else:
                self._print("Validating: iter = %d" % it)
                if not data:
                    return False
        self._print("  sec %s" % Progress.HeadStr())
        parts = [(f, k, conf.num_parts_per_file) for f in match_files(data) for k in range(conf.num_parts_per_file)]
        acc = [0.0] * 6
        last = [time.time()]
        stop = [False]

        def show():
            self.prog.data, acc[:] = list(acc), [0.0] * 6
            s = self.prog.PrintStr()
            if not s:
                return False
            self.rows.append((time.time() - self.start_time, s))
            self._print("%5.0f  %s" % (time.time() - self.start_time, s))
            return self.stop(self.prog, is_train)

        def report(p):
            for i in range(6):
                acc[i] += p[i]
            if is_train and time.time() - last[0] >= conf.print_sec:         # :165-171 print every print_sec
                last[0] = time.time()
                stop[0] = show() or stop[0]

        pred = None
        while parts and not stop[0]:
            f, k, n = parts.pop(_rand() % len(parts) if self.shuffle_parts else 0)
            if wl_type == PRED and conf.predict_out:                         # iter_solver.h:137-153 PredictStream
                pred = open(conf.predict_out + os.path.basename(f) + "_part-%d" % k, "w")
            self.worker.process(wl_type, f, k, n, report, pred)
            if pred:
                pred.close()
        return show() or stop[0]                                             # the rest of the pass (:174-177 for VAL)

    def Run(self):
        conf = self.conf
        self._print("Connected 1 servers and 1 workers")
        self.start_time = time.time()
        is_predict = bool(conf.predict_out)
        if is_predict and not conf.model_in:
            raise capi.WormholeError("should provide model_in for predicting")
        cur = 0
        if conf.model_in:
            if conf.load_iter > 0:
                self._print("Loading model from iter = %d" % conf.load_iter)
                cur = conf.load_iter
            else:
                self._print("Loading the last model")
                cur = -1
            self.server.load(self.model_name(conf.model_in, cur))
            self.iterate(cur, PRED)
            cur += 1
        if is_predict:
            self._print("Prediction is finished!")
            return True
        while cur < conf.max_data_pass:
            if self.iterate(cur, TRAIN) or self.iterate(cur, VAL):
                self._print("Hit stop critera")
                break
            if cur == conf.max_data_pass - 1:
                self._print("Hit max number of data passes %d" % conf.max_data_pass)
                break
            if conf.model_out and conf.save_iter > 0 and (cur + 1) % conf.save_iter == 0:
                self._print("Saving model for iter = %d" % cur)
                self.server.save(self.model_name(conf.model_out, cur))
            cur += 1
        if conf.model_out:
            self._print("Saving the final model")
            self.server.save(self.model_name(conf.model_out, -1))
        self._print("Training is finished!")
        return True


def main(argv=None):
    """linear.dmlc-style entry: python -m wormhole_b200.solver <conf> [key=value ...] [--max-key N]"""
    argv = list(sys.argv[1:] if argv is None else argv)
    max_key = capi.NO_MAX_KEY
    if "--max-key" in argv:
        i = argv.index("--max-key")
        max_key = int(argv[i + 1])
        del argv[i:i + 2]
    conf = Config.from_file(argv[0], argv[1:])
    MinibatchScheduler(conf, max_key=max_key).Run()


if __name__ == "__main__":
    main()
