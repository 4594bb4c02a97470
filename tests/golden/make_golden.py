"""Generates the golden vectors under tests/golden/ from oracle/_ref -- the reference's OWN headers
(/root/reference/learn/...) compiled unchanged (oracle/Makefile).  Run in the build container:

    python tests/golden/make_golden.py

The fixtures travel with the repo, so the CPU port and the CUDA path can be checked against the
reference's outputs where /root/reference does not exist (the GPU box).  Inputs are stored next to
the outputs; nothing here reads the reference at test time.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle as O  # noqa: E402
from tests.util import random_block, wide_ids  # noqa: E402
from wormhole_b200.rowblock import RowBlock, read_libsvm  # noqa: E402

REF_DATA = "/root/reference/learn/data/agaricus.txt.train"


def pack_block(prefix, blk, out):
    out[prefix + "offset"] = blk.offset
    out[prefix + "index"] = blk.index
    out[prefix + "label"] = blk.label
    if blk.value is not None:
        out[prefix + "value"] = blk.value


def pack_loc(prefix, r, out):
    out[prefix + "uniq"] = r.uniq
    out[prefix + "cnt"] = r.cnt
    out[prefix + "loffset"] = r.offset
    out[prefix + "lindex"] = r.index
    if r.value is not None:
        out[prefix + "lvalue"] = r.value


def main():
    ref = O.Ref()
    rng = np.random.default_rng(20260101)

    # --- agaricus (learn/data/agaricus.txt.train; learn/linear/guide/demo.conf) as a compact input fixture
    ag = read_libsvm(REF_DATA)
    assert ag.value is None and ag.index.max() < 256
    np.savez_compressed(os.path.join(HERE, "agaricus_train.npz"), offset=ag.offset.astype(np.uint32),
                        index=ag.index.astype(np.uint8), label=ag.label.astype(np.uint8))

    # --- localizer
    out = {}
    cases = []
    for i, (rows, mx, space, val) in enumerate([(1, 3, 10, False), (40, 8, 100, True), (700, 30, 2 ** 40, False),
                                                 (500, 12, 2 ** 62, True)]):
        blk = random_block(rng, rows, mx, space, val, empty_rows=(rows > 1))
        for max_key in (O.NO_MAX_KEY, 2 ** 24, 1000003, 7):
            name = "c%d_k%d_" % (i, [O.NO_MAX_KEY, 2 ** 24, 1000003, 7].index(max_key))
            pack_block(name, blk, out)
            out[name + "max_key"] = np.uint64(max_key)
            pack_loc(name, ref.localize(blk, max_key), out)
            cases.append(name)
    blk = RowBlock(np.arange(0, 2001, 4), wide_ids(rng, 2000), None, np.zeros(500))
    pack_block("wide_", blk, out); out["wide_max_key"] = np.uint64(O.NO_MAX_KEY)
    pack_loc("wide_", ref.localize(blk, O.NO_MAX_KEY), out)
    cases.append("wide_")
    # RemapIndex against a foreign dictionary
    blk = random_block(rng, 300, 10, 300, True)
    full = ref.localize(blk, 2 ** 24).uniq
    pack_block("dict_", blk, out)
    out["dict_keep"] = full[::2]
    r = ref.remap(blk, full[::2], 2 ** 24)
    out["dict_loffset"], out["dict_lindex"], out["dict_lvalue"] = r.offset, r.index, r.value
    out["cases"] = np.array(cases)
    np.savez_compressed(os.path.join(HERE, "golden_localize.npz"), **out)

    # --- SpMV / SpMM / loss / evaluation / handles
    out = {}
    for tag, val in (("b", False), ("v", True)):
        blk = random_block(rng, 600, 20, 3000, val)
        D = ref.localize(blk, 2 ** 24)
        U = len(D.uniq)
        pack_loc(tag + "_", D, out); out[tag + "_label"] = D.label
        x = rng.standard_normal(U).astype(np.float32)
        d = rng.standard_normal(D.rows).astype(np.float32)
        out[tag + "_x"], out[tag + "_d"] = x, d
        out[tag + "_times"] = ref.spmv_times(D, x)
        out[tag + "_trans"] = ref.spmv_trans_times(D, d, U)
        for dim in (1, 5, 16):
            X = rng.standard_normal(U * dim).astype(np.float32)
            Y = rng.standard_normal(D.rows * dim).astype(np.float32)
            Z = rng.standard_normal(U * dim).astype(np.float32)
            k = "%s_mm%d_" % (tag, dim)
            out[k + "X"], out[k + "Y"], out[k + "Z"] = X, Y, Z
            out[k + "times"] = ref.spmm_times(D, X, dim)
            out[k + "trans"] = ref.spmm_trans_times(D, Y, U * dim, dim)
            out[k + "trans_pz"] = ref.spmm_trans_times(D, Y, U * dim, dim, 0.5, Z)
        w = (rng.standard_normal(U) * 0.3).astype(np.float32)
        out[tag + "_w"] = w
        for loss in (O.LOSS_LOGIT, O.LOSS_SQUARE_HINGE):
            prog, grad = ref.linear_loss(loss, D, w, 2)
            out["%s_loss%d_prog" % (tag, loss)] = prog
            out["%s_loss%d_grad" % (tag, loss)] = grad
    label = (rng.random(5000) < 0.3).astype(np.float32)
    pred = (rng.standard_normal(5000) * 3).astype(np.float32)
    ev = ref.bin_eval(label, pred, 2)
    out["ev_label"], out["ev_pred"] = label, pred
    out["ev_out"] = np.array([ev["auc"], ev["acc"], ev["logloss"], ev["logit_objv"]], np.float32)
    n = 2000
    for algo in (O.ALGO_SGD, O.ALGO_ADAGRAD, O.ALGO_FTRL):
        w = (rng.standard_normal(n) * (rng.random(n) < 0.7)).astype(np.float32)
        z = rng.standard_normal(n).astype(np.float32)
        cg = np.abs(rng.standard_normal(n)).astype(np.float32)
        k = "h%d_" % algo
        out[k + "w0"], out[k + "z0"], out[k + "cg0"] = w.copy(), z.copy(), cg.copy()
        t = 1
        for step in range(2):
            g = (rng.standard_normal(n) * 2).astype(np.float32)
            out[k + "g%d" % step] = g
            dn, t = ref.handle_push(algo, 1.0, 0.1, 0.1, 1.0, g, w, z, cg, t)
            out[k + "w%d" % (step + 1)], out[k + "z%d" % (step + 1)], out[k + "cg%d" % (step + 1)] = w.copy(), z.copy(), cg.copy()
            out[k + "dn%d" % step] = np.int64(dn)
        out[k + "t"] = np.int64(t)
    np.savez_compressed(os.path.join(HERE, "golden_ops.npz"), **out)

    # --- the whole path on agaricus: demo.conf settings (FTRL, logit, lambda_l1 = 1, lr defaults,
    # minibatch 1000), one pass, 1 worker + 1 server, default ps::FLAGS_max_key (nibble-reversed keys)
    out = {}
    for name, algo, loss, max_key in (("ftrl_default", O.ALGO_FTRL, O.LOSS_LOGIT, O.NO_MAX_KEY),
                                      ("ftrl_hashed", O.ALGO_FTRL, O.LOSS_LOGIT, 2 ** 20),
                                      ("adagrad_hinge", O.ALGO_ADAGRAD, O.LOSS_SQUARE_HINGE, 2 ** 20),
                                      ("sgd_logit", O.ALGO_SGD, O.LOSS_LOGIT, 2 ** 20)):
        job = ref.linear_create(algo, loss, 1.0, 0.0, None, None, max_key)
        progs = []
        for b in range(0, ag.rows, 1000):
            progs.append(ref.linear_minibatch(job, ag.slice(b, min(ag.rows, b + 1000))))
        keys, w, z, cg = ref.linear_model(job)
        out[name + "_prog"] = np.array(progs)
        out[name + "_keys"], out[name + "_w"], out[name + "_z"], out[name + "_cg"] = keys, w, z, cg
        out[name + "_conf"] = np.array([algo, loss, max_key], dtype=np.uint64)
        ref.linear_destroy(job)
    np.savez_compressed(os.path.join(HERE, "golden_agaricus.npz"), **out)
    # --- DiFacto (learn/difacto): the loss on a ZVPull-shaped payload, and the guide's demo.conf
    # (embedding dim 5, threshold 5) on agaricus for two passes, 1 worker + 1 server
    from tests.test_oracle_vs_ref import make_model
    out = {}
    blk = random_block(rng, 400, 15, 900, True)
    D = ref.localize(blk, 2 ** 24)
    pack_loc("l_", D, out); out["l_label"] = D.label
    for dim in (0, 5):
        model, siz = make_model(rng, len(D.uniq), dim)
        prog, grad = ref.difacto_loss(D, model, siz, dim, 2, 0.0, 0.0)
        out["l%d_model" % dim], out["l%d_siz" % dim], out["l%d_prog" % dim], out["l%d_grad" % dim] = model, siz, prog, grad
    kw = dict(l1=1.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=5, threshold=5, v_l2=0.0, init_scale=.01,
              max_key=2 ** 20, seed=7)
    job = ref.difacto_create(**kw)
    progs = [ref.difacto_minibatch(job, ag.slice(b, min(ag.rows, b + 1000)), True, dp)
             for dp in range(2) for b in range(0, ag.rows, 1000)]
    keys, cnt, size, w, cg = ref.difacto_model(job, 5)
    ref.difacto_destroy(job)
    out["ag_prog"], out["ag_keys"], out["ag_cnt"], out["ag_size"], out["ag_w"], out["ag_cg"] = np.array(progs), keys, cnt, size, w, cg
    import ctypes
    libc = ctypes.CDLL("libc.so.6"); libc.srand(ctypes.c_uint(7))
    out["ag_rand"] = np.array([libc.rand() for _ in range(4096)], np.int32)   # the rand() values V was drawn from
    np.savez_compressed(os.path.join(HERE, "golden_difacto.npz"), **out)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
