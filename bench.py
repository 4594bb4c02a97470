#!/usr/bin/env python
"""Benchmark of the sparse minibatch training hot path (BASELINE.json: "train examples/sec at
Criteo shape, 1/2/4/8 B200; % of HBM roofline").

    python bench.py --gpus N --steps K --warmup W            our CUDA path (one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code

A "step" is one minibatch of the synthetic Criteo-shaped stream (100 000 rows x 39 binary
features, ids hashed into 2^24 keys, FTRL with L1+L2, logit loss) through
AsgdWorker::ProcessMinibatch's replacement: localize -> pull -> Xw -> loss + metrics ->
X^T dual -> FTRL push.  Rank 0 prints ONE JSON line.

  value     examples/s with the minibatches already resident in HBM (a pool larger than L2)
  e2e       the same through the C-ABI with HOST (pinned) buffers: H2D of every minibatch and the
            D2H of its progress numbers are inside the timed region
  roofline  the slowest kernel group of the step, timed with CUDA events on the launching stream
  cpu_baseline  the reference's own headers (oracle/_ref) on the host cores, bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, NNZ_PER_ROW, MAX_KEY = 100000, 39, 2 ** 24
FTRL = dict(l1=1.0, l2=0.1, eta=0.1, beta=1.0)
METRIC = "train examples/sec at Criteo shape"
WORKLOAD = "synthetic Criteo-shape sparse LR: 39 nnz/row, 2^24 hashed features, minibatch 100k, FTRL L1+L2, logit loss"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------ the reference arm ----------

def _ref_worker(rank, nprocs, rows, pool, warmup, steps, out_q, go):
    """One reference worker + server pair (AsgdWorker nt_=2 threads) on its own minibatch stream."""
    from oracle import oracle
    from wormhole_b200 import synth
    chk = oracle.best()
    job = chk.linear_create(oracle.ALGO_FTRL, oracle.LOSS_LOGIT, FTRL["l1"], FTRL["l2"], FTRL["eta"],
                            FTRL["beta"], MAX_KEY)
    blks = [synth.criteo_minibatch(rank * 1000 + m, rows=rows) for m in range(pool)]
    for i in range(warmup):
        chk.linear_minibatch(job, blks[i % pool])
    out_q.put(("ready", rank, chk.kind))
    go.wait()
    t0 = time.perf_counter()
    for i in range(steps):
        chk.linear_minibatch(job, blks[(warmup + i) % pool])
    out_q.put(("done", rank, time.perf_counter() - t0))


def run_reference(args):
    """Times the reference's own CPU implementation of the path (oracle/_ref = its headers compiled
    unchanged, driven through the real AsgdWorker::ProcessMinibatch against an in-process stand-in
    for ps-lite; the C port if _ref did not travel). All host cores: one worker+server pair per two
    cores, because the reference worker hard-codes nt_ = 2 (learn/linear/async_sgd.h:303)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 2
    nprocs = args.cpu_procs or max(1, min(32, cores // 2))
    # bounded sample: ~3.5 us per row per worker on this class of host; keep K+W steps within ~2 minutes
    budget_rows = int(120.0 / max(1, args.steps + args.warmup) / 3.5e-6)
    rows = max(2000, min(args.rows, budget_rows))
    # (load the checker library here, before forking: the children inherit it, and the parent's map shows what ran)
    from oracle import oracle as _O
    _O.best()
    from wormhole_b200 import synth as _synth
    _b = _synth.criteo_minibatch(0, rows=args.rows)
    _N = int(_b.offset[-1])
    _r = len(np.unique(_b.index % np.uint64(MAX_KEY))) / args.rows
    ctx = mp.get_context("fork")
    q, go = ctx.Queue(), ctx.Event()
    procs = [ctx.Process(target=_ref_worker, args=(r, nprocs, rows, 2, args.warmup, args.steps, q, go))
             for r in range(nprocs)]
    for p in procs:
        p.start()
    kind = "port"
    for _ in procs:
        _, _, kind = q.get()
    go.set()
    times = [q.get()[2] for _ in procs]
    for p in procs:
        p.join()
    secs = max(times)
    value = nprocs * rows * args.steps / secs
    kind = "reference" if kind == "reference" else "port"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "examples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config("criteo", "ftrl", "logit", MAX_KEY, args.rows, _N, _r, args.pool, max(1, args.gpus), args),
        "cpu_baseline": {"value": value, "unit": "examples/s", "cores": min(cores, 2 * nprocs), "kind": kind,
                         "sample": "%d worker+server pairs x %d minibatches of %d rows each (of the workload's %d), "
                                   "nt=2 per worker, in-process ps stand-in (no network, no serialisation)"
                                   % (nprocs, args.steps, rows, args.rows)},
        "e2e": {"value": value, "unit": "examples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------ our arm -------------------

class ClockSampler:
    """One `nvidia-smi -lms` process streaming clocks and throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc, self.rows = None, []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            pass

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        self.rows = [[x.strip() for x in l.split(",")] for l in out.splitlines() if l.count(",") >= 6]

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        # "under load": the samples whose power draw is in the upper half of what was seen
        if pw and len(pw) == len(sm):
            cut = (min(pw) + max(pw)) / 2
            load = [c for c, p in zip(sm, pw) if p >= cut] or sm
        else:
            load = sm
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows), "power_w_max": max(pw) if pw else None}


def segment_bytes(name, N, B, U, words):
    """Algorithmic bytes one launch group must move ("kernels")."""
    s = 12  # FTRL state bytes per key
    return {
        "loc_mark": 8 * N,                              # ids read once
        "loc_dir_scan": 8 * words + 4 * words + 8 * U,  # directory read, prefix written, keys emitted
        "loc_rank": 8 * N + 4 * N,                      # ids read, local ids written
        "loc_dir_clear": 8 * words,
        "pull": U * (8 + 4 + 4),                        # key read, weight gathered, weight written
        "spmv_times": 4 * N + 8 * B + 4 * B + 4 * U,    # local ids, offsets, Xw written, w gathered
        "spmv_loss": 4 * N + 8 * B + 12 * B + 4 * U,    # the same pass + label read, dual written
        "loss_forward": 12 * B,                         # label, Xw read; dual written
        "auc": 8 * B,                                   # label, Xw read
        "spmv_trans": 4 * N + 8 * B + 4 * B + 8 * U,    # local ids, offsets, dual, accumulators
        "spmv_fused": 4 * N + 8 * B + 8 * B + 4 * U + 8 * U,  # forward + backward in one launch: local ids once, offsets, label, Xw, w, accumulators
        "push": U * (8 + 8 + 2 * s),                    # key, accumulator, state read + written
        # the shared-GPU step (csrc/direct.cu): the 64-bit ids once, offsets, labels, Xw, one weight gathered and one
        # gradient accumulated per unique key -- SURVEY.md 8(d)'s per-example figure without the state update
        "fused_direct": 8 * N + 8 * B + 4 * B + 4 * B + U * (4 + 4),
    }.get(name, 0)


ALGOS = {"ftrl": 3, "adagrad": 2, "sgd": 1}
LOSSES = {"logit": 2, "sqhinge": 4, "square": 1}
STATE_BYTES = {"ftrl": 12, "adagrad": 8, "sgd": 4}


def workload_name(shape, algo, loss, max_key, rows):
    if (algo, loss, shape) == ("ftrl", "logit", "criteo"):
        return WORKLOAD
    return "synthetic %s-shape sparse LR, %s, %s loss, max_key %d, minibatch %d" % (shape, algo, loss, max_key, rows)


class Job:
    """One (worker, server) pair per rank for one configuration: rank g = worker g + server g of the key range
    [g*max_key/S, (g+1)*max_key/S). world == 1: worker and the one server share the GPU. world > 1, exchange p2p: every
    rank maps the other ranks' shards and lane outboxes over NVLink (CUDA IPC); exchange nccl: all-to-all-v of
    (key, weight) / (key, grad) messages through torch.distributed (wormhole_b200/dist.py)."""

    def __init__(self, ctx, rank, world, local, algo, loss, max_key, depth, exchange, max_uniq, max_delay=0):
        import torch
        import torch.distributed as dist
        from wormhole_b200 import capi
        self.world, self.depth, self.sw = world, depth, None
        if world == 1:
            self.srv = capi.Server(ctx, ALGOS[algo], FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], 0, max_key)
            self.wk = capi.Worker(ctx, self.srv, LOSSES[loss], max_key, depth=depth)
        else:
            from wormhole_b200.dist import DeviceShard, ShardedKV, ShardedWorker, key_range
            kb, ke = key_range(rank, world, max_key)
            self.srv = capi.Server(ctx, ALGOS[algo], FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], kb, ke)
            if exchange == "p2p":
                handles = [None] * world
                dist.all_gather_object(handles, self.srv.ipc_export())
                self.srv.ipc_attach(rank, world, max_key, handles)
                dist.barrier()
                self.wk = capi.Worker(ctx, self.srv, LOSSES[loss], max_key, depth=depth)
                boxes = [None] * world
                dist.all_gather_object(boxes, self.wk.ipc_export(max_uniq))
                self.wk.ipc_attach(boxes)
                dist.barrier()
            else:
                self.depth = 1
                self.wk = capi.Worker(ctx, None, LOSSES[loss], max_key)
                self.sw = ShardedWorker(ctx, self.wk, ShardedKV(DeviceShard(self.srv), max_key, device=torch.device("cuda", local)))
        if max_delay and self.sw is None:
            self.wk.set_max_delay(max_delay)

    def run(self, blocks, mem, first, count, progs=None):
        """`count` training minibatches, up to `depth` in flight (conf max_concurrency); every one is waited for, so
        its progress numbers have reached the host before the caller's clock stops."""
        n = len(blocks)
        if self.sw is not None:
            for i in range(count):
                b = blocks[(first + i) % n]
                p = self.sw.process_minibatch(b, True, mem, b.nnz)
                if progs is not None:
                    progs.append(p)
            return
        pending = []
        for i in range(count):
            if len(pending) == self.depth:
                p = self.wk.wait(pending.pop(0))
                if progs is not None:
                    progs.append(p)
            b = blocks[(first + i) % n]
            pending.append(self.wk.submit(b, True, mem, b.nnz))
        for t in pending:
            p = self.wk.wait(t)
            if progs is not None:
                progs.append(p)

    def close(self):
        self.wk.close(); self.srv.close()


def make_pool(shape, rank, pool_n, rows, want_host=True):
    """pool_n distinct minibatches of this rank's stream: in pinned host memory (e2e leg) and resident in HBM."""
    import torch
    from wormhole_b200 import synth
    host, dev, raw = [], [], []
    for m in range(pool_n):
        b = synth.criteo_minibatch(rank * 100000 + m, rows=rows) if shape == "criteo" else synth.kdd_minibatch(rank * 100000 + m, rows=rows)
        h = type("B", (), {})()
        h.rows, h.nnz = b.rows, b.nnz
        h.offset = torch.from_numpy(b.offset.view(np.int64))
        h.index = torch.from_numpy(b.index.view(np.int64))
        h.label = torch.from_numpy(b.label)
        if want_host:
            h.offset, h.index, h.label = h.offset.pin_memory(), h.index.pin_memory(), h.label.pin_memory()
        h.value = None
        d = type("B", (), {})()
        d.rows, d.nnz = b.rows, b.nnz
        d.offset, d.index, d.label, d.value = h.offset.cuda(), h.index.cuda(), h.label.cuda(), None
        host.append(h); dev.append(d); raw.append(b)
    return host, dev, raw


def timed_region(job, blocks, mem, steps, warmup, stream, world):
    """W untimed steps, then exactly K steps between two CUDA events on the launching stream, bracketed by a barrier and
    a device synchronisation on both sides; the maximum over the ranks."""
    import torch
    import torch.distributed as dist
    from wormhole_b200 import capi

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    job.run(blocks, mem, 0, warmup)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = capi.kernel_launches()
    t0 = time.perf_counter()
    e0.record(stream)
    job.run(blocks, mem, warmup, steps)
    e1.record(stream)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms, 1e3 * wall], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, wall = float(t[0].item()), float(t[1].item()) * 1e-3
    return ms, wall, capi.kernel_launches() - l0


def profile_segments(ctx, job, dev, steps):
    """Per-kernel-group timing with CUDA events on the launching stream (wh_profile_*), one minibatch at a time."""
    import ctypes as C
    from wormhole_b200 import capi
    lib = capi.lib()
    lib.wh_profile_enable(ctx.h, 1)
    for i in range(steps):
        b = dev[i % len(dev)]
        if job.sw is not None:
            job.sw.process_minibatch(b, True, capi.WH_DEVICE, b.nnz)
        else:
            job.wk.process_minibatch(b, True, capi.WH_DEVICE)
    segs = []
    for i in range(lib.wh_profile_count(ctx.h)):
        name, ms, cnt = C.c_char_p(), C.c_double(), C.c_uint64()
        lib.wh_profile_get(ctx.h, i, C.byref(name), C.byref(ms), C.byref(cnt))
        segs.append((name.value.decode(), ms.value / max(1, cnt.value)))
    lib.wh_profile_enable(ctx.h, 0)
    return segs


def parity_leg(ctx, rank, world, local, args, max_key, rows, shape, algo, loss, max_uniq, steps=3):
    """Outside every timed region: `steps` minibatches per rank of the bench configuration itself through a fresh job,
    then every rank's shard {w, z, sq_cum_grad} against the CPU checker (oracle/: the checker leg, the one place besides
    the CPU arm where bench.py runs it) driven in the same message order -- all pulls of a step see the model as of the
    previous step, each owner applies the pushes of a step one handle update per source rank in ascending rank order
    (tests/test_dist_gpu.py), and against the same simulation with X^T dual summed exactly. The device has to match the
    second one entry by entry (1e-5 relative + 1e-7; in practice bit for bit). Against the first, per field and shard, the
    relative L2 distance must stay within 1e-5 or within the distance the exact-sum simulation itself has from it -- a
    CPU-only number: the reference's own fp32 summation error, 1e-5 .. 3e-5 of a popular key's gradient at 100 000 rows,
    which compounds when several workers push per step; per-entry misses are counted and reported."""
    import torch.distributed as dist
    from wormhole_b200 import capi, synth
    from wormhole_b200.dist import key_range
    job = Job(ctx, rank, world, local, algo, loss, max_key, args.depth, args.exchange, max_uniq)
    _, dev, raw = make_pool(shape, rank, steps, rows, want_host=False)
    job.run(dev, capi.WH_DEVICE, 0, steps)
    import torch
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    # rank 0 regenerates every rank's minibatches, simulates, and tells each rank which keys to report
    out = None
    if rank == 0:
        from oracle import oracle as O
        chk = O.best()
        A = {"ftrl": O.ALGO_FTRL, "adagrad": O.ALGO_ADAGRAD, "sgd": O.ALGO_SGD}[algo]
        L = {"logit": O.LOSS_LOGIT, "sqhinge": O.LOSS_SQUARE_HINGE}[loss]
        gen = synth.criteo_minibatch if shape == "criteo" else synth.kdd_minibatch
        bounds = [key_range(r, world, max_key)[0] for r in range(world)] + [max_key]
        W, Z, CG = {}, {}, {}
        t = [1] * world
        touched = set()
        seen = np.zeros(max_key, np.float64)   # entries of each key so far: bounds sum |terms| of its gradients (|dual| <= 1)
        Wd = np.zeros(max_key, np.float32); Zd = np.zeros(max_key, np.float32); Cd = np.zeros(max_key, np.float32)
        # a second simulation with ONE change: X^T dual summed exactly (fp64 over the fp32 terms) instead of in fp32 row
        # order. The device adds integers, so it must agree with this one to rounding; the distance between the two
        # simulations is the reference's own summation error
        We = np.zeros(max_key, np.float32); Ze = np.zeros(max_key, np.float32); Ce = np.zeros(max_key, np.float32)
        te = [1] * world
        for step in range(steps):
            msgs = []
            emsgs = []
            for r in range(world):
                blk = raw[step] if r == 0 else gen(r * 100000 + step, rows=rows)
                D = chk.localize(blk, max_key)
                keys = D.uniq.astype(np.int64)
                _, grad = chk.linear_loss(L, D, Wd[keys], 2)
                seen[keys] += np.bincount(D.index, minlength=len(keys)) * (2.0 if loss == "sqhinge" else 1.0)
                msgs.append((keys, grad, np.searchsorted(keys, bounds)))
                xw = chk.spmv_times(D, We[keys])
                y = np.where(D.label > 0, 1.0, -1.0).astype(np.float32)
                if loss == "logit":
                    dual = (-y / (np.float32(1) + np.exp((y * xw).astype(np.float64)).astype(np.float32))).astype(np.float32)
                else:
                    dual = (y * ((y * xw) > np.float32(1.0))).astype(np.float32)
                rows_of = np.repeat(np.arange(D.rows), np.diff(D.offset.astype(np.int64)))
                ge = np.bincount(D.index, weights=dual[rows_of].astype(np.float64), minlength=len(keys))
                ge = (ge.astype(np.float32) * np.float32(-2.0 if loss == "sqhinge" else 1.0)).astype(np.float32)
                emsgs.append((keys, ge, msgs[-1][2]))
            for sidx in range(world):
                for r in range(world):
                    keys, grad, cut = msgs[r]
                    sl = slice(cut[sidx], cut[sidx + 1])
                    w, z, cg = Wd[keys[sl]].copy(), Zd[keys[sl]].copy(), Cd[keys[sl]].copy()
                    _, t[sidx] = chk.handle_push(A, FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], grad[sl].copy(), w, z, cg, t[sidx])
                    Wd[keys[sl]], Zd[keys[sl]], Cd[keys[sl]] = w, z, cg
                    keys, grad, cut = emsgs[r]
                    w, z, cg = We[keys[sl]].copy(), Ze[keys[sl]].copy(), Ce[keys[sl]].copy()
                    _, te[sidx] = chk.handle_push(A, FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], grad[sl].copy(), w, z, cg, te[sidx])
                    We[keys[sl]], Ze[keys[sl]], Ce[keys[sl]] = w, z, cg
            for keys, _, _ in msgs:
                touched.update(keys.tolist())
            if os.environ.get("WH_BENCH_DEBUG"):
                dd = np.abs(Wd.astype(np.float64) - We)
                sys.stderr.write("parity sim step %d: |W_ref - W_exact| max %g, count > 1e-5: %d\n" % (step, dd.max(), int((dd > 1e-5).sum())))
        allk = np.array(sorted(touched), np.uint64)
        cuts = np.searchsorted(allk, np.array(bounds, np.uint64))
        want_keys = [allk[cuts[r]:cuts[r + 1]] for r in range(world)]
        out = (want_keys, Wd, Zd, Cd, chk.kind, seen, We, Ze, Ce)
    if world > 1:
        box = [out[0] if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        my_keys = box[0][rank]
    else:
        my_keys = out[0][0]
    state = job.srv.get_state(my_keys)
    nnz = job.srv.nnz_w()
    if world > 1:
        got = [None] * world
        dist.gather_object((state, nnz), got if rank == 0 else None, dst=0)
    else:
        got = [(state, nnz)]
    job.close()
    if rank != 0:
        return None
    want_keys, Wd, Zd, Cd, kind, seen, We, Ze, Ce = out
    e_out = 0
    e_abs = e_rel = 0.0
    rel_l2 = []
    n = n_out = n_flip = 0
    max_abs = max_rel = 0.0
    ok = True
    for r in range(world):
        k = want_keys[r].astype(np.int64)
        (gw, gz, gc), _ = got[r]
        for a, b, ex in ((gw, Wd[k], We[k]), (gz, Zd[k], Ze[k]), (gc, Cd[k], Ce[k])):
            if algo == "sgd" and a is not gw:
                continue
            if algo == "adagrad" and a is gz:
                continue
            a = a.astype(np.float64); b = b.astype(np.float64); ex = ex.astype(np.float64)
            ee = np.abs(a - ex)
            emiss = ee > 1e-5 * np.abs(ex) + 1e-7
            e_out += int((emiss & ((a == 0) == (ex == 0))).sum())
            if a.size:
                e_abs = max(e_abs, float(ee.max())); e_rel = max(e_rel, float((ee / np.maximum(np.abs(ex), 1e-7))[~((a == 0) != (ex == 0))].max()))
            err = np.abs(a - b)
            miss = err > 1e-5 * np.abs(b) + 1e-7
            flip = miss & ((a == 0) != (b == 0))
            rest = miss & ~flip
            n += a.size; n_out += int(rest.sum()); n_flip += int(flip.sum())
            if a.size:
                max_abs = max(max_abs, float(err.max()))
                max_rel = max(max_rel, float((err / np.maximum(np.abs(b), 1e-7)).max()))
            if flip.any() and float(np.maximum(np.abs(a[flip]), np.abs(b[flip])).max()) > 1e-5:
                ok = False
            # Against the reference simulation a per-entry bar cannot hold once several workers push per step: the
            # reference adds each key's gradient in fp32 in row order (learn/base/spmv.h:103-117), exact to ~1e-5 of the
            # sum of |terms| only; that error enters w of the popular keys, through Xw every dual, and so every key's
            # next gradient. (The two CPU simulations differ from each other by exactly these amounts, with no GPU
            # involved.) Held to the bar in norm instead; the per-entry counts are reported.
            den = float((b ** 2).sum())
            if den > 0:
                d_dev = (float(((a - b) ** 2).sum()) / den) ** 0.5       # device vs reference simulation
                d_sim = (float(((ex - b) ** 2).sum()) / den) ** 0.5      # exact-sum simulation vs reference simulation (CPU only)
                rel_l2.append([round(d_dev, 9), round(d_sim, 9)])
                if d_dev > max(1e-5, 1.01 * d_sim):
                    ok = False
    if e_out > max(3, 1e-5 * n):     # (a handful of keys on the L1 threshold apart, the two must agree)
        ok = False
    nnz_dev = sum(g[1] for g in got)
    nnz_ref = int((Wd != 0).sum())
    if abs(nnz_dev - nnz_ref) > max(2, n_flip):
        ok = False
    return {"ranks": world, "minibatches_per_rank": steps, "rows": rows, "keys": int(sum(len(k) for k in want_keys)),
            "values_compared": n, "outside_1e-5_rel_1e-7_abs": n_out, "l1_flips": n_flip, "max_abs": max_abs,
            "max_rel": max_rel, "vs_exact_sum_simulation": {"outside_1e-5_rel_1e-7_abs": e_out, "max_abs": e_abs, "max_rel": e_rel},
            "rel_l2_vs_reference_per_field_[device,exact_sum_simulation]": rel_l2,
            "nnz_w": [int(nnz_dev), nnz_ref], "checker": kind, "ok": bool(ok)}


def side_config(ctx, rank, world, local, args, stream, name, algo, loss, shape, rows, max_key, nnz_row, steps, warmup, pool_n,
                max_delay=0):
    """A short device-resident measurement of one of BASELINE.json's other configurations (configs[2..3]); same timing
    rules as the headline, fewer steps. Returns its sub-record (rank 0) or None."""
    from wormhole_b200 import capi
    max_uniq = min(rows * nnz_row * (2 if shape == "kdd" else 1), max_key)
    job = Job(ctx, rank, world, local, algo, loss, max_key, args.depth, args.exchange, max_uniq, max_delay)
    _, dev, raw = make_pool(shape, rank, pool_n, rows, want_host=False)
    N = int(raw[0].offset[-1])
    U = len(np.unique(raw[0].index % np.uint64(max_key)))
    ms, _, _ = timed_region(job, dev, capi.WH_DEVICE, steps, warmup, stream, world)
    job.close()
    del dev
    import torch
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    pk, _ = peaks()
    r = U / rows
    bpe = N / rows * 8 + 12 + r * (8 + 2 * STATE_BYTES[algo])
    v = world * rows * steps / (ms * 1e-3)
    return {"workload": workload_name(shape, algo, loss, max_key, rows) + (", max_delay %d" % max_delay if max_delay else ""),
            "value": v, "unit": "examples/s", "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "rows_per_minibatch": rows, "nnz_per_minibatch": N,
            "unique_keys_per_row_r": round(r, 4), "pool_minibatches": pool_n,
            "step_roofline": {"bytes_per_example": bpe, "frac_of_hbm_peak": v * bpe / (world * pk["hbm_gbs"] * 1e9)}}


def difacto_config(ctx, rank, world, rows, steps, warmup):
    """BASELINE.json configs[4]: DiFacto, dim 16, on the Criteo shape (learn/difacto guide values). One
    GPU: wh_difacto_process_minibatch. Several: every rank is a worker and the shard of its key range, the minibatch in
    phases with the three exchanges over NCCL (dist.ShardedDifacto). The entry points are synchronous."""
    import torch
    import torch.distributed as dist
    from wormhole_b200 import capi
    from wormhole_b200.dist import ShardedDifacto, key_range
    max_key = MAX_KEY
    kb, ke = key_range(rank, world, max_key)
    d = capi.Difacto(ctx, l1=4.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=16, threshold=16, v_l2=1e-4, max_key=max_key,
                     key_begin=kb, key_end=ke)
    _, blks, raw = make_pool("criteo", 9 + rank, 3, rows, want_host=False)
    U = len(np.unique(raw[0].index % np.uint64(max_key)))
    if world > 1:
        sd = ShardedDifacto(ctx, d, max_key, device=torch.device("cuda", torch.cuda.current_device()))
        run = lambda b, dp: sd.process_minibatch(b, True, dp, capi.WH_DEVICE)
    else:
        run = lambda b, dp: d.process_minibatch(b, True, dp, capi.WH_DEVICE)
    for i in range(warmup):
        run(blks[i % 3], 0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    t0 = time.perf_counter()
    for i in range(steps):
        run(blks[(warmup + i) % 3], 1)
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ms = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # step roofline, SURVEY.md 8d config 5: 324 B/example of CSR + 416 B per unique key that carries an embedding
    # ((1+dim) floats pulled and pushed, entry {w, V, sqc_grad...} read and written) + 32 B per key without one
    keys = np.unique(raw[0].index % np.uint64(max_key))
    if world > 1:
        kt = torch.from_numpy(keys.astype(np.int64)).cuda()
        nk = torch.tensor([kt.numel()], dtype=torch.int64, device="cuda")
        dist.broadcast(nk, 0)
        if rank != 0:
            kt = torch.empty(int(nk.item()), dtype=torch.int64, device="cuda")
        dist.broadcast(kt, 0)
        keys = kt.cpu().numpy().astype(np.uint64)
    mine = keys[(keys >= np.uint64(kb)) & (keys < np.uint64(ke))]
    _, size, _, _ = d.get_state(mine)
    with_v = float((size > 1).sum())
    if world > 1:
        t = torch.tensor([with_v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        with_v = float(t.item())
    d.close()
    v = world * rows * steps / (ms * 1e-3)
    pk, _ = peaks()
    r, r_v = len(keys) / rows, with_v / rows
    bpe = 324 + 416 * r_v + 32 * (r - r_v)
    roof = {"bytes_per_example": bpe, "unique_keys_with_embedding_per_row": round(r_v, 4),
            "frac_of_hbm_peak": v * bpe / (world * pk["hbm_gbs"] * 1e9)}
    if rank != 0:
        return None
    return {"step_roofline": roof, "workload": "DiFacto dim 16 (threshold 16, lambda_l1 4, lr_eta .01) on the synthetic Criteo shape, minibatches resident in HBM; "
                        "the entry point is synchronous" + ("; %d key-range shards, count push / ZVPull / ZVPush as all-to-all-v over NCCL" % world if world > 1 else ""),
            "value": v, "unit": "examples/s", "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
            "rows_per_minibatch": rows, "unique_keys_per_row_r": round(U / rows, 4)}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from wormhole_b200 import capi, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != args.gpus and rank == 0:
        sys.stderr.write("bench.py: --gpus %d but WORLD_SIZE=%d; using WORLD_SIZE\n" % (args.gpus, world))

    # the bench line proper is configs[1] (ftrl / logit / criteo); --algo / --loss / --shape time another one instead
    rows, pool_n, shape, algo, loss = args.rows, args.pool, args.shape, args.algo, args.loss
    max_key, nnz_row = MAX_KEY, NNZ_PER_ROW
    if shape == "kdd":
        max_key, nnz_row = synth.KDD_VOCAB, 11
        rows = args.rows if args.rows != ROWS else 1000000
    default_cfg = (algo, loss, shape) == ("ftrl", "logit", "criteo") and rows == ROWS
    ctx = capi.Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    max_uniq = min(rows * nnz_row * (2 if shape == "kdd" else 1), max_key)
    job = Job(ctx, rank, world, local, algo, loss, max_key, args.depth, args.exchange, max_uniq, args.max_delay)

    # the minibatch pool: distinct minibatches, resident in HBM (device leg) and in pinned host memory (e2e leg).
    # pool_n x ~32 MB exceeds the 126 MB L2, so every step streams its ids from HBM.
    host, dev, raw = make_pool(shape, rank, pool_n, rows)
    N = int(raw[0].offset[-1])
    uniq_per_mb = len(np.unique(raw[0].index % np.uint64(max_key)))
    r = uniq_per_mb / rows
    torch.cuda.synchronize()

    sampler = ClockSampler(local)
    ms_dev, _, launches = timed_region(job, dev, capi.WH_DEVICE, args.steps, args.warmup, stream, world)
    ms_e2e, wall_e2e, _ = timed_region(job, host, capi.WH_HOST, args.steps, max(3, args.warmup // 2), stream, world)
    ms_e2e = max(ms_e2e, 1e3 * wall_e2e)  # host-side staging is part of the end-to-end time
    # A driver-sized run (20 steps of 0.1 ms) ends before nvidia-smi's 20 ms sampler has reported twice: keep the same
    # steps going, untimed, until the sampler has had half a second under this load
    t_more = time.perf_counter()
    while time.perf_counter() - t_more < 0.5:
        job.run(dev, capi.WH_DEVICE, 0, 50)
    sampler.stop()

    # the raw host-to-device ceiling of this rank's pinned pool (one cudaMemcpyAsync per minibatch array, no kernels)
    n_min = min(int(h.index.numel()) for h in host)     # (minibatches of the KDD shape differ in their entry counts)
    h2d_bytes = n_min * 8 + (rows + 1) * 8 + rows * 4
    scratch = [torch.empty(n_min, dtype=dev[0].index.dtype, device="cuda"), torch.empty_like(dev[0].offset), torch.empty_like(dev[0].label)]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(20):
        h = host[i % pool_n]
        scratch[0].copy_(h.index[:n_min], non_blocking=True); scratch[1].copy_(h.offset, non_blocking=True); scratch[2].copy_(h.label, non_blocking=True)
    torch.cuda.synchronize()
    h2d_gbs = 20 * h2d_bytes / (time.perf_counter() - t0) / 1e9
    if world > 1:
        t = torch.tensor([h2d_gbs], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        h2d_gbs = float(t.item())
    del scratch

    segs = profile_segments(ctx, job, dev, min(args.steps, 50))
    table_keys, census_ms, censuses = job.wk.table_keys() if job.sw is None else (np.zeros(0, np.uint32), 0.0, 0)
    hit = 0.0
    if len(table_keys):
        k = (raw[min(1, pool_n - 1)].index % np.uint64(max_key)).astype(np.uint32)
        hit = float(np.isin(k, table_keys).mean())
    job.close()
    del host, dev
    torch.cuda.empty_cache()

    # ---- legs outside the headline's timed regions -------------------------------------------------------------
    parity = None
    if not args.no_parity and loss != "square":
        parity = parity_leg(ctx, rank, world, local, args, max_key, min(rows, 100000), shape, algo, loss, max_uniq)
    others = {}
    if default_cfg and not args.no_others:
        K, Wm = max(5, min(args.steps, 40)), max(3, min(args.warmup, 5))
        for name, a2, l2 in (("adagrad_logit", "adagrad", "logit"), ("sgd_sqhinge", "sgd", "sqhinge")):
            others[name] = side_config(ctx, rank, world, local, args, stream, name, a2, l2, "criteo", ROWS, MAX_KEY, NNZ_PER_ROW, K, Wm, pool_n)
        if world > 1 and args.exchange == "p2p":
            # the reference's own default, max_concurrency = 2 (learn/linear/config.proto): a pull may overtake the worker's
            # previous push. Not the configuration parity is judged in (max_delay = 0, the headline); reported beside it.
            others["ftrl_logit_max_delay_1"] = side_config(ctx, rank, world, local, args, stream, "ftrl_logit_max_delay_1", "ftrl", "logit",
                                                           "criteo", ROWS, MAX_KEY, NNZ_PER_ROW, K, Wm, pool_n, max_delay=1)
        others["kdd_ftrl"] = side_config(ctx, rank, world, local, args, stream, "kdd_ftrl", "ftrl", "logit", "kdd", 1000000,
                                         synth.KDD_VOCAB, 11, max(8, K // 2), max(5, args.depth + 1), 3)   # (every lane warm)
        # (every rank takes part; a failure on one rank of several would leave the others in a collective, so it is
        #  only caught where there is a single rank)
        if world == 1:
            try:
                others["difacto_dim16"] = difacto_config(ctx, rank, world, ROWS, max(3, K // 4), 3)
            except Exception as e:  # noqa: BLE001
                others["difacto_dim16"] = {"unavailable": "failed: %r" % (e,)}
        else:
            others["difacto_dim16"] = difacto_config(ctx, rank, world, ROWS, max(3, K // 4), 3)
        ctx.set_stream(stream.cuda_stream)
    no_table = None
    if default_cfg and world == 1 and not args.no_others and len(table_keys):
        # the same device-resident leg with an empty table of frequent keys, in a process of its own (the switch is read
        # when the table is built)
        try:
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--steps", str(min(args.steps, 40)), "--warmup", "5",
                                  "--no-others", "--no-parity", "--no-cpu", "--pool", str(pool_n)], capture_output=True, text=True,
                                 timeout=300, env=dict(os.environ, WH_DX_TABLE="0", WH_HOT="0"))
            no_table = json.loads(out.stdout.strip().splitlines()[-1])["value"]
        except Exception:  # noqa: BLE001
            no_table = None

    if rank == 0:
        pk, pk_kind = peaks()
        words = max_key // 32
        ms_step = ms_dev / args.steps
        # the dominant kernel of the minibatch's dependent chain; "auc_tail" is the join on the AUC side stream (idle
        # time in the one-at-a-time pass, hidden behind the next minibatch otherwise). At N > 1 the one-at-a-time pass
        # also times the waits at the rank barriers inside its segments: those are reported, not ranked.
        chain = [x for x in segs if x[0] != "auc_tail" and (world == 1 or x[1] <= ms_step)]
        top = max(chain, key=lambda s: s[1]) if chain else ("none", 0.0)
        top_bytes = segment_bytes(top[0], N, rows, uniq_per_mb, words)
        achieved = top_bytes / (top[1] * 1e-3) / 1e9 if top[1] > 0 else 0.0
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
            kname = {"fused_direct": "fused_direct_kernel", "push": "push_sweep_kernel"}.get(top[0])
            traffic = tr.get(kname) if (default_cfg and world == 1) else None
        except Exception:
            pass
        bytes_per_ex = N / rows * 8 + 12 + r * (8 + 2 * STATE_BYTES[algo])
        value = world * rows * args.steps / (ms_dev * 1e-3)
        e2e = world * rows * args.steps / (ms_e2e * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": "examples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(shape, algo, loss, max_key, rows, N, r, pool_n, world, args),
            "e2e": {"value": e2e, "unit": "examples/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 72 + 16,
                    "h2d_ceiling_gbs": h2d_gbs, "h2d_achieved_gbs": e2e / world * h2d_bytes / rows / 1e9},
            "gpu_launches": int(launches),
            "clocks": dict(sampler.summary(), sampled_over="both timed regions + 0.5 s of the same steps"),
            "roofline": {"bound": "hbm", "kernel": top[0], "achieved": achieved, "peak": pk["hbm_gbs"],
                         "peak_kind": pk_kind, "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": traffic,
                         "kernel_ms": top[1], "kernel_share_of_step": top[1] / ms_step,
                         "kernel_algorithmic_bytes": top_bytes},
            "step_roofline": {"bytes_per_example": bytes_per_ex,
                              "frac_of_hbm_peak": value * bytes_per_ex / (world * pk["hbm_gbs"] * 1e9)},
            "kernel_ms": {n: round(ms, 5) for n, ms in segs if n != "auc_tail"},
            # (not a kernel of the chain: how long the one-at-a-time pass idles at the join with the AUC side stream; with
            #  several minibatches in flight the next minibatch runs in that time)
            "side_stream_join_ms": {n: round(ms, 5) for n, ms in segs if n == "auc_tail"},
            "kernel_ms_note": "one minibatch at a time (depth 1), CUDA events on the launching stream" +
                              ("; at N > 1 these segments include the waits at the rank barriers" if world > 1 else ""),
            "hot": {"hit_rate": round(hit, 4), "keys_placed": int(len(table_keys)), "census_ms": round(census_ms, 3),
                    "censuses": censuses, "value_with_empty_table": no_table},
            "parity_check": parity,
            "other_configs": others if others else None,
        }
        # CPU baseline on a bounded sample, rank 0 at N=1 only (subprocess: no fork after CUDA init)
        line["cpu_baseline"] = None
        if world == 1 and default_cfg and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg(rows)
        print(json.dumps(line))
        if parity is not None and not parity["ok"]:
            sys.stderr.write("bench.py: parity_check FAILED: %r\n" % (parity,))
    ctx.close()
    failed = parity is not None and rank == 0 and not parity["ok"]
    if world > 1:
        flag = torch.tensor([1 if failed else 0], device="cuda")
        dist.all_reduce(flag)
        failed = bool(flag.item())
        dist.destroy_process_group()
    if failed:
        raise SystemExit(3)


def bench_config(shape, algo, loss, max_key, rows, N, r, pool_n, world, args):
    return {"workload": workload_name(shape, algo, loss, max_key, rows), "rows_per_minibatch": rows, "nnz_per_minibatch": N,
            "max_key": max_key, "unique_keys_per_row_r": round(r, 4), "pool_minibatches": pool_n,
            "l2_policy": "inputs larger than L2 (pool of %d distinct minibatches, %.0f MB)" % (pool_n, pool_n * N * 8 / 1e6),
            "parallelism": "key-range shards x%d" % world + ("" if world == 1 else " (" + args.exchange + ")"), "updater": FTRL,
            "minibatches_in_flight": args.depth if (world == 1 or args.exchange == "p2p") else 1,
            "max_delay": args.max_delay if (world == 1 or args.exchange == "p2p") else 0}


def cpu_baseline_leg(rows):
    """The reference's CPU path on this box's host cores, bounded samples: one worker (nt = 2, the reference's
    hard-coded thread count) + one server, and as many worker + server pairs as the cores allow."""
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    res = {}
    for tag, procs, steps in (("nt2", "1", "4"), ("all_cores", "0", "3")):
        try:
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", steps, "--warmup", "1",
                                  "--cpu-procs", procs, "--rows", str(rows)], capture_output=True, text=True, timeout=600, env=env)
            res[tag] = json.loads(out.stdout.strip().splitlines()[-1])["cpu_baseline"]
        except Exception as e:  # noqa: BLE001
            res[tag] = {"value": None, "unit": "examples/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (e,)}
    cb = dict(res["nt2"])
    cb["sample"] = "1 worker (nt=2) + 1 server, 4 minibatches of %d rows after 1 warm-up" % rows
    cb["all_cores"] = res["all_cores"]
    try:
        model = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")]
        cb["cpu_model"], cb["nproc"] = (model[0] if model else "unknown"), os.cpu_count()
    except Exception:
        pass
    return cb


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS)
    ap.add_argument("--pool", type=int, default=8, help="distinct minibatches kept resident")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="N>1: how pull/push reach the owning GPU")
    ap.add_argument("--max-delay", type=int, default=0, help="bounded staleness of pulls (0 = sequential semantics)")
    ap.add_argument("--depth", type=int, default=4, help="minibatches in flight (conf max_concurrency)")
    ap.add_argument("--algo", default="ftrl", choices=["ftrl", "adagrad", "sgd"], help="updater (the bench line is ftrl)")
    ap.add_argument("--loss", default="logit", choices=["logit", "sqhinge", "square"])
    ap.add_argument("--shape", default="criteo", choices=["criteo", "kdd"], help="kdd: ~11 nnz/row, 54.7 M ids, 1 M-row minibatches")
    ap.add_argument("--cpu-procs", type=int, default=0, help="reference arm: worker+server pairs (0 = cores/2)")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity_check leg (checker vs device, outside the timed regions)")
    ap.add_argument("--no-others", action="store_true", help="skip the other_configs sub-records and the empty-table comparison")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
