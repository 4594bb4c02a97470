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
        "config": {"workload": WORKLOAD, "rows_per_minibatch": rows, "max_key": MAX_KEY},
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
    }.get(name, 0)


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

    rows, pool_n = args.rows, args.pool
    # parity-test configurations of BASELINE.json (configs[2..3]) can be timed with --algo / --loss / --shape; the
    # bench line proper is the default (configs[1])
    algo = {"ftrl": capi.ALGO_FTRL, "adagrad": capi.ALGO_ADAGRAD, "sgd": capi.ALGO_SGD}[args.algo]
    loss = {"logit": capi.LOSS_LOGIT, "sqhinge": capi.LOSS_SQUARE_HINGE}[args.loss]
    global MAX_KEY, WORKLOAD, NNZ_PER_ROW
    if args.shape == "kdd":
        MAX_KEY, NNZ_PER_ROW = synth.KDD_VOCAB, 11
        rows = args.rows if args.rows != ROWS else 1000000
    if (args.algo, args.loss, args.shape) != ("ftrl", "logit", "criteo"):
        WORKLOAD = "synthetic %s-shape sparse LR, %s, %s loss, max_key %d, minibatch %d" % (args.shape, args.algo, args.loss, MAX_KEY, rows)
    ctx = capi.Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    if world == 1:
        # worker and the one server share the GPU: the lane pipeline (minibatches in flight)
        srv = capi.Server(ctx, algo, FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], 0, MAX_KEY)
        wk = capi.Worker(ctx, srv, loss, MAX_KEY, depth=args.depth)
        sw = None
    else:
        # rank g = worker g + server g of the key range [g*max_key/S, (g+1)*max_key/S). Default: every rank maps
        # the other ranks' shards over NVLink (CUDA IPC); pull gathers from the owning GPU, push runs in S turns
        # with stream-ordered barriers (server.cu). --exchange nccl: all-to-all-v of (key, weight) / (key, grad)
        # messages through torch.distributed instead (wormhole_b200/dist.py). Weak scaling.
        from wormhole_b200.dist import DeviceShard, ShardedKV, ShardedWorker, key_range
        kb, ke = key_range(rank, world, MAX_KEY)
        srv = capi.Server(ctx, algo, FTRL["l1"], FTRL["l2"], FTRL["eta"], FTRL["beta"], kb, ke)
        if args.exchange == "p2p":
            handles = [None] * world
            dist.all_gather_object(handles, srv.ipc_export())
            srv.ipc_attach(rank, world, MAX_KEY, handles)
            dist.barrier()
            # lanes still apply: every server interaction of a rank is chained through one CUDA event, so the
            # barriers keep one global order; localize / staging / AUC of the next minibatch overlap
            wk = capi.Worker(ctx, srv, loss, MAX_KEY, depth=args.depth)
            boxes = [None] * world
            dist.all_gather_object(boxes, wk.ipc_export(min(rows * NNZ_PER_ROW * (2 if args.shape == "kdd" else 1), MAX_KEY)))
            wk.ipc_attach(boxes)
            dist.barrier()
            sw = None
        else:
            wk = capi.Worker(ctx, None, loss, MAX_KEY)
            sw = ShardedWorker(ctx, wk, ShardedKV(DeviceShard(srv), MAX_KEY, device=torch.device("cuda", local)))

    # the minibatch pool: distinct minibatches, resident in HBM (device leg) and in pinned host
    # memory (e2e leg). pool_n x ~32 MB exceeds the 126 MB L2, so every step streams from HBM.
    host, dev = [], []
    for m in range(pool_n):
        b = synth.criteo_minibatch(rank * 100000 + m, rows=rows) if args.shape == "criteo" else synth.kdd_minibatch(rank * 100000 + m, rows=rows)
        h = type("B", (), {})()
        h.rows = b.rows
        h.nnz = b.nnz
        h.offset = torch.from_numpy(b.offset.view(np.int64)).pin_memory()
        h.index = torch.from_numpy(b.index.view(np.int64)).pin_memory()
        h.label = torch.from_numpy(b.label).pin_memory()
        h.value = None
        d = type("B", (), {})()
        d.rows, d.nnz = b.rows, b.nnz
        d.offset, d.index, d.label, d.value = h.offset.cuda(), h.index.cuda(), h.label.cuda(), None
        host.append(h); dev.append(d)
    N = int(host[0].offset[-1])
    uniq_per_mb = len(np.unique(host[0].index.numpy().view(np.uint64) % np.uint64(MAX_KEY)))
    r = uniq_per_mb / rows
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(blocks, mem, first, count):
        if sw is not None:
            for i in range(count):
                sw.process_minibatch(blocks[(first + i) % pool_n], True, mem, blocks[(first + i) % pool_n].nnz)
            return
        # up to `depth` minibatches in flight (conf max_concurrency); every one is waited for,
        # so its progress numbers have reached the host before the clock stops
        pending = []
        for i in range(count):
            if len(pending) == args.depth:
                wk.wait(pending.pop(0))
            b = blocks[(first + i) % pool_n]
            pending.append(wk.submit(b, True, mem, b.nnz))
        for t in pending:
            wk.wait(t)

    def timed(blocks, mem, steps, warmup):
        run(blocks, mem, 0, warmup)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = capi.kernel_launches()
        t0 = time.perf_counter()
        e0.record(stream)
        run(blocks, mem, warmup, steps)
        e1.record(stream)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall, capi.kernel_launches() - l0

    sampler = ClockSampler(local)
    if args.max_delay and (world == 1 or args.exchange == "p2p"):
        wk.set_max_delay(args.max_delay)
    ms_dev, _, launches = timed(dev, capi.WH_DEVICE, args.steps, args.warmup)
    ms_e2e, wall_e2e, _ = timed(host, capi.WH_HOST, args.steps, max(3, args.warmup // 2))
    ms_e2e = max(ms_e2e, 1e3 * wall_e2e)  # host-side staging is part of the end-to-end time
    sampler.stop()

    # per-kernel timing (CUDA events on the launching stream), separate pass
    lib = capi.lib()
    import ctypes as C
    lib.wh_profile_enable(ctx.h, 1)
    for i in range(min(args.steps, 50)):
        if sw is not None:
            sw.process_minibatch(dev[i % pool_n], True, capi.WH_DEVICE, dev[i % pool_n].nnz)
        else:
            wk.process_minibatch(dev[i % pool_n], True, capi.WH_DEVICE)
    segs = []
    for i in range(lib.wh_profile_count(ctx.h)):
        name, ms, cnt = C.c_char_p(), C.c_double(), C.c_uint64()
        lib.wh_profile_get(ctx.h, i, C.byref(name), C.byref(ms), C.byref(cnt))
        segs.append((name.value.decode(), ms.value / max(1, cnt.value)))
    lib.wh_profile_enable(ctx.h, 0)

    if rank == 0:
        pk, pk_kind = peaks()
        words = MAX_KEY // 32
        total_seg = sum(ms for _, ms in segs) or 1.0
        # the dominant kernel of the minibatch's dependent chain; "auc_tail" is the join on the AUC side
        # stream (idle time in this one-at-a-time pass, hidden behind the next minibatch otherwise)
        chain = [x for x in segs if x[0] != "auc_tail"]
        top = max(chain, key=lambda s: s[1]) if chain else ("none", 0.0)
        top_bytes = segment_bytes(top[0], N, rows, uniq_per_mb, words)
        achieved = top_bytes / (top[1] * 1e-3) / 1e9 if top[1] > 0 else 0.0
        # DRAM bytes per launch of that kernel from the committed ncu --set full capture (profiles/), or null
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_d_traffic.json")))
            kname = {"spmv_fused": "spmv_fused_hot_kernel", "push": "server_push_kernel<1>", "loc_mark": "loc_mark_hot_plain_kernel<1>",
                     "pull": "server_pull_kernel",
                     "loc_dir_scan": "scan_coop_kernel<4, DirPopc, DirEmit, HotRankEpilogue>"}.get(top[0])
            default_cfg = rows == ROWS and world == 1 and (args.algo, args.loss, args.shape) == ("ftrl", "logit", "criteo")
            traffic = tr.get(kname) if default_cfg else None
        except Exception:
            pass
        bytes_per_ex = NNZ_PER_ROW * 8 + 12 + r * (8 + 2 * 12)
        value = world * rows * args.steps / (ms_dev * 1e-3)
        e2e = world * rows * args.steps / (ms_e2e * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": "examples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_minibatch": rows, "nnz_per_minibatch": N,
                       "max_key": MAX_KEY, "unique_keys_per_row_r": round(r, 4), "pool_minibatches": pool_n,
                       "l2_policy": "inputs larger than L2 (pool of %d distinct minibatches, %.0f MB)" % (pool_n, pool_n * N * 8 / 1e6),
                       "parallelism": "key-range shards x%d" % world + ("" if world == 1 else " (" + args.exchange + ")"), "updater": FTRL,
                       "minibatches_in_flight": args.depth if (world == 1 or args.exchange == "p2p") else 1,
                       "max_delay": args.max_delay if (world == 1 or args.exchange == "p2p") else 0},
            "e2e": {"value": e2e, "unit": "examples/s", "h2d_bytes_per_step": N * 8 + (rows + 1) * 8 + rows * 4,
                    "d2h_bytes_per_step": 72 + 16},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "roofline": {"bound": "hbm", "kernel": top[0], "achieved": achieved, "peak": pk["hbm_gbs"],
                         "peak_kind": pk_kind, "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": traffic,
                         "kernel_ms": top[1], "kernel_share_of_step": top[1] / total_seg,
                         "kernel_algorithmic_bytes": top_bytes},
            "step_roofline": {"bytes_per_example": bytes_per_ex,
                              "frac_of_hbm_peak": value * bytes_per_ex / (world * pk["hbm_gbs"] * 1e9)},
            "kernel_ms": {n: round(ms, 5) for n, ms in segs},
        }
        # CPU baseline on a bounded sample, rank 0 at N=1 only (subprocess: no fork after CUDA init)
        if world > 1:
            line["cpu_baseline"] = None
            print(json.dumps(line))
        try:
            if world > 1:
                raise StopIteration
            if (args.algo, args.loss, args.shape) != ("ftrl", "logit", "criteo"):
                line["cpu_baseline"] = None     # the CPU arm times the bench configuration only
                raise StopIteration
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "4",
                                  "--warmup", "1", "--cpu-procs", "1", "--rows", str(rows)],
                                 capture_output=True, text=True, timeout=600,
                                 env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")})
            cb = json.loads(out.stdout.strip().splitlines()[-1])["cpu_baseline"]
            cb["sample"] = "1 worker (nt=2) + 1 server, 4 minibatches of %d rows after 1 warm-up" % rows
            line["cpu_baseline"] = cb
        except StopIteration:
            pass
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": "examples/s", "cores": 0, "kind": "port",
                                    "sample": "failed: %r" % (e,)}
        if world == 1:
            print(json.dumps(line))
    wk.close(); srv.close(); ctx.close()
    if world > 1:
        dist.destroy_process_group()


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
    ap.add_argument("--loss", default="logit", choices=["logit", "sqhinge"])
    ap.add_argument("--shape", default="criteo", choices=["criteo", "kdd"], help="kdd: ~11 nnz/row, 54.7 M ids, 1 M-row minibatches")
    ap.add_argument("--cpu-procs", type=int, default=0, help="reference arm: worker+server pairs (0 = cores/2)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
