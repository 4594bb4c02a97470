"""Quick look at one GPU: per-segment kernel times (CUDA events, one minibatch at a time) and the throughput with
3 minibatches in flight. Not a bench line."""
import ctypes as C, sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth
ctx = capi.Context(0); srv = capi.Server(ctx, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, 2 ** 24)
depth = 3
wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, 2 ** 24, depth=depth)
blks = []
for m in range(6):
    b = synth.criteo_minibatch(m)
    d = type("B", (), {})(); d.rows = b.rows
    d.offset = torch.from_numpy(b.offset.view(np.int64)).cuda(); d.index = torch.from_numpy(b.index.view(np.int64)).cuda()
    d.label = torch.from_numpy(b.label).cuda(); d.value = None
    blks.append(d)
N = 3900000
for i in range(6): wk.process_minibatch(blks[i % 6], True, capi.WH_DEVICE) if False else wk.wait(wk.submit(blks[i % 6], True, capi.WH_DEVICE, N))
lib = capi.lib()
lib.wh_profile_enable(ctx.h, 1)
for i in range(30): wk.wait(wk.submit(blks[i % 6], True, capi.WH_DEVICE, N))
tot = 0
for i in range(lib.wh_profile_count(ctx.h)):
    name, ms, cnt = C.c_char_p(), C.c_double(), C.c_uint64()
    lib.wh_profile_get(ctx.h, i, C.byref(name), C.byref(ms), C.byref(cnt))
    us = 1e3 * ms.value / max(1, cnt.value); tot += us
    print("%-14s %8.1f us" % (name.value.decode(), us))
print("%-14s %8.1f us" % ("sum", tot))
lib.wh_profile_enable(ctx.h, 0)
def run(count):
    pend = []
    for i in range(count):
        if len(pend) == depth: wk.wait(pend.pop(0))
        pend.append(wk.submit(blks[i % 6], True, capi.WH_DEVICE, N))
    for t in pend: prog = wk.wait(t)
    return prog
run(20); ctx.sync()
t0 = time.perf_counter(); prog = run(300); ctx.sync(); dt = (time.perf_counter() - t0) / 300
print("depth %d: %.1f us/step  %.1f M ex/s   last prog %s" % (depth, dt * 1e6, 0.1 / dt, np.round(prog, 4)))
