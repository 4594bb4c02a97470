import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth
ctx = capi.Context(0); srv = capi.Server(ctx, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, 2 ** 24)
depth = 3
wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, 2 ** 24, depth=depth)
blks = []
for m in range(4):
    b = synth.criteo_minibatch(m)
    d = type("B", (), {})(); d.rows = b.rows
    d.offset = torch.from_numpy(b.offset.view(np.int64)).cuda(); d.index = torch.from_numpy(b.index.view(np.int64)).cuda()
    d.label = torch.from_numpy(b.label).cuda(); d.value = None
    blks.append(d)
N = 3900000
torch.cuda.synchronize()
def run(count):
    pend = []; t_sub = 0.0; t_wait = 0.0
    for i in range(count):
        if len(pend) == depth:
            t0 = time.perf_counter(); wk.wait(pend.pop(0)); t_wait += time.perf_counter() - t0
        t0 = time.perf_counter(); pend.append(wk.submit(blks[i % 4], True, capi.WH_DEVICE, N)); t_sub += time.perf_counter() - t0
    for t in pend:
        t0 = time.perf_counter(); wk.wait(t); t_wait += time.perf_counter() - t0
    return t_sub, t_wait
run(10)
t0 = time.perf_counter(); ts, tw = run(300); tot = time.perf_counter() - t0
print("per step: total %.1f us, in submit %.1f us, in wait %.1f us" % (tot / 300 * 1e6, ts / 300 * 1e6, tw / 300 * 1e6))
