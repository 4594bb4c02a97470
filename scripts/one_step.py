"""A handful of minibatches through the worker (for ncu captures)."""
import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth
ctx = capi.Context(0); srv = capi.Server(ctx, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, 2 ** 24)
wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, 2 ** 24, depth=1)
blks = []
for m in range(2):
    b = synth.criteo_minibatch(m)
    d = type("B", (), {})(); d.rows = b.rows
    d.offset = torch.from_numpy(b.offset.view(np.int64)).cuda(); d.index = torch.from_numpy(b.index.view(np.int64)).cuda()
    d.label = torch.from_numpy(b.label).cuda(); d.value = None
    blks.append(d)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    wk.wait(wk.submit(blks[i % 2], True, capi.WH_DEVICE, 3900000))
print("ok")
