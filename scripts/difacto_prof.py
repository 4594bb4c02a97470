"""DiFacto dim 16 at BASELINE.json configs[4]'s settings (lambda_l1 4, threshold 16, lr_eta .01, embedding lambda_l2 1e-4) on the
synthetic Criteo shape, a few synchronous minibatches: the command the DiFacto ncu launch list is taken over."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth
ctx = capi.Context(0)
rows, steps = 100000, int(sys.argv[1]) if len(sys.argv) > 1 else 6
d = capi.Difacto(ctx, l1=4.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=True, dim=16, threshold=16, v_l2=1e-4, max_key=2 ** 24)
blks = []
for m in range(3):
    b = synth.criteo_minibatch(9000 + m, rows=rows)
    x = type("B", (), {})(); x.rows = b.rows
    x.offset = torch.from_numpy(b.offset.view(np.int64)).cuda(); x.index = torch.from_numpy(b.index.view(np.int64)).cuda()
    x.label = torch.from_numpy(b.label).cuda(); x.value = None
    blks.append(x)
for i in range(3): d.process_minibatch(blks[i % 3], True, 0, capi.WH_DEVICE)
ctx.sync(); t0 = time.perf_counter()
for i in range(steps): p = d.process_minibatch(blks[i % 3], True, 1, capi.WH_DEVICE)
ctx.sync(); dt = (time.perf_counter() - t0) / steps
print("difacto dim16: %.3f ms/step  %.1f M ex/s  prog %s" % (dt * 1e3, rows / dt / 1e6, np.round(p, 3)))
