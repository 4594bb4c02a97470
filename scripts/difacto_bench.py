"""One-GPU timing of BASELINE.json configs[4]: DiFacto dim 16 on the Criteo shape (not a bench line)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth
ctx = capi.Context(0)
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
d = capi.Difacto(ctx, l1=1.0, l2=0.1, eta=0.1, beta=1.0, dim=16, threshold=2, v_l2=0.01, v_eta=0.01, max_key=2 ** 24)
blks = []
for m in range(4):
    b = synth.criteo_minibatch(m, rows=rows)
    x = type("B", (), {})(); x.rows = b.rows
    x.offset = torch.from_numpy(b.offset.view(np.int64)).cuda(); x.index = torch.from_numpy(b.index.view(np.int64)).cuda()
    x.label = torch.from_numpy(b.label).cuda(); x.value = None
    blks.append(x)
for i in range(4): d.process_minibatch(blks[i % 4], True, 0, capi.WH_DEVICE)
ctx.sync(); t0 = time.perf_counter()
n = 40
for i in range(n): p = d.process_minibatch(blks[i % 4], True, 0, capi.WH_DEVICE)
ctx.sync(); dt = (time.perf_counter() - t0) / n
print("difacto dim16: %.3f ms/step  %.1f M ex/s  prog %s" % (dt * 1e3, rows / dt / 1e6, np.round(p, 3)))
