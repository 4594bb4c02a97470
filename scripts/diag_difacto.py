import ctypes, sys
import numpy as np
sys.path.insert(0, ".")
from oracle import oracle as O
from wormhole_b200 import capi, synth
chk = O.best(); gpu = capi.Context(0)
max_key, seed = 2 ** 24, 5
kw = dict(l1=4.0, l2=0.0, eta=.01, beta=1.0, l1_shrk=False, dim=16, threshold=16, v_l2=1e-4, init_scale=.01)
libc = ctypes.CDLL("libc.so.6"); libc.srand(ctypes.c_uint(seed))
stream = np.array([libc.rand() for _ in range(1 << 20)], np.int32)
d = capi.Difacto(gpu, max_key=max_key, **kw)
d.set_init_stream(stream)
job = chk.difacto_create(max_key=max_key, seed=seed, **kw)
for m in range(2):
    blk = synth.criteo_minibatch(m)
    pg, pc = d.process_minibatch(blk, True, 0), chk.difacto_minibatch(job, blk, True, 0)
    print("prog gpu", pg); print("prog ref", pc)
    keys, cnt, size, w, cg = chk.difacto_model(job, 16)
    gcnt, gsize, gw, gcg = d.get_state(keys)
    w0r, w0g, zr, zg = w[:, 0], gw[:, 0], cg[:, 1], gcg[:, 1]
    flip = (w0r != 0) != (w0g != 0)
    print("keys", len(keys), "flips", flip.sum(), "max |z|-4 among flips", np.abs(np.abs(zr[flip]) - 4).max() if flip.any() else 0,
          "max |dz| all", np.abs(zr - zg).max(), "max |dw0|", np.abs(w0r - w0g).max(), "max|dV|", np.abs(w[:, 1:] - gw[:, 1:]).max(),
          "max |dcg0| rel", (np.abs(cg[:, 0] - gcg[:, 0]) / np.maximum(cg[:, 0], 1)).max())
    hv = size == 17
    dV = np.abs(w[hv, 1:] - gw[hv, 1:])
    print("V keys", hv.sum(), "frac entries |dV|>0.005:", (dV > 0.005).mean(), "rows affected:", (dV.max(axis=1) > 0.005).sum(),
          "first bad row idx:", np.where(dV.max(axis=1) > 0.005)[0][:5], "cnt of those", cnt[hv][np.where(dV.max(axis=1) > 0.005)[0][:5]])
    bad = np.where(dV.max(axis=1) > 0.005)[0][:3]
    for b in bad:
        print(" ref V", w[hv][b, 1:5], "gpu V", gw[hv][b, 1:5], "ref cgV", cg[hv][b, 2:6], "gpu cgV", gcg[hv][b, 2:6])
