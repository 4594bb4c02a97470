"""Throughput of the GPU reader on Criteo-shape text: parse alone (text resident in HBM), parse from pinned host
text, and text -> parse -> training step, next to the parsed-block e2e of bench.py."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth

rows = 100000
texts = [synth.criteo_text_minibatch(m, rows) for m in range(4)]
print("text bytes/row %.1f" % (len(texts[0]) / rows))
ctx = capi.Context(0)
dev = [torch.frombuffer(bytearray(t), dtype=torch.uint8).cuda() for t in texts]
pin = [torch.frombuffer(bytearray(t), dtype=torch.uint8).pin_memory() for t in texts]
torch.cuda.synchronize()

def timeit(fn, n=20):
    fn(0); ctx.sync()
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    ctx.sync()
    return (time.perf_counter() - t0) / n

R = capi.Rows(ctx)
def parse_dev(i):
    capi.parse_text(ctx, "criteo", dev[i % 4].data_ptr(), mem=capi.WH_DEVICE, length=len(texts[i % 4]), into=R)
def parse_pin(i):
    capi.parse_text(ctx, "criteo", pin[i % 4].numpy(), into=R)
t = timeit(parse_dev); print("parse, text in HBM   : %.3f ms  %.1f M rows/s  %.1f GB/s" % (t * 1e3, rows / t / 1e6, len(texts[0]) / t / 1e9))
t = timeit(parse_pin); print("parse, text in pinned: %.3f ms  %.1f M rows/s" % (t * 1e3, rows / t / 1e6))

srv = capi.Server(ctx, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, 2 ** 24)
wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, 2 ** 24, depth=3)
pend = []
ring = [capi.Rows(ctx) for _ in range(4)]
def step(i):
    if len(pend) == 3:
        wk.wait(pend.pop(0))
    r = capi.parse_text(ctx, "criteo", pin[i % 4].numpy(), into=ring[i % 4])
    pend.append(wk.submit(r, True, capi.WH_DEVICE, r.nnz))
t = timeit(step, 40)
for tk in pend: wk.wait(tk)
print("text -> parse -> step: %.3f ms  %.1f M ex/s" % (t * 1e3, rows / t / 1e6))
