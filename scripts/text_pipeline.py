"""Text -> GPU reader -> training step with the reading overlapped: reader threads (own context / stream each) parse
pinned text chunks into device row containers while the main thread trains on the previous ones. Prints examples/s
end to end from host TEXT, next to bench.py's e2e from host RowBlocks (not a bench line)."""
import queue, sys, threading, time
import numpy as np
sys.path.insert(0, ".")
import torch
from wormhole_b200 import capi, synth

rows, nread, depth = 100000, int(sys.argv[1]) if len(sys.argv) > 1 else 2, 4
texts = [synth.criteo_text_minibatch(m, rows) for m in range(4)]
pin = [torch.frombuffer(bytearray(t), dtype=torch.uint8).pin_memory() for t in texts]
ctx = capi.Context(0)
srv = capi.Server(ctx, capi.ALGO_FTRL, 1.0, 0.1, 0.1, 1.0, 0, 2 ** 24)
wk = capi.Worker(ctx, srv, capi.LOSS_LOGIT, 2 ** 24, depth=depth)
steps = 200
ready = queue.Queue(maxsize=depth + 2 * nread)
free = [queue.Queue() for _ in range(nread)]

def reader(r):
    rc = capi.Context(0)
    s = torch.cuda.Stream()
    rc.set_stream(s.cuda_stream)
    ring = [capi.Rows(rc) for _ in range(depth + 3)]
    for x in ring: free[r].put(x)
    for i in range(r, steps, nread):
        rowsobj = free[r].get()
        capi.parse_text(rc, "criteo", pin[i % 4].numpy(), into=rowsobj)     # H2D + parse + sync on this thread's stream
        ready.put((i, r, rowsobj))
    ready.put(None)

threads = [threading.Thread(target=reader, args=(r,)) for r in range(nread)]
t0 = time.perf_counter()
for t in threads: t.start()
pend, done, got = [], 0, 0
while done < nread:
    item = ready.get()
    if item is None:
        done += 1; continue
    i, r, ro = item
    if len(pend) == depth:
        tk, rr, old = pend.pop(0); wk.wait(tk); free[rr].put(old)
    pend.append((wk.submit(ro, True, capi.WH_DEVICE, ro.nnz), r, ro)); got += 1
for tk, rr, old in pend: wk.wait(tk)
dt = time.perf_counter() - t0
print("text bytes/row %.0f; %d reader threads: %.3f ms/step  %.1f M ex/s end to end from host text" % (len(texts[0]) / rows, nread, dt / got * 1e3, rows * got / dt / 1e6))
