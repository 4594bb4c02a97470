"""DiFacto's exchange on CPU: wormhole_b200.dist.ShardedDifacto (count push, ZVPull, ZVPush as all-to-all-v of sorted key
slices; one message per source rank, ascending) over gloo with a numpy stand-in for each rank's worker half and shard.
The stand-in restates learn/difacto/async_sgd.h:216-296 (Push / Pull / Resize / UpdateW / UpdateV) and uses the oracle
port for Localize and difacto::Loss; with one rank it must reproduce the port's own DiFacto job bit for bit (which
pins the stand-in), with two ranks the single-process replay of the same message schedule."""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O
from tests.util import random_block
from wormhole_b200.dist import ShardedDifacto, key_range

MAX_KEY = 2 ** 13
DIM, THR = 4, 3
L1, L2, ALPHA, BETA, V_L2, V_ALPHA, V_BETA, SCALE = 0.3, 0.01, 0.05, 1.0, 1e-4, 0.02, 1.0, 0.01
STEPS, PASSES = 4, 2
f32 = np.float32


def minibatch(rank, step):
    rng = np.random.default_rng(4000 + 100 * rank + step)
    return random_block(rng, 800, 12, 8000, with_value=(step % 2 == 1), empty_rows=False)


class CpuDifacto:
    """Worker half (oracle port) + the shard of [begin, end) as dense numpy arrays. `init`: "rand" draws new embeddings
    from glibc rand() in message order like the reference (async_sgd.h:254), "hash" from the key (rank-independent)."""

    def __init__(self, port, begin, end, init):
        self.port, self.kb, self.init, self.dim = port, begin, init, DIM
        n = end - begin
        self.cnt = np.zeros(n, np.uint32)
        self.w0, self.z0, self.cg0 = (np.zeros(n, f32) for _ in range(3))
        self.hasv = np.zeros(n, bool)
        self.V, self.cgV = np.zeros((n, DIM), f32), np.zeros((n, DIM), f32)
        self.new_w = self.new_V = 0
        self.libc = ctypes.CDLL("libc.so.6")

    # ---- worker half
    def begin(self, blk, mem=0):
        self.D = self.port.localize(blk, MAX_KEY)
        return torch.from_numpy(self.D.uniq.astype(np.int64)), torch.from_numpy(self.D.cnt.astype(np.int32))

    def finish(self, w, has, V, train):
        w, has, V = w.numpy(), has.numpy().astype(bool), V.numpy()
        siz = np.where(has, DIM + 1, 1).astype(np.int32)
        pos = np.cumsum(siz) - siz
        model = np.zeros(int(siz.sum()), f32)
        model[pos] = w
        for j in range(DIM):
            model[pos[has] + 1 + j] = V[has, j]
        prog, grad = self.port.difacto_loss(self.D, model, siz, DIM, 1, 0.0, 0.0, want_grad=train)
        gw, G = np.zeros(len(w), f32), np.zeros((len(w), DIM), f32)
        if train:
            gw = grad[pos].copy()
            for j in range(DIM):
                G[has, j] = grad[pos[has] + 1 + j]
        return prog, torch.from_numpy(gw), torch.from_numpy(G)

    # ---- the shard: one message each
    def _resize(self, k):
        """Resize, async_sgd.h:247-259, for the slots k (ascending): those that must grow an embedding now."""
        grow = k[(self.cnt[k] > THR) & ~self.hasv[k] & (self.w0[k] != 0)]      # l1_shrk = true
        for s in grow:
            for j in range(DIM):
                if self.init == "rand":
                    r = self.libc.rand()
                else:
                    r = int((int(s + self.kb) * 2654435761 + j * 40503) % 2147483647)
                self.V[s, j] = f32(r) / f32(2147483647) * f32(SCALE - (-SCALE)) + f32(-SCALE)
                self.cgV[s, j] = 0
            self.hasv[s] = True
            self.new_V += DIM

    def count_push(self, keys, cnt):
        k = keys.numpy().astype(np.int64) - self.kb
        assert ((k >= 0) & (k < len(self.cnt))).all(), "key outside this shard's range"
        self.cnt[k] += cnt.numpy().astype(np.uint32)
        self._resize(k)

    def pull(self, keys):
        k = keys.numpy().astype(np.int64) - self.kb
        assert ((k >= 0) & (k < len(self.cnt))).all(), "key outside this shard's range"
        has = self.hasv[k] & (self.w0[k] != 0)
        return torch.from_numpy(self.w0[k].copy()), torch.from_numpy(has.astype(np.uint8)), torch.from_numpy(self.V[k].copy())

    def push(self, keys, gw, G, has):
        k = keys.numpy().astype(np.int64) - self.kb
        gw, G, has = gw.numpy(), G.numpy(), has.numpy().astype(bool)
        w = self.w0[k].copy()
        g = gw + f32(L2) * w                                              # UpdateW, :262-286
        cg = self.cg0[k]
        cg_new = np.sqrt(cg * cg + g * g, dtype=f32)
        self.cg0[k] = cg_new
        self.z0[k] -= g - (cg_new - cg) / f32(ALPHA) * w
        z = self.z0[k]
        eta = (f32(BETA) + cg_new) / f32(ALPHA)
        nw = np.where((z <= f32(L1)) & (z >= -f32(L1)), f32(0), np.where(z > 0, z - f32(L1), z + f32(L1)) / eta).astype(f32)
        self.w0[k] = nw
        self.new_w += int(((w == 0) & (nw != 0)).sum()) - int(((w != 0) & (nw == 0)).sum())
        self._resize(k[(w == 0) & (nw != 0)])
        kv = k[has]                                                      # UpdateV, :289-296, where the worker pulled V
        grad = G[has] + f32(V_L2) * self.V[kv]
        c2 = np.sqrt(self.cgV[kv] * self.cgV[kv] + grad * grad, dtype=f32)
        self.cgV[kv] = c2
        self.V[kv] -= f32(V_ALPHA) / (c2 + f32(V_BETA)) * grad

    def progress(self):
        out = (float(self.new_w), float(self.new_V))
        self.new_w = self.new_V = 0
        return out


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_one_rank_reproduces_the_ports_difacto_job():
    port = O.Port()
    kw = dict(l1=L1, l2=L2, eta=ALPHA, beta=BETA, l1_shrk=True, dim=DIM, threshold=THR, v_l2=V_L2, v_eta=V_ALPHA, v_beta=V_BETA,
              init_scale=SCALE, nt=1, max_key=MAX_KEY, seed=55)
    job = port.difacto_create(**kw)
    want = np.array([port.difacto_minibatch(job, minibatch(0, s), True, dp) for dp in range(PASSES) for s in range(STEPS)])
    keys, cnt, size, w, cg = port.difacto_model(job, DIM)
    port.difacto_destroy(job)
    ctypes.CDLL("libc.so.6").srand(55)
    d = CpuDifacto(port, 0, MAX_KEY, "rand")
    sd = ShardedDifacto(None, d, MAX_KEY)
    got = np.array([sd.process_minibatch(minibatch(0, s), True, dp) for dp in range(PASSES) for s in range(STEPS)])
    k = keys.astype(np.int64)
    assert np.array_equal(d.cnt[k], cnt) and np.array_equal(np.where(d.hasv[k], DIM + 1, 1), size)
    assert (size == DIM + 1).any()
    assert np.array_equal(got[:, 4:], want[:, 4:])                       # count, new_ex, new_w, new_V
    assert np.array_equal(d.w0[k], w[:, 0]) and np.array_equal(d.cg0[k], cg[:, 0]) and np.array_equal(d.z0[k], cg[:, 1])
    hv = d.hasv[k]
    assert np.array_equal(d.V[k][hv], w[hv, 1:]) and np.array_equal(d.cgV[k][hv], cg[hv, 2:])
    assert np.array_equal(got[:, 0], want[:, 0])


def run_rank(rank, world, port_no, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    port = O.Port()
    b, e = key_range(rank, world, MAX_KEY)
    d = CpuDifacto(port, b, e, "hash")
    sd = ShardedDifacto(None, d, MAX_KEY)
    progs = [sd.process_minibatch(minibatch(rank, s), True, dp) for dp in range(PASSES) for s in range(STEPS)]
    progs.append(sd.process_minibatch(minibatch(rank, 0), False, 1))
    np.savez(os.path.join(out_dir, "dfc%d.npz" % rank), cnt=d.cnt, w0=d.w0, z0=d.z0, cg0=d.cg0, hasv=d.hasv, V=d.V, cgV=d.cgV,
             prog=np.array(progs))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_over_gloo_match_the_replayed_schedule(tmp_path):
    world = 2
    mp.spawn(run_rank, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    # the same schedule in one process: `world` worker halves against one shard that owns every key
    port = O.Port()
    server = CpuDifacto(port, 0, MAX_KEY, "hash")
    workers = [CpuDifacto(port, 0, 1, "hash") for _ in range(world)]
    progs = [[] for _ in range(world)]
    new = []

    def step(s, train, dp):
        began = [workers[r].begin(minibatch(r, s)) for r in range(world)]
        if train and dp == 0:
            for k, c in began:
                server.count_push(k, c)
        pulled = [server.pull(k) for k, _ in began]
        grads = []
        for r in range(world):
            p, gw, G = workers[r].finish(*pulled[r], train)
            progs[r].append(p)
            grads.append((gw, G))
        if train:
            for (k, _), (_, has, _), (gw, G) in zip(began, pulled, grads):
                server.push(k, gw, G, has)
        new.append(server.progress())

    for dp in range(PASSES):
        for s in range(STEPS):
            step(s, True, dp)
    step(0, False, 1)
    got_new = np.zeros((len(new), 2))
    for r in range(world):
        got = np.load(tmp_path / ("dfc%d.npz" % r))
        b, e = key_range(r, world, MAX_KEY)
        for name in ("cnt", "w0", "z0", "cg0", "hasv", "V", "cgV"):
            assert np.array_equal(got[name], getattr(server, name)[b:e]), name
        assert got["hasv"].any()
        assert np.array_equal(got["prog"][:, :6], np.array(progs[r])[:, :6])
        got_new += got["prog"][:, 6:8]
    assert np.array_equal(got_new, np.array(new))
