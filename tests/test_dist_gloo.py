"""The N>1 path on CPU: two processes over gloo run wormhole_b200.dist.ShardedKV (key-range split of
the sorted unique keys, all-to-all-v of keys / weights / gradients, one push message per source rank in
ascending order) with a CPU stand-in for each rank's shard and worker maths (the oracle port), and must
end with exactly the model a single-process simulation of the same message order produces."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O
from tests.util import random_block
from wormhole_b200.dist import ShardedKV, key_range

MAX_KEY = 1000003          # not a power of two: uneven ranges, general modulo
ALGOS = [(O.ALGO_FTRL, O.LOSS_LOGIT), (O.ALGO_SGD, O.LOSS_SQUARE_HINGE), (O.ALGO_ADAGRAD, O.LOSS_LOGIT)]
L1, L2, ETA, BETA = 0.5, 0.1, 0.1, 1.0
STEPS = 5


class CpuShard:
    """Dense {w, z, cg} arrays for one key range; updates by the oracle's handle (test stand-in)."""

    def __init__(self, port, algo, begin, end):
        self.port, self.algo, self.begin = port, algo, begin
        n = end - begin
        self.w, self.z, self.cg = (np.zeros(n, np.float32) for _ in range(3))
        self.t, self.nnz = 1, 0

    def pull(self, keys):
        k = keys.numpy().astype(np.int64) - self.begin
        assert (k >= 0).all() and (k < len(self.w)).all(), "key outside this shard's range"
        return torch.from_numpy(self.w[k].copy())

    def push(self, keys, grads):
        k = keys.numpy().astype(np.int64) - self.begin
        w, z, cg = self.w[k].copy(), self.z[k].copy(), self.cg[k].copy()
        dn, self.t = self.port.handle_push(self.algo, L1, L2, ETA, BETA, grads.numpy().copy(), w, z, cg, self.t)
        self.w[k], self.z[k], self.cg[k] = w, z, cg
        self.nnz += dn


def minibatch(rank, step):
    rng = np.random.default_rng(1000 * rank + step)
    return random_block(rng, 600, 12, 5000, with_value=(step % 2 == 1), empty_rows=False)


def worker_step(port, loss, blk, kv):
    D = port.localize(blk, MAX_KEY)
    keys = torch.from_numpy(D.uniq.astype(np.int64))
    w = kv.pull(keys).numpy()
    _, grad = port.linear_loss(loss, D, w, 2)
    kv.push(torch.from_numpy(grad))


def run_rank(rank, world, port_no, algo, loss, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    port = O.Port()
    b, e = key_range(rank, world, MAX_KEY)
    shard = CpuShard(port, algo, b, e)
    kv = ShardedKV(shard, MAX_KEY)
    for step in range(STEPS):
        worker_step(port, loss, minibatch(rank, step), kv)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), w=shard.w, z=shard.z, cg=shard.cg, t=shard.t, nnz=shard.nnz)
    dist.barrier()
    dist.destroy_process_group()


def simulate(world, algo, loss):
    """Same job in one process: per step every worker pulls from the model as of the end of the previous
    step, then the pushes are applied per (owner, source) in ascending source order."""
    port = O.Port()
    shards = [CpuShard(port, algo, *key_range(r, world, MAX_KEY)) for r in range(world)]
    bounds = [key_range(r, world, MAX_KEY)[0] for r in range(world)] + [MAX_KEY]
    for step in range(STEPS):
        msgs = []
        for r in range(world):
            D = port.localize(minibatch(r, step), MAX_KEY)
            keys = D.uniq.astype(np.int64)
            w = np.empty(len(keys), np.float32)
            cut = np.searchsorted(keys, bounds)
            for s in range(world):
                sl = slice(cut[s], cut[s + 1])
                w[sl] = shards[s].pull(torch.from_numpy(keys[sl])).numpy()
            _, grad = port.linear_loss(loss, D, w, 2)
            msgs.append((keys, grad, cut))
        for s in range(world):
            for r in range(world):
                keys, grad, cut = msgs[r]
                sl = slice(cut[s], cut[s + 1])
                shards[s].push(torch.from_numpy(keys[sl]), torch.from_numpy(grad[sl]))
    return shards


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_key_ranges_cover_the_key_space():
    for world in (1, 2, 3, 8):
        for mk in (7, 2 ** 24, 1000003):
            r = [key_range(g, world, mk) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == mk
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))


def test_single_rank_degenerates_to_local_push_pull():
    port = O.Port()
    shard = CpuShard(port, O.ALGO_FTRL, 0, MAX_KEY)
    kv = ShardedKV(shard, MAX_KEY)
    for step in range(3):
        worker_step(port, O.LOSS_LOGIT, minibatch(0, step), kv)
    job = port.linear_create(O.ALGO_FTRL, O.LOSS_LOGIT, L1, L2, ETA, BETA, MAX_KEY)
    for step in range(3):
        port.linear_minibatch(job, minibatch(0, step))
    keys, w, z, cg = port.linear_model(job)
    assert np.array_equal(shard.w[keys.astype(np.int64)], w)
    assert np.array_equal(shard.z[keys.astype(np.int64)], z)
    assert shard.nnz == int((w != 0).sum())


@pytest.mark.parametrize("algo,loss", ALGOS)
def test_two_ranks_over_gloo_match_the_simulation(tmp_path, algo, loss):
    world = 2
    mp.spawn(run_rank, args=(world, free_port(), algo, loss, str(tmp_path)), nprocs=world, join=True)
    want = simulate(world, algo, loss)
    for r in range(world):
        got = np.load(tmp_path / ("rank%d.npz" % r))
        assert np.array_equal(got["w"], want[r].w)
        assert np.array_equal(got["z"], want[r].z)
        assert np.array_equal(got["cg"], want[r].cg)
        assert int(got["t"]) == want[r].t            # SGD: t advances once per push MESSAGE per shard
        assert int(got["nnz"]) == want[r].nnz
    if algo == O.ALGO_SGD:
        assert want[0].t == 1 + world * STEPS
