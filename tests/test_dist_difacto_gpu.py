"""DiFacto (learn/difacto/async_sgd.h) over key-range shards on several GPUs. The minibatch in phases on one GPU must
equal wh_difacto_process_minibatch bit for bit; S ranks over NCCL must equal the same message schedule -- all count
pushes, all pulls, all pushes of a step, each in ascending source order -- replayed on ONE GPU against one shard that
holds every key (the embeddings are initialised from the (seed, key, j) hash, which does not depend on who owns a key)."""
import os
import socket

import numpy as np
import pytest

from tests.util import random_block

pytestmark = pytest.mark.gpu
MAX_KEY = 2 ** 13
STEPS, PASSES = 4, 2
KW = dict(l1=0.3, l2=0.01, eta=0.05, beta=1.0, l1_shrk=True, threshold=3, v_l2=1e-4, v_eta=0.02, init_scale=0.01, seed=77)


def minibatch(rank, step):
    rng = np.random.default_rng(9000 + 100 * rank + step)
    return random_block(rng, 2000, 12, 8000, with_value=(step % 2 == 1), empty_rows=False)


def all_keys(world):
    return np.unique(np.concatenate([minibatch(r, s).index % np.uint64(MAX_KEY) for r in range(world) for s in range(STEPS)]))


@pytest.mark.parametrize("dim", [0, 4, 16])
@pytest.mark.parametrize("init", ["hash", "rand"])
def test_phases_equal_process_minibatch(dim, init):
    import torch
    from wormhole_b200 import capi
    from wormhole_b200.dist import ShardedDifacto
    ctx = capi.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    keys = all_keys(1)
    rand = np.random.default_rng(5).integers(0, 2 ** 31 - 1, 1 << 20).astype(np.int32)
    out = []
    for phased in (False, True):
        d = capi.Difacto(ctx, dim=dim, max_key=MAX_KEY, **KW)
        if init == "rand":
            d.set_init_stream(rand)
        if phased:
            sd = ShardedDifacto(ctx, d, MAX_KEY, device=torch.device("cuda", 0))
            progs = [sd.process_minibatch(minibatch(0, s), True, dp) for dp in range(PASSES) for s in range(STEPS)]
            progs.append(sd.process_minibatch(minibatch(0, 0), False, 1))
        else:
            progs = [d.process_minibatch(minibatch(0, s), True, dp) for dp in range(PASSES) for s in range(STEPS)]
            progs.append(d.process_minibatch(minibatch(0, 0), False, 1))
        torch.cuda.synchronize()
        out.append((np.array(progs), d.get_state(keys)))
        d.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert out[0][0][:, 6].sum() > 0 and (dim == 0 or out[0][0][:, 7].sum() > 0)
    for a, b in zip(out[0][1], out[1][1]):
        assert np.array_equal(a, b)
    ctx.close()


def _rank_main(rank, world, port_no, out_dir, dim):
    import torch
    import torch.distributed as dist
    from wormhole_b200 import capi
    from wormhole_b200.dist import ShardedDifacto, key_range
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port_no)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ctx = capi.Context(rank)
    b, e = key_range(rank, world, MAX_KEY)
    d = capi.Difacto(ctx, dim=dim, max_key=MAX_KEY, key_begin=b, key_end=e, **KW)
    sd = ShardedDifacto(ctx, d, MAX_KEY, device=torch.device("cuda", rank))
    progs = [sd.process_minibatch(minibatch(rank, s), True, dp) for dp in range(PASSES) for s in range(STEPS)]
    progs.append(sd.process_minibatch(minibatch(rank, 0), False, 1))
    torch.cuda.synchronize()
    keys = all_keys(world)
    keys = keys[(keys >= b) & (keys < e)]
    cnt, size, w, cg = d.get_state(keys)
    np.savez(os.path.join(out_dir, "dfc%d.npz" % rank), keys=keys, cnt=cnt, size=size, w=w, cg=cg, prog=np.array(progs))
    dist.barrier()
    d.close(); ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,dim", [(2, 4), (2, 0), (4, 8), (8, 4)])
def test_gpus_match_the_schedule_on_one_gpu(tmp_path, world, dim):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (gpurun --gpus %d)" % (world, world))
    import torch.multiprocessing as mp
    from wormhole_b200 import capi
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port_no = s.getsockname()[1]; s.close()
    mp.spawn(_rank_main, args=(world, port_no, str(tmp_path), dim), nprocs=world, join=True)

    # the same schedule on one GPU: `world` workers (their own shards unused) against one shard with every key
    ctx = capi.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    dev, sd = torch.device("cuda", 0), max(dim, 1)
    server = capi.Difacto(ctx, dim=dim, max_key=MAX_KEY, **KW)
    workers = [capi.Difacto(ctx, dim=dim, max_key=MAX_KEY, **KW) for _ in range(world)]
    from wormhole_b200.dist import _wrap_device, _wrap_device_i64
    progs = [[] for _ in range(world)]
    new = []

    def step(s, train, dp):
        began = []
        for r in range(world):
            pu, pc, U = workers[r].begin(minibatch(r, s))
            began.append((_wrap_device_i64(pu, U, dev), _wrap_device(pc, U, "<i4", torch.int32, dev), U))
        if train and dp == 0 and dim > 0:
            for k, c, U in began:
                server.shard_count_push(k.data_ptr(), c.data_ptr(), U)
        pulled = []
        for k, c, U in began:
            w = torch.empty(U, dtype=torch.float32, device=dev); has = torch.empty(U, dtype=torch.uint8, device=dev)
            V = torch.empty((U, sd), dtype=torch.float32, device=dev)
            server.shard_pull(k.data_ptr(), U, w.data_ptr(), has.data_ptr(), V.data_ptr())
            pulled.append((w, has, V))
        grads = []
        for r, ((k, c, U), (w, has, V)) in enumerate(zip(began, pulled)):
            gw = torch.empty(U, dtype=torch.float32, device=dev); G = torch.empty((U, sd), dtype=torch.float32, device=dev)
            progs[r].append(workers[r].finish(w.data_ptr(), has.data_ptr(), V.data_ptr(), train, gw.data_ptr(), G.data_ptr()))
            grads.append((gw, G))
        if train:
            for (k, c, U), (w, has, V), (gw, G) in zip(began, pulled, grads):
                server.shard_push(k.data_ptr(), U, gw.data_ptr(), G.data_ptr(), has.data_ptr())
        new.append(server.shard_progress())

    for dp in range(PASSES):
        for s in range(STEPS):
            step(s, True, dp)
    step(0, False, 1)
    torch.cuda.synchronize()
    got_new = np.zeros((len(new), 2))
    for r in range(world):
        got = np.load(tmp_path / ("dfc%d.npz" % r))
        cnt, size, w, cg = server.get_state(got["keys"])
        assert len(got["keys"]) > 0
        assert np.array_equal(got["cnt"], cnt) and np.array_equal(got["size"], size)
        assert np.array_equal(got["w"], w) and np.array_equal(got["cg"], cg)
        assert np.array_equal(got["prog"][:, :6], np.array(progs[r])[:, :6])
        got_new += got["prog"][:, 6:8]
    assert np.array_equal(got_new, np.array(new))
    assert got_new[:, 0].sum() > 0 and (dim == 0 or got_new[:, 1].sum() > 0)
    server.close()
    for wk in workers:
        wk.close()
    ctx.close()
