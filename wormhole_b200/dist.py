"""Key-range sharded push / pull across the GPUs of one box: the ps-lite part of the path.

In the reference every worker cuts its SORTED unique key list into contiguous slices, one per
server, and sends each slice to the server that owns that key range (ps::KVWorker::ZPull / ZPush,
call sites learn/linear/async_sgd.h:277-278,287; ps-lite itself is not vendored). Here rank g is
both worker g (its own minibatch stream) and server g, owner of the transformed keys in
[floor(g*max_key/S), floor((g+1)*max_key/S)). One pull = all-to-all-v of key slices, a Handle::Pull
per received key on the owner, all-to-all-v of the weights back. One push = all-to-all-v of the
gradients (the key slices of the pull are reused: the reference's KEY_CACHING filter,
learn/linear/async_sgd.h:295-297), then on each owner one push message PER SOURCE RANK applied in
ascending source order -- every (worker, minibatch) gradient is its own handle update, as in the
asynchronous parameter server; nothing is summed across workers.

Only torch.distributed is used for the exchange (NCCL over NVLink on GPUs, gloo in the CPU tests);
the shard itself is anything with pull(keys)->weights and push(keys, grads): the CUDA shard
(DeviceShard below) in production, a CPU stand-in in the world_size-2 gloo tests.
"""
import torch
import torch.distributed as dist


def key_range(rank, world, max_key):
    """[begin, end) of the transformed keys that server `rank` of `world` owns."""
    return (rank * max_key) // world, ((rank + 1) * max_key) // world


def range_boundaries(world, max_key, device):
    """The world+1 range boundaries as an int64 tensor (max_key <= 2^62 so they fit)."""
    return torch.tensor([(r * max_key) // world for r in range(world + 1)], dtype=torch.int64, device=device)


class DeviceShard:
    """Adapter: a capi.Server (CUDA shard) behind the pull / push interface ShardedKV expects."""

    def __init__(self, server):
        self.server = server

    def pull(self, keys):
        vals = torch.empty(keys.numel(), dtype=torch.float32, device=keys.device)
        if keys.numel():
            self.server.pull_dev(keys.data_ptr(), keys.numel(), vals.data_ptr())
        return vals

    def push(self, keys, grads):
        # an empty slice is still a push message for this server: SGDHandle::Start advances t
        # (learn/linear/async_sgd.h:85-90); wh_server_push handles n == 0 that way
        self.server.push_dev(keys.data_ptr() if keys.numel() else 0, keys.numel(),
                             grads.data_ptr() if grads.numel() else 0)


class ShardedKV:
    """ZPull / ZPush of one worker against `world` key-range shards (one per rank)."""

    def __init__(self, shard, max_key, group=None, device=None):
        self.shard, self.max_key, self.group = shard, max_key, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device if device is not None else torch.device("cpu")
        self.bounds = range_boundaries(self.world, max_key, self.device)
        self._cache = None   # (send_counts, recv_counts, recv_keys) of the last pull

    def _split(self, keys):
        """Entries of the ascending key list per destination shard."""
        cut = torch.searchsorted(keys, self.bounds)          # keys are int64 views of values < 2^62
        return (cut[1:] - cut[:-1]).tolist()                 # host sync: the split sizes

    def _a2a(self, out, inp, out_split, in_split):
        if self.world == 1:
            out.copy_(inp)
        else:
            dist.all_to_all_single(out, inp, out_split, in_split, group=self.group)

    def pull(self, keys):
        """keys: ascending unique transformed keys (int64 tensor). Returns their weights (float32)."""
        send = self._split(keys)
        if self.world == 1:
            recv = list(send)
        else:
            s = torch.tensor(send, dtype=torch.int64, device=self.device)
            r = torch.empty_like(s)
            dist.all_to_all_single(r, s, group=self.group)
            recv = r.tolist()
        recv_keys = torch.empty(sum(recv), dtype=torch.int64, device=self.device)
        self._a2a(recv_keys, keys, recv, send)
        # one pull message per source; Handle::Pull is stateless, so serve them as one batch
        recv_vals = self.shard.pull(recv_keys)
        vals = torch.empty(keys.numel(), dtype=torch.float32, device=self.device)
        self._a2a(vals, recv_vals, send, recv)
        self._cache = (send, recv, recv_keys)
        return vals

    def push(self, grads):
        """grads: one float per key of the last pull. Applied on the owners, one message per source."""
        send, recv, recv_keys = self._cache
        recv_grads = torch.empty(sum(recv), dtype=torch.float32, device=self.device)
        self._a2a(recv_grads, grads, recv, send)
        o = 0
        for src in range(self.world):           # ascending source rank: one valid async interleaving
            n = recv[src]
            self.shard.push(recv_keys[o:o + n], recv_grads[o:o + n])
            o += n


class ShardedWorker:
    """AsgdWorker::ProcessMinibatch on rank g of a key-range sharded job (CUDA path):
    wh_worker_begin (Localize) -> ShardedKV.pull -> wh_worker_finish (Init/Evaluate/CalcGrad) ->
    ShardedKV.push."""

    def __init__(self, ctx, worker, kv):
        self.ctx, self.worker, self.kv = ctx, worker, kv
        # The exchange interleaves torch work (searchsorted, all_to_all, empty) with library launches on the context's
        # stream, and nothing else orders the two: bind the context to the stream torch is using, here and again at
        # every minibatch (torch's current stream may change under a torch.cuda.stream(...) block).
        self._bind()

    def _bind(self):
        s = torch.cuda.current_stream(self.kv.device).cuda_stream
        if getattr(self, "_bound", None) != s:
            self.ctx.set_stream(s)
            self._bound = s

    def process_minibatch(self, blk, train=True, mem=0, nnz=0):
        self._bind()
        ptr, U = self.worker.begin(blk, mem, nnz)
        keys = _wrap_device_i64(ptr, U, self.kv.device)
        w = self.kv.pull(keys)
        grads = torch.empty(U, dtype=torch.float32, device=self.kv.device)
        prog = self.worker.finish(w.data_ptr() if U else 0, grads.data_ptr() if U else 0, train)
        if train:
            self.kv.push(grads)
        return prog


class DeviceDifacto:
    """Adapter: a capi.Difacto (CUDA worker half + shard) behind the tensor-level interface ShardedDifacto drives. A CPU
    stand-in with the same methods runs the exchange in the world_size-2 gloo tests."""

    def __init__(self, difacto, device):
        self.d, self.device, self.dim = difacto, device, difacto.dim

    @staticmethod
    def _p(t):
        return t.data_ptr() if t.numel() else 0

    def begin(self, blk, mem=0):
        """Localize with counts -> (ascending unique keys int64[U], counts int32[U]), valid until the next begin."""
        pu, pc, U = self.d.begin(blk, mem)
        return _wrap_device_i64(pu, U, self.device), _wrap_device(pc, U, "<i4", torch.int32, self.device)

    def count_push(self, keys, cnt):
        self.d.shard_count_push(self._p(keys), self._p(cnt), keys.numel())

    def pull(self, keys):
        n, sd = keys.numel(), max(self.dim, 1)
        w = torch.empty(n, dtype=torch.float32, device=self.device)
        has = torch.empty(n, dtype=torch.uint8, device=self.device)
        V = torch.empty((n, sd), dtype=torch.float32, device=self.device)
        self.d.shard_pull(self._p(keys), n, self._p(w), self._p(has), self._p(V))
        return w, has, V

    def finish(self, w, has, V, train):
        U, sd = w.numel(), max(self.dim, 1)
        gw = torch.empty(U, dtype=torch.float32, device=self.device)
        G = torch.empty((U, sd), dtype=torch.float32, device=self.device)
        prog = self.d.finish(self._p(w), self._p(has), self._p(V), train, self._p(gw), self._p(G))
        return prog, gw, G

    def push(self, keys, gw, G, has):
        self.d.shard_push(self._p(keys), keys.numel(), self._p(gw), self._p(G), self._p(has))

    def progress(self):
        return self.d.shard_progress()


class ShardedDifacto:
    """AsyncWorker::ProcessMinibatch of learn/difacto/async_sgd.h:363-425 on rank g of a key-range sharded job: rank g
    is worker g and the AsyncServer shard of [g*max_key/S, (g+1)*max_key/S). The three exchanges of the minibatch --
    count push (:374-381), ZVPull (:424), ZVPush (:408-411) -- are all-to-all-v's of the sorted key slices; the owners
    apply one message per source rank in ascending source order (one valid interleaving of the asynchronous server).
    The variable-length payloads travel as fixed-width rows {w, has, V[dim]} (has = the length flag).

    `difacto` is a capi.Difacto (wrapped in DeviceDifacto) or anything with DeviceDifacto's methods."""

    def __init__(self, ctx, difacto, max_key, group=None, device=None):
        self.ctx, self.max_key, self.group = ctx, max_key, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if ctx is not None else torch.device("cpu")
        self.device = device
        self.d = difacto if hasattr(difacto, "count_push") else DeviceDifacto(difacto, device)
        self.dim = self.d.dim
        self.bounds = range_boundaries(self.world, max_key, self.device)
        self._bound = None
        self._bind()

    def _bind(self):
        if self.ctx is None:
            return
        s = torch.cuda.current_stream(self.device).cuda_stream
        if self._bound != s:
            self.ctx.set_stream(s)
            self._bound = s

    def _a2a(self, inp, out_split, in_split):
        out = torch.empty((sum(out_split),) + tuple(inp.shape[1:]), dtype=inp.dtype, device=self.device)
        if self.world == 1:
            out.copy_(inp)
        else:
            dist.all_to_all_single(out, inp.contiguous(), out_split, in_split, group=self.group)
        return out

    def _sizes(self, keys):
        cut = torch.searchsorted(keys, self.bounds)
        send = (cut[1:] - cut[:-1]).tolist()
        if self.world == 1:
            return send, list(send)
        s = torch.tensor(send, dtype=torch.int64, device=self.device)
        r = torch.empty_like(s)
        dist.all_to_all_single(r, s, group=self.group)
        return send, r.tolist()

    def process_minibatch(self, blk, train=True, data_pass=0, mem=0):
        self._bind()
        dim = self.dim
        keys, cnt = self.d.begin(blk, mem)
        send, recv = self._sizes(keys)
        rkeys = self._a2a(keys, recv, send)
        if train and data_pass == 0 and dim > 0:
            rcnt = self._a2a(cnt, recv, send)
            o = 0
            for n in recv:                       # one count-push message per source, ascending
                self.d.count_push(rkeys[o:o + n], rcnt[o:o + n])
                o += n
        # ZVPull: Handle::Pull does not change the shard, so the S messages are served as one batch
        rw, rhas, rV = self.d.pull(rkeys)
        w = self._a2a(rw, send, recv)
        has = self._a2a(rhas, send, recv)
        V = self._a2a(rV, send, recv) if dim > 0 else rV[:0]
        prog, gw, G = self.d.finish(w, has, V, train)
        if train:
            rgw = self._a2a(gw, recv, send)
            rG = self._a2a(G, recv, send) if dim > 0 else rV
            # the push carries the lengths the worker pulled (rhas is unchanged since the pull of these very keys)
            o = 0
            for n in recv:                       # one gradient-push message per source, ascending
                self.d.push(rkeys[o:o + n], rgw[o:o + n], rG[o:o + n], rhas[o:o + n])
                o += n
        prog[6], prog[7] = self.d.progress()
        return prog


def _wrap_device(ptr, n, typestr, dtype, device):
    if n == 0:
        return torch.empty(0, dtype=dtype, device=device)

    class _Arr:
        __cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}

    return torch.as_tensor(_Arr(), device=device)


def _wrap_device_i64(ptr, n, device):
    """A torch int64 view of n uint64 keys at a raw device address (keys < 2^63 by construction)."""
    if n == 0:
        return torch.empty(0, dtype=torch.int64, device=device)

    class _Arr:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 2}

    return torch.as_tensor(_Arr(), device=device)
