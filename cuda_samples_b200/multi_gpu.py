"""Sharded (one process per GPU) versions of the primitives over torch.distributed / NCCL.

Only the exchange step each primitive really has is a collective:
  histogram   shard by bytes          -> all_reduce(SUM) of the 256+64 u32 bins (wrap-add is associative:
                                         bit-exact whatever algorithm NCCL picks)
  reduce      shard by elements       -> int: all_reduce(SUM) of one i32 (wraps like reduceCPU<int>);
                                         float/double: all_gather of the G shard sums, added in rank
                                         order in fp64 on every rank (deterministic, same on all ranks)
  scan        one long array, contiguous shards -> shard total, all_gather(G x u32), exclusive prefix of
                                         the lower ranks' totals is this shard's carry, chained scan
                                         seeded with it (batched scans shard by array: no collective)
  sort        sample sort             -> all_gather of samples, splitters, local stable partition,
                                         all_gather of bucket counts, all_to_all of keys and values laid
                                         out in source-rank order, local stable sort.  Stable globally:
                                         equal keys share a bucket, sources arrive in rank (= index) order.

With enable_p2p() (CUDA only, one NVSwitch node) the three small exchanges -- bins, shard sums, shard totals --
leave NCCL altogether: the last CTA of the producing kernel stores its payload into every peer's window over
NVLink, waits for the peers' flags and finishes the combine (csrc/p2p.cuh).  NCCL then only carries the control
plane (IPC handles, barriers) and the sort's all-to-all.

The local compute is an injected `ops` object: `CudaOps` (libb200prim.so, the product path) by default.
The CPU test-suite (gloo, world_size 2) injects a numpy stand-in that lives in tests/ -- there is no
CPU path in this package.
"""
import torch
import torch.distributed as dist

from . import api


_P2P = {"world": 0}


def enable_p2p(group=None):
    """Create this rank's exchange window, swap the CUDA IPC handles over the process group and map the peers.
    Collective: every rank of `group` must call it.  -> True when the fused kernels will be used."""
    rank, world = _world(group)
    if world == 1 or world > 16:
        return False
    handles = [None] * world
    dist.all_gather_object(handles, api.p2p_window_create(), group=group)
    api.p2p_connect(rank, world, handles)
    dist.barrier(group=group)
    _P2P["world"] = world
    return True


def disable_p2p(group=None):
    if _P2P["world"]:
        torch.cuda.synchronize()
        dist.barrier(group=group)
        for b in _RECV.values():
            for r, q in enumerate(b["pairs"]):
                if r != b["rank"]:
                    api.p2p_buffer_close(q)
        dist.barrier(group=group)
        for b in _RECV.values():
            api.p2p_buffer_free(b["pairs"][b["rank"]])
        _RECV.clear()
        api.p2p_disconnect()
        _P2P["world"] = 0


_RECV = {}  # n_local -> receive buffers of the fused sort exchange, mapped on every rank


def _recv_buffers(n_local, factor, rank, world, group):
    """Peer-mapped receive buffers for shards of n_local pairs.  The first call for a given n_local is a host
    collective (max shard size, IPC handles): every rank must reach it together (same sequence of shapes)."""
    key = (n_local, factor)
    b = _RECV.get(key)
    if b is None:
        biggest = torch.tensor([n_local], dtype=torch.int64, device="cuda")
        dist.all_reduce(biggest, op=dist.ReduceOp.MAX, group=group)
        capacity = (int(int(biggest.item()) * factor) + 2 * 8192 - 1) // 8192 * 8192
        pp, ph = api.p2p_buffer_create(8 * capacity)  # (key, value) pairs: one 8-byte peer store per pair
        handles = [None] * world
        dist.all_gather_object(handles, ph, group=group)
        b = {"rank": rank, "capacity": capacity,
             "pairs": [pp if r == rank else api.p2p_buffer_open(handles[r]) for r in range(world)]}
        dist.barrier(group=group)
        _RECV[key] = b
    return b


def p2p_enabled():
    return _P2P["world"] > 1


class CudaOps:
    """Local steps on the current CUDA device through the C ABI."""

    device_type = "cuda"

    def histogram_fused_allreduce(self, data, nbytes):
        bins = torch.empty(320, dtype=torch.int32, device=data.device)
        api.histogram_fused_allreduce(bins[:256], bins[256:], data, nbytes)
        return bins

    def reduce_partial_exchange(self, x, n, slots, index, count, gathered, sums):
        api.reduce_partial_exchange(x, n, slots, index, count, gathered, sums)

    def scan_shard(self, dst, src, n):
        carry = torch.empty(1, dtype=torch.int32, device=src.device)
        work = torch.empty(2 + 16, dtype=torch.float64, device=src.device)
        api.scan_exclusive_shard(dst, src, n, work, carry)

    def scan_total_exchange(self, src, n):
        carry = torch.empty(1, dtype=torch.int32, device=src.device)
        work = torch.empty(2 + 16, dtype=torch.float64, device=src.device)
        api.scan_total_exchange(carry, src, n, work)
        return carry

    def histogram_fused(self, data, nbytes):
        bins = torch.empty(320, dtype=torch.int32, device=data.device)
        api.histogram_fused(bins[:256], bins[256:], data, nbytes)
        return bins

    def reduce(self, x, n):
        out = torch.empty(1, dtype=x.dtype, device=x.device)
        api.reduce(n, 256, 1, 7, x, out)
        return out

    def reduce_partial(self, x, n, slot):
        api.reduce_partial(x, n, slot)

    def scan_total(self, src, n):
        tot = torch.empty(1, dtype=torch.int32, device=src.device)
        api.scan_total(tot, src, n)
        return tot

    def scan_with_carry(self, dst, src, n, carry):
        api.scanExclusiveLarge(dst, src, 1, n, d_carry=carry)

    def partition(self, keys, vals, n, splitters, buckets, direction):
        dk, dv = torch.empty_like(keys), torch.empty_like(vals)
        counts = torch.empty(buckets, dtype=torch.int32, device=keys.device)
        api.partition_kv(dk, dv, keys, vals, n, splitters, buckets, direction, counts)
        return dk, dv, counts

    def sort(self, keys, vals, n, direction):
        dk, dv = torch.empty_like(keys), torch.empty_like(vals)
        bk, bv = torch.empty_like(keys), torch.empty_like(vals)
        api.mergeSort(dk, dv, bk, bv, keys, vals, n, direction)
        return dk, dv


def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def _u32_as_i64(t):
    return t.to(torch.int64) & 0xFFFFFFFF


# ---------------------------------------------------------------------------------------- histogram
def histogram_sharded(data, nbytes, ops=None, group=None):
    """-> (hist256, hist64) of the concatenation of all ranks' shards, on every rank (int32 = u32 bits)."""
    ops = ops or CudaOps()
    if p2p_enabled() and isinstance(ops, CudaOps) and _world(group)[1] == _P2P["world"]:
        bins = ops.histogram_fused_allreduce(data, nbytes)  # one kernel: histogram + exchange over NVLink
        return bins[:256], bins[256:]
    bins = ops.histogram_fused(data, nbytes)
    if _world(group)[1] > 1:
        dist.all_reduce(bins, op=dist.ReduceOp.SUM, group=group)
    return bins[:256], bins[256:]


# ------------------------------------------------------------------------------------------- reduce
def reduce_sharded(x, n, ops=None, group=None):
    """-> 1-element tensor with the sum over all shards (same bits on every rank)."""
    ops = ops or CudaOps()
    part = ops.reduce(x, n)
    rank, world = _world(group)
    if world == 1:
        return part
    if part.dtype == torch.int32:
        dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
        return part
    parts = torch.empty(world, dtype=part.dtype, device=part.device)
    dist.all_gather_into_tensor(parts, part, group=group)
    return parts.to(torch.float64).sum().to(part.dtype).reshape(1)  # fixed rank order


class PendingReduce:
    """Result of reduce_many_sharded: the collective may still be in flight (it runs on NCCL's own stream,
    overlapping whatever the caller enqueues next); result() makes the current stream wait for it."""

    def __init__(self, work, gathered, world, dtypes):
        self.work, self.gathered, self.world, self.dtypes = work, gathered, world, dtypes

    def result(self):
        """-> float64 tensor [k]: entry i is the sum over ranks (rank order, fp64) of the i-th shard partial.
        Integer entries are exact integers; apply the 32-bit wrap with typed()."""
        if self.work is not None:
            self.work.wait()
            self.work = None
        if self.world == 1:
            return self.gathered
        return self.gathered.view(self.world, -1).sum(0)

    def typed(self):
        """-> list of python scalars in the inputs' own types (int: wrapped to int32 like reduceCPU<int>)."""
        vals = self.result().cpu().tolist()
        check_p2p()
        out = []
        for v, dt in zip(vals, self.dtypes):
            if dt == torch.int32:
                out.append(((int(v) + 2**31) % 2**32) - 2**31)
            elif dt == torch.float32:
                out.append(float(torch.tensor(v, dtype=torch.float64).to(torch.float32)))
            else:
                out.append(float(v))
        return out


def reduce_many_sharded(tensors, ns, ops=None, group=None, overlap=False):
    """Several sharded reductions with ONE collective: every local kernel writes its shard partial (widened
    exactly to a double) into one small buffer, the buffer is all_gathered once, and the per-rank partials are
    added in rank order in fp64.  With overlap=True the all_gather is asynchronous (NCCL stream); measured on
    behind the caller's next kernels.  -> PendingReduce"""
    ops = ops or CudaOps()
    rank, world = _world(group)
    k = len(tensors)
    buf = torch.empty(k, dtype=torch.float64, device=tensors[0].device)
    dtypes = [x.dtype for x in tensors]
    if p2p_enabled() and isinstance(ops, CudaOps) and world == _P2P["world"]:
        # the last local kernel also carries the exchange and the rank-order sums: no collective call at all
        gathered = torch.empty(world * k, dtype=torch.float64, device=buf.device)
        sums = torch.empty(k, dtype=torch.float64, device=buf.device)
        for i, (x, n) in enumerate(zip(tensors, ns)):
            if i < k - 1:
                ops.reduce_partial(x, n, buf[i:i + 1])
            else:
                ops.reduce_partial_exchange(x, n, buf, i, k, gathered, sums)
        return PendingReduce(None, sums, 1, dtypes)
    for i, (x, n) in enumerate(zip(tensors, ns)):
        ops.reduce_partial(x, n, buf[i:i + 1])
    if world == 1:
        return PendingReduce(None, buf, 1, dtypes)
    gathered = torch.empty(world * k, dtype=torch.float64, device=buf.device)
    work = dist.all_gather_into_tensor(gathered, buf, group=group, async_op=overlap)
    return PendingReduce(work if overlap else None, gathered, world, dtypes)


# --------------------------------------------------------------------------------------------- scan
def shard_carry(totals, rank):
    """Exclusive prefix (mod 2^32) of the shard totals below `rank`; totals: int32 tensor of G u32."""
    if rank == 0:
        return torch.zeros(1, dtype=torch.int32, device=totals.device)
    s = _u32_as_i64(totals[:rank]).sum() & 0xFFFFFFFF
    s = torch.where(s >= 2**31, s - 2**32, s)
    return s.to(torch.int32).reshape(1)


def scan_sharded(dst, src, n_local, ops=None, group=None):
    """Exclusive scan of ONE array split into contiguous shards (rank r holds the r-th slice)."""
    ops = ops or CudaOps()
    rank, world = _world(group)
    if world == 1:
        ops.scan_with_carry(dst, src, n_local, None)
        return dst
    if p2p_enabled() and isinstance(ops, CudaOps) and world == _P2P["world"]:
        ops.scan_shard(dst, src, n_local)  # 2 kernels, no collective call (tile sums + exchange, scan of independent tiles)
        return dst
    total = ops.scan_total(src, n_local)
    totals = torch.empty(world, dtype=torch.int32, device=total.device)
    dist.all_gather_into_tensor(totals, total, group=group)
    ops.scan_with_carry(dst, src, n_local, shard_carry(totals, rank))
    return dst


# --------------------------------------------------------------------------------------------- sort
def choose_splitters(samples_ordered, world):
    """samples_ordered: 1-D int64 tensor of ALL ranks' samples in ordered key space (key ^ flip).
    -> world-1 ascending splitters (int64); bucket(o) = #splitters <= o."""
    s, _ = torch.sort(samples_ordered)
    per = s.numel() // world
    idx = torch.arange(1, world, device=s.device) * per
    return s[idx]


def sort_exchange_only(keys, vals, n_local, direction=1, group=None, capacity_factor=1.25):
    """The exchange phase of sort_sharded alone (p2p path): splitters, bucket counts + their exchange, and the
    partition pass that stores the bucket runs into the peers' receive buffers, closed by the node-wide barrier.
    For bench.py's NVLink-roofline row.  -> device tensor meta [world*world counts | pairs received | overflow]."""
    rank, world = _world(group)
    b = _recv_buffers(n_local, capacity_factor, rank, world, group)
    if "work" not in b:
        sort_sharded(keys, vals, n_local, direction, group=group, capacity_factor=capacity_factor)
    w = b["work"]
    api.sort_splitters(keys, n_local, direction, w["spl"])
    api.partition_exchange_kv(keys, vals, n_local, w["spl"], direction, b["pairs"], None, b["capacity"],
                              w["meta"][:world * world], w["meta"][world * world:])
    return w["meta"]


def nvlink_probe(src, nbytes, n_local, group=None, capacity_factor=1.25):
    """Measured NVLink roofline for the exchange: every rank writes nbytes from `src` into the receive buffer of rank
    (rank+1) % world at the same time.  -> {mode name: GB/s per direction per GPU on this rank}."""
    rank, world = _world(group)
    b = _recv_buffers(n_local, capacity_factor, rank, world, group)
    peer = b["pairs"][(rank + 1) % world]
    nbytes = min(nbytes, 8 * b["capacity"])
    out = {}
    for mode, name in ((3, "cudaMemcpyAsync"), (1, "16-byte lane stores"), (0, "4-byte lane stores"),
                       (2, "cp.async.bulk from shared memory")):
        torch.cuda.synchronize()
        dist.barrier(group=group)
        out[name] = api.p2p_probe_store(peer, src, nbytes, mode, 5)
    torch.cuda.synchronize()
    dist.barrier(group=group)
    return out


def check_p2p():
    """Raise if any fused exchange on this rank ever lost a peer (sticky device-side count, csrc/p2p.cuh): the results
    of that call are void.  Synchronises the device."""
    if p2p_enabled():
        n = api.p2p_errors()
        if n:
            raise api.B200Error("NVLink peer-window exchange: %d word(s) never arrived within the timeout -- a rank "
                                "is missing or lagging; results of the affected calls are void" % n)


def sort_sharded(keys, vals, n_local, direction=1, ops=None, group=None, samples_per_rank=1024, capacity_factor=1.25,
                 sync=True):
    """Stable global key-value sort.  Rank r ends up with the r-th key range (ascending ranges for
    direction=1, descending for 0).  -> (keys_out, vals_out, n_out, counts_matrix).
    With enable_p2p() nothing but kernels of libb200prim.so run: splitters are chosen on the device (samples exchanged
    through the peer windows), the partition pass stores every bucket straight into its destination rank's receive
    buffer over NVLink (no all-to-all, no host round trip before the local sort), the local sort reads its pair count
    on the device.  A receive buffer holds capacity_factor x the largest shard; a distribution that overflows it falls
    back to the NCCL path.  The output tensors of the p2p path belong to the library and are reused by the next call
    with the same shard size.  sync=False (p2p path only) returns (keys_out, vals_out, meta, None) without any host
    synchronisation: meta is the device tensor [world*world counts | pairs received | overflow flag]."""
    ops = ops or CudaOps()
    rank, world = _world(group)
    if world == 1:
        dk, dv = ops.sort(keys, vals, n_local, direction)
        return dk, dv, n_local, None
    flip = 0 if direction else 0xFFFFFFFF
    spl = None
    if p2p_enabled() and isinstance(ops, CudaOps) and world == _P2P["world"]:
        b = _recv_buffers(n_local, capacity_factor, rank, world, group)
        cap = b["capacity"]
        if "work" not in b:  # per-shape work buffers live with the receive buffers: nothing is allocated per call
            mk = lambda n: torch.empty(n, dtype=torch.int32, device=keys.device)  # noqa: E731
            b["work"] = {"spl": mk(max(world - 1, 1)), "meta": mk(world * world + 2), "dk": mk(cap), "dv": mk(cap)}
            api.sort_reserve(max(cap, n_local))  # library scratch too: no allocation between kernels that wait on peers
            torch.cuda.synchronize()
            dist.barrier(group=group)
        w = b["work"]
        spl, meta = w["spl"], w["meta"]
        # 1. splitters on the device: regular samples -> peer windows -> bitonic sort -> same splitters everywhere
        api.sort_splitters(keys, n_local, direction, spl)
        # 2-4 fused: count + exchange of counts, partition pass that scatters into the peers' receive buffers
        api.partition_exchange_kv(keys, vals, n_local, spl, direction, b["pairs"], None, cap,
                                  meta[:world * world], meta[world * world:])
        # 5. local stable sort of the (pair-format) receive buffer, received count read on the device: no host round trip
        api.sort_kv_dyn(w["dk"], w["dv"], None, None, b["pairs"][rank], None, cap, meta[world * world:], direction)
        if not sync:
            return w["dk"], w["dv"], meta, None
        host = meta.cpu()  # the only synchronisation: after everything is queued
        check_p2p()
        if int(host[-1]) == 0:
            return w["dk"], w["dv"], int(host[-2]), host[:world * world].reshape(world, world)
        # some rank would receive more than `cap` pairs: nothing was sent, every rank takes the NCCL path below
    if spl is None:
        # 1. regular samples of the local keys, in ordered space
        ns = min(samples_per_rank, max(n_local, 1))
        step = max(n_local // ns, 1)
        local = _u32_as_i64(keys[:n_local:step][:ns]) ^ flip
        if local.numel() < ns:  # tiny shards: pad by repetition so every rank contributes ns samples
            pad = local[-1:].expand(ns - local.numel()) if local.numel() else torch.zeros(ns, dtype=torch.int64,
                                                                                        device=keys.device)
            local = torch.cat([local, pad])
        gathered = torch.empty(world * ns, dtype=torch.int64, device=keys.device)
        dist.all_gather_into_tensor(gathered, local.contiguous(), group=group)
        spl64 = choose_splitters(gathered, world)
        spl = torch.where(spl64 >= 2**31, spl64 - 2**32, spl64).to(torch.int32).contiguous()
    # 2. local stable partition into `world` buckets
    pk, pv, counts = ops.partition(keys, vals, n_local, spl, world, direction)
    # 3. everyone learns everyone's bucket sizes
    matrix = torch.empty(world * world, dtype=torch.int32, device=counts.device)
    dist.all_gather_into_tensor(matrix, counts, group=group)
    m = matrix.reshape(world, world).cpu()
    send = m[rank].tolist()
    recv = m[:, rank].tolist()
    n_out = int(sum(recv))
    rk = torch.empty(max(n_out, 1), dtype=keys.dtype, device=keys.device)
    rv = torch.empty(max(n_out, 1), dtype=vals.dtype, device=vals.device)
    # 4. all-to-all: the receive buffer is laid out in source-rank order (= original index order)
    dist.all_to_all_single(rk[:n_out], pk[:n_local], recv, send, group=group)
    dist.all_to_all_single(rv[:n_out], pv[:n_local], recv, send, group=group)
    # 5. local stable sort of the received range
    dk, dv = ops.sort(rk, rv, n_out, direction)
    return dk, dv, n_out, m
