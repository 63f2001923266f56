"""Host-side logic of the multi-GPU path: one process per GPU, torch.distributed for plumbing only.

The distance path shards by query rows with a replicated index (src/dist.cpp:174-258 / :560-692 have no cross-row
state), so the only exchange is the all-gather of the per-rank sketch shards before the index is built.  Everything
here is plain torch on whatever device the tensors live on, so the same code runs under NCCL on B200s and under
gloo on CPUs (tests/test_multirank_gloo.py)."""
import os
import time

import torch

# diagnostic: KSSD_MG_TRACE=1 -> every sub-step is followed by a device synchronize and its wall time accumulated here
TRACE = {} if os.environ.get("KSSD_MG_TRACE") else None


def _tick(name, t0):
    if TRACE is None:
        return t0
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    TRACE.setdefault(name, []).append((t1 - t0) * 1e3)
    return t1


def triangle_bounds(n, parts, row_cost=0.0):
    """Row boundaries b[0..parts] that give every part (nearly) the same cost of the n x n triangle: row i costs
    n-1-i cells (src/dist.cpp:207 scans j in (i, n)) plus row_cost cell-equivalents of per-row work (the posting-range
    lookups of the row's hashes: list_lookup_kernel).  row_cost = 0 balances cells alone."""
    if row_cost <= 0:
        total = n * (n - 1) / 2.0
        b = [0]
        for p in range(1, parts):
            target = total * p / parts
            # cells in rows [0, r) = r*n - r(r+1)/2  -> solve for r
            r = int(round(((2 * n - 1) - ((2 * n - 1) ** 2 - 8 * target) ** 0.5) / 2))
            b.append(min(max(r, b[-1]), n))
        b.append(n)
        return b
    # cost of rows [0, r) = r*(n + row_cost) - r(r+1)/2: same quadratic with n + row_cost in place of n
    m = n + float(row_cost)
    total = n * m - n * (n + 1) / 2.0
    b = [0]
    for p in range(1, parts):
        target = total * p / parts
        r = int(round(((2 * m - 1) - ((2 * m - 1) ** 2 - 8 * target) ** 0.5) / 2))
        b.append(min(max(r, b[-1]), n))
    b.append(n)
    return b


def rect_bounds(n_query, parts):
    """dist (query x ref rectangle): contiguous blocks of query rows, sizes differing by at most one."""
    base, extra = divmod(n_query, parts)
    b = [0]
    for p in range(parts):
        b.append(b[-1] + base + (1 if p < extra else 0))
    return b


class PeerGather:
    """Sketch gather over peer memory (NVLink through CUDA IPC) instead of an NCCL transfer: every rank publishes its
    shard in its own buffer and reads the others' in place while concatenating (kssd_b200_sketch_gather_peers).
    torch.distributed only carries the 128-byte IPC handles once; the per-step barrier is a flag every rank stores into its
    peers' memory after publishing (no collective in the step at all)."""

    def __init__(self, dist, ctx, n_local, cap, rank, world, nccl_barrier=False):
        self.dist, self.ctx, self.world, self.nccl_barrier = dist, ctx, world, nccl_barrier
        dev = torch.device("cuda", torch.cuda.current_device())
        mine = torch.frombuffer(bytearray(ctx.peer_init(n_local, cap)), dtype=torch.uint8).to(dev)
        every = torch.empty(128 * world, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(every, mine)
        ctx.peer_connect(bytes(every.cpu().numpy().tobytes()), world, rank)
        self.token = torch.zeros(1, dtype=torch.int32, device=dev)

    def gather(self):
        """After ctx.sketch_finish_device(): -> (d_hashes ptr, d_genome_off ptr, total) or None on overflow (every rank alike)."""
        t0 = _tick("", time.perf_counter()) if TRACE is not None else 0.0
        self.ctx.sketch_publish()
        t0 = _tick("peer.publish", t0)
        # the library's concat kernels wait on flags the peers store over NVLink when their shard is complete (segsort.cu:
        # peer_wait); round 1 passed a one-element all-reduce here instead (kept behind nccl_barrier for comparison)
        if self.nccl_barrier:
            self.dist.all_reduce(self.token)
            t0 = _tick("peer.barrier", t0)
        ph, po, total, overflow = self.ctx.sketch_gather_peers()
        _tick("peer.pull", t0)
        return None if overflow else (ph, po, total)


def gather_sketches_device(dist, ctx, hashes_local, sizes_local, world, cap):
    """CUDA fast path of gather_sketches: ONE all-gather of equally sized blocks [n_local sizes | hashes | padding to
    cap] and the library's concat kernels (kssd_b200_sketch_concat_device) -- no host round trip for the shard sizes,
    no padded copy + cat.  cap: hashes per rank every rank agrees on beforehand (e.g. 1.25 x the expected sketch
    size of its genomes).  Returns (d_hashes ptr, d_genome_off ptr, total): library-owned device memory, or None
    when some rank's shard exceeded cap (the same answer on every rank: callers fall back to gather_sketches)."""
    n_local = sizes_local.numel()
    tot = hashes_local.numel()
    dev = hashes_local.device
    t0 = _tick("", time.perf_counter()) if TRACE is not None else 0.0
    send = torch.empty(n_local + cap, dtype=torch.int32, device=dev)
    send[:n_local] = sizes_local.to(torch.int32)
    send[n_local:n_local + min(tot, cap)] = hashes_local[:cap]
    recv = torch.empty(world * (n_local + cap), dtype=torch.int32, device=dev)
    t0 = _tick("gather.pack", t0)
    dist.all_gather_into_tensor(recv, send)
    t0 = _tick("gather.nccl", t0)
    ph, po, total, overflow = ctx.sketch_concat_device(recv.data_ptr(), n_local + cap, n_local, world)
    t0 = _tick("gather.concat", t0)
    if overflow:
        return None
    return ph, po, total  # the concat call synchronised the stream: recv may be freed


def build_index_sharded(dist, ctx, d_hashes, d_genome_off, n_genomes, total, rank, world):
    """Index build sharded by hash range + all-gather of the pieces over NCCL (north_star: "the reference index
    replicated by NCCL over NVLink"): rank r builds the CSR of hash range r from the sketches of all genomes
    (kssd_b200_index_build_range, ~1/world of the replicated build), the piece sizes are exchanged, blocks of the
    largest size are all-gathered and every rank adopts them (kssd_b200_index_adopt_parts).  Returns False when a
    piece could not be built -- the same answer on every rank -- and the caller builds the whole index locally
    (ctx.index_build_device)."""
    # a piece holds P ~ total/world postings and U <= P distinct hashes: 5 + 2U + P words; generous local capacity
    cap_words = 3 * (2 * (total // world) + 1024) + 8
    dev = torch.device("cuda", torch.cuda.current_device())
    t0 = _tick("", time.perf_counter()) if TRACE is not None else 0.0
    send = torch.empty(cap_words, dtype=torch.int32, device=dev)
    overflow, used = ctx.index_build_range(d_hashes, d_genome_off, n_genomes, rank, world, send.data_ptr(), cap_words)
    t0 = _tick("index.build_range", t0)
    mine = torch.tensor([-1 if overflow else used], dtype=torch.int64).to(dev, non_blocking=True)
    sizes = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(sizes, mine)
    sizes = sizes.cpu()
    t0 = _tick("index.sizes", t0)
    if int(sizes.min()) < 0:
        return False
    block_words = min(cap_words, (int(sizes.max()) + 1023) // 1024 * 1024)
    recv = torch.empty(world * block_words, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(recv, send[:block_words])
    t0 = _tick("index.nccl", t0)
    ok = not ctx.index_adopt_parts(recv.data_ptr(), block_words, world, d_genome_off, n_genomes)
    _tick("index.adopt", t0)
    return ok


def gather_sketches(dist, hashes_local, sizes_local, world):
    """All-gather of CSR sketch shards.  hashes_local: int32[total_local] (the u32 hashes, bit-cast), sizes_local:
    int64[n_local] per-genome sketch sizes; every rank holds the same number of genomes.  Returns
    (hashes int32[total], off int64[n_total+1]) in rank order, identical on every rank."""
    n_local = sizes_local.numel()
    dev = hashes_local.device
    if world == 1:
        off = torch.zeros(n_local + 1, dtype=torch.int64, device=dev)
        off[1:] = torch.cumsum(sizes_local, 0)
        return hashes_local, off
    sizes_all = torch.empty(world * n_local, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(sizes_all, sizes_local.contiguous())
    tots = sizes_all.view(world, n_local).sum(1).cpu()
    mx = int(tots.max())
    pad = torch.zeros(mx, dtype=torch.int32, device=dev)
    pad[:hashes_local.numel()] = hashes_local
    allh = torch.empty(world * mx, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(allh, pad)
    hashes = torch.cat([allh[i * mx:i * mx + int(tots[i])] for i in range(world)])
    off = torch.zeros(world * n_local + 1, dtype=torch.int64, device=dev)
    off[1:] = torch.cumsum(sizes_all, 0)
    return hashes, off


def replicate_index(dist, ctx, rank, root=0):
    """The reference index replicated by NCCL broadcast over NVLink (north_star): the root's CSR (distinct
    hashes, list starts, postings, reference sizes: kssd_b200_index_device_views) is broadcast array by array and adopted
    by every other rank's context (kssd_b200_index_adopt_device)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    v = ctx.index_device_views() if rank == root else None
    shape = torch.tensor([v["n_uniq"], v["total"], v["n_ref"]] if v else [0, 0, 0], dtype=torch.int64, device=dev)
    dist.broadcast(shape, root)
    n_uniq, total, n_ref = (int(x) for x in shape.tolist())

    def view(ptr, n, dtype, ts):
        class V:
            __cuda_array_interface__ = {"shape": (n,), "typestr": ts, "data": (ptr, False), "version": 2}
        return torch.as_tensor(V(), device=dev) if n else torch.zeros(0, dtype=dtype, device=dev)
    spec = [("d_uniq_hash", n_uniq, torch.int32, "<i4"), ("d_uniq_off", n_uniq + 1, torch.int64, "<i8"),
            ("d_dict", total, torch.int32, "<i4"), ("d_ref_sizes", n_ref, torch.int32, "<i4")]
    bufs = {}
    for name, n, dtype, ts in spec:
        t = view(v[name], n, dtype, ts).clone() if rank == root else torch.empty(n, dtype=dtype, device=dev)
        if n:
            dist.broadcast(t, root)
        bufs[name] = t
    if rank != root:
        torch.cuda.current_stream().synchronize()
        ctx.index_adopt_device(n_uniq, total, n_ref, bufs["d_uniq_hash"].data_ptr(), bufs["d_uniq_off"].data_ptr(),
                               bufs["d_dict"].data_ptr(), bufs["d_ref_sizes"].data_ptr())
    return n_uniq, total, n_ref
