"""CPU, world_size 2 and 3 over gloo: the host-side logic of the multi-GPU path (rabbitkssd_b200/multigpu.py) --
sketch-shard all-gather and the query-row partitions -- with the ORACLE standing in for the per-rank device work,
so that the N>1 orchestration bench.py uses on B200s is exercised without a GPU."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from rabbitkssd_b200 import multigpu, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_clusters, cluster, size, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h, off = synth.synth_sketches(4711, n_clusters, cluster, size, 24)
        n = len(off) - 1
        n_local = n // world
        g0, g1 = rank * n_local, (rank + 1) * n_local
        h_loc = torch.from_numpy(h[int(off[g0]):int(off[g1])].view(np.int32).copy())
        sizes_loc = torch.from_numpy(np.diff(off[g0:g1 + 1]).astype(np.int64))
        gh, goff = multigpu.gather_sketches(dist, h_loc, sizes_loc, world)
        ok_gather = np.array_equal(gh.numpy().view(np.uint32), h) and np.array_equal(goff.numpy().astype(np.uint64), off)
        # alldist: this rank's block of triangle rows against the full (replicated) index
        gh_np, goff_np = gh.numpy().view(np.uint32), goff.numpy().astype(np.uint64)
        b = multigpu.triangle_bounds(n, world)
        hr, hc, hm = oracle.alldist(gh_np, goff_np, 20, 0.08, rows=(b[rank], b[rank + 1]))
        tri = np.stack([hr, hc, hm], axis=1).astype(np.int64)
        # dist: a block of query rows (queries = other members of the same clusters)
        qh, qoff = synth.synth_sketches(4711, n_clusters, 3, size, 24, survival_step=0.1)
        rb = multigpu.rect_bounds(len(qoff) - 1, world)
        dr, dc, dm = oracle.dist(qh, qoff, gh_np, goff_np, 20, 0.1, rows=(rb[rank], rb[rank + 1]))
        rect = np.stack([dr, dc, dm], axis=1).astype(np.int64)
        outs = [None] * world
        dist.all_gather_object(outs, (ok_gather, tri, rect))
        if rank == 0:
            q.put(outs)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_row_sharded_alldist_and_dist_over_gloo(world):
    n_clusters, cluster, size = 12, 6, 200  # 72 genomes: divisible by 2 and 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_clusters, cluster, size, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(o[0] for o in outs), "gathered sketches differ from the unsharded ones"
    h, off = synth.synth_sketches(4711, n_clusters, cluster, size, 24)
    hr, hc, hm = oracle.alldist(h, off, 20, 0.08)
    want = np.stack([hr, hc, hm], axis=1).astype(np.int64)
    got = np.concatenate([o[1] for o in outs])
    got = got[np.lexsort((got[:, 1], got[:, 0]))]
    assert len(want) > 50 and np.array_equal(got, want)
    qh, qoff = synth.synth_sketches(4711, n_clusters, 3, size, 24, survival_step=0.1)
    dr, dc, dm = oracle.dist(qh, qoff, h, off, 20, 0.1)
    want = np.stack([dr, dc, dm], axis=1).astype(np.int64)
    got = np.concatenate([o[2] for o in outs])
    got = got[np.lexsort((got[:, 1], got[:, 0]))]
    assert len(want) > 20 and np.array_equal(got, want)


def test_partitions_cover_and_balance():
    for n in (1, 2, 7, 100, 10000, 100001):
        for parts in (1, 2, 3, 4, 8):
            b = multigpu.triangle_bounds(n, parts)
            assert b[0] == 0 and b[-1] == n and len(b) == parts + 1 and all(x <= y for x, y in zip(b, b[1:]))
            r = multigpu.rect_bounds(n, parts)
            assert r[0] == 0 and r[-1] == n and max(np.diff(r)) - min(np.diff(r)) <= 1
    b = multigpu.triangle_bounds(10000, 8)
    cells = [sum(10000 - 1 - i for i in range(b[p], b[p + 1])) for p in range(8)]
    assert max(cells) / (sum(cells) / 8) < 1.01


def test_partitions_with_row_cost():
    """row_cost > 0: every part gets the same cells + row_cost * rows; row_cost = 0 is the cells-only split."""
    n, parts, rc = 10000, 8, 6000.0
    b = multigpu.triangle_bounds(n, parts, row_cost=rc)
    assert b[0] == 0 and b[-1] == n and all(x <= y for x, y in zip(b, b[1:]))
    cost = [sum(n - 1 - i + rc for i in range(b[p], b[p + 1])) for p in range(parts)]
    assert max(cost) / (sum(cost) / parts) < 1.01
    assert multigpu.triangle_bounds(n, parts, row_cost=0.0) == multigpu.triangle_bounds(n, parts)
    for nn in (1, 2, 7, 100):
        for pp in (1, 2, 3, 8):
            bb = multigpu.triangle_bounds(nn, pp, row_cost=3.0)
            assert bb[0] == 0 and bb[-1] == nn and len(bb) == pp + 1 and all(x <= y for x, y in zip(bb, bb[1:]))
