"""GPU: the distance path beyond the golden cases -- device-side top-N (dist -N, src/dist.cpp:625-640), the loud refusal
of a triangle run whose queries are not the indexed sketches, replication of an index into another context
(kssd_b200_index_device_views / _index_adopt_device), tiles whose first-touch list overflows, and the cross-process
peer-memory gather (CUDA IPC) with two real processes."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import oracle
import util
from rabbitkssd_b200 import _lib, api, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _sorted(h):
    h = np.asarray(h).reshape(-1, 3)
    return h[np.lexsort((h[:, 1], h[:, 0]))].astype(np.int64)


def _exact_d(hits, qs, rs, kmer, containment=0):
    out = np.empty(len(hits))
    for i, (r, c, m) in enumerate(hits):
        out[i] = oracle.distance(int(m), int(rs[c]), int(qs[r]), kmer, containment)[1]
    return out


@pytest.mark.parametrize("n_top,max_dist,containment", [(1, 0.2, 0), (3, 0.2, 0), (5, 1.0, 0), (2, 0.15, 1), (40, 1.0, 0)])
def test_device_top_n_is_superset_of_exact_top_n(ctx, n_top, max_dist, containment):
    """With kssd_b200_set_max_neighbor the device reports, per query row, the n_top smallest distances (+ ties) among
    the cells with common > 0 that pass max_dist: a superset of what the reference's bounded heap keeps, and never
    more than the unselected hit list."""
    rh, roff = synth.synth_sketches(909, 60, 10, 300, 20)  # 600 refs, 2^20 hash space: many random common = 1 cells
    qh, qoff = synth.synth_sketches(909, 60, 4, 300, 20, survival_step=0.12)
    rs, qs = np.diff(roff).astype(np.int64), np.diff(qoff).astype(np.int64)
    nq = len(qs)
    ctx.index_build(rh, roff)
    ctx.set_emit_zero_cells(False)
    try:
        ctx.set_max_neighbor(0)
        full = _sorted(ctx.dist_rows(qh, qoff, False, 18, max_dist, containment))
        ctx.set_max_neighbor(n_top)
        sel = _sorted(ctx.dist_rows(qh, qoff, False, 18, max_dist, containment))
    finally:
        ctx.set_max_neighbor(0)
        ctx.set_emit_zero_cells(True)
    full_set = {tuple(x) for x in full}
    assert {tuple(x) for x in sel} <= full_set
    d_full = _exact_d(full, qs, rs, 18, containment)
    ok = d_full <= max_dist  # the device threshold is conservative: drop what the exact test drops
    full, d_full = full[ok], d_full[ok]
    sel_set = {tuple(x) for x in sel}
    n_rows_cut = 0
    for r in range(nq):
        m = full[:, 0] == r
        if not m.any():
            continue
        d = d_full[m]
        cells = full[m]
        kth = np.sort(d)[min(n_top, len(d)) - 1]
        must = cells[d <= kth]  # the top-N and everything tied with the N-th
        for x in must:
            assert tuple(x) in sel_set, (r, x)
        n_sel = sum(1 for x in sel if x[0] == r)
        assert n_sel >= min(n_top, len(d))
        if len(d) > n_top and (d > kth).any():
            n_rows_cut += n_sel < len(d)
    if n_top < 40:
        assert n_rows_cut > 0 and len(sel) < len(full_set)  # the selection really cut something


def test_triangle_needs_the_indexed_sketches(ctx):
    h, off = synth.synth_sketches(11, 20, 5, 200, 24)
    ctx.index_build(h, off)
    with pytest.raises(_lib.KssdError) as e:  # fewer queries than indexed genomes
        ctx.dist_rows(h[:int(off[50])], off[:51], True, 20, 0.1)
    assert e.value.code == -2
    # same count, other sketches (rows with more posting ranges than hashes are detected on the device)
    off2 = off.copy()
    sizes = np.diff(off).astype(np.int64)
    sizes[:] = 1
    off2 = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    with pytest.raises(_lib.KssdError) as e:
        ctx.dist_rows(h[:int(off2[-1])], off2, True, 20, 0.1)
    assert e.value.code == -2 and "indexed sketches" in str(e.value)
    # the context stays usable
    hits = ctx.dist_rows(h, off, True, 20, 0.1)
    hr, hc, hm = oracle.alldist(h, off, 20, 0.1)
    sz = np.diff(off).astype(np.int64)
    hits, _, _ = api.refine_hits(hits, sz, sz, 20, 0.1, True)
    assert np.array_equal(_sorted(hits), np.stack([hr, hc, hm], axis=1).astype(np.int64))


def test_dense_tile_first_touch_list_overflow(ctx):
    """One cluster of 300 near-identical sketches: every cell of a tile is non-zero, far more than the 4096-entry
    first-touch list of a tile holds -- the epilogue must fall back to scanning the counters and still report every cell."""
    h, off = synth.synth_sketches(77, 1, 300, 500, 28, survival_step=0.001)
    sizes = np.diff(off).astype(np.int64)
    ctx.index_build(h, off)
    hits = ctx.dist_rows(h, off, True, 20, 0.3)
    hits, _, _ = api.refine_hits(hits, sizes, sizes, 20, 0.3, True)
    hr, hc, hm = oracle.alldist(h, off, 20, 0.3)
    assert len(hr) == 300 * 299 // 2
    assert np.array_equal(_sorted(hits), np.stack([hr, hc, hm], axis=1).astype(np.int64))
    q = ctx.dist_rows(h, off, False, 20, 0.3)
    q, _, _ = api.refine_hits(q, sizes, sizes, 20, 0.3, False)
    dr, dc, dm = oracle.dist(h, off, h, off, 20, 0.3)
    assert np.array_equal(_sorted(q), np.stack([dr, dc, dm], axis=1).astype(np.int64))


def test_index_replicated_into_second_context(ctx):
    """kssd_b200_index_device_views + _index_adopt_device: what a host with one context per GPU does to replicate the
    reference index -- here both contexts sit on device 0."""
    rh, roff = synth.synth_sketches(5150, 40, 8, 250, 24)
    qh, qoff = synth.synth_sketches(5150, 40, 3, 250, 24, survival_step=0.1)
    rs, qs = np.diff(roff).astype(np.int64), np.diff(qoff).astype(np.int64)
    ctx.index_build(rh, roff)
    ctx.sync()
    v = ctx.index_device_views()
    c2 = api.Context(0)
    try:
        c2.index_adopt_device(v["n_uniq"], v["total"], v["n_ref"], v["d_uniq_hash"], v["d_uniq_off"], v["d_dict"], v["d_ref_sizes"])
        for a, b in zip(c2.index_export(), ctx.index_export()):
            assert np.array_equal(a, b)
        got = c2.dist_rows(qh, qoff, False, 18, 0.1)
        got, _, _ = api.refine_hits(got, qs, rs, 18, 0.1, False)
        dr, dc, dm = oracle.dist(qh, qoff, rh, roff, 18, 0.1)
        assert len(dr) > 50 and np.array_equal(_sorted(got), np.stack([dr, dc, dm], axis=1).astype(np.int64))
        tri = c2.dist_rows(rh, roff, True, 18, 0.1)
        tri, _, _ = api.refine_hits(tri, rs, rs, 18, 0.1, True)
        hr, hc, hm = oracle.alldist(rh, roff, 18, 0.1)
        assert np.array_equal(_sorted(tri), np.stack([hr, hc, hm], axis=1).astype(np.int64))
    finally:
        c2.close()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_peer_gather_across_processes(world):
    """The real cross-process gather: `world` processes (all on cuda:0 -- CUDA IPC works between processes of one device
    exactly as between devices), each sketches its own genomes, publishes the shard in its own buffer and pulls the others'
    through IPC mappings.  Every rank must see exactly the concatenation of the shards, over several steps (the double
    buffer flips), and the sharded index build pulled over the same mappings must equal the full build."""
    port = _free_port()
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE=str(world))
    procs = []
    for r in range(world):
        e = dict(env, RANK=str(r), LOCAL_RANK="0")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "peer_worker.py")], env=e, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out.decode(errors="replace"))
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r}:\n{o[-3000:]}"
        assert "PEER_OK" in o, f"rank {r}:\n{o[-3000:]}"


def test_row_restricted_index_equals_full_index():
    """kssd_b200_index_build_rows: for every block of alldist rows the restricted index must give exactly the hits of the
    full index for those rows, and hold fewer postings; rows outside the block and dist (rectangle) are refused."""
    from rabbitkssd_b200 import _lib
    h, off = synth.synth_sketches(991, 60, 8, 300, 28)
    n = len(off) - 1
    c = api.Context(0)
    try:
        c.set_params(10, 6, 3, util.shuffled_dim(10, 6, 3))
        c.index_build(h, off)
        full = c.dist_rows(h, off, True, 20, 0.1)
        full = full[np.lexsort((full[:, 1], full[:, 0]))]
        _, _, full_dict = c.index_export()
        assert len(full) > 500
        parts = []
        for (r0, r1) in [(0, 100), (100, 101), (101, 333), (333, n)]:
            c.index_build_rows(h, off, (r0, r1))
            _, _, d = c.index_export()
            assert len(d) < len(full_dict) and (r0 == 0 or d.min() >= r0)
            got = c.dist_rows(h, off, True, 20, 0.1, rows=(r0, r1))
            parts.append(got)
            with pytest.raises(_lib.KssdError):
                c.dist_rows(h, off, False, 20, 0.1, rows=(r0, r1))
            if r0 > 0:
                with pytest.raises(_lib.KssdError):
                    c.dist_rows(h, off, True, 20, 0.1, rows=(r0 - 1, r1))
        got = np.concatenate(parts)
        got = got[np.lexsort((got[:, 1], got[:, 0]))]
        assert np.array_equal(got, full)
        c.index_build_rows(h, off, (0, n))  # the whole triangle: a plain full build again
        _, _, d = c.index_export()
        assert np.array_equal(d, full_dict)
    finally:
        c.close()


@pytest.mark.parametrize("rows", [None, (1000, 3777)])
def test_packed_triangle_tiles_equal_oracle(ctx, rows):
    """alldist with one column tile packs the triangle: a tile holds only the columns right of its first row, so tiles
    further down take more rows (4 ... 32).  4,800 small sketches give tiles of every height; whole triangle and a row block."""
    h, off = synth.synth_sketches(7321, 800, 6, 48, 28)
    n = len(off) - 1
    sizes = np.diff(off).astype(np.int64)
    ctx.set_params(10, 6, 3, util.shuffled_dim(10, 6, 3))
    ctx.index_build(h, off)
    for D, cont in ((0.12, 0), (0.3, 1)):
        hits = ctx.dist_rows(h, off, True, 20, D, cont, rows=rows)
        hits, _, _ = api.refine_hits(hits, sizes, sizes, 20, D, True, cont)
        hr, hc, hm = oracle.alldist(h, off, 20, D, cont, rows=rows)
        o = np.lexsort((hits[:, 1], hits[:, 0]))
        assert len(hr) > 2000 and np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1)), (D, cont)


@pytest.mark.parametrize("dtype,bits", [(np.uint32, 28), (np.uint64, 40)])
def test_set_union_and_subtract(ctx, dtype, bits):
    """kssd_b200_set_union / _set_subtract (the arithmetic of `rabbit_kssd union` / `sub`, src/subCommand.cpp:307-794)
    against numpy: union = sorted distinct hashes; subtract = query sketches without the hashes of the reference set, order
    inside a genome kept; empty inputs and empty genomes included."""
    rng = np.random.default_rng(11)
    ref = rng.integers(0, 1 << bits, 300000, dtype=np.uint64).astype(dtype)
    ref[1000:2000] = ref[:1000]  # duplicates
    got = ctx.set_union(ref)
    assert got.dtype == dtype and np.array_equal(got, np.unique(ref))
    assert len(ctx.set_union(np.zeros(0, dtype))) == 0
    sizes = [0, 5, 1221, 0, 70000, 333]
    qs = []
    for i, n in enumerate(sizes):
        own = rng.integers(0, 1 << bits, n, dtype=np.uint64).astype(dtype)
        take = rng.choice(got, n // 2) if n else np.zeros(0, dtype)
        q = np.concatenate([own, take])
        rng.shuffle(q)  # unsorted on purpose: the order of the survivors must be kept
        qs.append(q)
    off = np.concatenate([[0], np.cumsum([len(q) for q in qs])]).astype(np.uint64)
    h, goff = ctx.set_subtract(np.concatenate(qs), off, got)
    for i, q in enumerate(qs):
        want = q[~np.isin(q, got)]
        assert np.array_equal(h[int(goff[i]):int(goff[i + 1])], want), i
    h, goff = ctx.set_subtract(np.concatenate(qs), off, np.zeros(0, dtype))  # nothing to subtract
    assert np.array_equal(h, np.concatenate(qs)) and np.array_equal(goff, off)
