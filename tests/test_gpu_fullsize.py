"""GPU: BASELINE.json configs 3-5 at (or near) their full shape -- the parts of the path that small fixtures do not
reach: column-tiled counters (n_ref > 98,304), > 2^30-byte inputs cut into several launches, the G = 4 filter of
s = 7 shuffle files, 50,000-reference posting lists.  Checked against the oracle (fully where the oracle finishes in
seconds, on sampled row blocks otherwise) plus size-independent properties (partition sums, sortedness)."""
import numpy as np
import pytest

import oracle
import util
from rabbitkssd_b200 import api, multigpu, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _sorted_hits(h):
    return h[np.lexsort((h[:, 1], h[:, 0]))].astype(np.int64)


def test_config3_dist_1000_queries_vs_50000_refs_L3K9(ctx):
    rh, roff = synth.synth_sketches(3003, 5000, 10, 1221, 24)
    qh, qoff = synth.synth_sketches(3003, 5000, 3, 1221, 24, survival_step=0.1)
    qoff = qoff[:1001]  # the first 1000 queries (members of the first 334 clusters)
    qh = qh[:int(qoff[-1])]
    rs, qs = np.diff(roff).astype(np.int64), np.diff(qoff).astype(np.int64)
    assert len(rs) == 50000 and len(qs) == 1000
    ctx.index_build(rh, roff)
    b = multigpu.rect_bounds(1000, 8)
    parts = [ctx.dist_rows(qh, qoff, False, 18, 0.1, 0, rows=(b[i], b[i + 1])) for i in range(8)]
    hits, _, _ = api.refine_hits(np.concatenate(parts), qs, rs, 18, 0.1, False)
    dr, dc, dm = oracle.dist(qh, qoff, rh, roff, 18, 0.1)
    want = np.stack([dr, dc, dm], axis=1).astype(np.int64)
    assert len(want) > 5000
    assert np.array_equal(_sorted_hits(hits), want)


def test_config5_alldist_100000_sketches_column_tiles(ctx):
    n_clusters, cluster, size = 10000, 10, 400
    h, off = synth.synth_sketches(5005, n_clusters, cluster, size, 28)
    n = len(off) - 1
    sizes = np.diff(off).astype(np.int64)
    assert n == 100000
    ctx.index_build(h, off)
    # size-independent property: any partition of the rows gives the same hit set as one call
    full = ctx.dist_rows(h, off, True, 20, 0.05)
    b = multigpu.triangle_bounds(n, 8)
    parts = [ctx.dist_rows(h, off, True, 20, 0.05, rows=(b[p], b[p + 1])) for p in range(8)]
    assert np.array_equal(_sorted_hits(full), _sorted_hits(np.concatenate(parts)))
    kept, _, _ = api.refine_hits(full, sizes, sizes, 20, 0.05, True)
    assert len(kept) > 100000
    assert np.all(kept[:, 1] > kept[:, 0]) and np.all(kept[:, 1] // cluster == kept[:, 0] // cluster)
    # sampled row blocks against the oracle (first rows see both column tiles; last rows only the second)
    kept = _sorted_hits(kept)
    for lo, hi in [(0, 400), (49900, 50300), (98200, 98600), (99700, 100000)]:
        hr, hc, hm = oracle.alldist(h, off, 20, 0.05, rows=(lo, hi))
        want = np.stack([hr, hc, hm], axis=1).astype(np.int64)
        got = kept[(kept[:, 0] >= lo) & (kept[:, 0] < hi)]
        assert len(want) > 100 and np.array_equal(got, want), (lo, hi)


def test_config4_shape_multi_launch_L4K10(ctx):
    """2.4 Gbp in 24 'genomes' of 100 Mbp, 3 contigs each, resident in HBM: the scan is cut into three launches
    (positions are 32-bit inside a launch) with 32-byte context overlap; s = 7 selects the G = 4 filter."""
    import torch
    k, s, l = 10, 7, 4
    dim = util.shuffled_dim(k, s, l)
    ctx.set_params(k, s, l, dim)
    n_g, glen, n_contig = 24, 100_000_000, 3
    stride = glen + n_contig  # one separator after every contig
    gen = torch.Generator(device="cuda")
    gen.manual_seed(404)
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device="cuda")
    buf = torch.empty(n_g * stride, dtype=torch.uint8, device="cuda")
    clen = glen // n_contig
    for g in range(n_g):
        codes = torch.randint(0, 4, (glen,), dtype=torch.uint8, device="cuda", generator=gen)
        asc = lut[codes.long()]
        o = g * stride
        cuts = [0, clen, 2 * clen, glen]
        for c in range(n_contig):
            seg = asc[cuts[c]:cuts[c + 1]]
            buf[o:o + len(seg)] = seg
            buf[o + len(seg)] = 10  # '\n'
            o += len(seg) + 1
    # a few N runs and lowercase stretches across launch borders
    for pos in (1 << 30, (1 << 30) - 7, 2 * (1 << 30) + 3):
        buf[pos - 40:pos - 30] = ord("N")
        buf[pos + 100:pos + 400] |= 0x20
    seg_off = (np.arange(n_g + 1, dtype=np.uint64) * np.uint64(stride))
    d_so = torch.from_numpy(seg_off.astype(np.int64)).cuda()
    d_sg = torch.arange(n_g, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.sketch_begin(n_g)
    ctx.scan_device(buf.data_ptr(), buf.numel(), d_so.data_ptr(), d_sg.data_ptr(), n_g)
    got_h, got_off = ctx.sketch_finish()
    host = buf.cpu().numpy()
    del buf
    # oracle: records = contigs (split at the separators)
    p = oracle.init_params(k, s, l)
    rec_off, rec_first = [0], [0]
    seqs = []
    for g in range(n_g):
        o = g * stride
        cuts = [0, clen, 2 * clen, glen]
        for c in range(n_contig):
            ln = cuts[c + 1] - cuts[c]
            seqs.append(host[o:o + ln])
            rec_off.append(rec_off[-1] + ln)
            o += ln + 1
        rec_first.append(len(rec_off) - 1)
    want = oracle.sketch_batch(p, dim, np.concatenate(seqs), np.asarray(rec_off, np.uint64), np.asarray(rec_first, np.uint64),
                               out_cap=200000)
    for g in range(n_g):
        a = got_h[int(got_off[g]):int(got_off[g + 1])].astype(np.uint64)
        assert abs(len(a) - glen / 65536) < 200
        assert np.array_equal(a, want[g]), g
