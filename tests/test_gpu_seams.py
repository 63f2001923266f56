"""GPU: every seam of the sketch path -- the chunk borders of kssd_b200_sketch_scan_host (the call behind bench.py's
`e2e`), the launch pieces of sketch_scan_device, a record continued over two pinned batches (kssd_batch.first_emit),
the emit-buffer regrow + replay on repeat-rich input -- checked per genome for SET EQUALITY against the oracle, with
N runs, lowercase, record borders and retained k-mers placed to straddle each border (reference semantics:
src/sketch.cpp:487-532, the valid-run counter restarts at every record and at every non-ACGT byte)."""
import numpy as np
import pytest

import oracle
import util
from rabbitkssd_b200 import api, synth

pytestmark = pytest.mark.gpu

K, S, L = 10, 6, 3


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    c.set_params(K, S, L, util.shuffled_dim(K, S, L))
    yield c
    c.close()


def retained_kmers(k, s, l, n_want=4, seed=99):
    """2k-base strings that are retained k-mers on their own (one oracle call over many one-record genomes)."""
    n = 40000
    bases = synth.random_bases(seed, n * 2 * k).reshape(n, 2 * k)
    got = util.oracle_sketch(k, s, l, [[bases[i]] for i in range(n)])
    idx = [i for i in range(n) if len(got[i])][:n_want]
    assert len(idx) == n_want
    return [bases[i].copy() for i in idx]


def seam_genomes(chunk, n_chunks, kmers, with_qual=False, seed=1):
    """Three genomes packed into ~n_chunks * chunk bytes whose interesting features sit ON the chunk borders of the
    packed buffer: for border b (byte b*chunk of the batch) a retained k-mer placed to end 0..19 bytes after it, an N run
    ending exactly at it, a lowercase stretch across it, a record border at / next to it."""
    total = chunk * n_chunks
    g_len = [total // 2 - 64, total // 3, total - (total // 2 - 64) - total // 3 - 64]
    genomes, quals = [], []
    pos = 0
    rng = np.random.default_rng(seed)
    for gi, n in enumerate(g_len):
        seq = synth.random_bases(seed * 17 + gi, n)
        q = ((synth.splitmix64(seed * 31 + gi, n) % np.uint64(41)) + 33).astype(np.uint8)
        start = (pos + 15) // 16 * 16  # pack_batch aligns genomes to 16
        cuts = [0]
        first_border = (start + chunk - 1) // chunk
        for b in range(first_border, (start + n) // chunk + 1):
            o = b * chunk - start  # offset of the border inside this genome
            if o < 64 or o > n - 64:
                continue
            kind = b % 6
            km = kmers[b % len(kmers)]
            if kind == 0:    # retained k-mer straddling the border, every alignment over the borders of this kind
                e = o - 20 + 1 + (b // 6) % 20
                seq[e:e + 20] = km
                q[e:e + 20] = 73
            elif kind == 1:  # N run ending exactly at the border, retained k-mer right behind it
                seq[o - 9:o] = ord("N")
                seq[o:o + 20] = km
                q[o:o + 20] = 73
            elif kind == 2:  # lowercase across the border (valid bases)
                seq[o - 30:o + 30] |= 0x20
                seq[o - 10:o + 10] = km | 0x20
                q[o - 10:o + 10] = 73
            elif kind == 3:  # record border exactly at the chunk border: the k-mer before it ends at o-1, one starts at o
                cuts.append(o)
                seq[o - 20:o] = km
                seq[o:o + 20] = kmers[(b + 1) % len(kmers)]
                q[o - 20:o + 20] = 73
            elif kind == 4:  # record border one byte before / after
                cuts.append(o + (1 if (b // 6) % 2 else -1))
                seq[o - 25:o - 5] = km
                q[o - 25:o - 5] = 73
            else:            # low quality base inside a retained k-mer across the border (FASTQ), byte >= 0x80 nearby
                seq[o - 10:o + 10] = km
                q[o - 10:o + 10] = 73
                q[o - 3] = 34
                seq[o + 40] = 0xC1
        cuts.append(n)
        cuts = sorted(set(cuts))
        recs, rq = [], []
        for a, b_ in zip(cuts[:-1], cuts[1:]):
            # a record border consumes no byte in the oracle but one separator in the packed batch: drop one base
            recs.append(seq[a:b_ - 1] if b_ != n else seq[a:b_])
            rq.append(q[a:b_ - 1] if b_ != n else q[a:b_])
        genomes.append(recs)
        quals.append(rq)
        pos = start + n + 1
    return genomes, (quals if with_qual else None)


def _sets(h, goff):
    return [h[int(goff[g]):int(goff[g + 1])].astype(np.uint64) for g in range(len(goff) - 1)]


@pytest.mark.parametrize("fastq", [False, True])
@pytest.mark.parametrize("chunk", [1 << 16, 100000 // 16 * 16])
def test_scan_host_chunk_borders_equal_oracle_and_device(ctx, fastq, chunk):
    import torch
    kmers = retained_kmers(K, S, L)
    n_chunks = 9
    genomes, quals = seam_genomes(chunk, n_chunks, kmers, with_qual=fastq, seed=3 + fastq)
    Q, nmin = (35, 1) if fastq else (0, 1)  # qualities are uniform in 33..73: ~37 % of the k-mers survive -Q 35
    want = util.oracle_sketch(K, S, L, genomes, quals, Q, nmin)
    assert sum(len(w) for w in want) > (25 if fastq else 100)
    seq, qual, off = synth.pack_batch(genomes, quals)
    assert len(seq) > (n_chunks - 1) * chunk  # at least n_chunks - 1 borders inside the buffer
    n = len(genomes)
    seg_off = off.copy()
    seg_gen = np.arange(n, dtype=np.uint32)
    # (1) the host-streamed scan: chunks of `chunk` bytes with a 32-byte context prefix each
    ctx.sketch_begin(n, fastq, Q, nmin)
    ctx.scan_host(seq.ctypes.data, len(seq), seg_off, seg_gen, qual.ctypes.data if fastq else 0, chunk_bytes=chunk)
    got_host = _sets(*ctx.sketch_finish())
    # (2) the resident scan of the same bytes
    d_seq = torch.from_numpy(seq).cuda()
    d_qual = torch.from_numpy(qual).cuda() if fastq else None
    d_so = torch.from_numpy(seg_off.astype(np.int64)).cuda()
    d_sg = torch.from_numpy(seg_gen.astype(np.int32)).cuda()
    torch.cuda.synchronize()
    ctx.sketch_begin(n, fastq, Q, nmin)
    ctx.scan_device(d_seq.data_ptr(), len(seq), d_so.data_ptr(), d_sg.data_ptr(), n, d_qual.data_ptr() if fastq else 0)
    got_dev = _sets(*ctx.sketch_finish())
    for g in range(n):
        assert np.array_equal(got_host[g], want[g]), ("host", g)
        assert np.array_equal(got_dev[g], want[g]), ("device", g)
    # the planted k-mers are really there: the sketch of each lone k-mer is contained in some genome's sketch
    lone = util.oracle_sketch(K, S, L, [[k] for k in kmers])
    union = np.unique(np.concatenate(want))
    for x in lone:
        assert np.isin(x, union).all()


def test_scan_host_one_call_per_genome_and_tiny_chunks(ctx):
    """Several scan_host calls in one session (one genome each), chunk = 4 KiB: hundreds of borders."""
    fam = synth.family(606, 3, 200000, 0.02, n_contigs=4)
    fam = [[synth.sprinkle(c, 9 + gi * 5 + ci, n_runs=20, run_len=11, lowercase_frac=0.05) for ci, c in enumerate(g)]
           for gi, g in enumerate(fam)]
    want = util.oracle_sketch(K, S, L, fam)
    ctx.sketch_begin(len(fam))
    keep = []
    for gi, g in enumerate(fam):
        seq, _, off = synth.pack_batch([g])
        so = np.array([0, len(seq)], dtype=np.uint64)
        sg = np.array([gi], dtype=np.uint32)
        keep.append((seq, so, sg))  # buffers stay valid until sketch_finish (include/kssd_b200.h)
        ctx.scan_host(seq.ctypes.data, len(seq), so, sg, chunk_bytes=4096)
    got = _sets(*ctx.sketch_finish())
    for g in range(len(fam)):
        assert np.array_equal(got[g], want[g]), g


def test_record_continued_over_two_batches_first_emit(ctx):
    """kssd_batch.first_emit: a record split over two pinned batches.  The second batch starts with the last 2k-1 bytes of
    the first part (context only: k-mers ending before first_emit are not emitted again) -- what a big-file parser
    (RabbitFX consumer, src/sketch.cpp:173-238) does when a chromosome does not fit one batch."""
    kmers = retained_kmers(K, S, L, n_want=3, seed=7)
    ctxlen = 2 * K - 1
    rec = synth.random_bases(4040, 300000)
    for cut, km in zip((100000, 200000), kmers):  # retained k-mers straddling both cuts, 7 bytes before the cut
        rec[cut - 7:cut - 7 + 20] = km
    rec[99990:99993] = ord("N")                    # an N run inside the context of the first cut
    other = synth.random_bases(4041, 50000)
    want = util.oracle_sketch(K, S, L, [[rec], [other]])
    ctx.sketch_begin(2)
    parts = [(0, 100000), (100000, 200000), (200000, 300000)]
    for pi, (a, b) in enumerate(parts):
        pre = ctxlen if pi else 0
        body = rec[a - pre:b]
        if pi == 1:  # the middle batch also carries the whole second genome behind the record part
            seq = np.concatenate([body, [10], other, [10]]).astype(np.uint8)
            seg_off = np.array([0, len(body) + 1, len(seq)], dtype=np.uint64)
            seg_gen = np.array([0, 1], dtype=np.uint32)
        else:
            seq = np.concatenate([body, [10]]).astype(np.uint8) if pi == 2 else body.copy()
            seg_off = np.array([0, len(seq)], dtype=np.uint64)
            seg_gen = np.array([0], dtype=np.uint32)
        ctx.submit_host(seq, seg_off, seg_gen, first_emit=pre)
    got = _sets(*ctx.sketch_finish())
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    # negative control: without the context prefix the k-mers across the cuts are lost
    ctx.sketch_begin(2)
    for a, b in parts:
        seq = np.concatenate([rec[a:b], [10]]).astype(np.uint8)
        ctx.submit_host(seq, np.array([0, len(seq)], dtype=np.uint64), np.array([0], dtype=np.uint32))
    lost = _sets(*ctx.sketch_finish())[0]
    assert len(lost) < len(want[0]) or not np.array_equal(lost, want[0])


def _retained_repeat_unit(k, s, l, seed=5):
    """A 2k-base unit whose tandem repeat emits (one of its rotations is a retained k-mer)."""
    for t in range(5000):
        unit = synth.random_bases(seed * 100003 + t, 2 * k)
        if len(util.oracle_sketch(k, s, l, [[np.tile(unit, 4)]])[0]):
            return unit
    raise AssertionError("no retained repeat unit found")


@pytest.mark.parametrize("path", ["batches", "device", "host"])
def test_emit_overflow_regrows_and_replays(ctx, path):
    """400 Mbp of one tandem repeat whose k-mer is retained: one emit per period = 20 M keys against an emit capacity of
    8 x bases / 4096 + 4 Mi = 5 M.  The reference's unordered_set simply holds one hash (src/sketch.cpp:526-529); the
    library must return the same sketch (grow the buffer, run the affected scans again), not KSSD_E_OVERFLOW."""
    import torch
    unit = _retained_repeat_unit(K, S, L)
    tail = synth.random_bases(912, 300000)
    small = [[np.tile(unit, 500), tail]]
    want_rep = util.oracle_sketch(K, S, L, small)[0]           # hashes of the repeat + the tail
    want_tail = util.oracle_sketch(K, S, L, [[tail]])[0]
    # four genomes = four pinned batches on the batch path: the overflowing one is the second, so both the per-batch cursor
    # snapshot (retire_stage) and the replay of two unverified batches are exercised
    genomes = [[tail[:100000]], [np.tile(unit, 20_000_000), tail], [tail[:50000]], [tail[:70000]]]
    want_small = util.oracle_sketch(K, S, L, [genomes[0], genomes[2], genomes[3]])
    seq, _, off = synth.pack_batch(genomes)
    n = len(genomes)
    if path == "batches":
        h, goff = ctx.sketch(seq, off, batch_bytes=60000)
    else:
        seg_gen = np.arange(n, dtype=np.uint32)
        ctx.sketch_begin(n)
        if path == "device":
            d_seq = torch.from_numpy(seq).cuda()
            d_so = torch.from_numpy(off.astype(np.int64)).cuda()
            d_sg = torch.from_numpy(seg_gen.astype(np.int32)).cuda()
            torch.cuda.synchronize()
            ctx.scan_device(d_seq.data_ptr(), len(seq), d_so.data_ptr(), d_sg.data_ptr(), n)
        else:
            ctx.scan_host(seq.ctypes.data, len(seq), off, seg_gen, chunk_bytes=96 << 20)
        h, goff = ctx.sketch_finish()
    got = _sets(h, goff)
    assert len(want_rep) > len(want_tail)  # the repeat contributes at least one hash
    assert np.array_equal(got[1], want_rep)
    for g, w in zip((0, 2, 3), want_small):
        assert np.array_equal(got[g], w), g
    # the context stays usable, and a FASTQ count filter sees every occurrence exactly once after a replay
    reads = [np.tile(unit, 5)] * 7 + [tail[:3000]]
    quals = [np.full(len(r), 70, np.uint8) for r in reads]
    for nmin in (1, 7 * 4, 7 * 5 + 1):
        w = util.oracle_sketch(K, S, L, [reads], [quals], 0, nmin)[0]
        s2, q2, o2 = synth.pack_batch([reads], [quals])
        h2, g2 = ctx.sketch(s2, o2, q2, 0, nmin)
        assert np.array_equal(h2.astype(np.uint64), w), nmin


def test_fastq_count_filter_survives_overflow_replay():
    """-n counts must not double (or vanish) when scans are undone and run again: 130,000 reads of a retained tandem repeat
    emit ~6.4 M keys against a capacity of ~4.5 M, through several pinned batches; thresholds just below / above the number
    of occurrences must give the oracle's answer (the count filter of src/sketch.cpp:704-706 sees every occurrence once)."""
    unit = _retained_repeat_unit(K, S, L)
    c = api.Context(0)
    try:
        c.set_params(K, S, L, util.shuffled_dim(K, S, L))
        read = np.tile(unit, 50)  # 1000 bases: ~49 occurrences of every rotation's k-mer
        n_reads = 130000
        reads = [read] * n_reads
        quals = [np.full(len(read), 70, np.uint8)] * n_reads
        seq, qual, off = synth.pack_batch([reads], [quals])
        sizes = []
        for nmin in (n_reads * 49, n_reads * 49 + 1, n_reads * 50 + 1):
            want = util.oracle_sketch(K, S, L, [reads], [quals], 0, nmin)[0]
            h, goff = c.sketch(seq, off, qual, 0, nmin, batch_bytes=32 << 20)
            assert np.array_equal(h.astype(np.uint64), want), nmin
            sizes.append(len(want))
        assert sizes[0] >= 1 and sizes[-1] == 0  # the thresholds bracket the true occurrence counts
    finally:
        c.close()
