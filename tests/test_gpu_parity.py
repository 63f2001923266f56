"""GPU: the CUDA path through the C-ABI against (a) the reference binary's golden fixtures and (b) the oracle on
seeded inputs -- bit-exact hash sets, index, intersection counts and passing pair sets."""
import numpy as np
import pytest

import oracle
import util
from rabbitkssd_b200 import api, synth

pytestmark = pytest.mark.gpu
CASES = sorted(util.golden_meta().keys())  # includes the 64-bit hash configuration (fa_k12s6l3: half_k - drlevel = 9)


# every test runs on both implementations of the sorts: the shared-memory segment sorts (default) and the
# device-wide radix sort they fall back to
@pytest.fixture(scope="module", params=["segsort", "radix"])
def ctx(request):
    c = api.Context(0)
    c.set_force_radix(request.param == "radix")
    yield c
    c.close()


def _gpu_sketch(ctx, k, s, l, genomes, quals=None, least_qual=0, least_num_kmer=1, batch_bytes=None):
    ctx.set_params(k, s, l, util.shuffled_dim(k, s, l))
    seq, qual, off = synth.pack_batch(genomes, quals)
    h, goff = ctx.sketch(seq, off, qual, least_qual, least_num_kmer, batch_bytes)
    return [h[int(goff[g]):int(goff[g + 1])] for g in range(len(genomes))], h, goff


@pytest.mark.parametrize("case", CASES)
def test_sketch_equals_reference_golden(ctx, case):
    m = util.golden_meta()[case]
    c = m["recipe"]
    names, genomes, quals = util.case_genomes(case)
    sets, _, _ = _gpu_sketch(ctx, c["k"], c["s"], c["l"], genomes, quals, c.get("Q", 0), c.get("n", 1))
    z = util.golden_npz(case)
    for n, s in zip(names, sets):
        assert np.array_equal(s.astype(np.uint64), z["set_" + n]), n


@pytest.mark.parametrize("case", CASES)
def test_index_and_distances_equal_reference_golden(ctx, case):
    m = util.golden_meta()[case]
    c = m["recipe"]
    z = util.golden_npz(case)
    names = m["names"]
    sets = [z["set_" + n] for n in names]
    wide = util.use64(case)
    h, off = util.flatten_sets(sets, np.uint64 if wide else np.uint32)
    sizes = np.diff(off).astype(np.int64)
    ctx.set_params(c["k"], c["s"], c["l"], util.shuffled_dim(c["k"], c["s"], c["l"]))
    ctx.index_build(h, off)
    uh, uc, dict_ = ctx.index_export()
    if wide:
        # sparse .index in the iteration order of the reference's hash map (src/sketch.cpp:937-958): same MAPPING hash ->
        # posting list; ours comes out in ascending hash order, which is what the oracle builds on the hashes' ranks
        assert uh.dtype == np.uint64 and np.all(np.diff(uh.astype(np.float64)) > 0)
        want, start = {}, 0
        for hv, cnt in zip(z["index_hash"], z["index_count"]):
            want[int(hv)] = sorted(int(x) for x in z["dict"][start:start + cnt])  # fixture ids are in `names` order, lists in the reference's push order
            start += cnt
        got, start = {}, 0
        for hv, cnt in zip(uh, uc):
            got[int(hv)] = [int(x) for x in dict_[start:start + cnt]]
            start += cnt
        assert got == want
        rk, roff, _ = util.compress_hashes(sets)
        o_dict, _, _, _ = oracle.index_build(rk, roff)
    else:
        assert np.array_equal(uh.astype(np.uint64), z["index_hash"])
        assert np.array_equal(uc, z["index_count"])
        o_dict, _, _, _ = oracle.index_build(h, off)
    assert np.array_equal(dict_, o_dict)
    for tag in ["alldist", "alldist_M1", "dist", "dist_M1", "dist_D1"]:
        cont = 1 if tag.endswith("M1") else 0
        D = 1.0 if tag.endswith("D1") else c["D"]
        tri = tag.startswith("alldist")
        hits = ctx.dist_rows(h, off, tri, 2 * c["k"], D, cont)
        hits, jorc, d = api.refine_hits(hits, sizes, sizes, 2 * c["k"], D, tri, cont)
        got = []
        for (r, cl, cm), j_, d_ in zip(hits, jorc, d):
            if tri:
                got.append((names[cl], names[r], int(cm), int(sizes[r]), int(sizes[cl]), "%f" % j_, "%f" % d_))
            else:
                got.append((names[r], names[cl], int(cm), int(sizes[cl]), int(sizes[r]), "%f" % j_, "%f" % d_))
        want = sorted(tuple(x) for x in m["lines"][tag])
        assert sorted(got) == want, tag


def test_index_load_from_reference_files_layout(ctx):
    """index_load takes the dense count array + postings exactly as the reference's .index/.dict hold them."""
    m = util.golden_meta()["fa_k9s6l3"]
    z = util.golden_npz("fa_k9s6l3")
    sets = [z["set_" + n] for n in m["names"]]
    h, off = util.flatten_sets(sets)
    sizes = np.diff(off).astype(np.uint32)
    counts = np.zeros(m["hash_size"], dtype=np.uint32)
    counts[z["index_hash"].astype(np.int64)] = z["index_count"]
    o_dict, _, _, _ = oracle.index_build(h, off)
    ctx.index_load(counts, o_dict, sizes)
    hits = ctx.dist_rows(h, off, True, 18, 0.15)
    hr, hc, hm = oracle.alldist(h, off, 18, 0.15)
    hits, _, _ = api.refine_hits(hits, sizes, sizes, 18, 0.15, True)
    o = np.lexsort((hits[:, 1], hits[:, 0]))
    assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1))


@pytest.mark.parametrize("k,s,l,length,n", [(10, 6, 3, 1200000, 6), (9, 6, 3, 700000, 5), (8, 5, 2, 90000, 7)])
def test_sketch_equals_oracle_random(ctx, k, s, l, length, n):
    fam = synth.family(777 + k * 10 + l, n, length, 0.03, n_contigs=3)
    fam = [[synth.sprinkle(c, 5 + gi * 7 + ci, n_runs=8, run_len=7, lowercase_frac=0.03) for ci, c in enumerate(g)]
           for gi, g in enumerate(fam)]
    want = util.oracle_sketch(k, s, l, fam)
    # small batch size => several pinned batches through the double buffer
    got, _, _ = _gpu_sketch(ctx, k, s, l, fam, batch_bytes=max(length * 2, 1 << 20))
    for g, (a, b) in enumerate(zip(got, want)):
        assert np.array_equal(a.astype(np.uint64), b), g


def test_sketch_edge_cases(ctx):
    k, s, l = 10, 6, 3
    base = synth.random_bases(4242, 300000)
    genomes = [
        [np.zeros(0, np.uint8)],                                   # empty genome
        [base[:19]],                                               # shorter than one k-mer
        [base[:20]],                                               # exactly one k-mer
        [np.full(5000, ord("N"), np.uint8)],                       # nothing valid
        [base[:100000], np.zeros(0, np.uint8), base[100000:100019], base[100019:]],  # ragged records
        [np.concatenate([base[:50000], np.frombuffer(b"\xff\x80\x00*", np.uint8), base[50000:]])],  # bytes >= 0x80
        [base],
    ]
    want = util.oracle_sketch(k, s, l, genomes)
    got, _, _ = _gpu_sketch(ctx, k, s, l, genomes)
    for g, (a, b) in enumerate(zip(got, want)):
        assert np.array_equal(a.astype(np.uint64), b), g
    assert len(got[0]) == 0 and len(got[1]) == 0 and len(got[3]) == 0


def test_fastq_quality_and_count_filter(ctx):
    k, s, l = 10, 6, 3
    seq = synth.random_bases(99, 400000)
    n_reads, rl = 9000, 150
    r = synth.splitmix64(1234, n_reads)
    starts = (r % np.uint64(len(seq) - rl)).astype(np.int64)
    reads = [seq[a:a + rl] for a in starts]
    quals = [((synth.splitmix64(555 + i, rl) % np.uint64(41)) + 33).astype(np.uint8) for i in range(n_reads)]
    for Q, n in [(0, 1), (40, 1), (38, 2), (50, 3)]:
        want = util.oracle_sketch(k, s, l, [reads], [quals], Q, n)
        got, _, _ = _gpu_sketch(ctx, k, s, l, [reads], [quals], Q, n)
        assert np.array_equal(got[0].astype(np.uint64), want[0]), (Q, n)


@pytest.mark.parametrize("containment", [0, 1])
def test_alldist_equals_oracle_synthetic_sketches(ctx, containment):
    h, off = synth.synth_sketches(31337, 150, 10, 300, 28)
    sizes = np.diff(off).astype(np.int64)
    ctx.index_build(h, off)
    hits = ctx.dist_rows(h, off, True, 20, 0.05, containment)
    hits, _, _ = api.refine_hits(hits, sizes, sizes, 20, 0.05, True, containment)
    hr, hc, hm = oracle.alldist(h, off, 20, 0.05, containment)
    o = np.lexsort((hits[:, 1], hits[:, 0]))
    assert len(hr) > 100
    assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1))


def test_dist_row_sharding_equals_oracle(ctx):
    rh, roff = synth.synth_sketches(2024, 120, 10, 250, 24)
    qh, qoff = synth.synth_sketches(2024, 120, 3, 250, 24, survival_step=0.1)  # same roots, other members
    rs, qs = np.diff(roff).astype(np.int64), np.diff(qoff).astype(np.int64)
    ctx.index_build(rh, roff)
    nq = len(qoff) - 1
    parts = []
    for lo, hi in [(0, nq // 3), (nq // 3, nq // 2), (nq // 2, nq)]:  # three "GPUs"
        parts.append(ctx.dist_rows(qh, qoff, False, 18, 0.08, 0, rows=(lo, hi)))
    hits = np.concatenate(parts)
    hits, _, _ = api.refine_hits(hits, qs, rs, 18, 0.08, False)
    hr, hc, hm = oracle.dist(qh, qoff, rh, roff, 18, 0.08)
    o = np.lexsort((hits[:, 1], hits[:, 0]))
    assert len(hr) > 50
    assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1))


def test_full_size_properties_c1(ctx):
    """BASELINE config 1 shape (100 x 5 Mbp is too slow for the oracle in a unit test: 12 x 5 Mbp here, oracle on 3)
    plus size-independent properties: sorted, unique, idempotent, reverse-complement invariant."""
    k, s, l = 10, 6, 3
    fam = synth.family(20261104, 12, 5000000, 0.04, n_contigs=2)
    sets, h, goff = _gpu_sketch(ctx, k, s, l, fam)
    for x in sets:
        assert np.all(np.diff(x.astype(np.int64)) > 0)
        assert abs(len(x) - 5000000 / 4096) < 200
    want = util.oracle_sketch(k, s, l, fam[:3])
    for a, b in zip(sets[:3], want):
        assert np.array_equal(a.astype(np.uint64), b)
    comp = np.zeros(256, np.uint8)
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    rc = [[comp[c[::-1]].copy() for c in g] for g in fam[:4]]
    sets_rc, _, _ = _gpu_sketch(ctx, k, s, l, rc)
    for a, b in zip(sets[:4], sets_rc):
        assert np.array_equal(a, b)  # canonical k-mers: strand does not matter
    sets2, _, _ = _gpu_sketch(ctx, k, s, l, fam[:4])
    for a, b in zip(sets[:4], sets2):
        assert np.array_equal(a, b)


def test_sort_fallbacks_when_shared_memory_is_too_small():
    """A genome with more emitted keys than one CTA's shared memory holds (> 57,344) takes the radix path on its own;
    an index whose hashes crowd one bucket (> 8,192 pairs) does too.  Both must equal the oracle."""
    c = api.Context(0)
    try:
        k, s, l = 8, 5, 2
        g = [[synth.random_bases(8181, 16000000)], [synth.random_bases(8182, 200000)]]
        c.set_params(k, s, l, util.shuffled_dim(k, s, l))
        seq, qual, off = synth.pack_batch(g)
        h, goff = c.sketch(seq, off)
        assert goff[1] > 57344
        want = util.oracle_sketch(k, s, l, g)
        for i in range(2):
            assert np.array_equal(h[int(goff[i]):int(goff[i + 1])].astype(np.uint64), want[i]), i
        rng = np.random.default_rng(5)
        sets = [np.unique(rng.integers(0, 4096, 400)).astype(np.uint64) for _ in range(40)]
        sh, soff = util.flatten_sets(sets)
        assert soff[-1] > 8192
        c.index_build(sh, soff)
        uh, uc, dict_ = c.index_export()
        o_dict, o_uh, o_uc, _ = oracle.index_build(sh, soff)
        assert np.array_equal(dict_, o_dict)
        sizes = np.diff(soff).astype(np.int64)
        hits = c.dist_rows(sh, soff, True, 16, 0.3)
        hits, _, _ = api.refine_hits(hits, sizes, sizes, 16, 0.3, True)
        hr, hc, hm = oracle.alldist(sh, soff, 16, 0.3)
        o = np.lexsort((hits[:, 1], hits[:, 0]))
        assert len(hr) > 10
        assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1))
    finally:
        c.close()


def test_sketch_concat_device_matches_numpy():
    """Multi-GPU gather epilogue (kssd_b200_sketch_concat_device): blocks [sizes | hashes | padding] of 3 simulated ranks
    -> concatenated CSR; a shard larger than its block reports overflow."""
    import torch
    rng = np.random.default_rng(7)
    world, n_local, cap = 3, 5, 4000
    c = api.Context(0)
    try:
        sizes = rng.integers(0, 700, size=(world, n_local)).astype(np.uint32)
        sizes[1, 2] = 0
        parts = np.zeros((world, n_local + cap), dtype=np.uint32)
        want = []
        for r in range(world):
            h = rng.integers(0, 2 ** 28, size=int(sizes[r].sum()), dtype=np.uint32)
            parts[r, :n_local] = sizes[r]
            parts[r, n_local:n_local + len(h)] = h
            parts[r, n_local + len(h):] = 0xDEADBEEF
            want.append(h)
        want = np.concatenate(want)
        d = torch.from_numpy(parts.view(np.int32)).cuda()
        ph, po, total, overflow = c.sketch_concat_device(d.data_ptr(), n_local + cap, n_local, world)
        assert not overflow and total == len(want)
        class View:  # raw device pointer -> torch, no copy
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}
        got_h = torch.as_tensor(View(ph, total, "<i4"), device="cuda")
        got_o = torch.as_tensor(View(po, world * n_local + 1, "<i8"), device="cuda")
        assert np.array_equal(got_h.cpu().numpy().view(np.uint32), want)
        assert np.array_equal(got_o.cpu().numpy(), np.concatenate([[0], np.cumsum(sizes.reshape(-1).astype(np.int64))]))
        # rank 1 claims more hashes than a block holds
        parts[1, 0] = cap + 1
        d = torch.from_numpy(parts.view(np.int32)).cuda()
        _, _, _, overflow = c.sketch_concat_device(d.data_ptr(), n_local + cap, n_local, world)
        assert overflow
    finally:
        c.close()


@pytest.mark.parametrize("parts", [1, 2, 3, 8])
def test_index_sharded_by_hash_range_equals_full_build(parts):
    """Multi-GPU index path on one GPU: every hash-range piece (kssd_b200_index_build_range) packed into its block, the
    blocks adopted (kssd_b200_index_adopt_parts) -> the same CSR as the full build, and the same alldist hits."""
    import torch
    k, s, l = 10, 6, 3
    fams = [g for f in range(6) for g in synth.family(100 + f, 5, 120000, 0.03, n_contigs=2)]
    c = api.Context(0)
    try:
        c.set_params(k, s, l, util.shuffled_dim(k, s, l))
        seq, _, off = synth.pack_batch(fams)
        h, goff = c.sketch(seq, off)
        n = len(fams)
        c.index_build(h, goff)
        want = [np.array(x, copy=True) for x in c.index_export()]
        want_hits = c.dist_rows(h, goff, True, 2 * k, 0.2)
        d_h = torch.from_numpy(h.view(np.int32)).cuda()
        d_o = torch.from_numpy(goff.astype(np.int64)).cuda()
        cap = 3 * len(h) + 64
        blocks = torch.zeros(parts * cap, dtype=torch.int32, device="cuda")
        used_all = []
        for p in range(parts):
            ov, used = c.index_build_range(d_h.data_ptr(), d_o.data_ptr(), n, p, parts, blocks.data_ptr() + p * cap * 4, cap)
            assert not ov and 5 <= used <= cap
            used_all.append(used)
        # the pieces hold comparable numbers of postings (ranges follow the hash density)
        posts = [int(blocks[p * cap + 1]) for p in range(parts)]
        assert sum(posts) == len(h)
        if parts > 1:
            assert max(posts) < 1.5 * len(h) / parts
        # ship only the used prefix, as the all-gather does
        stride = (max(used_all) + 1023) // 1024 * 1024
        packed = torch.cat([blocks[p * cap:p * cap + stride] for p in range(parts)]) if stride <= cap else blocks
        assert not c.index_adopt_parts(packed.data_ptr(), min(stride, cap), parts, d_o.data_ptr(), n)
        got = c.index_export()
        for a, b in zip(got, want):
            assert np.array_equal(np.asarray(a), np.asarray(b))
        hits = c.dist_rows(h, goff, True, 2 * k, 0.2)
        key = lambda x: x[np.lexsort((x[:, 2], x[:, 1], x[:, 0]))]
        assert np.array_equal(key(np.asarray(hits)), key(np.asarray(want_hits)))
        # a block that is too small reports overflow, and adopting it reports overflow too
        small = torch.zeros(16, dtype=torch.int32, device="cuda")
        ov, _ = c.index_build_range(d_h.data_ptr(), d_o.data_ptr(), n, 0, 1, small.data_ptr(), 16)
        assert ov and c.index_adopt_parts(small.data_ptr(), 16, 1, d_o.data_ptr(), n)
    finally:
        c.close()


def test_peer_gather_single_rank():
    """Peer-memory gather with world = 1 (the own buffer stands in for the peers; the IPC mappings themselves need
    two processes and are exercised by bench.py --gpus N): publish + pull returns the session's sketch, twice
    (double buffer flip), and a shard larger than the buffer reports overflow."""
    import torch
    k, s, l = 10, 6, 3
    fams = synth.family(31, 6, 150000, 0.03, n_contigs=2)
    c = api.Context(0)
    try:
        c.set_params(k, s, l, util.shuffled_dim(k, s, l))
        seq, _, off = synth.pack_batch(fams)
        c.peer_connect(c.peer_init(len(fams), 4096), 1, 0)

        class View:
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}
        for _ in range(3):
            h, goff = c.sketch(seq, off)
            h, goff = np.array(h, copy=True), np.array(goff, copy=True)
            c.sketch_publish()
            ph, po, total, overflow = c.sketch_gather_peers()
            assert not overflow and total == len(h)
            got_h = torch.as_tensor(View(ph, total, "<i4"), device="cuda").cpu().numpy().view(np.uint32)
            got_o = torch.as_tensor(View(po, len(fams) + 1, "<i8"), device="cuda").cpu().numpy()
            assert np.array_equal(got_h, h) and np.array_equal(got_o.astype(np.uint64), goff.astype(np.uint64))
    finally:
        c.close()
    c = api.Context(0)
    try:
        c.set_params(k, s, l, util.shuffled_dim(k, s, l))
        c.peer_connect(c.peer_init(len(fams), 16), 1, 0)  # far too small
        c.sketch(seq, off)
        c.sketch_publish()
        assert c.sketch_gather_peers()[3]
    finally:
        c.close()


def _retained_repeat_unit(k, s, l, seed=5):
    """A 2k-base unit whose tandem repeat emits (one of its rotations is a retained k-mer)."""
    for t in range(5000):
        unit = synth.random_bases(seed * 100003 + t, 2 * k)
        if len(util.oracle_sketch(k, s, l, [[np.tile(unit, 4)]])[0]):
            return unit
    raise AssertionError("no retained repeat unit found")


def test_tandem_repeats_duplicates(ctx):
    """Repeat-rich input: a tandem repeat of a retained k-mer emits the same hash once per period -- thousands of
    duplicates that the per-genome sort + unique must collapse exactly like the reference's unordered_set
    (src/sketch.cpp:526-529).  (Beyond the emit buffer's headroom the library regrows and replays:
    tests/test_gpu_seams.py::test_emit_overflow_regrows_and_replays.)"""
    k, s, l = 10, 6, 3
    unit = _retained_repeat_unit(k, s, l)
    base = synth.random_bases(911, 400000)
    genomes = [
        [np.concatenate([base[:150000], np.tile(unit, 20000), base[150000:]])],       # 400 kbp of repeat: ~20 k duplicate emits
        [np.tile(unit, 3000), base[:50000], np.tile(unit[::-1].copy(), 3000)],        # repeats at record borders
        [base],
    ]
    want = util.oracle_sketch(k, s, l, genomes)
    got, _, _ = _gpu_sketch(ctx, k, s, l, genomes)
    for g, (a, b) in enumerate(zip(got, want)):
        assert np.array_equal(a.astype(np.uint64), b), g


def test_64bit_hashes_equal_oracle_random(ctx):
    """half_k - drlevel > 8 (src/sketch.cpp:336): 36- and 40-bit hashes, several genomes, N runs / lowercase, against the
    oracle (pinned for this configuration by the golden case fa_k12s6l3); index + alldist + dist on the u64 hashes against
    the oracle on their ranks; sparse-index round trip through index_load64 in shuffled order."""
    for (k, s, l, length, n) in [(12, 6, 3, 900000, 5), (11, 5, 1, 30000, 4)]:
        fam = synth.family(99 + k, n, length, 0.02, n_contigs=3)
        genomes = [[synth.sprinkle(c, 7 + i, n_runs=3, lowercase_frac=0.02) for c in g] for i, g in enumerate(fam)]
        want = util.oracle_sketch(k, s, l, genomes)
        got, h, goff = _gpu_sketch(ctx, k, s, l, genomes)
        assert h.dtype == np.uint64
        for g, (a, b) in enumerate(zip(got, want)):
            assert np.array_equal(a, b), (k, g)
        assert max(int(x.max()) for x in want if len(x)) >= 1 << 32  # the case really needs more than 32 bits
        sizes = np.diff(goff).astype(np.int64)
        rk, roff, uniq = util.compress_hashes(want)
        ctx.index_build(h, goff)
        uh, uc, dict_ = ctx.index_export()
        o_dict, o_uh, o_uc, _ = oracle.index_build(rk, roff)
        assert np.array_equal(uh, uniq[o_uh]) and np.array_equal(uc, o_uc) and np.array_equal(dict_, o_dict)
        for tri, D in [(True, 0.2), (False, 0.1), (False, 1.0)]:
            hits = ctx.dist_rows(h, goff, tri, 2 * k, D)
            hits, _, _ = api.refine_hits(hits, sizes, sizes, 2 * k, D, tri)
            hr, hc, hm = oracle.alldist(rk, roff, 2 * k, D) if tri else oracle.dist(rk, roff, rk, roff, 2 * k, D)
            o = np.lexsort((hits[:, 1], hits[:, 0]))
            assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1)), (k, tri, D)
        # the reference's files hold the lists in hash-map order: load them shuffled
        perm = np.random.default_rng(5).permutation(len(uh))
        starts = np.concatenate([[0], np.cumsum(uc.astype(np.int64))]).astype(np.int64)
        shuffled = np.concatenate([dict_[starts[u]:starts[u + 1]] for u in perm]) if len(perm) else dict_
        ctx.index_load64(uh[perm], uc[perm], shuffled, sizes.astype(np.uint32))
        hits = ctx.dist_rows(h, goff, True, 2 * k, 0.2)
        hits, _, _ = api.refine_hits(hits, sizes, sizes, 2 * k, 0.2, True)
        hr, hc, hm = oracle.alldist(rk, roff, 2 * k, 0.2)
        o = np.lexsort((hits[:, 1], hits[:, 0]))
        assert np.array_equal(hits[o], np.stack([hr, hc, hm], axis=1))


def test_hash_width_mismatch_is_refused():
    from rabbitkssd_b200 import _lib
    import ctypes as C
    c = api.Context(0)
    try:
        c.set_params(12, 6, 3, util.shuffled_dim(12, 6, 3))
        c.sketch_begin(1)
        hp, op = _lib.u32p(), _lib.u64p()
        assert c.lib.kssd_b200_sketch_finish(c.h, C.byref(hp), C.byref(op)) == -2  # KSSD_E_ARG: 64-bit session, 32-bit entry point
    finally:
        c.close()


def test_relaxed_guard_L5K10(ctx):
    """BASELINE config 4 names L5K10 (half_subk 7, drlevel 5), which the reference refuses (common.cpp:37-39); the library
    refuses it too unless kssd_b200_set_relax_guard is on.  Parity UNPINNED (no reference output exists): checked against
    the oracle with the same guard relaxed.  Run here at s = 5, l = 3 (the same s - l = 2, 1 MiB table)."""
    from rabbitkssd_b200 import _lib
    k, s, l = 10, 5, 3
    dim = oracle.shuffle_dim(k, s, l)
    with pytest.raises(_lib.KssdError) as e:
        ctx.set_params(k, s, l, dim)
    assert e.value.code == -4
    ctx.set_relax_guard(True)
    try:
        ctx.set_params(k, s, l, dim)
        genomes = synth.family(31, 3, 400000, 0.02, n_contigs=2)
        seq, _, off = synth.pack_batch(genomes)
        h, goff = ctx.sketch(seq, off)
        p = oracle.init_params(k, s, l, relax_guard=True)
        oseq, rec_off, rec_first = synth.oracle_records(genomes)
        want = oracle.sketch_batch(p, dim, oseq, rec_off, rec_first)
        for g in range(3):
            assert np.array_equal(h[int(goff[g]):int(goff[g + 1])].astype(np.uint64), want[g]), g
        assert sum(len(x) for x in want) > 100
    finally:
        ctx.set_relax_guard(False)
