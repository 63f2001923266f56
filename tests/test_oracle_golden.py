"""CPU: the oracle restatement against the fixtures produced by the unmodified reference binary."""
import numpy as np
import pytest

import oracle
import util

CASES = sorted(util.golden_meta().keys())


@pytest.mark.parametrize("case", CASES)
def test_shuffle_table_matches_reference_file(case):
    m = util.golden_meta()[case]
    c = m["recipe"]
    dim = util.shuffled_dim(c["k"], c["s"], c["l"])
    assert util.shuf_sha256(c["k"], c["s"], c["l"], dim) == m["shuf_sha256"]


def test_masks_l3k10():
    # values printed by the reference for L3K10
    p = oracle.init_params(10, 6, 3)
    assert (p.tupmask, p.domask, p.undomask0, p.undomask1) == (0xFFFFFFFFFF, 0xFFFFFF00, 0xFF00000000, 0xFF)
    assert p.dim_end == 4096 and p.kmer_size == 20 and p.rev_add_move == 38
    with pytest.raises(ValueError):
        oracle.init_params(10, 6, 4)  # init_shuffle.sh's L4 with -s 6: reference aborts (common.cpp:37-39)


@pytest.mark.parametrize("case", CASES)
def test_sketch_sets_equal_reference(case):
    m = util.golden_meta()[case]
    c = m["recipe"]
    names, genomes, quals = util.case_genomes(case)
    sets = util.oracle_sketch(c["k"], c["s"], c["l"], genomes, quals, c.get("Q", 0), c.get("n", 1))
    z = util.golden_npz(case)
    for n, s in zip(names, sets):
        assert np.array_equal(s, z["set_" + n]), n


@pytest.mark.parametrize("case", CASES)
def test_index_equals_reference_files(case):
    m = util.golden_meta()[case]
    z = util.golden_npz(case)
    # NOTE posting ids follow the genome order of the reference's .sketch (completion order); the
    # fixture's dict is compared as a multiset of (hash, count) only; posting order is checked on canon order
    sets = [z["set_" + n] for n in m["names"]]
    if util.use64(case):
        # sparse index in the iteration order of the reference's hash map (src/sketch.cpp:904-969): compared as a
        # mapping hash -> posting set; the oracle's index is built on the ranks of the 64-bit hashes
        h, off, uniq = util.compress_hashes(sets)
        dict_, uh, uc, _ = oracle.index_build(h, off)
        got, start = {}, 0
        for u, cnt in zip(uh, uc):
            got[int(uniq[u])] = sorted(int(x) for x in dict_[start:start + cnt])
            start += cnt
        want, start = {}, 0
        for hv, cnt in zip(z["index_hash"], z["index_count"]):
            want[int(hv)] = sorted(int(x) for x in z["dict"][start:start + cnt])
            start += cnt
        assert m["hash_size"] == len(want) == len(uniq) and got == want
        return
    h, off = util.flatten_sets(sets)
    dict_, uh, uc, _ = oracle.index_build(h, off)
    assert np.array_equal(uh.astype(np.uint64), z["index_hash"])
    assert np.array_equal(uc, z["index_count"])
    assert len(dict_) == len(z["dict"])
    # postings ascending within each list (transSketches pushes genome ids in ascending order)
    start = 0
    for cnt in uc:
        seg = dict_[start:start + cnt]
        assert np.all(np.diff(seg.astype(np.int64)) > 0)
        start += cnt


def _lines_from_hits(names, sizes, hr, hc, hm, k, containment, alldist):
    out = []
    for r, c_, common in zip(hr, hc, hm):
        if alldist:
            s0, s1 = sizes[r], sizes[c_]
            a, b = names[c_], names[r]
        else:
            s0, s1 = sizes[c_], sizes[r]
            a, b = names[r], names[c_]
        jorc, d = oracle.distance(common, s0, s1, 2 * k, containment)
        out.append((a, b, int(common), int(s0), int(s1), "%f" % jorc, "%f" % d))
    return sorted(out)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("tag", ["alldist", "alldist_M1", "dist", "dist_M1", "dist_D1"])
def test_distance_lines_equal_reference(case, tag):
    m = util.golden_meta()[case]
    c = m["recipe"]
    z = util.golden_npz(case)
    sets = [z["set_" + n] for n in m["names"]]
    if util.use64(case):
        h, off, _ = util.compress_hashes(sets)  # counts depend on hash equality only
    else:
        h, off = util.flatten_sets(sets)
    sizes = np.diff(off).astype(np.int64)
    cont = 1 if tag.endswith("M1") else 0
    D = 1.0 if tag.endswith("D1") else c["D"]
    if tag.startswith("alldist"):
        hr, hc, hm = oracle.alldist(h, off, 2 * c["k"], D, cont)
        got = _lines_from_hits(m["names"], sizes, hr, hc, hm, c["k"], cont, True)
    else:
        hr, hc, hm = oracle.dist(h, off, h, off, 2 * c["k"], D, cont)
        got = _lines_from_hits(m["names"], sizes, hr, hc, hm, c["k"], cont, False)
    want = sorted(tuple(x) for x in m["lines"][tag])
    assert got == want
