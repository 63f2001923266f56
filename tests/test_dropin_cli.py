"""GPU: the drop-in CLI (reference main/subCommand/parsers/file I/O + the B200 shims behind the five hot functions)
against the UNMODIFIED reference binary on identical inputs: .sketch hash sets, .dict/.index bytes, .alldist/.dist
line sets (incl. -M 1, -N, FASTQ -Q/-n).  Both binaries are built by __graft_entry__.build() where the reference
tree exists and travel to the GPU box; nothing here reads /root/reference."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from rabbitkssd_b200 import formats, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")
NEW = os.path.join(ROOT, "rabbitkssd_b200", "_dropin", "rabbit_kssd_b200")


def _run(binary, args, cwd):
    p = subprocess.run([binary] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    assert p.returncode == 0, f"{os.path.basename(binary)} {' '.join(args)} failed:\n{p.stdout.decode()[-3000:]}"
    return p.stdout.decode()


def _sketch_sets(path):
    _, names, sets = formats.read_sketch(path)
    return {n: np.sort(s) for n, s in zip(names, sets)}


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    if not (os.access(REF, os.X_OK) and os.access(NEW, os.X_OK)):
        pytest.skip("reference / drop-in binaries not built (run __graft_entry__.build() where the reference tree exists)")
    d = str(tmp_path_factory.mktemp("dropin"))
    data = os.path.join(d, "data")
    os.makedirs(data)
    # two families of related genomes + N runs / lowercase, 2-3 contigs each
    refs, queries = [], []
    for f in range(2):
        fam = synth.family(4100 + f, 5, 400000, 0.05, n_contigs=2 + f)
        for m, g in enumerate(fam):
            g = [synth.sprinkle(c, 17 * f + m + ci, n_runs=4, run_len=9, lowercase_frac=0.02) for ci, c in enumerate(g)]
            p = os.path.join(data, f"f{f}m{m}.fna")
            synth.write_fasta(p, g, name=f"f{f}m{m}")
            (refs if m < 4 else queries).append(p)
    for i, p in enumerate(refs[:2]):  # queries that equal a reference and a fresh unrelated genome
        queries.append(p)
    extra = os.path.join(data, "lonely.fna")
    synth.write_fasta(extra, [synth.random_bases(999, 250000)], name="lonely")
    queries.append(extra)
    # FASTQ reads of one genome
    base = synth.random_bases(4242, 200000)
    fq = []
    for i in range(2):
        n_reads, rl = 6000, 150
        starts = (synth.splitmix64(77 + i, n_reads) % np.uint64(len(base) - rl)).astype(np.int64)
        reads = [base[a:a + rl] for a in starts]
        quals = [((synth.splitmix64(900 + 13 * i + r, rl) % np.uint64(41)) + 33).astype(np.uint8) for r in range(n_reads)]
        p = os.path.join(data, f"reads{i}.fq")
        synth.write_fastq(p, reads, quals, name=f"r{i}")
        fq.append(p)
    for name, lst in (("refs.list", refs), ("queries.list", queries), ("fq.list", fq)):
        with open(os.path.join(d, name), "w") as f:
            f.write("\n".join(lst) + "\n")
    _run(REF, ["shuffle", "-k", "10", "-s", "6", "-l", "3", "-o", "L3K10.shuf"], d)
    dirs = {}
    for tag in ("ref", "new"):
        dirs[tag] = os.path.join(d, tag)
        os.makedirs(dirs[tag])
    return {"d": d, "dirs": dirs, "bins": {"ref": REF, "new": NEW}}


def _both(work, args):
    for tag in ("ref", "new"):
        _run(work["bins"][tag], args, work["dirs"][tag])


def test_sketch_fasta_and_index_files(work):
    d = work["d"]
    _both(work, ["sketch", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"])
    a = _sketch_sets(os.path.join(work["dirs"]["ref"], "refs.sketch"))
    b = _sketch_sets(os.path.join(work["dirs"]["new"], "refs.sketch"))
    assert sorted(a) == sorted(b)
    for n in a:
        assert len(a[n]) > 50 and np.array_equal(a[n], b[n]), n
    # genome order inside the files may differ (the reference appends in thread-completion order): compare the
    # inverted index as hash -> set of genome names
    def inverted(tag):
        _, names, _ = formats.read_sketch(os.path.join(work["dirs"][tag], "refs.sketch"))
        counts, postings = formats.read_index_dict(os.path.join(work["dirs"][tag], "refs.sketch.index"),
                                                   os.path.join(work["dirs"][tag], "refs.sketch.dict"))
        nz = np.nonzero(counts)[0]
        off = np.concatenate([[0], np.cumsum(counts[nz].astype(np.int64))]).astype(np.int64)
        return {int(h): frozenset(names[g] for g in postings[off[i]:off[i + 1]]) for i, h in enumerate(nz)}, len(counts)
    ia, na = inverted("ref")
    ib, nb = inverted("new")
    assert na == nb == 1 << 28
    assert ia == ib


def test_sketch_fastq_quality_and_count(work):
    d = work["d"]
    _both(work, ["sketch", "-q", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "fq.list"), "-o", "fq", "-Q", "35", "-n", "2",
                 "-t", "2"])
    a = _sketch_sets(os.path.join(work["dirs"]["ref"], "fq.sketch"))
    b = _sketch_sets(os.path.join(work["dirs"]["new"], "fq.sketch"))
    assert sorted(a) == sorted(b)
    for n in a:
        assert len(a[n]) > 5 and np.array_equal(a[n], b[n]), n


@pytest.mark.parametrize("D,extra", [("0.08", []), ("0.08", ["-M", "1"]), ("1.5", [])])
def test_alldist_from_same_sketch(work, D, extra):
    """Same .sketch file in both directories: .dict/.index must be byte-identical, .alldist line sets equal."""
    d = work["d"]
    src = os.path.join(work["dirs"]["ref"], "refs.sketch")
    if not os.path.exists(src):
        _run(REF, ["sketch", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"], work["dirs"]["ref"])
    for tag in ("ref", "new"):
        for f in ("x.sketch", "x.sketch.index", "x.sketch.dict", "x.alldist"):
            p = os.path.join(work["dirs"][tag], f)
            if os.path.exists(p):
                os.remove(p)
        shutil.copy(src, os.path.join(work["dirs"][tag], "x.sketch"))
    _both(work, ["alldist", "-i", "x.sketch", "-D", D, "-o", "x.alldist", "-t", "3"] + extra)  # -D 1.5: every pair, also common = 0
    for f in ("x.sketch.dict", "x.sketch.index"):
        with open(os.path.join(work["dirs"]["ref"], f), "rb") as fa, open(os.path.join(work["dirs"]["new"], f), "rb") as fb:
            assert fa.read() == fb.read(), f
    la = sorted(formats.parse_dist_text(os.path.join(work["dirs"]["ref"], "x.alldist")))
    lb = sorted(formats.parse_dist_text(os.path.join(work["dirs"]["new"], "x.alldist")))
    assert len(la) >= 8
    assert la == lb
    with open(os.path.join(work["dirs"]["new"], "x.alldist")) as f:
        assert f.readline() == " genome0\tgenome1\tcommon|size0|size1\tjaccard\tmashD\n"
    assert not os.path.exists(os.path.join(work["dirs"]["new"], "x.alldist.dir"))


@pytest.mark.parametrize("extra", [["-D", "0.1"], ["-N", "3"], ["-D", "0.3", "-N", "2", "-M", "1"], []])
def test_dist_from_same_sketches(work, extra):
    d = work["d"]
    rdir = work["dirs"]["ref"]
    if not os.path.exists(os.path.join(rdir, "refs.sketch")):
        _run(REF, ["sketch", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"], rdir)
    if not os.path.exists(os.path.join(rdir, "queries.sketch")):
        _run(REF, ["sketch", "-q", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "queries.list"), "-o", "queries", "-t", "4"], rdir)
    for tag in ("ref", "new"):
        for f in ("r.sketch", "r.sketch.index", "r.sketch.dict", "q.sketch", "r.dist"):
            p = os.path.join(work["dirs"][tag], f)
            if os.path.exists(p):
                os.remove(p)
        shutil.copy(os.path.join(rdir, "refs.sketch"), os.path.join(work["dirs"][tag], "r.sketch"))
        shutil.copy(os.path.join(rdir, "queries.sketch"), os.path.join(work["dirs"][tag], "q.sketch"))
    _both(work, ["dist", "-r", "r.sketch", "-q", "q.sketch", "-o", "r.dist", "-t", "3"] + extra)
    la = sorted(formats.parse_dist_text(os.path.join(work["dirs"]["ref"], "r.dist")))
    lb = sorted(formats.parse_dist_text(os.path.join(work["dirs"]["new"], "r.dist")))
    assert len(la) >= 4
    assert la == lb


def test_alldist_straight_from_genome_list(work):
    """alldist -i <genome list>: sketch + index + distances in one command, each binary on its own sketches.
    Line orientation (which genome is printed first) depends on the genome order inside the .sketch, which the
    reference leaves to thread timing: compare unordered name pairs."""
    d = work["d"]
    _both(work, ["alldist", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "refs.list"), "-D", "0.08", "-o", "g.alldist", "-t", "4"])
    def canon(tag):
        out = []
        for a, b, c, s0, s1, j, dd in formats.parse_dist_text(os.path.join(work["dirs"][tag], "g.alldist")):
            out.append((frozenset([(a, s1), (b, s0)]), c, j, dd))
        return sorted(out, key=lambda x: (sorted(x[0]), x[1]))
    la, lb = canon("ref"), canon("new")
    assert len(la) >= 8
    assert la == lb


def test_64bit_hash_configuration_k12(work):
    """-k 12 -l 3: half_k - drlevel = 9 > 8 selects the 64-bit hash path (src/sketch.cpp:336): hashSet64 in .sketch, sparse
    .index (hash -> count, in the reference's hash-map order; ascending here), alldist / dist on the 64-bit index files
    -- each binary also READS the other's index files (same .sketch, swapped .index/.dict)."""
    d = work["d"]
    _run(REF, ["shuffle", "-k", "12", "-s", "6", "-l", "3", "-o", "L3K12.shuf"], d)
    shuf = os.path.join(d, "L3K12.shuf")
    _both(work, ["sketch", "-L", shuf, "-i", os.path.join(d, "refs.list"), "-o", "w", "-t", "4"])
    _both(work, ["sketch", "-q", "-L", shuf, "-i", os.path.join(d, "queries.list"), "-o", "wq", "-t", "4"])
    a = _sketch_sets(os.path.join(work["dirs"]["ref"], "w.sketch"))
    b = _sketch_sets(os.path.join(work["dirs"]["new"], "w.sketch"))
    assert sorted(a) == sorted(b)
    for n in a:
        assert a[n].dtype == np.uint64 and len(a[n]) > 20 and np.array_equal(a[n], b[n]), n
    assert max(int(x.max()) for x in a.values()) >= 1 << 32

    def mapping(tag, stem):
        _, names, _ = formats.read_sketch(os.path.join(work["dirs"][tag], stem + ".sketch"))
        hashes, counts, postings = formats.read_index_dict64(os.path.join(work["dirs"][tag], stem + ".sketch.index"),
                                                             os.path.join(work["dirs"][tag], stem + ".sketch.dict"))
        off = np.concatenate([[0], np.cumsum(counts.astype(np.int64))])
        return {int(h): frozenset(names[g] for g in postings[off[i]:off[i + 1]]) for i, h in enumerate(hashes)}
    assert mapping("ref", "w") == mapping("new", "w")
    # same .sketch everywhere; the drop-in reads the REFERENCE's index files (hash-map order) and vice versa
    rdir, ndir = work["dirs"]["ref"], work["dirs"]["new"]
    for f in ("w.sketch", "wq.sketch"):
        shutil.copy(os.path.join(rdir, f), os.path.join(ndir, f))
    for ext in (".index", ".dict"):  # built for the drop-in's own genome order: stale now
        os.remove(os.path.join(ndir, "w.sketch" + ext))
    _run(NEW, ["alldist", "-i", "w.sketch", "-D", "0.3", "-o", "own.alldist", "-t", "3"], ndir)   # builds its own index files
    for ext in (".index", ".dict"):
        os.replace(os.path.join(ndir, "w.sketch" + ext), os.path.join(ndir, "mine" + ext))
        shutil.copy(os.path.join(rdir, "w.sketch" + ext), os.path.join(ndir, "w.sketch" + ext))
    _both(work, ["alldist", "-i", "w.sketch", "-D", "0.3", "-o", "w.alldist", "-t", "3"])
    la = sorted(formats.parse_dist_text(os.path.join(rdir, "w.alldist")))
    assert len(la) >= 8
    assert la == sorted(formats.parse_dist_text(os.path.join(ndir, "w.alldist")))
    assert la == sorted(formats.parse_dist_text(os.path.join(ndir, "own.alldist")))
    for extra in (["-D", "0.3"], ["-N", "2"]):
        _both(work, ["dist", "-r", "w.sketch", "-q", "wq.sketch", "-o", "w.dist", "-t", "3"] + extra)
        lr = sorted(formats.parse_dist_text(os.path.join(rdir, "w.dist")))
        assert len(lr) >= 4
        assert lr == sorted(formats.parse_dist_text(os.path.join(ndir, "w.dist")))
    for ext in (".index", ".dict"):  # the reference reads OUR index files
        shutil.copy(os.path.join(ndir, "mine" + ext), os.path.join(rdir, "w.sketch" + ext))
    _run(REF, ["alldist", "-i", "w.sketch", "-D", "0.3", "-o", "w2.alldist", "-t", "3"], rdir)
    assert la == sorted(formats.parse_dist_text(os.path.join(rdir, "w2.alldist")))


def _run_env(binary, args, cwd, env):
    e = dict(os.environ)
    e.update(env)
    p = subprocess.run([binary] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600, env=e)
    assert p.returncode == 0, f"{os.path.basename(binary)} {' '.join(args)} failed:\n{p.stdout.decode()[-3000:]}"
    return p.stdout.decode()


def test_index_stays_on_device_and_files_are_still_written(work):
    """alldist -i <genome list> in one process: the distance step uses the CSR transSketches left on the device (no read-back
    of the 1 GiB .index), while the .index / .dict files are still produced -- a second run that starts from the .sketch
    reads them from disk and must print the same lines; the dense .index is what the reference's reader expects."""
    d, ndir = work["d"], work["dirs"]["new"]
    lst = os.path.join(ndir, "mylist")
    shutil.copy(os.path.join(d, "refs.list"), lst)
    out = _run(NEW, ["alldist", "-L", os.path.join(d, "L3K10.shuf"), "-i", "mylist", "-D", "0.08", "-o", "c1.alldist", "-t", "4"], ndir)
    assert "is not read back" in out
    counts, postings = formats.read_index_dict(os.path.join(ndir, "mylist.sketch.index"), os.path.join(ndir, "mylist.sketch.dict"))
    assert len(counts) == 1 << 28 and int(counts.sum()) == len(postings)
    out2 = _run(NEW, ["alldist", "-i", "mylist.sketch", "-D", "0.08", "-o", "c2.alldist", "-t", "4"], ndir)
    assert "is not read back" not in out2
    l1 = sorted(formats.parse_dist_text(os.path.join(ndir, "c1.alldist")))
    assert len(l1) >= 8 and l1 == sorted(formats.parse_dist_text(os.path.join(ndir, "c2.alldist")))
    # and the reference reads the files the drop-in wrote
    rdir = work["dirs"]["ref"]
    for f in ("mylist.sketch", "mylist.sketch.index", "mylist.sketch.dict"):
        shutil.copy(os.path.join(ndir, f), os.path.join(rdir, f))
    _run(REF, ["alldist", "-i", "mylist.sketch", "-D", "0.08", "-o", "c3.alldist", "-t", "4"], rdir)
    assert l1 == sorted(formats.parse_dist_text(os.path.join(rdir, "c3.alldist")))


def test_output_split_into_subfiles(work):
    """The > 4 GiB branch of the result-file policy (src/dist.cpp:314-336): sub-files kept in <out>.dir, <out>.index maps
    every genome to its sub-file.  Taken here with KSSD_B200_MERGE_LIMIT (the real 4 GiB case is run by bench C5)."""
    d, ndir, rdir = work["d"], work["dirs"]["new"], work["dirs"]["ref"]
    if not os.path.exists(os.path.join(rdir, "refs.sketch")):
        _run(REF, ["sketch", "-L", os.path.join(d, "L3K10.shuf"), "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"], rdir)
    for f in ("s.sketch.index", "s.sketch.dict"):
        if os.path.exists(os.path.join(ndir, f)):
            os.remove(os.path.join(ndir, f))
    shutil.copy(os.path.join(rdir, "refs.sketch"), os.path.join(ndir, "s.sketch"))
    _run(NEW, ["alldist", "-i", "s.sketch", "-D", "0.5", "-o", "whole.alldist", "-t", "3"], ndir)
    _run_env(NEW, ["alldist", "-i", "s.sketch", "-D", "0.5", "-o", "split.alldist", "-t", "3"], ndir, {"KSSD_B200_MERGE_LIMIT": "100"})
    assert not os.path.exists(os.path.join(ndir, "split.alldist"))
    sub = sorted(os.listdir(os.path.join(ndir, "split.alldist.dir")))
    assert sub == ["split.alldist.0", "split.alldist.1", "split.alldist.2"]
    lines = []
    for f in sub:
        lines += formats.parse_dist_text(os.path.join(ndir, "split.alldist.dir", f))
    whole = sorted(formats.parse_dist_text(os.path.join(ndir, "whole.alldist")))
    assert len(whole) >= 8 and sorted(lines) == whole
    with open(os.path.join(ndir, "split.alldist.index")) as f:
        idx = f.read().splitlines()
    _, names, _ = formats.read_sketch(os.path.join(ndir, "s.sketch"))
    assert idx[0] == "genomeName\tdistFileName"
    got = dict(x.split("\t") for x in idx[1:])
    assert sorted(got) == sorted(names)
    for a, b, *_ in lines:  # a line "name_j name_i ..." lives in the sub-file of its row genome (name_i)
        pass
    for f in sub:
        rows_here = {b for a, b, *_ in formats.parse_dist_text(os.path.join(ndir, "split.alldist.dir", f))}
        assert all(got[r].endswith(f) for r in rows_here)


def test_multi_gpu_cli_equals_reference(work):
    """One context per visible GPU inside the drop-in (files dealt over the GPUs for sketching, query rows sharded for
    alldist / dist with the index copied device to device): same files as the reference.  Needs >= 2 GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    d, ndir, rdir = work["d"], work["dirs"]["new"], work["dirs"]["ref"]
    shuf = os.path.join(d, "L3K10.shuf")
    out = _run_env(NEW, ["sketch", "-L", shuf, "-i", os.path.join(d, "refs.list"), "-o", "mg", "-t", "4"], ndir, {"KSSD_B200_DEVICES": "all", "KSSD_B200_SKETCH_BYTES_PER_GPU": "1000000"})
    one = _run_env(NEW, ["sketch", "-L", shuf, "-i", os.path.join(d, "refs.list"), "-o", "sg", "-t", "4"], ndir, {"KSSD_B200_DEVICE": "0"})
    a, b = _sketch_sets(os.path.join(ndir, "mg.sketch")), _sketch_sets(os.path.join(ndir, "sg.sketch"))
    assert sorted(a) == sorted(b) and all(np.array_equal(a[n], b[n]) for n in a)
    if not os.path.exists(os.path.join(rdir, "refs.sketch")):
        _run(REF, ["sketch", "-L", shuf, "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"], rdir)
    if not os.path.exists(os.path.join(rdir, "queries.sketch")):
        _run(REF, ["sketch", "-q", "-L", shuf, "-i", os.path.join(d, "queries.list"), "-o", "queries", "-t", "4"], rdir)
    for tag in ("ref", "new"):
        for f in ("m.sketch.index", "m.sketch.dict"):
            if os.path.exists(os.path.join(work["dirs"][tag], f)):
                os.remove(os.path.join(work["dirs"][tag], f))
        shutil.copy(os.path.join(rdir, "refs.sketch"), os.path.join(work["dirs"][tag], "m.sketch"))
        shutil.copy(os.path.join(rdir, "queries.sketch"), os.path.join(work["dirs"][tag], "mq.sketch"))
    _run(REF, ["alldist", "-i", "m.sketch", "-D", "0.5", "-o", "m.alldist", "-t", "4"], rdir)
    _run_env(NEW, ["alldist", "-i", "m.sketch", "-D", "0.5", "-o", "m.alldist", "-t", "4"], ndir, {"KSSD_B200_DEVICES": "all", "KSSD_B200_CELLS_PER_GPU": "8"})
    assert sorted(formats.parse_dist_text(os.path.join(rdir, "m.alldist"))) == sorted(formats.parse_dist_text(os.path.join(ndir, "m.alldist")))
    for extra in (["-D", "0.3"], ["-N", "2"], []):
        _run(REF, ["dist", "-r", "m.sketch", "-q", "mq.sketch", "-o", "m.dist", "-t", "4"] + extra, rdir)
        _run_env(NEW, ["dist", "-r", "m.sketch", "-q", "mq.sketch", "-o", "m.dist", "-t", "4"] + extra, ndir, {"KSSD_B200_DEVICES": "all", "KSSD_B200_CELLS_PER_GPU": "8"})
        assert sorted(formats.parse_dist_text(os.path.join(rdir, "m.dist"))) == sorted(formats.parse_dist_text(os.path.join(ndir, "m.dist")))


def test_raw_fasta_text_path_equals_reference(work, tmp_path):
    """Plain FASTA files go to the GPU as raw text (sliced into pieces by the shim, tokenised by fasta.cu); gzip files and
    anything that does not start with '>' go through kseq.  Small pieces (KSSD_B200_RAW_PIECE_BYTES) force many cuts; the
    hash sets must equal the reference's (kseq on the host) for one-line records, CRLF, blank lines, tiny records; a file
    with a '+' line makes the shim fall back to kseq for the whole job."""
    import gzip
    import test_gpu_fasta as tf
    d = work["d"]
    texts = tf._texts()
    texts.append(b">h1\n" + bytes(synth.random_bases(5, 200000)) + b"\nAC\nGT\n>h2 x\nA\n\n\nC\n>h3\n" + bytes(synth.random_bases(6, 90000)) + b"\r\nGG\r\n")
    files = []
    for i, t in enumerate(texts):
        p = str(tmp_path / f"raw{i}.fa")
        with open(p, "wb") as f:
            f.write(t)
        files.append(p)
    lst = str(tmp_path / "raw.list")
    with open(lst, "w") as f:
        f.write("\n".join(files) + "\n")
    shuf = os.path.join(d, "L3K10.shuf")
    # a list of gzip files goes through kseq as before (the reference wants a list to be all plain or all gzip, sketch.cpp:68-110)
    gzs = []
    for i in (0, 4):
        gz = str(tmp_path / f"z{i}.fa.gz")
        with gzip.open(gz, "wb") as f:
            f.write(texts[i])
        gzs.append(gz)
    gzl = str(tmp_path / "gz.list")
    with open(gzl, "w") as f:
        f.write("\n".join(gzs) + "\n")
    _both(work, ["sketch", "-q", "-L", shuf, "-i", gzl, "-o", "gzs", "-t", "2"])
    a, b = _sketch_sets(os.path.join(work["dirs"]["ref"], "gzs.sketch")), _sketch_sets(os.path.join(work["dirs"]["new"], "gzs.sketch"))
    assert sorted(a) == sorted(b) and all(np.array_equal(a[n], b[n]) for n in a)
    _run(REF, ["sketch", "-q", "-L", shuf, "-i", lst, "-o", "rawref", "-t", "3"], work["dirs"]["ref"])
    want = _sketch_sets(os.path.join(work["dirs"]["ref"], "rawref.sketch"))
    for env in ({}, {"KSSD_B200_RAW_PIECE_BYTES": "4099"}, {"KSSD_B200_RAW_PIECE_BYTES": "70001"}, {"KSSD_B200_NO_RAW_FASTA": "1"}):
        out = _run_env(NEW, ["sketch", "-q", "-L", shuf, "-i", lst, "-o", "rawnew", "-t", "3"], work["dirs"]["new"], env)
        got = _sketch_sets(os.path.join(work["dirs"]["new"], "rawnew.sketch"))
        assert sorted(got) == sorted(want), env
        for n in want:
            assert np.array_equal(got[n], want[n]), (env, n)
        assert "parsing all files with kseq" not in out
    # '+' line: not plain FASTA
    bad = str(tmp_path / "bad.fa")
    t = texts[0]
    cut = t.find(b"\n", 3000) + 1
    with open(bad, "wb") as f:
        f.write(t[:cut] + b"+\n" + t[cut:])
    with open(lst, "a") as f:
        f.write(bad + "\n")
    _run(REF, ["sketch", "-q", "-L", shuf, "-i", lst, "-o", "rawref2", "-t", "3"], work["dirs"]["ref"])
    out = _run(NEW, ["sketch", "-q", "-L", shuf, "-i", lst, "-o", "rawnew2", "-t", "3"], work["dirs"]["new"])
    assert "parsing all files with kseq" in out
    want = _sketch_sets(os.path.join(work["dirs"]["ref"], "rawref2.sketch"))
    got = _sketch_sets(os.path.join(work["dirs"]["new"], "rawnew2.sketch"))
    assert sorted(got) == sorted(want) and all(np.array_equal(got[n], want[n]) for n in want)


def test_raw_fastq_text_path_equals_reference(work, tmp_path):
    """Plain four-line FASTQ files go to the GPU as raw text (cut at record boundaries by the shim, tokenised by fastq.cu):
    ragged and empty reads, CRLF, many records per 32-byte chunk, a last line without a line end, trailing blank lines, tiny pieces
    (KSSD_B200_RAW_PIECE_BYTES) -- the hash sets must equal the reference's (kseq on the host; -t 1 keeps the reference off its
    RabbitFX big-file branch, src/sketch.cpp:644-650).  A multi-line record makes the shim fall back to kseq for the whole job."""
    import test_gpu_fastq as tq
    d = work["d"]
    fl, texts = tq._files()
    r, q = tq.make_reads(11, [100] * 3000)
    texts.append(tq.fastq_text(r, q)[:-1])         # the last line has no line end
    r, q = tq.make_reads(12, [75] * 3000)
    texts.append(tq.fastq_text(r, q) + b"\n\n")    # blank lines after the last record
    files = []
    for i, t in enumerate(texts):
        p = str(tmp_path / f"raw{i}.fq")
        with open(p, "wb") as f:
            f.write(t)
        files.append(p)
    lst = str(tmp_path / "rawfq.list")
    with open(lst, "w") as f:
        f.write("\n".join(files) + "\n")
    shuf = os.path.join(d, "L3K10.shuf")
    args = ["sketch", "-q", "-L", shuf, "-i", lst, "-Q", "50", "-n", "1"]
    _run(REF, args + ["-o", "fqref", "-t", "1"], work["dirs"]["ref"])
    want = _sketch_sets(os.path.join(work["dirs"]["ref"], "fqref.sketch"))
    assert sum(len(v) for v in want.values()) > 300
    for env in ({}, {"KSSD_B200_RAW_PIECE_BYTES": "4099"}, {"KSSD_B200_RAW_PIECE_BYTES": "70001"}, {"KSSD_B200_NO_RAW_FASTA": "1"}):
        out = _run_env(NEW, args + ["-o", "fqnew", "-t", "3"], work["dirs"]["new"], env)
        got = _sketch_sets(os.path.join(work["dirs"]["new"], "fqnew.sketch"))
        assert sorted(got) == sorted(want), env
        for n in want:
            assert np.array_equal(got[n], want[n]), (env, n)
        assert "parsing all files with kseq" not in out
    # a record with its sequence on two lines: kseq reads it, the device tokeniser refuses it
    good, bad = tq._bad_texts()
    badp = str(tmp_path / "multi.fq")
    with open(badp, "wb") as f:
        f.write(bad["two_line_seq"] * 40)
    with open(lst, "a") as f:
        f.write(badp + "\n")
    _run(REF, args + ["-o", "fqref2", "-t", "1"], work["dirs"]["ref"])
    out = _run(NEW, args + ["-o", "fqnew2", "-t", "3"], work["dirs"]["new"])
    assert "parsing all files with kseq" in out
    want = _sketch_sets(os.path.join(work["dirs"]["ref"], "fqref2.sketch"))
    got = _sketch_sets(os.path.join(work["dirs"]["new"], "fqnew2.sketch"))
    assert sorted(got) == sorted(want) and all(np.array_equal(got[n], want[n]) for n in want)


def test_union_and_sub_commands(work):
    """`rabbit_kssd union` / `sub` (set algebra, src/subCommand.cpp:307-794) on the GPU: same .sketch content as the
    reference -- union: one sketch "<file> merged sketches" with the ascending distinct hashes; sub: every query genome
    without the hashes of the reference sketches (compared per genome as sets: the reference writes genomes in the order its
    threads finish)."""
    d, rdir, ndir = work["d"], work["dirs"]["ref"], work["dirs"]["new"]
    shuf = os.path.join(d, "L3K10.shuf")
    if not os.path.exists(os.path.join(rdir, "refs.sketch")):
        _run(REF, ["sketch", "-L", shuf, "-i", os.path.join(d, "refs.list"), "-o", "refs", "-t", "4"], rdir)
    if not os.path.exists(os.path.join(rdir, "queries.sketch")):
        _run(REF, ["sketch", "-q", "-L", shuf, "-i", os.path.join(d, "queries.list"), "-o", "queries", "-t", "4"], rdir)
    for tag in ("ref", "new"):
        shutil.copy(os.path.join(rdir, "refs.sketch"), os.path.join(work["dirs"][tag], "u_in.sketch"))
        shutil.copy(os.path.join(rdir, "queries.sketch"), os.path.join(work["dirs"][tag], "u_q.sketch"))
    _both(work, ["union", "-i", "u_in.sketch", "-o", "u_out.sketch", "-t", "3"])
    ia, na, sa = formats.read_sketch(os.path.join(rdir, "u_out.sketch"))
    ib, nb, sb = formats.read_sketch(os.path.join(ndir, "u_out.sketch"))
    assert ia == ib and na == nb == ["u_in.sketch merged sketches"]
    assert len(sa[0]) > 200 and np.array_equal(sa[0], sb[0])
    _both(work, ["sub", "--rs", "u_out.sketch", "--qs", "u_q.sketch", "-o", "s_out.sketch", "-t", "3"])
    a = _sketch_sets(os.path.join(rdir, "s_out.sketch"))
    b = _sketch_sets(os.path.join(ndir, "s_out.sketch"))
    assert sorted(a) == sorted(b) and any(len(v) for v in a.values())
    for n in a:
        assert np.array_equal(a[n], b[n]), n
    # subtracting the multi-genome reference file directly gives the same result as subtracting its union
    _run(NEW, ["sub", "--rs", "u_in.sketch", "--qs", "u_q.sketch", "-o", "s_out2.sketch", "-t", "3"], ndir)
    c = _sketch_sets(os.path.join(ndir, "s_out2.sketch"))
    assert sorted(c) == sorted(b) and all(np.array_equal(c[n], b[n]) for n in b)
