#!/usr/bin/env python
"""Generate the golden fixtures in this directory by running the UNMODIFIED reference binary
(oracle/_ref/rabbit_kssd, built by oracle/build_ref.sh from /root/reference).

The reference ships no golden vectors; these files pin parity to the reference's own
CPU output.  Inputs are NOT stored: they are regenerated from the seeds in cases() by rabbitkssd_b200.synth
(counter-based splitmix64, machine independent).  Run from the repo root:  python tests/golden/make_golden.py
[case ...] (no names = all cases).
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from rabbitkssd_b200 import formats, synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")


def cases():
    """name -> recipe.  Kept small so the CPU test-suite stays fast."""
    return {
        "fa_k10s6l3": dict(kind="fasta", k=10, s=6, l=3, seed=20261104, n_fam=2, fam=4, length=300000,
                           contigs=2, max_div=0.04, n_runs=6, lower=0.02, D=0.2, N=2),
        "fa_k9s6l3": dict(kind="fasta", k=9, s=6, l=3, seed=20261105, n_fam=2, fam=3, length=250000,
                          contigs=1, max_div=0.03, n_runs=3, lower=0.0, D=0.15, N=1),
        "fa_k8s5l2": dict(kind="fasta", k=8, s=5, l=2, seed=20261106, n_fam=3, fam=3, length=40000,
                          contigs=3, max_div=0.05, n_runs=4, lower=0.05, D=0.3, N=3),
        "fa_k10s7l4": dict(kind="fasta", k=10, s=7, l=4, seed=20261107, n_fam=1, fam=3, length=1500000,
                           contigs=2, max_div=0.02, n_runs=5, lower=0.01, D=0.3, N=1),
        # half_k - drlevel = 9 > 8: the reference's 64-bit hash path (36-bit hashes, sparse .index).
        # Pins the ORACLE for the next round; the CUDA library refuses this configuration (KSSD_E_UNSUPPORTED).
        "fa_k12s6l3": dict(kind="fasta", k=12, s=6, l=3, seed=20261109, n_fam=2, fam=3, length=250000,
                           contigs=2, max_div=0.03, n_runs=4, lower=0.02, D=0.3, N=2),
        "fq_k10s6l3": dict(kind="fastq", k=10, s=6, l=3, seed=20261108, n_fam=2, fam=2, length=120000,
                           read_len=150, coverage=3, Q=40, n=2, D=0.5, N=1),
    }


def make_genomes(c):
    """-> list of (name, contigs) in list-file order."""
    out = []
    for f in range(c["n_fam"]):
        fam = synth.family(c["seed"] + 101 * f, c["fam"], c["length"], c["max_div"] if "max_div" in c else 0.03,
                           c.get("contigs", 1))
        for m, g in enumerate(fam):
            g = [synth.sprinkle(x, c["seed"] + 17 * f + m + ci, c.get("n_runs", 0), 10, c.get("lower", 0.0))
                 for ci, x in enumerate(g)]
            out.append((f"g{f}_{m}.fna", g))
    return out


def make_reads(c):
    """-> list of (name, reads, quals): shotgun reads (both strands) off each genome of make_genomes()."""
    out = []
    comp = np.zeros(256, np.uint8)
    for a, b in zip(b"ACGTNacgtn", b"TGCANtgcan"):
        comp[a] = b
    for gi, (name, g) in enumerate(make_genomes(dict(c, contigs=1))):
        seq = g[0]
        n_reads = c["coverage"] * len(seq) // c["read_len"]
        r = synth.splitmix64(c["seed"] + 999 + gi, 2 * n_reads)
        starts = (r[:n_reads] % np.uint64(len(seq) - c["read_len"])).astype(np.int64)
        strand = (r[n_reads:] & np.uint64(1)).astype(bool)
        reads, quals = [], []
        for i in range(n_reads):
            x = seq[starts[i]:starts[i] + c["read_len"]]
            if strand[i]:
                x = comp[x[::-1]]
            q = (synth.splitmix64(c["seed"] * 7 + gi * 100003 + i, c["read_len"]) % np.uint64(39)).astype(np.uint8) + 35
            low = (synth.splitmix64(c["seed"] * 11 + gi * 100003 + i, c["read_len"]) % np.uint64(8)) != 0
            q = np.where(low, np.maximum(q, 45), q).astype(np.uint8)  # ~1/8 of bases may fall under Q=40
            reads.append(np.ascontiguousarray(x))
            quals.append(q)
        out.append((name.replace(".fna", ".fq"), reads, quals))
    return out


def run(cmd, cwd):
    p = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    if p.returncode != 0:
        raise RuntimeError(f"{cmd} failed: {p.stderr.decode()[-2000:]}")
    return p


def write_inputs(c, d):
    names = []
    if c["kind"] == "fasta":
        for name, g in make_genomes(c):
            synth.write_fasta(os.path.join(d, name), g, name=name)
            names.append(name)
    else:
        for name, reads, quals in make_reads(c):
            synth.write_fastq(os.path.join(d, name), reads, quals, name=name)
            names.append(name)
    with open(os.path.join(d, "list.txt"), "w") as f:
        f.write("\n".join(names) + "\n")
    return names


def main():
    if not os.access(REF, os.X_OK):
        sys.exit("reference binary missing: run oracle/build_ref.sh first")
    meta = {}
    only = sys.argv[1:]  # optional: names of the cases to (re)generate; the others are kept as they are in golden.json
    if only and os.path.exists(os.path.join(HERE, "golden.json")):
        meta = json.load(open(os.path.join(HERE, "golden.json")))
    for cname, c in cases().items():
        if only and cname not in only:
            continue
        with tempfile.TemporaryDirectory() as d:
            names = write_inputs(c, d)
            shuf = "x.shuf"
            run([REF, "shuffle", "-k", str(c["k"]), "-s", str(c["s"]), "-l", str(c["l"]), "-o", shuf], d)
            sha = hashlib.sha256(open(os.path.join(d, shuf), "rb").read()).hexdigest()
            extra = []
            if c["kind"] == "fastq":
                extra = ["-Q", str(c["Q"]), "-n", str(c["n"])]
            run([REF, "sketch", "-L", shuf, "-i", "list.txt", "-o", "ref", "-t", "4"] + extra, d)
            info, snames, sets = formats.read_sketch(os.path.join(d, "ref.sketch"))
            by_name = {n: np.sort(s.astype(np.uint64)) for n, s in zip(snames, sets)}
            use64 = c["k"] - c["l"] > 8
            if use64:  # sparse index in hash-map iteration order: stored as the file has it
                ihash, icount, postings = formats.read_index_dict64(os.path.join(d, "ref.sketch.index"),
                                                                    os.path.join(d, "ref.sketch.dict"))
                # posting ids follow the order of ref.sketch (completion order): remap to list order
                file_to_list = np.array([names.index(n) for n in snames], dtype=np.uint32)
                postings = file_to_list[postings]
            else:
                counts, postings = formats.read_index_dict(os.path.join(d, "ref.sketch.index"),
                                                           os.path.join(d, "ref.sketch.dict"))
            # canonical sketch file in list order so that alldist/dist orientation is reproducible
            hashes = np.concatenate([by_name[n] for n in names])
            off = np.concatenate([[0], np.cumsum([len(by_name[n]) for n in names])]).astype(np.uint64)
            formats.write_sketch(os.path.join(d, "canon.sketch"), c["k"], c["s"], c["l"], names, hashes, off)
            out = {}
            for tag, args in {
                "alldist": ["alldist", "-i", "canon.sketch", "-D", str(c["D"]), "-o", "o_alldist"],
                "alldist_M1": ["alldist", "-i", "canon.sketch", "-D", str(c["D"]), "-M", "1", "-o", "o_alldist_m1"],
                "dist": ["dist", "-r", "canon.sketch", "-q", "canon.sketch", "-D", str(c["D"]), "-o", "o_dist"],
                "dist_M1": ["dist", "-r", "canon.sketch", "-q", "canon.sketch", "-D", str(c["D"]), "-M", "1",
                            "-o", "o_dist_m1"],
                "dist_D1": ["dist", "-r", "canon.sketch", "-q", "canon.sketch", "-o", "o_dist_d1"],
            }.items():
                run([REF] + args + ["-t", "3"], d)
                out[tag] = sorted(formats.parse_dist_text(os.path.join(d, args[-1])))
            if use64:
                index_hash, index_count, hash_size = ihash, icount, int(len(ihash))
            else:
                nz = np.nonzero(counts)[0]
                index_hash, index_count, hash_size = nz.astype(np.uint64), counts[nz], int(len(counts))
            np.savez_compressed(os.path.join(HERE, cname + ".npz"),
                                **{"set_" + n: by_name[n] for n in names},
                                index_hash=index_hash, index_count=index_count, dict=postings)
            meta[cname] = dict(recipe=c, shuf_sha256=sha, names=names, info=info, hash_size=hash_size,
                               lines=out)
            print(cname, "ok:", {n: len(by_name[n]) for n in names}, {k: len(v) for k, v in out.items()})
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=0, sort_keys=True)


if __name__ == "__main__":
    main()
