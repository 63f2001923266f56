#!/usr/bin/env python
"""Wall clock of `sketch -Q -n` on plain FASTQ files: the drop-in (raw text tokenised on the device, fastq.cu), the drop-in with
KSSD_B200_NO_RAW_FASTA=1 (kseq on the host, the path before fastq.cu) and the unmodified reference, on identical files.
One file of fixed-width four-line records is generated and copied --files times (same work per file, same hash sets).
Prints one JSON object."""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from rabbitkssd_b200 import formats, synth  # noqa: E402


def make_fastq(path, n_reads, read_len=150, seed=1):
    genome = synth.random_bases(seed, 5_000_000)
    width = 11 + (read_len + 1) + 2 + (read_len + 1)
    rec = np.empty((n_reads, width), dtype=np.uint8)
    idx = np.arange(n_reads, dtype=np.int64)
    rec[:, 0] = ord("@")
    for d in range(9):
        rec[:, 9 - d] = ord("0") + (idx // 10 ** d) % 10
    rec[:, 10] = 10
    starts = (synth.splitmix64(seed + 1, n_reads) % np.uint64(len(genome) - read_len)).astype(np.int64)
    rec[:, 11:11 + read_len] = genome[starts[:, None] + np.arange(read_len)]
    rec[:, 11 + read_len] = 10
    rec[:, 12 + read_len] = ord("+")
    rec[:, 13 + read_len] = 10
    x = synth.splitmix64(seed + 2, n_reads * (read_len // 8 + 1)).view(np.uint8)[:n_reads * read_len].reshape(n_reads, read_len)
    rec[:, 14 + read_len:14 + 2 * read_len] = np.where(x < 8, 33 + x % 41, 73)  # 'I', 3 % of the positions anywhere in '!'..'I'
    rec[:, 14 + 2 * read_len] = 10
    with open(path, "wb") as f:
        f.write(rec.tobytes())
    return width * n_reads


def run(binary, argv, cwd, env=None):
    e = dict(os.environ)
    e.update(env or {})
    t0 = time.perf_counter()
    p = subprocess.run([binary] + argv, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=e)
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        raise RuntimeError(f"{binary} {' '.join(argv)} failed:\n{p.stdout.decode(errors='replace')[-2000:]}")
    return wall, p.stdout.decode(errors="replace")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=16)
    ap.add_argument("--reads", type=int, default=2_000_000, help="reads per file (315 bytes each at the default length)")
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--repeat", type=int, default=2)
    args = ap.parse_args()
    ref = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")
    new = os.path.join(ROOT, "rabbitkssd_b200", "_dropin", "rabbit_kssd_b200")
    base = "/dev/shm" if os.access("/dev/shm", os.W_OK) else None
    d = tempfile.mkdtemp(prefix="kssd_fq_", dir=base)
    try:
        st = os.statvfs(d)
        room = st.f_bavail * st.f_frsize
        per_file = 315 * args.reads
        while args.files > 1 and per_file * args.files > 0.6 * room:
            args.files //= 2
        first = os.path.join(d, "reads0.fq")
        t0 = time.perf_counter()
        nbytes = make_fastq(first, args.reads)
        files = [first]
        for i in range(1, args.files):
            p = os.path.join(d, f"reads{i}.fq")
            shutil.copyfile(first, p)
            files.append(p)
        gen_s = time.perf_counter() - t0
        with open(os.path.join(d, "fq.list"), "w") as f:
            f.write("\n".join(files) + "\n")
        run(ref, ["shuffle", "-k", "10", "-s", "6", "-l", "3", "-o", "L3K10.shuf"], d)
        argv = ["sketch", "-q", "-L", "L3K10.shuf", "-i", "fq.list", "-Q", "50", "-n", "2", "-t", str(args.threads)]
        out = {"files": args.files, "bytes_per_file": nbytes, "total_GB": nbytes * args.files / 1e9, "threads": args.threads, "generation_s": gen_s,
               "command": "sketch -q -Q 50 -n 2 -t %d" % args.threads}
        walls = {"reference": [], "dropin_raw": [], "dropin_kseq": []}
        for r in range(args.repeat):
            walls["dropin_raw"].append(run(new, argv + ["-o", "new_raw"], d)[0])
            walls["dropin_kseq"].append(run(new, argv + ["-o", "new_kseq"], d, {"KSSD_B200_NO_RAW_FASTA": "1"})[0])
        walls["reference"].append(run(ref, argv + ["-o", "ref"], d)[0])
        trace = run(new, argv + ["-o", "new_raw"], d, {"KSSD_B200_TRACE": "1"})[1]
        out["wall_s"] = {k: min(v) for k, v in walls.items()}
        out["all_walls_s"] = walls
        out["speedup_vs_reference"] = out["wall_s"]["reference"] / out["wall_s"]["dropin_raw"]
        out["trace"] = [l for l in trace.splitlines() if "kssd_b200 trace" in l]
        sets = {}
        for tag in ("ref", "new_raw", "new_kseq"):
            _, names, hs = formats.read_sketch(os.path.join(d, tag + ".sketch"))
            sets[tag] = {os.path.basename(n): np.sort(h) for n, h in zip(names, hs)}
        out["hashes_per_file"] = int(np.mean([len(v) for v in sets["ref"].values()]))
        out["sets_equal_reference"] = {tag: bool(sorted(sets[tag]) == sorted(sets["ref"]) and all(np.array_equal(sets[tag][n], sets["ref"][n]) for n in sets["ref"]))
                                       for tag in ("new_raw", "new_kseq")}
        print(json.dumps(out))
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
