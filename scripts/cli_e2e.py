#!/usr/bin/env python
"""Wall clock of the real drop-in: `rabbit_kssd_b200 alldist -L .shuf -i <FASTA list>` next to the unmodified `rabbit_kssd`
on identical files (BASELINE config 1 shape at --genomes 100, 1/10 of config 2 at --genomes 1000), with the reference's own
stderr timers for the phase split (subCommand.cpp:192, sketch.cpp:588, dist.cpp:134, :273, :340).  Prints one JSON object."""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import oracle  # noqa: E402  (test / bench infrastructure: generates the FASTA files)
from rabbitkssd_b200 import formats  # noqa: E402

TIMERS = {"sketch_s": r"time of computing sketches and save sketches into file is ([0-9.eE+-]+)",
          "transSketches_s": r"the time of transSketches is: ([0-9.eE+-]+)",
          "index_load_s": r"time of read index and offset sketch file is: ([0-9.eE+-]+)",
          "distance_s": r"time of multiple threads distance computing and save the subFile is: ([0-9.eE+-]+)",
          "merge_s": r"time of merge the subFiles into final files is: ([0-9.eE+-]+)"}


def run(binary, argv, cwd, env=None):
    e = dict(os.environ)
    e.update(env or {})
    t0 = time.perf_counter()
    p = subprocess.run([binary] + argv, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=e)
    wall = time.perf_counter() - t0
    txt = p.stdout.decode(errors="replace")
    if p.returncode != 0:
        raise RuntimeError(f"{binary} {' '.join(argv)} failed:\n{txt[-2000:]}")
    out = {"wall_s": wall}
    for k, pat in TIMERS.items():
        m = re.search(pat, txt)
        if m:
            out[k] = float(m.group(1))
    return out


def lines_of(path):
    out = set()
    for a, b, c, s0, s1, j, d in formats.parse_dist_text(path):
        out.add((frozenset([(a, s1), (b, s0)]), c, j, d))  # orientation depends on the genome order inside the .sketch
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=1000)
    ap.add_argument("--genome-len", type=int, default=5000000)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--repeat", type=int, default=2)
    args = ap.parse_args()
    ref = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")
    new = os.path.join(ROOT, "rabbitkssd_b200", "_dropin", "rabbit_kssd_b200")
    base = "/dev/shm" if os.access("/dev/shm", os.W_OK) else None
    d = tempfile.mkdtemp(prefix="kssd_cli_", dir=base)
    try:
        os.makedirs(os.path.join(d, "fa"))
        oracle.synth_fasta_files(os.path.join(d, "fa"), 20261106, 0, args.genomes, 10, args.genome_len, 0.04, 2)
        with open(os.path.join(d, "list"), "w") as f:
            f.write("".join(f"fa/g{i}.fna\n" for i in range(args.genomes)))
        subprocess.run([ref, "shuffle", "-k", "10", "-s", "6", "-l", "3", "-o", "L3K10.shuf"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        res = {"workload": f"alldist -L L3K10.shuf -i <{args.genomes} FASTA files of {args.genome_len / 1e6:g} Mbp> -D 0.05 -t {args.threads}, files in /dev/shm",
               "threads": args.threads}
        for tag, binary, env in (("reference", ref, {}), ("b200", new, {}), ("b200_kseq_parse", new, {"KSSD_B200_NO_RAW_FASTA": "1"})):
            runs = []
            for r in range(args.repeat):
                for f in ("list.sketch", "list.sketch.index", "list.sketch.dict", f"{tag}.alldist"):
                    if os.path.exists(os.path.join(d, f)):
                        os.remove(os.path.join(d, f))
                runs.append(run(binary, ["alldist", "-L", "L3K10.shuf", "-i", "list", "-D", "0.05", "-o", f"{tag}.alldist", "-t", str(args.threads)], d, env))
            res[tag] = min(runs, key=lambda x: x["wall_s"])
            res[tag]["all_walls_s"] = [x["wall_s"] for x in runs]
        a, b, c = (lines_of(os.path.join(d, f"{t}.alldist")) for t in ("reference", "b200", "b200_kseq_parse"))
        res["parity"] = {"lines": len(a), "set_equal": a == b, "set_equal_kseq_parse": a == c}
        res["speedup_wall"] = res["reference"]["wall_s"] / res["b200"]["wall_s"]
        print(json.dumps(res))
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
