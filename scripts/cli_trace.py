#!/usr/bin/env python
"""Phase trace of the drop-in's sketch step (KSSD_B200_TRACE) on FASTA files in /dev/shm; raw-text path vs kseq path."""
import os, shutil, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import oracle
d = tempfile.mkdtemp(prefix="kssd_trace_", dir="/dev/shm")
try:
    os.makedirs(os.path.join(d, "fa"))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    oracle.synth_fasta_files(os.path.join(d, "fa"), 20261106, 0, n, 10, 5000000, 0.04, 2)
    for name, k in (("list100", min(100, n)), ("list", n)):
        open(os.path.join(d, name), "w").write("".join(f"fa/g{i}.fna\n" for i in range(k)))
    ref = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")
    new = os.path.join(ROOT, "rabbitkssd_b200", "_dropin", "rabbit_kssd_b200")
    subprocess.run([ref, "shuffle", "-k", "10", "-s", "6", "-l", "3", "-o", "L3K10.shuf"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for lst in ("list100", "list", "list"):
        for mode, env in (("raw", {}), ("kseq", {"KSSD_B200_NO_RAW_FASTA": "1"}), ("reference", None)):
            e = dict(os.environ, KSSD_B200_TRACE="1", **(env or {}))
            t0 = time.perf_counter()
            p = subprocess.run([ref if env is None else new, "sketch", "-q", "-L", "L3K10.shuf", "-i", lst, "-o", "out_" + lst, "-t", str(os.cpu_count())], cwd=d,
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=e)
            wall = time.perf_counter() - t0
            keep = [l for l in p.stdout.decode(errors="replace").splitlines() if "trace" in l or "time of" in l]
            print(f"== {lst} {mode}: wall {wall:.3f} s rc={p.returncode}")
            for l in keep:
                print("   ", l)
finally:
    shutil.rmtree(d, ignore_errors=True)
