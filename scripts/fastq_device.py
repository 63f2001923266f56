#!/usr/bin/env python
"""One session of raw FASTQ text through the C-ABI (kssd_batch.raw_fasta = KSSD_RAW_FASTQ): --mb megabytes of fixed-width
four-line records in 48 MiB batches.  Prints the number of hashes; run under ncu for the per-kernel times of the tokeniser."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "scripts")]
import cli_fastq  # noqa: E402
from rabbitkssd_b200 import api  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=480)
    args = ap.parse_args()
    path = "/dev/shm/kssd_fq_device.fq"
    n_reads = args.mb * (1 << 20) // 315
    cli_fastq.make_fastq(path, n_reads)
    text = open(path, "rb").read()
    os.unlink(path)
    per_batch = (48 << 20) // 315 * 315
    ctx = api.Context(0)
    ctx.set_params(10, 6, 3, api.shuffle_table(10, 6, 3))
    for rep in range(2):
        t0 = time.perf_counter()
        ctx.sketch_begin(1, True, 50, 2)
        for a in range(0, len(text), per_batch):
            ctx.submit_raw_fastq([(0, text[a:a + per_batch])])
        h, off = ctx.sketch_finish()
        dt = time.perf_counter() - t0
    print(json.dumps({"bytes": len(text), "batches": (len(text) + per_batch - 1) // per_batch, "hashes": int(off[-1]), "wall_s_second_session": dt}))
    ctx.close()


if __name__ == "__main__":
    main()
