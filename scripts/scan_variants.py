#!/usr/bin/env python
"""Scan-kernel tuning run: times kssd_scan_kernel alone (CUDA events inside the library, device-resident input larger
than L2) with the libkssd_b200.so that is in place; --tag labels the output line (scripts/run_variants.sh swaps in
differently compiled libraries).
usage: python scripts/scan_variants.py [--genomes 2000] [--steps 5] [--warmup 2] [--tag name]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=2000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--tag", default="default")
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench
    from rabbitkssd_b200 import api
    length, family = 5_000_000, 10
    stride = (length + 2 + 15) // 16 * 16  # as bench.py
    n = args.genomes
    buf = bench.gen_genomes_device(torch, n, 0, length, family, stride)
    d_seg_off = torch.from_numpy((np.arange(n + 1, dtype=np.int64) * stride)).cuda()
    d_seg_gen = torch.arange(n, dtype=torch.int32, device="cuda")
    dim = api.shuffle_table(bench.K, bench.S, bench.L_)
    ref_tot = None
    for v in [args.tag]:
        ctx = api.Context(0)
        ctx.set_params(bench.K, bench.S, bench.L_, dim)
        ctx.set_profiling(True)
        for i in range(args.warmup + args.steps):
            if i == args.warmup:
                ctx.sync()
                ctx.scan_kernel_time()
            ctx.sketch_begin(n)
            ctx.scan_device(buf.data_ptr(), n * stride, d_seg_off.data_ptr(), d_seg_gen.data_ptr(), n)
            _, _, tot = ctx.sketch_finish_device()
        ms, launches, nbytes = ctx.scan_kernel_time()
        if ref_tot is None:
            ref_tot = tot
        print(json.dumps({"lib": v, "ms_per_launch": ms / max(launches, 1), "launches": launches, "GBps": nbytes / ms / 1e6 if ms else None,
                          "sketch_total": int(tot), "same_as_first": bool(tot == ref_tot)}), flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
