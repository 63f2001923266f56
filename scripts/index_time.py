#!/usr/bin/env python
"""Index-build tuning run: times kssd_b200_index_build (device input) and the alldist row pass on synthetic sketches of
the config-2 shape (10,000 sketches of ~1221 hashes, families of 10) with the libkssd_b200.so that is in place.
usage: python scripts/index_time.py [--tag name] [--reps 10]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", default="default")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    import numpy as np
    import torch
    from rabbitkssd_b200 import api, synth
    hashes, off = synth.synth_sketches(20261106, 1000, 10, 1221, 28)
    n = len(off) - 1
    d_h = torch.from_numpy(np.ascontiguousarray(hashes, dtype=np.uint32).view(np.int32)).cuda()
    d_o = torch.from_numpy(np.ascontiguousarray(off, dtype=np.uint64).view(np.int64)).cuda()
    ctx = api.Context(0)
    ctx.set_params(10, 6, 3, api.shuffle_table(10, 6, 3))
    t_ix, t_d = [], []
    for _ in range(args.reps + 2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.index_build_device(d_h.data_ptr(), d_o.data_ptr(), n)
        ctx.sync()
        t1 = time.perf_counter()
        _, n_hits = ctx.dist_rows_device(d_h.data_ptr(), d_o.data_ptr(), n, True, 20, 0.05, 0)
        ctx.sync()
        t2 = time.perf_counter()
        t_ix.append((t1 - t0) * 1e3)
        t_d.append((t2 - t1) * 1e3)
    print(json.dumps({"lib": args.tag, "postings": int(off[-1]), "index_ms_median": round(sorted(t_ix[2:])[args.reps // 2], 4),
                      "index_ms_min": round(min(t_ix[2:]), 4), "dist_ms_median": round(sorted(t_d[2:])[args.reps // 2], 4), "hits": int(n_hits)}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
