#!/usr/bin/env python
"""Diagnostic: alldist of large clusters (dense tiles) against the oracle; prints what differs."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import oracle
from rabbitkssd_b200 import api, synth

def check(n_clusters, cluster, size, rows=None):
    h, off = synth.synth_sketches(20261109, n_clusters, cluster, size, 28, survival_step=0.55 / max(cluster - 1, 1))
    n = len(off) - 1
    sizes = np.diff(off).astype(np.int64)
    ctx = api.Context(0)
    ctx.set_params(10, 6, 3, api.shuffle_table(10, 6, 3))
    ctx.index_build(h, off)
    rows = rows or (0, n)
    hits = ctx.dist_rows(h, off, True, 20, 0.05, rows=rows)
    kept = api.refine_hits(hits, sizes, sizes, 20, 0.05, True)[0].astype(np.int64)
    kept = kept[np.lexsort((kept[:, 1], kept[:, 0]))]
    hr, hc, hm = oracle.alldist(h, off, 20, 0.05, rows=rows)
    want = np.stack([hr, hc, hm], axis=1).astype(np.int64)
    eq = np.array_equal(kept, want)
    print(f"clusters={n_clusters} x {cluster} size={size} rows={rows}: device {len(hits)} kept {len(kept)} want {len(want)} equal={eq}")
    if not eq:
        a = {(r, c): m for r, c, m in kept.tolist()}
        b = {(r, c): m for r, c, m in want.tolist()}
        miss = sorted(set(b) - set(a)); extra = sorted(set(a) - set(b)); diff = [(k, a[k], b[k]) for k in a if k in b and a[k] != b[k]]
        print("  missing", len(miss), miss[:8], "extra", len(extra), extra[:8], "common differs", len(diff), diff[:8])
        if miss:
            rr = np.array([m[0] for m in miss]); cc = np.array([m[1] for m in miss])
            print("  missing rows range", rr.min(), rr.max(), "cols range", cc.min(), cc.max(), "distinct rows", len(set(rr.tolist())))
    ctx.close()

check(2, 1250, 1221)
check(8, 1250, 1221, rows=(0, 64))
check(8, 1250, 1221, rows=(5000, 5064))
check(40, 1250, 1221, rows=(30000, 30064))
check(80, 1250, 1221, rows=(50000, 50064))
check(80, 1250, 1221, rows=(99936, 100000))
