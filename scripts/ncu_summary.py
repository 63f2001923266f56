#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + source page) -> text: key metrics, per-region instruction counts, top stalls."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode()
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[-1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct', 'sm__warps_active.avg.pct_of_peak',
        'launch__registers_per_thread', 'sm__inst_executed.sum', 'sm__inst_executed_pipe_alu.avg.pct', 'sm__inst_executed_pipe_fma.avg.pct',
        'sm__inst_executed_pipe_lsu', 'sm__inst_executed_pipe_xu', 'sm__inst_executed_pipe_uniform', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts', 'sm__cycles_elapsed.max', 'smsp__issue_active.avg.pct', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__average_warp', 'smsp__warp_issue_stalled', 'sm__inst_executed_pipe_adu', 'sm__inst_executed_pipe_cbu', 'launch__grid_size', 'launch__block_size',
        'smsp__inst_executed.sum', 'sm__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__throughput.avg.pct', 'smsp__average_warps_issue_stalled']
for i, n in enumerate(h):
    if any(n.startswith(w) for w in want):
        print(f"{n} = {v[i]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode()
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
h = rows[hi]; body = rows[hi + 1:]
ie = h.index('Instructions Executed'); isamp = h.index('# Samples')
tot = sum(int(r[ie]) for r in body); ts = sum(int(r[isamp]) for r in body)
print("total warp instrs", tot, "samples", ts)
stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
agg = {s: sum(int(r[h.index(s)] or 0) for r in body) for s in stalls}
print("stall samples:", {k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
if len(sys.argv) > 2:
    print("--- top instructions by samples")
    top = sorted(range(len(body)), key=lambda i: -int(body[i][isamp]))[:int(sys.argv[2])]
    for i in sorted(top):
        r = body[i]
        st = {s[6:]: int(r[h.index(s)] or 0) for s in stalls if int(r[h.index(s)] or 0)}
        print(i, r[1].strip()[:60], 'exec', r[ie], 'samp', r[isamp], st)
# regions with similar exec count
cur = None; start = 0; acc = 0; n = 0; sacc = 0
print("--- regions")
for i, r in enumerate(body):
    e = int(r[ie])
    if cur is None or abs(e - cur) > 0.02 * max(cur, 1):
        if cur is not None and acc / max(tot, 1) > 0.004: print(f"rows {start}-{i-1}: n={n} exec_each~{cur} share={acc/tot:.3f} samples={sacc/ts:.3f}")
        cur = e; start = i; acc = 0; n = 0; sacc = 0
    acc += e; n += 1; sacc += int(r[isamp])
print(f"rows {start}-{len(body)-1}: n={n} exec_each~{cur} share={acc/tot:.3f} samples={sacc/ts:.3f}")
