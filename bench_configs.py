"""bench.py --config C3 | C4 | C5: the BASELINE.json configs beyond the default (configs[2..4]); same JSON contract.

C3  dist of 1,000 query sketches vs a 50,000-genome reference sketch index (L3K9), query rows sharded over the ranks, the
    index replicated by one NCCL broadcast of the CSR.  step = posting accumulation + epilogue of the rank's
    rows; metric genome-pairs/s = Q x R / t.
C4  sketch of 20 synthetic 3 Gbp genomes (>= 24 contigs each) with L4K10 (s = 7: the G = 4 filter template), bases resident in
    HBM; metric sketch Gbp/s; roofline = the scan kernel.  --extra shuffle=L5K10 runs the configuration the reference
    refuses (common.cpp:37-39) behind kssd_b200_set_relax_guard.
C5  alldist of 100,000 synthesised sketches (L3K10, S ~ 1221): step = index build + alldist of the rank's row block.
"""
import json
import os
import shutil
import subprocess
import sys
import time

import numpy as np

import bench as B


def _extra(args):
    out = {}
    for kv in filter(None, args.extra.split(",")):
        k, _, v = kv.partition("=")
        out[k.strip()] = v.strip()
    return out


def _setup(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    all_cpus = os.sched_getaffinity(0)
    numa = B.bind_to_gpu_numa(local_rank)
    torch.cuda.set_device(local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from rabbitkssd_b200 import api
    ctx = api.Context(local_rank)
    stream = torch.cuda.Stream(device=local_rank)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_profiling(True)
    return torch, dist, world, rank, local_rank, ctx, stream, all_cpus, numa


def _timed(torch, dist, world, stream, warmup, steps, fn):
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(max(warmup, 3)):
        fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        out = fn()
    e1.record(stream)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) / steps, out


def _finish(dist, world):
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _dist_lines(hits_kept, jorc, dd, q_sizes, r_sizes):
    """dist result lines (src/dist.cpp:642: query ref common|size_ref|size_query jorc dist) with indices for names."""
    out = set()
    for (r, c, m), jv, dv in zip(hits_kept.tolist(), jorc.tolist(), dd.tolist()):
        out.add(f"{r}\t{c}\t{m}|{int(r_sizes[c])}|{int(q_sizes[r])}\t{'%f' % jv}\t{'%f' % dv}")
    return out


# ---------------------------------------------------------------------------------------------------------------------
def c3_inputs(n_ref_clusters, cluster, n_query, size, seed):
    from rabbitkssd_b200 import synth
    rh, roff = synth.synth_sketches(seed, n_ref_clusters, cluster, size, 24)
    # queries: fresh variants of random reference genomes: keep ~85 % of the hashes, top up with private ones
    rng = np.random.default_rng(seed + 1)
    picks = rng.integers(0, len(roff) - 1, n_query)
    qs, qoff = [], [0]
    for g in picks:
        src = rh[int(roff[g]):int(roff[g + 1])]
        keep = src[rng.random(len(src)) < 0.85]
        priv = rng.integers(0, 1 << 24, max(0, size - len(keep)), dtype=np.uint32)
        q = np.unique(np.concatenate([keep, priv]))
        qs.append(q)
        qoff.append(qoff[-1] + len(q))
    return rh, roff, np.concatenate(qs).astype(np.uint32), np.asarray(qoff, dtype=np.uint64)


def run_c3(args):
    ex = _extra(args)
    n_clusters, cluster = int(ex.get("ref_clusters", 5000)), int(ex.get("cluster", 10))
    n_query, size, D = int(ex.get("queries", 1000)), int(ex.get("size", 1221)), float(ex.get("D", 0.05))
    K = 9
    cfg = {"workload": f"dist of {n_query} query sketches vs a {n_clusters * cluster}-genome reference sketch index, L3K9, -D {D} (BASELINE.json configs[2])",
           "queries": n_query, "refs": n_clusters * cluster, "sketch_size": size, "max_dist": D, "seed": B.SEED + 3,
           "l2": "the index (postings + directory, ~0.3 GB) exceeds the 126 MB L2; queries stream through it"}
    rh, roff, qh, qoff = c3_inputs(n_clusters, cluster, n_query, size, B.SEED + 3)
    n_ref = len(roff) - 1
    pairs = float(n_query) * n_ref
    if args.impl == "reference":
        return ref_c3(args, cfg, rh, roff, qh, qoff, D, pairs)
    torch, dist, world, rank, local_rank, ctx, stream, all_cpus, numa = _setup(args)
    from rabbitkssd_b200 import api, multigpu
    ctx.set_params(K, 6, 3, api.shuffle_table(K, 6, 3))
    t0 = time.perf_counter()
    if rank == 0:
        ctx.index_build(rh, roff)
    ctx.sync()
    t_index = time.perf_counter() - t0
    t0 = time.perf_counter()
    if world > 1:
        multigpu.replicate_index(dist, ctx, rank)
        torch.cuda.synchronize()
    t_bcast = time.perf_counter() - t0
    d_qh = torch.from_numpy(qh.view(np.int32)).cuda()
    d_qo = torch.from_numpy(qoff.astype(np.int64)).cuda()
    rb = multigpu.rect_bounds(n_query, world)
    r0, r1 = rb[rank], rb[rank + 1]
    with B.ClockSampler(local_rank) as clk:
        ctx.count_kernel_time()
        launches0 = ctx.launch_count()
        ms, (hp, nh) = _timed(torch, dist, world, stream, args.warmup, args.steps,
                              lambda: ctx.dist_rows_device(d_qh.data_ptr(), d_qo.data_ptr(), n_query, False, 2 * K, D, 0, rows=(r0, r1)))
    launches = (ctx.launch_count() - launches0) // (max(args.warmup, 3) + args.steps) * args.steps
    cnt_ms, cnt_n, cnt_post, cnt_atom = ctx.count_kernel_time()
    mine = B.dev_tensor(torch, hp, int(nh) * 3, torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 3)

    # e2e: host query arrays -> C-ABI (H2D of the queries, kernels, D2H of the hits through the sink)
    def e2e_step():
        return ctx.dist_rows(qh, qoff, False, 2 * K, D, 0, rows=(r0, r1))
    e2e_ms, e2e_hits = _timed(torch, dist, world, stream, 1, args.e2e_steps, e2e_step)
    os.sched_setaffinity(0, all_cpus)
    if world > 1:
        outs = [None] * world
        dist.all_gather_object(outs, (mine, e2e_hits))
        mine = np.concatenate([o[0] for o in outs])
        e2e_hits = np.concatenate([o[1] for o in outs])
    if rank != 0:
        return _finish(dist, world)
    q_sizes, r_sizes = np.diff(qoff).astype(np.int64), np.diff(roff).astype(np.int64)
    lines = _dist_lines(*api.refine_hits(mine, q_sizes, r_sizes, 2 * K, D, False), q_sizes, r_sizes)
    lines_e2e = _dist_lines(*api.refine_hits(e2e_hits, q_sizes, r_sizes, 2 * K, D, False), q_sizes, r_sizes)
    peak, peak_src = B.measured_hbm_peak()
    # algorithmic bytes: 4 B per query hash + 8 B offset pair + 4 B per posting streamed + 12 B per hit
    alg = 12.0 * float(qoff[r1] - qoff[r0]) + 4.0 * cnt_post / max(cnt_n, 1) + 12.0 * nh
    k_ms = cnt_ms / max(cnt_n, 1)
    out = {"metric": "dist genome-pairs/s (query x reference cells through the inverted index)", "value": pairs / (ms / 1e3), "unit": "genome-pairs/s",
           "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "u32 hashes, u16 smem counters, f64 distances", "data": "synthetic sketches (splitmix64 / PCG64 clusters)",
           "config": cfg, "gpu_launches": int(launches), "clocks": clk.summary(), "hits": len(lines),
           "roofline": {"bound": "hbm", "kernel": "dist_tile_kernel (rectangle: fused lookup)", "achieved": alg / (k_ms / 1e3) / 1e9 if k_ms else None,
                        "peak": peak, "unit": "GB/s", "frac": alg / (k_ms / 1e3) / 1e9 / peak if k_ms else None, "traffic": None, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": alg, "ms_per_launch": k_ms,
                        "note": "latency / shared-atomic bound, not HBM bound: the index is L2 + HBM resident and only ~6 MB of query hashes and their lists move"},
           "dist_kernel": {"postings_per_s": cnt_post / (cnt_ms / 1e3) if cnt_ms else None, "posting_GBps": 4 * cnt_post / (cnt_ms / 1e3) / 1e9 if cnt_ms else None,
                           "smem_atomics_per_s": cnt_atom / (cnt_ms / 1e3) if cnt_ms else None},
           "index": {"build_ms_rank0_incl_h2d": t_index * 1e3, "nccl_broadcast_ms": t_bcast * 1e3 if world > 1 else None,
                     "postings": int(roff[-1]), "note": "built once before the steps: the reference sketch index is a resident database"},
           "e2e": {"value": pairs / (e2e_ms / 1e3), "unit": "genome-pairs/s", "ms_per_step": e2e_ms, "steps": args.e2e_steps,
                   "h2d_bytes_per_step": int(qh.nbytes + qoff.nbytes), "d2h_bytes_per_step": int(len(e2e_hits)) * 12, "numa_rank0": numa},
           "parity": {"pairs_passing": len(lines), "sha256": B.lines_sha256(lines), "e2e_set_equal": lines == lines_e2e}}
    if world == 1 and not args.no_cpu_baseline:
        cb = ref_c3(args, cfg, rh, roff, qh, qoff, D, pairs, quiet=True)
        ref_lines = cb.pop("lines")
        out["cpu_baseline"] = cb
        out["parity"].update(set_equal=ref_lines == lines, reference_pairs_passing=len(ref_lines), only_in_reference=len(ref_lines - lines),
                             only_in_gpu=len(lines - ref_lines), checked_against="reference binary: r.dist line set of all query x reference cells")
    print(json.dumps(out))
    _finish(dist, world)


def ref_c3(args, cfg, rh, roff, qh, qoff, D, pairs, quiet=False):
    """`rabbit_kssd dist -r r.sketch -q q.sketch -D`: index files written by a first (untimed) run, the timed run loads
    them and computes the distances (all host cores)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    from rabbitkssd_b200 import formats
    ref = B.ref_binary()
    threads = os.cpu_count() or 1
    d = B.scratch_dir(1 << 30)
    try:
        if ref:
            formats.write_sketch(os.path.join(d, "r.sketch"), 9, 6, 3, [f"r{i}" for i in range(len(roff) - 1)], rh, roff)
            formats.write_sketch(os.path.join(d, "q.sketch"), 9, 6, 3, [f"q{i}" for i in range(len(qoff) - 1)], qh, qoff)
            argv = ["dist", "-r", "r.sketch", "-q", "q.sketch", "-D", str(D), "-o", "r.dist", "-t", str(threads)]
            B.run_ref(ref, argv, d)  # writes r.sketch.index / .dict
            times = [B.run_ref(ref, argv, d) for _ in range(max(1, min(args.steps, 3)))]
            wall = min(times)
            lines = set()
            for a, b, c, s0, s1, jorc, dd in formats.parse_dist_text(os.path.join(d, "r.dist")):
                lines.add(f"{int(a[1:])}\t{int(b[1:])}\t{c}|{s0}|{s1}\t{jorc}\t{dd}")
            kind = "reference"
            how = f"dist -r r.sketch -q q.sketch -D {D} -t {threads} with the index files in place: best of {len(times)} runs (sketch + index load, distance loop, merge)"
        else:
            import oracle
            t0 = time.perf_counter()
            hr, hc, hm = oracle.dist(qh, qoff, rh, roff, 18, D)
            wall = time.perf_counter() - t0
            q_sizes, r_sizes = np.diff(qoff).astype(np.int64), np.diff(roff).astype(np.int64)
            lines = set()
            for r, c, m in zip(hr.tolist(), hc.tolist(), hm.tolist()):
                jv, dv = oracle.distance(m, int(r_sizes[c]), int(q_sizes[r]), 18)
                lines.add(f"{r}\t{c}\t{m}|{int(r_sizes[c])}|{int(q_sizes[r])}\t{'%f' % jv}\t{'%f' % dv}")
            kind, threads = "port", oracle.max_threads()
            how = "oracle port (index build + distance loop)"
    finally:
        shutil.rmtree(d, ignore_errors=True)
    cb = {"value": pairs / wall, "unit": "genome-pairs/s", "cores": threads, "kind": kind, "sample": "the whole workload: " + how, "wall_s": wall}
    if quiet:
        cb["lines"] = lines
        return cb
    out = {"impl": "reference", "metric": "dist genome-pairs/s (query x reference cells through the inverted index)", "value": pairs / wall,
           "unit": "genome-pairs/s", "n_gpus": args.gpus, "steps": 1, "warmup": 1, "ms_per_step": wall * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int counters, f64 distances", "data": "synthetic sketches", "config": cfg,
           "cpu_baseline": cb, "parity": {"pairs_passing": len(lines), "sha256": B.lines_sha256(lines)},
           "e2e": {"value": pairs / wall, "unit": "genome-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------------------------------
def run_c4(args):
    ex = _extra(args)
    shuffle = ex.get("shuffle", "L4K10")
    k, s, l = {"L4K10": (10, 7, 4), "L5K10": (10, 7, 5), "L3K10": (10, 6, 3)}[shuffle]
    n_g, glen, n_contig = int(ex.get("genomes", 20)), int(float(ex.get("genome_len", 3e9))), int(ex.get("contigs", 24))
    sample_len = int(float(ex.get("sample_len", 2.4e8)))
    cfg = {"workload": f"sketch of {n_g} synthetic {glen / 1e9:g} Gbp genomes ({n_contig} contigs each), {shuffle} (BASELINE.json configs[3])",
           "genomes": n_g, "genome_len": glen, "contigs": n_contig, "shuffle": f"{shuffle} (k={k},s={s},l={l})", "seed": B.SEED + 4,
           "l2": f"no flush needed: every step streams {n_g * glen / 1e9:.0f} GB of sequence, far more than the 126 MB L2"}
    if args.impl == "reference":
        return ref_c4(args, cfg, k, s, l, sample_len, n_contig, n_g * glen)
    torch, dist, world, rank, local_rank, ctx, stream, all_cpus, numa = _setup(args)
    from rabbitkssd_b200 import api, synth
    if s - l < 3:
        ctx.set_relax_guard(True)
    dim = api.shuffle_table(k, s, l)
    ctx.set_params(k, s, l, dim)
    del dim
    mine = list(range(rank, n_g, world))  # genomes dealt round-robin over the ranks (no exchange: sketches are per genome)
    clen = glen // n_contig
    stride = (glen + n_contig + 15) // 16 * 16
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device="cuda")
    buf = torch.full((len(mine) * stride,), 10, dtype=torch.uint8, device="cuda")
    gen = torch.Generator(device="cuda")
    for gi, g in enumerate(mine):
        gen.manual_seed((B.SEED + 4) * 1000003 + g)
        o = gi * stride
        for c in range(n_contig):
            ln = clen if c < n_contig - 1 else glen - clen * (n_contig - 1)
            for b0 in range(0, ln, 1 << 28):
                nb = min(1 << 28, ln - b0)
                buf[o + b0:o + b0 + nb] = lut[torch.randint(0, 4, (nb,), dtype=torch.uint8, device="cuda", generator=gen).long()]
            o += ln + 1  # one separator byte after every contig
    # genome 0 of rank 0 starts with the deterministic sample the CPU baseline sketches (same bytes as FASTA)
    sample = None
    if rank == 0 and sample_len:
        sample = synth.family_device(torch, B.SEED + 4, 1, sample_len)[0]
    n_local = len(mine)
    seg_off_np = np.arange(n_local + 1, dtype=np.uint64) * np.uint64(stride)
    seg_gen_np = np.arange(n_local, dtype=np.uint32)
    d_so = torch.from_numpy(seg_off_np.astype(np.int64)).cuda()
    d_sg = torch.from_numpy(seg_gen_np.astype(np.int32)).cuda()
    n_bytes = n_local * stride

    def step():
        ctx.sketch_begin(n_local)
        ctx.scan_device(buf.data_ptr(), n_bytes, d_so.data_ptr(), d_sg.data_ptr(), n_local)
        return ctx.sketch_finish_device()

    with B.ClockSampler(local_rank) as clk:
        ctx.scan_kernel_time()
        launches0 = ctx.launch_count()
        ms, (dh, do, tot) = _timed(torch, dist, world, stream, args.warmup, args.steps, step)
    launches = (ctx.launch_count() - launches0) // (max(args.warmup, 3) + args.steps) * args.steps
    scan_ms, scan_n, scan_bytes = ctx.scan_kernel_time()
    scan_ms, scan_n, scan_bytes = (scan_ms * args.steps / (max(args.warmup, 3) + args.steps), scan_n * args.steps // (max(args.warmup, 3) + args.steps),
                                   scan_bytes * args.steps // (max(args.warmup, 3) + args.steps))
    off_dev = B.dev_tensor(torch, do, n_local + 1, torch.int64).cpu().numpy()
    h_dev = B.dev_tensor(torch, dh, int(tot), torch.int32).cpu().numpy().view(np.uint32).copy()
    # size-independent properties at full size: per genome sorted + unique, size ~ bases / 16^l
    ok_sorted = all(np.all(np.diff(h_dev[off_dev[g]:off_dev[g + 1]].astype(np.int64)) > 0) for g in range(n_local))
    sizes = np.diff(off_dev)
    expect = glen / 16.0 ** l
    ok_size = bool(np.all(np.abs(sizes - expect) < 6 * expect ** 0.5 + 0.002 * expect))

    # e2e: pinned host bytes -> scan_host (double-buffered H2D inside the timed region) -> sketches read back
    host = torch.empty(n_bytes, dtype=torch.uint8, pin_memory=True)
    host.copy_(buf)
    torch.cuda.synchronize()

    def e2e_step():
        ctx.sketch_begin(n_local)
        ctx.scan_host(host.data_ptr(), n_bytes, seg_off_np, seg_gen_np)
        return ctx.sketch_finish()
    e2e_ms, (eh, eo) = _timed(torch, dist, world, stream, 1, max(1, min(args.e2e_steps, 3)), e2e_step)
    e2e_equal = bool(np.array_equal(eh, h_dev) and np.array_equal(eo.astype(np.int64), off_dev))
    del host
    os.sched_setaffinity(0, all_cpus)
    sample_sets = None
    if sample is not None:  # the deterministic sample through the same kernels, for the CPU baseline's parity check
        cuts = [sample_len * c // n_contig for c in range(n_contig + 1)]
        sbuf = torch.full(((sample_len + n_contig + 15) // 16 * 16,), 10, dtype=torch.uint8, device="cuda")
        o = 0
        for c in range(n_contig):
            sbuf[o:o + cuts[c + 1] - cuts[c]] = sample[cuts[c]:cuts[c + 1]]
            o += cuts[c + 1] - cuts[c] + 1
        so = torch.tensor([0, sbuf.numel()], dtype=torch.int64, device="cuda")
        sg = torch.zeros(1, dtype=torch.int32, device="cuda")
        ctx.sketch_begin(1)
        ctx.scan_device(sbuf.data_ptr(), sbuf.numel(), so.data_ptr(), sg.data_ptr(), 1)
        sample_sets, _ = ctx.sketch_finish()
    if rank != 0:
        return _finish(dist, world)
    peak, peak_src = B.measured_hbm_peak()
    alg = scan_bytes / max(scan_n, 1) + 8.0 * float(tot) / max(scan_n / max(args.steps, 1), 1)
    per = scan_ms / max(scan_n, 1)
    ach = alg / (per / 1e3) / 1e9 if per else 0.0
    t_ratio, t_src = B.measured_traffic_ratio("dram_bytes_per_input_byte_g4")
    gbp = n_g * glen / 1e9
    out = {"metric": "sketch Gbp/s (Kssd sketch of whole genomes: scan + per-genome sort + unique)", "value": gbp / (ms / 1e3), "unit": "Gbp/s",
           "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "u8 bases -> u32 hashes (u64 k-mer arithmetic)", "data": "synthetic (uniform ACGT generated in HBM)",
           "config": cfg, "gpu_launches": int(launches), "clocks": clk.summary(), "hashes": int(tot) if world == 1 else None,
           "roofline": {"bound": "hbm", "kernel": f"kssd_scan_kernel<G={max(1, min(4, 2 * s - 9))}>", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                        "traffic": (t_ratio * scan_bytes / max(scan_n, 1)) if t_ratio else None, "traffic_source": t_src, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": alg, "ms_per_launch": per, "share_of_step": scan_ms / (ms * args.steps) if ms else None},
           "e2e": {"value": gbp / (e2e_ms / 1e3), "unit": "Gbp/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(n_bytes) * world,
                   "d2h_bytes_per_step": int(eh.nbytes + eo.nbytes), "numa_rank0": numa},
           "parity": {"full_size_properties": {"sorted_unique_per_genome": bool(ok_sorted), "sizes_within_6_sigma_of_bases_over_16^l": ok_size,
                                               "e2e_equals_device_resident": e2e_equal}}}
    if s - l < 3:
        out["parity"]["note"] = "reference: unsupported configuration (common.cpp:37-39); parity unpinned, oracle = restatement with the guard relaxed"
    if world == 1 and not args.no_cpu_baseline and sample_sets is not None:
        cb = ref_c4(args, cfg, k, s, l, sample_len, n_contig, n_g * glen, quiet=True)
        got = cb.pop("sets")
        out["cpu_baseline"] = cb
        out["parity"].update(sample_set_equal=None if got is None else bool(np.array_equal(np.sort(got).astype(np.uint32), sample_sets)),
                             checked_against=f"{cb['kind']}: hash set of a {sample_len / 1e6:g} Mbp sample genome ({n_contig} contigs) sketched from the same bytes")
    print(json.dumps(out))
    _finish(dist, world)


def ref_c4(args, cfg, k, s, l, sample_len, n_contig, total_bases, quiet=False):
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    import oracle
    from rabbitkssd_b200 import formats
    ref = B.ref_binary() if s - l >= 3 else None  # the reference refuses s - l < 3
    threads = os.cpu_count() or 1
    d = B.scratch_dir(sample_len * 1.1 + (5 << 30 if s == 7 else 1 << 28))
    sets = None
    try:
        os.makedirs(os.path.join(d, "x"))
        oracle.synth_fasta_files(os.path.join(d, "x"), B.SEED + 4, 0, 1, 1, sample_len, 0.0, n_contig)
        if ref:
            B.run_ref(ref, ["shuffle", "-k", str(k), "-s", str(s), "-l", str(l), "-o", "L.shuf"], d)
            with open(os.path.join(d, "list.txt"), "w") as f:
                f.write("x/g0.fna\n")
            wall = min(B.run_ref(ref, ["sketch", "-q", "-L", "L.shuf", "-i", "list.txt", "-o", "s", "-t", str(threads)], d) for _ in range(2))
            sets = formats.read_sketch(os.path.join(d, "s.sketch"))[2][0]
            kind = "reference"
            how = f"sketch -q -t {threads} of one {sample_len / 1e6:g} Mbp genome ({n_contig} contigs, FASTA; best of 2, incl. the {4 * 16 ** s / 2 ** 20:.0f} MiB .shuf load)"
        else:
            from rabbitkssd_b200 import synth
            p = oracle.init_params(k, s, l, relax_guard=True)
            dim = oracle.shuffle_dim(k, s, l)
            g = oracle.synth_genome(B.SEED + 4, 0, 1, sample_len, 0.0)
            cuts = [sample_len * c // n_contig for c in range(n_contig + 1)]
            seq, rec_off, rec_first = synth.oracle_records([[g[cuts[c]:cuts[c + 1]] for c in range(n_contig)]])
            t0 = time.perf_counter()
            sets = oracle.sketch_batch(p, dim, seq, rec_off, rec_first)[0]
            wall = time.perf_counter() - t0
            kind, threads = "port", 1
            how = f"oracle port, one thread, one {sample_len / 1e6:g} Mbp genome (the reference refuses this configuration)"
    finally:
        shutil.rmtree(d, ignore_errors=True)
    gbps = sample_len / 1e9 / wall
    cb = {"value": gbps, "unit": "Gbp/s", "cores": threads, "kind": kind, "sample": how, "wall_s": wall}
    if quiet:
        cb["sets"] = sets
        return cb
    out = {"impl": "reference", "metric": "sketch Gbp/s (Kssd sketch of whole genomes: scan + per-genome sort + unique)", "value": gbps, "unit": "Gbp/s",
           "n_gpus": args.gpus, "steps": 1, "warmup": 1, "ms_per_step": total_bases / 1e9 / gbps * 1e3, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "u64 k-mer arithmetic", "data": "synthetic", "config": cfg, "cpu_baseline": cb,
           "e2e": {"value": gbps, "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------------------------------
def run_c5(args):
    ex = _extra(args)
    n_clusters, cluster, size, D = int(ex.get("clusters", 80)), int(ex.get("cluster", 1250)), int(ex.get("size", 1221)), float(ex.get("D", 0.05))
    sample_clusters = int(ex.get("sample_clusters", 8))
    K = 10
    n = n_clusters * cluster
    cfg = {"workload": f"alldist of {n} synthesised sketches ({n_clusters} clusters x {cluster}, S ~ {size}), L3K10, -D {D} (BASELINE.json configs[4])",
           "sketches": n, "clusters": n_clusters, "cluster": cluster, "sketch_size": size, "max_dist": D, "seed": B.SEED + 5,
           "l2": "postings (~0.5 GB) + work lists exceed the 126 MB L2"}
    from rabbitkssd_b200 import synth
    pairs = n * (n - 1) / 2.0
    if args.impl == "reference":
        return ref_c5(args, cfg, n_clusters, cluster, size, D, sample_clusters, pairs)
    torch, dist, world, rank, local_rank, ctx, stream, all_cpus, numa = _setup(args)
    from rabbitkssd_b200 import _lib, api, multigpu
    h, off = synth.synth_sketches(B.SEED + 5, n_clusters, cluster, size, 28, survival_step=0.55 / max(cluster - 1, 1))
    ctx.set_params(K, 6, 3, api.shuffle_table(K, 6, 3))
    d_h = torch.from_numpy(h.view(np.int32)).cuda()
    d_o = torch.from_numpy(off.astype(np.int64)).cuda()
    b = multigpu.triangle_bounds(n, world, row_cost=40.0 * size)
    r0, r1 = b[rank], b[rank + 1]

    def step():
        if world == 1:
            ctx.index_build_device(d_h.data_ptr(), d_o.data_ptr(), n)
        else:
            try:
                ctx.index_build_rows_device(d_h.data_ptr(), d_o.data_ptr(), n, (r0, r1))
            except _lib.KssdError as e:
                if e.code != -5:
                    raise
                ctx.index_build_device(d_h.data_ptr(), d_o.data_ptr(), n)
        return ctx.dist_rows_device(d_h.data_ptr(), d_o.data_ptr(), n, True, 2 * K, D, 0, rows=(r0, r1))

    with B.ClockSampler(local_rank) as clk:
        ctx.count_kernel_time()
        launches0 = ctx.launch_count()
        ms, (hp, nh) = _timed(torch, dist, world, stream, args.warmup, args.steps, step)
    launches = (ctx.launch_count() - launches0) // (max(args.warmup, 3) + args.steps) * args.steps
    cnt_ms, cnt_n, cnt_post, cnt_atom = ctx.count_kernel_time()
    nh_t = torch.tensor([nh], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(nh_t)
    # e2e: host sketch arrays -> index build (H2D) -> alldist rows -> hits through the sink (D2H); one step (hundreds of MB of hits)
    def e2e_step():
        if world == 1:
            ctx.index_build(h, off)
        else:
            ctx.index_build_rows(h, off, (r0, r1))
        return ctx.dist_rows(h, off, True, 2 * K, D, 0, rows=(r0, r1))
    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e2e_hits = e2e_step()
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    os.sched_setaffinity(0, all_cpus)
    # parity at full size, this rank's rows: (a) the device-resident and the host path report the same cells; (b) sampled row
    # blocks of rank 0 against the oracle (exact refine: the reference's double expressions)
    mine = B.dev_tensor(torch, hp, int(nh) * 3, torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 3)
    same = bool(np.array_equal(mine[np.lexsort((mine[:, 1], mine[:, 0]))], e2e_hits[np.lexsort((e2e_hits[:, 1], e2e_hits[:, 0]))]))
    same_t = torch.tensor([int(same)], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(same_t, op=dist.ReduceOp.MIN)
    if rank != 0:
        return _finish(dist, world)
    import oracle
    sizes = np.diff(off).astype(np.int64)
    blocks_ok, n_checked = True, 0
    for lo in (r0, r0 + (r1 - r0) // 2, max(r0, r1 - 64)):  # the last rows of a cluster have no partner under the threshold: empty blocks count
        hi = min(r1, lo + 64)
        hr, hc, hm = oracle.alldist(h, off, 2 * K, D, rows=(lo, hi))
        want = np.stack([hr, hc, hm], axis=1).astype(np.int64)
        g = api.refine_hits(mine[(mine[:, 0] >= lo) & (mine[:, 0] < hi)], sizes, sizes, 2 * K, D, True)[0].astype(np.int64)
        g = g[np.lexsort((g[:, 1], g[:, 0]))]
        blocks_ok = blocks_ok and np.array_equal(g, want)
        n_checked += len(want)
    blocks_ok = blocks_ok and n_checked > 1000
    peak, peak_src = B.measured_hbm_peak()
    k_ms = cnt_ms / max(cnt_n, 1)
    alg = 4.0 * cnt_post / max(cnt_n, 1) + 12.0 * nh + 12.0 * float(off[r1] - off[r0])
    out = {"metric": "alldist genome-pairs/s (index build + alldist of pre-computed sketches)", "value": pairs / (ms / 1e3), "unit": "genome-pairs/s",
           "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "u32 hashes, u16 smem counters, f64 distances", "data": "synthetic sketches (splitmix64 clusters)",
           "config": cfg, "gpu_launches": int(launches), "clocks": clk.summary(), "hits_device_superset": int(nh_t.item()),
           "text_output_bytes_estimate": int(nh_t.item()) * 46,
           "roofline": {"bound": "hbm", "kernel": "row_work_kernel + dist_tile_kernel (triangle)", "achieved": alg / (k_ms / 1e3) / 1e9 if k_ms else None, "peak": peak,
                        "unit": "GB/s", "frac": alg / (k_ms / 1e3) / 1e9 / peak if k_ms else None, "traffic": None, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": alg, "ms_per_launch": k_ms, "share_of_step": k_ms / ms if ms else None,
                        "note": "shared-memory-atomic bound (one increment per posting streamed), see smem_atomics_per_s"},
           "alldist_kernel": {"postings_per_s": cnt_post / (cnt_ms / 1e3) if cnt_ms else None, "posting_GBps": 4 * cnt_post / (cnt_ms / 1e3) / 1e9 if cnt_ms else None,
                              "smem_atomics_per_s": cnt_atom / (cnt_ms / 1e3) if cnt_ms else None},
           "e2e": {"value": pairs / (e2e_ms / 1e3), "unit": "genome-pairs/s", "ms_per_step": e2e_ms, "steps": 1,
                   "h2d_bytes_per_step": int(h.nbytes + off.nbytes) * world, "d2h_bytes_per_step": int(nh_t.item()) * 12, "numa_rank0": numa},
           "parity": {"device_and_host_path_sets_equal_every_rank": bool(same_t.item()), "sampled_row_blocks_equal_oracle": bool(blocks_ok), "sampled_pairs": int(n_checked),
                      "checked_against": "oracle (pinned to the reference binary) on 3 blocks of 64 rows of rank 0 at full size"}}
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = ref_c5(args, cfg, n_clusters, cluster, size, D, sample_clusters, pairs, quiet=True)
    print(json.dumps(out))
    _finish(dist, world)


def ref_c5(args, cfg, n_clusters, cluster, size, D, sample_clusters, pairs, quiet=False):
    """Bounded sample: `alldist -i x.sketch` on the first sample_clusters clusters (cells scale with n^2: extrapolated on cells)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    from rabbitkssd_b200 import formats, synth
    ref = B.ref_binary()
    threads = os.cpu_count() or 1
    sc = min(sample_clusters, n_clusters)
    h, off = synth.synth_sketches(B.SEED + 5, sc, cluster, size, 28, survival_step=0.55 / max(cluster - 1, 1))
    ns = len(off) - 1
    d = B.scratch_dir(6 << 30)
    try:
        if ref:
            formats.write_sketch(os.path.join(d, "x.sketch"), 10, 6, 3, [f"g{i}" for i in range(ns)], h, off)
            wall = B.run_ref(ref, ["alldist", "-i", "x.sketch", "-D", str(D), "-o", "x.alldist", "-t", str(threads)], d)
            n_lines = sum(1 for _ in open(os.path.join(d, "x.alldist"), "rb")) - 1 if os.path.exists(os.path.join(d, "x.alldist")) else None
            kind = "reference"
        else:
            import oracle
            t0 = time.perf_counter()
            n_lines = oracle.alldist(h, off, 20, D, want_hits=False)
            wall = time.perf_counter() - t0
            kind, threads = "port", oracle.max_threads()
    finally:
        shutil.rmtree(d, ignore_errors=True)
    cells = ns * (ns - 1) / 2.0
    cb = {"value": cells / wall, "unit": "genome-pairs/s", "cores": threads, "kind": kind, "wall_s": wall, "pairs_passing_in_sample": n_lines,
          "sample": f"alldist -i x.sketch -D {D} -t {threads} on the first {sc} clusters ({ns} sketches, {cells:.3g} cells): index build (fixed ~15 s for 2^28 "
                    f"buckets) + load + distance loop + text output; value = sample cells / wall"}
    if quiet:
        return cb
    out = {"impl": "reference", "metric": "alldist genome-pairs/s (index build + alldist of pre-computed sketches)", "value": cells / wall, "unit": "genome-pairs/s",
           "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": pairs / (cells / wall) * 1e3, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "int counters, f64 distances", "data": "synthetic sketches", "config": cfg, "cpu_baseline": cb,
           "e2e": {"value": cells / wall, "unit": "genome-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out))


def main(args):
    return {"C3": run_c3, "C4": run_c4, "C5": run_c5}[args.config](args)
