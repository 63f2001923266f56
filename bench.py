#!/usr/bin/env python
"""bench.py -- BASELINE.json config[1]: alldist of 10,000 synthetic 5 Mbp bacterial genomes (L3K10, -D 0.05).

One step = the whole hot path over the whole workload: Kssd sketch scan of every genome (parsed bytes resident in
HBM) -> per-genome radix sort + unique -> [N>1: all-gather of the sketch shards over NCCL] -> inverted-index build
-> alldist posting accumulation + distance/threshold epilogue for this rank's query-row block -> compacted hits.
metric = genome-pairs/s = N(N-1)/2 / step time (whole job, max over ranks); strong scaling over GPUs.
`e2e` runs the same step from PINNED HOST buffers through the C-ABI (H2D inside the timed region, hits read back).
`roofline` is the sketch scan kernel (the dominant kernel) against MEASURED_PEAKS.json's HBM copy bandwidth.
`cpu_baseline` / `--impl reference` time the unmodified reference binary (oracle/_ref/rabbit_kssd) on the host cores.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

K, S, L_ = 10, 6, 3
MAX_DIST = 0.05
SEED = 20261104 + 2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--genomes", type=int, default=10000)
    ap.add_argument("--genome-len", type=int, default=5000000)
    ap.add_argument("--family", type=int, default=10)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--nccl-gather", action="store_true", help="N > 1: gather the sketch shards with one NCCL all-gather instead of peer-memory reads")
    ap.add_argument("--index-sharded", action="store_true", help="N > 1: build the index sharded by hash range and all-gather the pieces")
    ap.add_argument("--phases", action="store_true", help="diagnostic: per-phase wall times (synchronised) on stderr after the timed runs")
    ap.add_argument("--cpu-sample-genomes", type=int, default=256)
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.stop = threading.Event()
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                     timeout=5).stdout.decode().strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.03)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].startswith("Active") for s in self.samples)]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


def measured_traffic_ratio():
    """DRAM bytes per input byte of the scan kernel from the committed `ncu --set full` capture (profiles/)."""
    p = os.path.join(ROOT, "profiles", "scan_traffic.json")
    try:
        d = json.load(open(p))
        return float(d["dram_bytes_per_input_byte"]), d["capture"]
    except Exception:
        return None, None


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------------------
# reference binary helpers (CPU legs)
def ref_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "rabbit_kssd")
    if os.access(p, os.X_OK):
        try:
            if subprocess.run([p, "--help"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=20).returncode == 0:
                return p
        except Exception:
            pass
    return None


def scratch_dir():
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    return tempfile.mkdtemp(prefix="kssd_bench_", dir=base)


def host_genomes(n, length, family, seed):
    from rabbitkssd_b200 import synth
    out = []
    f = 0
    while len(out) < n:
        for g in synth.family(seed + 101 * f, family, length, 0.04, 2):
            if len(out) < n:
                out.append(g)
        f += 1
    return out


def ref_sketch_rate(ref, d, n_genomes, length, family, threads):
    """`rabbit_kssd sketch -q` on n_genomes FASTA files -> (Gbp/s incl. .shuf load, wall s)."""
    from rabbitkssd_b200 import synth
    names = []
    for i, g in enumerate(host_genomes(n_genomes, length, family, SEED)):
        nm = f"g{i}.fna"
        synth.write_fasta(os.path.join(d, nm), g, name=nm)
        names.append(nm)
    open(os.path.join(d, "list.txt"), "w").write("\n".join(names) + "\n")
    if not os.path.exists(os.path.join(d, "L3K10.shuf")):
        subprocess.run([ref, "shuffle", "-k", str(K), "-s", str(S), "-l", str(L_), "-o", "L3K10.shuf"], cwd=d, check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)

    def once():
        t0 = time.perf_counter()
        subprocess.run([ref, "sketch", "-q", "-L", "L3K10.shuf", "-i", "list.txt", "-o", "s", "-t", str(threads)], cwd=d,
                       check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return time.perf_counter() - t0
    return once


def ref_alldist_from_sketch(ref, d, hashes, off, threads):
    """`rabbit_kssd alldist -i x.sketch -D 0.05`: index build + load + distance loop + merge -> (wall s, #lines)."""
    from rabbitkssd_b200 import formats
    n = len(off) - 1
    formats.write_sketch(os.path.join(d, "x.sketch"), K, S, L_, [f"g{i}" for i in range(n)], hashes, off)
    for f in ("x.sketch.index", "x.sketch.dict"):
        if os.path.exists(os.path.join(d, f)):
            os.remove(os.path.join(d, f))
    t0 = time.perf_counter()
    subprocess.run([ref, "alldist", "-i", "x.sketch", "-D", str(MAX_DIST), "-o", "x.alldist", "-t", str(threads)], cwd=d,
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    wall = time.perf_counter() - t0
    lines = 0
    with open(os.path.join(d, "x.alldist"), "rb") as f:
        for _ in f:
            lines += 1
    return wall, lines - 1  # header


def port_rates(n_sample, length, family, hashes, off, threads):
    """Fallback when the reference binary is unavailable: the oracle port under OpenMP."""
    import oracle
    from rabbitkssd_b200 import synth
    p = oracle.init_params(K, S, L_)
    dim = oracle.shuffle_dim(K, S, L_)
    seq, rec_off, rec_first = synth.oracle_records(host_genomes(n_sample, length, family, SEED))
    t0 = time.perf_counter()
    oracle.sketch_batch(p, dim, seq, rec_off, rec_first)
    t_sk = time.perf_counter() - t0
    t0 = time.perf_counter()
    n_hit = oracle.alldist(hashes, off, 2 * K, MAX_DIST, want_hits=False)
    t_ad = time.perf_counter() - t0
    return t_sk, t_ad, n_hit


# ---------------------------------------------------------------------------------------------------------------
class DevView:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def dev_tensor(torch, ptr, n, dtype):
    if n == 0:
        return torch.zeros(0, dtype=dtype, device="cuda")
    ts = {torch.int32: "<i4", torch.int64: "<i8", torch.uint8: "|u1"}[dtype]
    return torch.as_tensor(DevView(ptr, n, ts), device="cuda")


def gen_genomes_device(torch, n_local, first_genome, length, family, stride):
    """Synthetic bacterial families straight into HBM: base = uniform ACGT, member m = base with i.i.d. substitutions
    at rate 0.04*m/(family-1); two contigs separated by one '\\n'.  Layout: genome g at bytes [g*stride, g*stride+length+1)."""
    buf = torch.full((n_local * stride,), 10, dtype=torch.uint8, device="cuda")
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device="cuda")
    half = length // 2
    for f0 in range(0, n_local, family):
        fam_id = (first_genome + f0) // family
        gen = torch.Generator(device="cuda")
        gen.manual_seed(SEED * 1000003 + fam_id)
        base = torch.randint(0, 4, (length,), dtype=torch.uint8, device="cuda", generator=gen)
        for m in range(min(family, n_local - f0)):
            codes = base
            if m:
                rate = 0.04 * m / max(1, family - 1)
                mut = torch.rand(length, device="cuda", generator=gen) < rate
                shift = torch.randint(1, 4, (length,), dtype=torch.uint8, device="cuda", generator=gen)
                codes = (base + shift * mut) & 3
            o = (f0 + m) * stride
            asc = lut[codes.long()]
            buf[o:o + half] = asc[:half]
            buf[o + half + 1:o + length + 1] = asc[half:]
    return buf


def run_b200(args):
    import torch
    import torch.distributed as dist
    from rabbitkssd_b200 import api
    from rabbitkssd_b200.multigpu import PeerGather, build_index_sharded, gather_sketches, gather_sketches_device, triangle_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_total, length, family = args.genomes, args.genome_len, args.family
    assert n_total % (world * family) == 0, "genomes must divide evenly into families per rank"
    n_local = n_total // world
    first = rank * n_local
    stride = (length + 2 + 15) // 16 * 16

    ctx = api.Context(local_rank)
    # ONE explicit (non-default) stream for torch, NCCL and the library: handle 0 (the legacy default stream) means
    # "library-owned stream" to kssd_b200_set_stream, and torch ops on stream 0 would then race with the kernels
    stream = torch.cuda.Stream(device=local_rank)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx.set_stream(stream.cuda_stream)
    dim = api.shuffle_table(K, S, L_)  # the .shuf table is an INPUT of the path (glibc-rand permutation, host code)
    ctx.set_params(K, S, L_, dim)
    ctx.set_profiling(True)

    buf = gen_genomes_device(torch, n_local, first, length, family, stride)
    seg_off_np = (np.arange(n_local + 1, dtype=np.uint64) * np.uint64(stride))
    seg_gen_np = np.arange(n_local, dtype=np.uint32)
    d_seg_off = torch.from_numpy(seg_off_np.astype(np.int64)).cuda()
    d_seg_gen = torch.from_numpy(seg_gen_np.astype(np.int32)).cuda()
    n_bytes = n_local * stride
    sk_size = length / 16.0 ** L_                      # expected hashes per genome
    gather_cap = int(1.25 * n_local * sk_size) + 1024  # block size of the sketch all-gather, the same on every rank
    # a row costs its cells plus per-hash work (posting-range lookup, posting streaming).  Fitted to the dist phase of
    # the 1- and 8-GPU runs (profiles/r01x): ~50 cell-equivalents per hash, i.e. the per-row part dominates
    bounds = triangle_bounds(n_total, world, row_cost=40.0 * sk_size)
    r0, r1 = bounds[rank], bounds[rank + 1]
    state = {}
    if world > 1 and not args.nccl_gather:
        state["peers"] = PeerGather(dist, ctx, n_local, gather_cap, rank, world)

    phase_t = {}

    def mark(name, t0):
        if args.phases and state.get("diag"):
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            phase_t[name] = phase_t.get(name, 0.0) + (t1 - t0) * 1e3
            return t1
        return t0

    def finish_and_dist(keep=False):
        t0 = mark("scan", state.get("t_step0", 0.0))
        d_h, d_o, tot = ctx.sketch_finish_device()
        if state.setdefault("tot0", tot) != tot:  # the sketch of a fixed input must not change between steps
            raise RuntimeError(f"rank {rank}: sketch size changed between steps: {state['tot0']} -> {tot}")
        t0 = mark("sort_unique", t0)
        if world > 1:
            def shard():
                o_loc = dev_tensor(torch, d_o, n_local + 1, torch.int64)
                return dev_tensor(torch, d_h, tot, torch.int32), o_loc[1:] - o_loc[:-1]
            got = state["peers"].gather() if not args.nccl_gather else gather_sketches_device(dist, ctx, *shard(), world, gather_cap)
            if got is not None:
                ph, po, state["total_all"] = got
            else:  # some rank's shard exceeded the agreed block size: two-step gather with exact sizes
                hashes, off = gather_sketches(dist, *shard(), world)
                ph, po = hashes.data_ptr(), off.data_ptr()
                state["keep"] = (hashes, off)
                state["total_all"] = hashes.numel()
        else:
            ph, po = d_h, d_o
        t0 = mark("gather", t0)
        # every rank builds the whole index from the gathered sketches (0.69 ms for 12.2 M postings); the build sharded by
        # hash range + all-gather of the pieces (--index-sharded) measured the same at 4 GPUs (profiles/r01u): its
        # fixed costs (two more collectives, adopt pass) eat what the smaller sort saves
        if world == 1 or not args.index_sharded or not build_index_sharded(dist, ctx, ph, po, n_total, state["total_all"], rank, world):
            ctx.index_build_device(ph, po, n_total)
        t0 = mark("index", t0)
        d_hits, n_hits = ctx.dist_rows_device(ph, po, n_total, True, 2 * K, MAX_DIST, 0, rows=(r0, r1))
        t0 = mark("dist", t0)
        state.update(d_hits=d_hits, n_hits=n_hits, ph=ph, po=po)
        return n_hits

    def step_device():
        if args.phases and state.get("diag"):
            torch.cuda.synchronize()
            state["t_step0"] = time.perf_counter()
        ctx.sketch_begin(n_local)
        ctx.scan_device(buf.data_ptr(), n_bytes, d_seg_off.data_ptr(), d_seg_gen.data_ptr(), n_local)
        return finish_and_dist()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # clocks are sampled over warm-up + timed steps (the same load): the timed region alone can be shorter than one
    # nvidia-smi poll
    with ClockSampler(local_rank) as clk:
        for _ in range(max(args.warmup, 3)):
            step_device()
        ctx.scan_kernel_time()
        ctx.count_kernel_time()
        launches0 = ctx.launch_count()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            n_hits = step_device()
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - launches0
    scan_ms, scan_n, scan_bytes = ctx.scan_kernel_time()
    cnt_ms, cnt_n, cnt_post, cnt_atom = ctx.count_kernel_time()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    hits_t = torch.tensor([n_hits], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(hits_t, op=dist.ReduceOp.SUM)
    ms = float(t.item())
    ms_per_step = ms / args.steps
    pairs = n_total * (n_total - 1) / 2.0
    value = pairs / (ms_per_step / 1e3)

    if args.phases:
        state["diag"] = True
        for _ in range(3):
            step_device()
        state["diag"] = False
        sys.stderr.write("phases rank %d (ms/step, synchronised): %s\n" % (rank, json.dumps({k: round(v / 3, 3) for k, v in phase_t.items()})))
        from rabbitkssd_b200 import multigpu as _mg
        if _mg.TRACE:
            sys.stderr.write("multigpu trace rank %d (median ms per call, synchronised): %s\n" % (rank, json.dumps({k: round(sorted(v)[len(v) // 2], 3) for k, v in _mg.TRACE.items() if k})))

    # ---- end to end: pinned host buffers -> C-ABI (H2D inside the timed region) -> hits read back ----
    host = torch.empty(n_bytes, dtype=torch.uint8, pin_memory=True)
    host.copy_(buf)
    torch.cuda.synchronize()
    hits_host = torch.empty(max(int(state["n_hits"]) * 2, 1 << 20) * 3, dtype=torch.int32, pin_memory=True)

    def step_e2e():
        ctx.sketch_begin(n_local)
        ctx.scan_host(host.data_ptr(), n_bytes, seg_off_np, seg_gen_np)
        n = finish_and_dist()
        hv = dev_tensor(torch, state["d_hits"], n * 3, torch.int32)
        hits_host[:n * 3].copy_(hv, non_blocking=True)
        torch.cuda.synchronize()
        return n

    step_e2e()
    barrier()
    e0.record(stream)
    for _ in range(args.e2e_steps):
        n_e2e = step_e2e()
    e1.record(stream)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item()) / args.e2e_steps
    e2e_val = pairs / (e2e_ms / 1e3)
    ctx.scan_kernel_time()
    ctx.count_kernel_time()

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_hbm_peak()
    cells_rank0 = sum(n_total - 1 - i for i in range(r0, r1))
    hashes_total = int(dev_tensor(torch, state["po"], n_total + 1, torch.int64)[-1].item())
    # algorithmic bytes of the scan kernel: 1 B per input byte + 8 B per emitted (genome,hash) key;
    # emitted keys ~ unique hashes (duplicates inside a 5 Mbp genome are negligible)
    alg_bytes = scan_bytes / max(scan_n, 1) + 8.0 * (hashes_total / world) / max(scan_n / max(args.steps, 1), 1)
    scan_ms_per = scan_ms / max(scan_n, 1)
    t_ratio, t_src = measured_traffic_ratio()
    achieved = alg_bytes / (scan_ms_per / 1e3) / 1e9 if scan_ms_per > 0 else 0.0
    out = {
        "metric": "alldist genome-pairs/s (sketch + index + alldist of the whole genome set)", "value": value,
        "unit": "genome-pairs/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8 bases -> u32 hashes (u64 k-mer arithmetic), u16 smem counters, f64 distances",
        "data": "synthetic (generated in HBM with torch: uniform ACGT families + substitution variants)",
        "config": {"workload": f"alldist of {n_total} synthetic {length / 1e6:g} Mbp genomes, L3K10, -D {MAX_DIST} "
                               "(BASELINE.json configs[1])", "genomes": n_total, "genome_len": length,
                   "family": family, "shuffle": "L3K10 (k=10,s=6,l=3)", "max_dist": MAX_DIST,
                   "partition": f"genomes and triangle query rows sharded over {world} rank(s)",
                   "l2": "inputs (%.1f GB per rank) exceed the 126 MB L2" % (n_bytes / 1e9)},
        "sketch_gbps": (n_total * length) / ((scan_ms / max(args.steps, 1)) / 1e3) / 1e9 if scan_ms > 0 else None,
        "sketch_gbps_note": "bases of ALL ranks / rank-0 scan-kernel time per step (ranks run concurrently)",
        "hits": int(hits_t.item()),
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "roofline": {"bound": "hbm", "kernel": "kssd_scan_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak,
                     "traffic": (t_ratio * scan_bytes / max(scan_n, 1)) if t_ratio else None, "traffic_source": t_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": scan_ms_per,
                     "share_of_step": scan_ms / ms if ms else None},
        "alldist_kernel": {"ms_per_launch": cnt_ms / max(cnt_n, 1), "postings_per_s": cnt_post / (cnt_ms / 1e3) if cnt_ms else None,
                           "posting_GBps": 4 * cnt_post / (cnt_ms / 1e3) / 1e9 if cnt_ms else None,
                           "smem_atomics_per_s": cnt_atom / (cnt_ms / 1e3) if cnt_ms else None,
                           "pairs_per_s_kernel_only": cells_rank0 / ((cnt_ms / max(cnt_n, 1)) / 1e3) if cnt_ms else None},
        "e2e": {"value": e2e_val, "unit": "genome-pairs/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(n_bytes) * world,
                "d2h_bytes_per_step": int(n_e2e) * 12},
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, torch, state, n_total, length, family, pairs, int(hits_t.item()), ctx)
    print(json.dumps(out))
    sys.stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(args, torch, state, n_total, length, family, pairs, gpu_hits_device, ctx):
    """Reference CPU path on this box's host cores, bounded: (a) `sketch -q` on a sample of genomes -> Gbp/s,
    extrapolated linearly to all genomes; (b) `alldist -i x.sketch` on the sketches of ALL genomes (the ones the
    GPU just produced, so the reference's line count doubles as a full-size parity check)."""
    from rabbitkssd_b200 import api
    threads = os.cpu_count() or 1
    off = dev_tensor(torch, state["po"], n_total + 1, torch.int64).cpu().numpy().astype(np.uint64)
    hashes = dev_tensor(torch, state["ph"], int(off[-1]), torch.int32).cpu().numpy().view(np.uint32)
    sizes = np.diff(off).astype(np.int64)
    hv = dev_tensor(torch, state["d_hits"], int(state["n_hits"]) * 3, torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 3)
    kept, _, _ = api.refine_hits(hv, sizes, sizes, 2 * K, MAX_DIST, True)
    ref = ref_binary()
    n_s = min(args.cpu_sample_genomes, n_total)
    d = scratch_dir()
    try:
        if ref:
            once = ref_sketch_rate(ref, d, n_s, length, family, threads)
            t_sk = min(once(), once())
            t_ad, lines = ref_alldist_from_sketch(ref, d, hashes, off, threads)
            kind = "reference"
        else:
            t_sk, t_ad, lines = port_rates(n_s, length, family, hashes, off, threads)
            kind = "port"
    finally:
        shutil.rmtree(d, ignore_errors=True)
    gbps = n_s * length / t_sk / 1e9
    t_full = (n_total * length / 1e9) / gbps + t_ad
    return {"value": pairs / t_full, "unit": "genome-pairs/s", "cores": threads, "kind": kind,
            "sample": f"sketch -q of {n_s} x {length / 1e6:g} Mbp FASTA files: {t_sk:.2f} s = {gbps:.3f} Gbp/s (incl. .shuf load), "
                      f"extrapolated linearly to {n_total} genomes; alldist -i x.sketch -D {MAX_DIST} on the sketches of all "
                      f"{n_total} genomes: {t_ad:.2f} s (index build + load + distance loop + merge)",
            "sketch_gbps": gbps, "alldist_from_sketch_s": t_ad, "projected_full_job_s": t_full,
            "parity": {"reference_pairs_passing": int(lines), "gpu_pairs_passing_after_host_refine": int(len(kept)),
                       "gpu_pairs_device_superset": int(gpu_hits_device), "equal": bool(lines == len(kept))}}


def run_reference(args):
    """Reference arm: the UNMODIFIED reference binary on the host cores.  A step = `sketch -q` on a bounded sample of
    genomes (time extrapolated linearly to the whole set); the fixed-size `alldist -i x.sketch` part (index build +
    load + distance loop on sketches of all genomes) is measured once before the steps and added to every step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from rabbitkssd_b200 import synth
    n_total, length, family = args.genomes, args.genome_len, args.family
    threads = os.cpu_count() or 1
    pairs = n_total * (n_total - 1) / 2.0
    ref = ref_binary()
    n_s = min(args.cpu_sample_genomes, n_total)
    sk_size = max(1, int(length / 16 ** L_))
    hashes, off = synth.synth_sketches(SEED, n_total // family, family, sk_size, 4 * (K - L_))
    d = scratch_dir()
    try:
        if ref:
            once = ref_sketch_rate(ref, d, n_s, length, family, threads)
            t_ad, lines = ref_alldist_from_sketch(ref, d, hashes, off, threads)
            kind = "reference"
        else:
            import oracle
            p = oracle.init_params(K, S, L_)
            dim = oracle.shuffle_dim(K, S, L_)
            seq, rec_off, rec_first = synth.oracle_records(host_genomes(n_s, length, family, SEED))

            def once():
                t0 = time.perf_counter()
                oracle.sketch_batch(p, dim, seq, rec_off, rec_first)
                return time.perf_counter() - t0
            t0 = time.perf_counter()
            lines = oracle.alldist(hashes, off, 2 * K, MAX_DIST, want_hits=False)
            t_ad = time.perf_counter() - t0
            kind = "port"
        for _ in range(args.warmup):
            once()
        times = [once() for _ in range(args.steps)]
    finally:
        shutil.rmtree(d, ignore_errors=True)
    t_sk = sum(times) / len(times)
    gbps = n_s * length / t_sk / 1e9
    t_full = (n_total * length / 1e9) / gbps + t_ad
    value = pairs / t_full
    sample = (f"step = sketch -q of {n_s} x {length / 1e6:g} Mbp FASTA files ({t_sk:.2f} s, {gbps:.3f} Gbp/s) extrapolated to "
              f"{n_total} genomes + alldist -i x.sketch -D {MAX_DIST} on {n_total} synthetic sketches measured once ({t_ad:.2f} s)")
    out = {"impl": "reference", "metric": "alldist genome-pairs/s (sketch + index + alldist of the whole genome set)",
           "value": value, "unit": "genome-pairs/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": t_full * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "u64 k-mer arithmetic, int counters, f64 distances", "data": "synthetic",
           "config": {"workload": f"alldist of {n_total} synthetic {length / 1e6:g} Mbp genomes, L3K10, -D {MAX_DIST} "
                                  "(BASELINE.json configs[1])", "genomes": n_total, "genome_len": length, "family": family,
                      "shuffle": "L3K10 (k=10,s=6,l=3)", "max_dist": MAX_DIST},
           "cpu_baseline": {"value": value, "unit": "genome-pairs/s", "cores": threads, "kind": kind, "sample": sample,
                            "sketch_gbps": gbps, "alldist_from_sketch_s": t_ad, "pairs_passing": int(lines)},
           "e2e": {"value": value, "unit": "genome-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
