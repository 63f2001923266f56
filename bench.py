#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's configs; the default is configs[1] (SURVEY.md "C2"):
alldist of 10,000 synthetic 5 Mbp bacterial genomes (L3K10, -D 0.05).

One step = the whole hot path over the whole workload: Kssd sketch scan of every genome (parsed bytes resident in
HBM) -> per-genome sort + unique -> [N>1: gather of the sketch shards over peer memory] -> inverted-index build
-> alldist posting accumulation + distance/threshold epilogue for this rank's query-row block -> compacted hits.
metric = genome-pairs/s = N(N-1)/2 / step time (whole job, max over ranks); strong scaling over GPUs.
`e2e` runs the same step from PINNED HOST buffers through the C-ABI (H2D inside the timed region, hits read back).
`roofline` is the sketch scan kernel (the dominant kernel) against MEASURED_PEAKS.json's HBM copy bandwidth.
`cpu_baseline` / `--impl reference` time the unmodified reference binary (oracle/_ref/rabbit_kssd) on the host cores.
`parity` compares the SET of result lines (names, common|size|size, %f Jaccard, %f distance) with the reference's.

Other workloads (bench_configs.py): --config C1 (100 genomes), C3 (dist 1,000 x 50,000, L3K9), C4 (sketch of
20 x 3 Gbp, L4K10 / L5K10), C5 (alldist of 100,000 sketches).
"""
import argparse
import hashlib
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
METRIC = "alldist genome-pairs/s (sketch + index + alldist of the whole genome set)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5"],
                    help="SURVEY.md 8 workload: C1/C2 = BASELINE.json configs[0]/[1] (this file), C3..C5 = configs[2..4] (bench_configs.py)")
    ap.add_argument("--genomes", type=int, default=None)
    ap.add_argument("--genome-len", type=int, default=5000000)
    ap.add_argument("--family", type=int, default=10)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--nccl-gather", action="store_true", help="N > 1: gather the sketch shards with one NCCL all-gather instead of peer-memory reads")
    ap.add_argument("--peer-nccl-barrier", action="store_true", help="N > 1: one-element all-reduce between publish and pull (round 1) instead of the peer-memory flags")
    ap.add_argument("--index-sharded", action="store_true", help="N > 1: build the index sharded by hash range and all-gather the pieces")
    ap.add_argument("--index-replicated", action="store_true", help="N > 1: every rank builds the whole index (default: the index restricted to its own row block)")
    ap.add_argument("--phases", action="store_true", help="diagnostic: per-phase wall times (synchronised) on stderr after the timed runs")
    ap.add_argument("--cpu-sample-genomes", type=int, default=256)
    ap.add_argument("--ref-batch-genomes", type=int, default=1000, help="reference arm: FASTA files on disk at a time")
    ap.add_argument("--extra", default="", help="bench_configs.py: comma-separated key=value overrides of a config's shape")
    args = ap.parse_args()
    if args.genomes is None:
        args.genomes = 100 if args.config == "C1" else 10000
    return args


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


def measured_traffic_ratio(key="dram_bytes_per_input_byte"):
    """DRAM bytes per input byte of the scan kernel from the committed `ncu --set full` capture (profiles/)."""
    p = os.path.join(ROOT, "profiles", "scan_traffic.json")
    try:
        d = json.load(open(p))
        return float(d[key]), d.get("capture_" + key, d["capture"])
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


def bind_to_gpu_numa(local_rank):
    """Pin this process to the CPUs of the GPU's NUMA node BEFORE pinned host memory is allocated, so that the staging
    pages are local to the root complex the GPU hangs off (first touch).  Best effort; returns what was done."""
    info = {"node": None, "cpus": None}
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(local_rank), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=10).stdout.decode().strip()
        bus = out.lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read().strip())
        info["node"] = node
        if node >= 0:
            cpulist = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
            cpus = set()
            for part in cpulist.split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                info["cpus"] = len(cpus)
    except Exception as e:  # no sysfs entry, restricted container...: keep the default placement
        info["error"] = str(e)[:80]
    return info


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


def scratch_dir(need_bytes=0):
    """/dev/shm when it has the room (the reference then reads its inputs from memory, as with a warm page cache)."""
    for base in ("/dev/shm", None):
        if base and not (os.path.isdir(base) and os.access(base, os.W_OK)):
            continue
        d = tempfile.mkdtemp(prefix="kssd_bench_", dir=base)
        if shutil.disk_usage(d).free > need_bytes * 1.1 + (1 << 30) or base is None:
            return d
        shutil.rmtree(d, ignore_errors=True)
    return tempfile.mkdtemp(prefix="kssd_bench_")


def run_ref(ref, argv, cwd):
    t0 = time.perf_counter()
    subprocess.run([ref] + argv, cwd=cwd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return time.perf_counter() - t0


def ref_sketch_genomes(ref, d, first, n, length, family, threads, tag):
    """FASTA files of genomes [first, first+n) (oracle's C generator: the bytes the GPU arm generates in HBM) -> timed
    `rabbit_kssd sketch -q` -> (wall s, {genome index: u32 hash array}).  The FASTA files are removed again."""
    import oracle
    from rabbitkssd_b200 import formats
    oracle.synth_fasta_files(d, SEED, first, n, family, length, 0.04, 2)
    names = [f"g{i}.fna" for i in range(first, first + n)]
    with open(os.path.join(d, "list.txt"), "w") as f:
        f.write("\n".join(names) + "\n")
    if not os.path.exists(os.path.join(d, "L3K10.shuf")):
        run_ref(ref, ["shuffle", "-k", str(K), "-s", str(S), "-l", str(L_), "-o", "L3K10.shuf"], d)
    wall = run_ref(ref, ["sketch", "-q", "-L", "L3K10.shuf", "-i", "list.txt", "-o", tag, "-t", str(threads)], d)
    _, got_names, sets = formats.read_sketch(os.path.join(d, tag + ".sketch"))
    out = {int(nm[1:-4]): s for nm, s in zip(got_names, sets)}
    for nm in names:
        os.remove(os.path.join(d, nm))
    os.remove(os.path.join(d, tag + ".sketch"))
    return wall, out


def ref_alldist_from_sketch(ref, d, hashes, off, threads):
    """`rabbit_kssd alldist -i x.sketch -D 0.05`: index build + load + distance loop + merge -> (wall s, result lines)."""
    from rabbitkssd_b200 import formats
    n = len(off) - 1
    formats.write_sketch(os.path.join(d, "x.sketch"), K, S, L_, [f"g{i}" for i in range(n)], hashes, off)
    for f in ("x.sketch.index", "x.sketch.dict"):
        if os.path.exists(os.path.join(d, f)):
            os.remove(os.path.join(d, f))
    wall = run_ref(ref, ["alldist", "-i", "x.sketch", "-D", str(MAX_DIST), "-o", "x.alldist", "-t", str(threads)], d)
    lines = set()
    for a, b, c, s0, s1, jorc, dd in formats.parse_dist_text(os.path.join(d, "x.alldist")):
        lines.add(canonical_line(int(b[1:]), int(a[1:]), c, s0, s1, jorc, dd))
    return wall, lines


# ---------------------------------------------------------------------------------------------------------------
# parity: the SET of result lines, in a form both arms can produce
def canonical_line(i, j, common, size_i, size_j, jorc_text, dist_text):
    """One alldist result line (src/dist.cpp:233: name[j] name[i] common|size_i|size_j jaccard mashD, j > i) with the
    genome names replaced by their indices."""
    return f"{j}\t{i}\t{common}|{size_i}|{size_j}\t{jorc_text}\t{dist_text}"


def lines_from_hits(hits_kept, jorc, dist_, sizes):
    out = set()
    for (r, c, m), jv, dv in zip(hits_kept.tolist(), jorc.tolist(), dist_.tolist()):
        out.add(canonical_line(r, c, m, int(sizes[r]), int(sizes[c]), "%f" % jv, "%f" % dv))  # std::to_string(double) is "%f"
    return out


def lines_sha256(lines):
    h = hashlib.sha256()
    for ln in sorted(lines):
        h.update(ln.encode())
        h.update(b"\n")
    return h.hexdigest()


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
    """Synthetic bacterial families straight into HBM with rabbitkssd_b200.synth.family_device -- byte-identical to
    synth.family() on the host and to the FASTA files the reference arm sketches: base = uniform ACGT, member m =
    base with i.i.d. substitutions at rate 0.04*m/(family-1); two contigs separated by one '\\n'.
    Layout: genome g at bytes [g*stride, g*stride+length+1)."""
    from rabbitkssd_b200 import synth
    buf = torch.full((n_local * stride,), 10, dtype=torch.uint8, device="cuda")
    half = length // 2
    assert first_genome % family == 0
    for f0 in range(0, n_local, family):
        fam_id = (first_genome + f0) // family
        members = synth.family_device(torch, SEED + 101 * fam_id, family, length, 0.04)
        for m in range(min(family, n_local - f0)):
            o = (f0 + m) * stride
            buf[o:o + half] = members[m][:half]
            buf[o + half + 1:o + length + 1] = members[m][half:]
        del members
    return buf


def run_b200(args):
    import torch
    import torch.distributed as dist
    from rabbitkssd_b200 import _lib, api
    from rabbitkssd_b200.multigpu import PeerGather, build_index_sharded, gather_sketches, gather_sketches_device, triangle_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    all_cpus = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa(local_rank)
    torch.cuda.set_device(local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # NCCL would print its version on stdout, next to the JSON line
            os.environ["NCCL_DEBUG"] = "WARN"
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
    # a row costs its cells plus per-hash work (posting ranges, posting streaming).  Fitted to the dist phase of
    # the 1- and 8-GPU runs (profiles/r01x): ~40 cell-equivalents per hash, i.e. the per-row part dominates
    bounds = triangle_bounds(n_total, world, row_cost=40.0 * sk_size)
    r0, r1 = bounds[rank], bounds[rank + 1]
    state = {}
    if world > 1 and not args.nccl_gather:
        state["peers"] = PeerGather(dist, ctx, n_local, gather_cap, rank, world, nccl_barrier=args.peer_nccl_barrier)

    phase_t = {}

    def mark(name, t0):
        if args.phases and state.get("diag"):
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            phase_t[name] = phase_t.get(name, 0.0) + (t1 - t0) * 1e3
            return t1
        return t0

    def finish_and_dist(rows=None):
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
        # N > 1: every rank indexes only what its own block of triangle rows streams (kssd_b200_index_build_rows: ~1/N of the
        # postings).  Alternatives kept for comparison: the whole index on every rank (--index-replicated, round 1) and the
        # build sharded by hash range + all-gather of the pieces (--index-sharded; same time as replicated at 4 GPUs, r01u)
        rr = rows or (r0, r1)
        if world == 1 or rr == (0, n_total) or args.index_replicated:
            ctx.index_build_device(ph, po, n_total)
        elif args.index_sharded:
            if not build_index_sharded(dist, ctx, ph, po, n_total, state["total_all"], rank, world):
                ctx.index_build_device(ph, po, n_total)
        else:
            try:
                ctx.index_build_rows_device(ph, po, n_total, rr)
            except _lib.KssdError as e:
                if e.code != -5:  # KSSD_E_OVERFLOW: a crowded bucket -> full build
                    raise
                ctx.index_build_device(ph, po, n_total)
        t0 = mark("index", t0)
        d_hits, n_hits = ctx.dist_rows_device(ph, po, n_total, True, 2 * K, MAX_DIST, 0, rows=rows or (r0, r1))
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

    def hits_numpy():
        n = int(state["n_hits"])
        return dev_tensor(torch, state["d_hits"], n * 3, torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 3)

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
    my_hits = hits_numpy()  # this rank's rows, from the last timed step

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

    barrier()  # pinned allocation takes a different time on every rank: the peer gather waits for the slowest at most ~10 s
    step_e2e()
    barrier()
    e0.record(stream)
    for _ in range(args.e2e_steps):
        n_e2e = step_e2e()
    e1.record(stream)
    barrier()
    my_e2e_ms = e0.elapsed_time(e1) / args.e2e_steps
    t = torch.tensor([my_e2e_ms], dtype=torch.float64, device="cuda")
    per_rank = torch.zeros(world, dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_gather_into_tensor(per_rank, t)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    else:
        per_rank[0] = t[0]
    e2e_ms = float(t.item())
    e2e_val = pairs / (e2e_ms / 1e3)
    e2e_hits = hits_host[:n_e2e * 3].numpy().view(np.uint32).reshape(-1, 3).copy()
    ctx.scan_kernel_time()
    ctx.count_kernel_time()
    del host
    os.sched_setaffinity(0, all_cpus)  # the CPU legs below use every host core

    # ---- parity of the result SET: hits of all ranks (device-timed step AND e2e step) against one full-triangle run ----
    def gather_all(h):
        if world == 1:
            return h
        n_max = torch.tensor([len(h)], dtype=torch.int64, device="cuda")
        dist.all_reduce(n_max, op=dist.ReduceOp.MAX)
        pad = torch.full((int(n_max.item()) * 3 + 1,), -1, dtype=torch.int32, device="cuda")
        pad[0] = len(h)
        pad[1:1 + h.size] = torch.from_numpy(h.view(np.int32).reshape(-1)).cuda()
        every = torch.empty(world * pad.numel(), dtype=torch.int32, device="cuda")
        dist.all_gather_into_tensor(every, pad)
        every = every.cpu().numpy().reshape(world, -1)
        return np.concatenate([every[r, 1:1 + 3 * every[r, 0]].view(np.uint32).reshape(-1, 3) for r in range(world)])

    all_hits = gather_all(my_hits)
    all_e2e = gather_all(e2e_hits)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    off = dev_tensor(torch, state["po"], n_total + 1, torch.int64).cpu().numpy().astype(np.uint64)
    hashes = dev_tensor(torch, state["ph"], int(off[-1]), torch.int32).cpu().numpy().view(np.uint32).copy()
    sizes = np.diff(off).astype(np.int64)
    gpu_lines = lines_from_hits(*api.refine_hits(all_hits, sizes, sizes, 2 * K, MAX_DIST, True), sizes)
    e2e_lines = lines_from_hits(*api.refine_hits(all_e2e, sizes, sizes, 2 * K, MAX_DIST, True), sizes)
    parity = {"pairs_passing": len(gpu_lines), "sha256": lines_sha256(gpu_lines), "e2e_set_equal": gpu_lines == e2e_lines,
              "device_hits_superset": int(len(all_hits))}
    if world > 1:  # every rank holds all sketches: rank 0 can run the whole triangle alone and compare the sets
        ctx.index_build_device(state["ph"], state["po"], n_total)
        hp, n1 = ctx.dist_rows_device(state["ph"], state["po"], n_total, True, 2 * K, MAX_DIST, 0, rows=(0, n_total))
        one = dev_tensor(torch, hp, int(n1) * 3, torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 3)
        one_lines = lines_from_hits(*api.refine_hits(one, sizes, sizes, 2 * K, MAX_DIST, True), sizes)
        parity.update(set_equal=gpu_lines == one_lines, checked_against="the full triangle run on rank 0 alone (full index built from the gathered sketches); "
                      "its sha256 equals the N=1 line's, which is checked against the reference binary")

    peak, peak_src = measured_hbm_peak()
    cells_rank0 = sum(n_total - 1 - i for i in range(r0, r1))
    hashes_total = int(off[-1])
    # algorithmic bytes of the scan kernel: 1 B per input byte + 8 B per emitted (genome,hash) key;
    # emitted keys ~ unique hashes (duplicates inside a 5 Mbp genome are negligible)
    alg_bytes = scan_bytes / max(scan_n, 1) + 8.0 * (hashes_total / world) / max(scan_n / max(args.steps, 1), 1)
    scan_ms_per = scan_ms / max(scan_n, 1)
    t_ratio, t_src = measured_traffic_ratio()
    achieved = alg_bytes / (scan_ms_per / 1e3) / 1e9 if scan_ms_per > 0 else 0.0
    per_rank_ms = [float(x) for x in per_rank.cpu().tolist()]
    out = {
        "metric": METRIC, "value": value,
        "unit": "genome-pairs/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8 bases -> u32 hashes (u64 k-mer arithmetic), u16 smem counters, f64 distances",
        "data": "synthetic (splitmix64 families generated in HBM: uniform ACGT base genomes + substitution variants; the reference arm sketches the same bytes as FASTA)",
        "config": workload_config(n_total, length, family),
        "partition": f"genomes and triangle query rows sharded over {world} rank(s), {n_bytes / 1e9:.1f} GB of sequence per rank",
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
                           "pairs_per_s_kernel_only": cells_rank0 / ((cnt_ms / max(cnt_n, 1)) / 1e3) if cnt_ms else None,
                           "note": "timed region = row_work_kernel + dist_tile_kernel of rank 0's row block"},
        "e2e": {"value": e2e_val, "unit": "genome-pairs/s", "ms_per_step": e2e_ms, "steps": args.e2e_steps,
                "h2d_bytes_per_step": int(n_bytes) * world, "d2h_bytes_per_step": int(len(all_e2e)) * 12,
                "per_rank_ms": per_rank_ms, "per_rank_h2d_GBps": [n_bytes / (m / 1e3) / 1e9 for m in per_rank_ms],
                "numa_rank0": numa},
        "parity": parity,
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, hashes, off, gpu_lines, n_total, length, family, pairs)
        par = out["cpu_baseline"].pop("parity")
        parity.update(par)
    print(json.dumps(out))
    sys.stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def workload_config(n_total, length, family):
    """The same dict in both arms (the driver compares them)."""
    which = "configs[0] shape" if n_total == 100 else "configs[1]" if n_total == 10000 else "custom size"
    return {"workload": f"alldist of {n_total} synthetic {length / 1e6:g} Mbp genomes, L3K10, -D {MAX_DIST} (BASELINE.json {which})",
            "genomes": n_total, "genome_len": length, "family": family, "shuffle": "L3K10 (k=10,s=6,l=3)", "max_dist": MAX_DIST,
            "seed": SEED, "l2": f"no flush needed: every step streams {n_total * length / 1e9:.1f} GB of sequence (>= 1/8 of it per "
                                "rank at up to 8 ranks), far more than the 126 MB L2"}


def cpu_baseline(args, hashes, off, gpu_lines, n_total, length, family, pairs):
    """Reference CPU path on this box's host cores, bounded: (a) `sketch -q` on a sample of the SAME genomes -> Gbp/s,
    extrapolated linearly to all genomes, and the sample's hash sets compared with the GPU's; (b) `alldist -i x.sketch`
    on the sketches of ALL genomes (the ones the GPU just produced): its result lines are compared with the GPU's as a
    SET (names, common|size|size and the %f texts)."""
    threads = os.cpu_count() or 1
    ref = ref_binary()
    n_s = min(args.cpu_sample_genomes, n_total)
    d = scratch_dir(n_s * length)
    sample_equal = None
    try:
        if ref:
            t_sk, got = ref_sketch_genomes(ref, d, 0, n_s, length, family, threads, "s0")
            t_sk2, _ = ref_sketch_genomes(ref, d, 0, n_s, length, family, threads, "s1")
            t_sk = min(t_sk, t_sk2)
            sample_equal = all(np.array_equal(np.sort(got[g]), hashes[int(off[g]):int(off[g + 1])]) for g in range(n_s))
            t_ad, ref_lines = ref_alldist_from_sketch(ref, d, hashes, off, threads)
            kind = "reference"
        else:
            import oracle
            from rabbitkssd_b200 import synth
            p = oracle.init_params(K, S, L_)
            dim = oracle.shuffle_dim(K, S, L_)
            gen = [[oracle.synth_genome(SEED + 101 * (g // family), g % family, family, length)] for g in range(n_s)]
            seq, rec_off, rec_first = synth.oracle_records([[g[0][:length // 2], g[0][length // 2:]] for g in gen])
            t0 = time.perf_counter()
            got = oracle.sketch_batch(p, dim, seq, rec_off, rec_first)
            t_sk = time.perf_counter() - t0
            sample_equal = all(np.array_equal(got[g].astype(np.uint32), hashes[int(off[g]):int(off[g + 1])]) for g in range(n_s))
            t0 = time.perf_counter()
            hr, hc, hm = oracle.alldist(hashes, off, 2 * K, MAX_DIST)
            t_ad = time.perf_counter() - t0
            sizes = np.diff(off).astype(np.int64)
            ref_lines = set()
            for r, c, m in zip(hr.tolist(), hc.tolist(), hm.tolist()):
                jv, dv = oracle.distance(m, int(sizes[r]), int(sizes[c]), 2 * K)
                ref_lines.add(canonical_line(r, c, m, int(sizes[r]), int(sizes[c]), "%f" % jv, "%f" % dv))
            kind = "port"
            threads = oracle.max_threads()
    finally:
        shutil.rmtree(d, ignore_errors=True)
    gbps = n_s * length / t_sk / 1e9
    t_full = (n_total * length / 1e9) / gbps + t_ad
    return {"value": pairs / t_full, "unit": "genome-pairs/s", "cores": threads, "kind": kind,
            "sample": f"sketch -q of the first {n_s} of the {n_total} genomes as FASTA files: {t_sk:.2f} s = {gbps:.3f} Gbp/s (incl. .shuf load), "
                      f"extrapolated linearly to {n_total} genomes; alldist -i x.sketch -D {MAX_DIST} on the sketches of all "
                      f"{n_total} genomes: {t_ad:.2f} s (index build + load + distance loop + merge)",
            "sketch_gbps": gbps, "alldist_from_sketch_s": t_ad, "projected_full_job_s": t_full,
            "parity": {"set_equal": bool(ref_lines == gpu_lines), "reference_pairs_passing": len(ref_lines),
                       "only_in_reference": len(ref_lines - gpu_lines), "only_in_gpu": len(gpu_lines - ref_lines),
                       "reference_sha256": lines_sha256(ref_lines), "sketch_sample_sets_equal": sample_equal,
                       "checked_against": f"reference binary: x.alldist line set of all {n_total} genomes + .sketch hash sets of {n_s} genomes"}}


def run_reference(args):
    """Reference arm: the UNMODIFIED reference binary on the host cores, on the workload of the GPU arm -- the same
    genomes, written as FASTA files by the oracle's C generator.  ONE measured step = `sketch -q` over all genomes (in
    batches of --ref-batch-genomes files so that the FASTA text fits the scratch space; wall times summed, sketches
    concatenated) + `alldist -i x.sketch -D 0.05` on the resulting sketches.  A full pass takes about a minute, so one
    pass is measured whatever --steps says, and the line says so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_total, length, family = args.genomes, args.genome_len, args.family
    threads = os.cpu_count() or 1
    pairs = n_total * (n_total - 1) / 2.0
    ref = ref_binary()
    batch = max(family, min(args.ref_batch_genomes, n_total) // family * family)
    d = scratch_dir(batch * length * 1.02)
    free = shutil.disk_usage(d).free
    while batch > family and batch * length * 1.05 > free * 0.8:
        batch = max(family, batch // 2 // family * family)
    t_wall0 = time.perf_counter()
    t_gen = 0.0
    try:
        if ref:
            sets, t_sk = {}, 0.0
            for b0 in range(0, n_total, batch):
                nb = min(batch, n_total - b0)
                tg = time.perf_counter()
                w, got = ref_sketch_genomes(ref, d, b0, nb, length, family, threads, f"s{b0}")
                t_gen += time.perf_counter() - tg - w
                t_sk += w
                sets.update(got)
            off = np.zeros(n_total + 1, dtype=np.uint64)
            off[1:] = np.cumsum([len(sets[g]) for g in range(n_total)])
            hashes = np.concatenate([np.sort(sets[g]) for g in range(n_total)]).astype(np.uint32)
            t_ad, lines = ref_alldist_from_sketch(ref, d, hashes, off, threads)
            kind = "reference"
            how = (f"sketch -q -t {threads} over all {n_total} genomes as FASTA files ({(n_total + batch - 1) // batch} runs of <= {batch} files, walls summed: "
                   f"{t_sk:.2f} s) + alldist -i x.sketch -D {MAX_DIST} -t {threads} on their sketches ({t_ad:.2f} s: index build + load + distance loop + merge)")
        else:
            import oracle
            from rabbitkssd_b200 import synth
            p = oracle.init_params(K, S, L_)
            dim = oracle.shuffle_dim(K, S, L_)
            sets, t_sk = [], 0.0
            for b0 in range(0, n_total, batch):
                gen = [oracle.synth_genome(SEED + 101 * (g // family), g % family, family, length) for g in range(b0, min(n_total, b0 + batch))]
                seq, rec_off, rec_first = synth.oracle_records([[g[:length // 2], g[length // 2:]] for g in gen])
                t0 = time.perf_counter()
                sets += oracle.sketch_batch(p, dim, seq, rec_off, rec_first)
                t_sk += time.perf_counter() - t0
            off = np.zeros(n_total + 1, dtype=np.uint64)
            off[1:] = np.cumsum([len(x) for x in sets])
            hashes = np.concatenate(sets).astype(np.uint32)
            sizes = np.diff(off).astype(np.int64)
            t0 = time.perf_counter()
            hr, hc, hm = oracle.alldist(hashes, off, 2 * K, MAX_DIST)
            t_ad = time.perf_counter() - t0
            lines = set()
            for r, c, m in zip(hr.tolist(), hc.tolist(), hm.tolist()):
                jv, dv = oracle.distance(m, int(sizes[r]), int(sizes[c]), 2 * K)
                lines.add(canonical_line(r, c, m, int(sizes[r]), int(sizes[c]), "%f" % jv, "%f" % dv))
            kind = "port"
            threads = oracle.max_threads()
            how = f"oracle port (OpenMP, {threads} threads): sketch of all {n_total} genomes {t_sk:.2f} s + alldist {t_ad:.2f} s"
    finally:
        shutil.rmtree(d, ignore_errors=True)
    t_full = t_sk + t_ad
    value = pairs / t_full
    out = {"impl": "reference", "metric": METRIC,
           "value": value, "unit": "genome-pairs/s", "n_gpus": args.gpus, "steps": 1, "warmup": 0,
           "steps_requested": args.steps, "warmup_requested": args.warmup,
           "ms_per_step": t_full * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "u64 k-mer arithmetic, int counters, f64 distances", "data": "synthetic (the GPU arm's genomes, as FASTA files)",
           "config": workload_config(n_total, length, family),
           "cpu_baseline": {"value": value, "unit": "genome-pairs/s", "cores": threads, "kind": kind,
                            "sample": "the whole workload, measured once: " + how,
                            "sketch_s": t_sk, "sketch_gbps": n_total * length / t_sk / 1e9, "alldist_from_sketch_s": t_ad,
                            "untimed_input_generation_s": t_gen, "wall_s": time.perf_counter() - t_wall0},
           "parity": {"pairs_passing": len(lines), "sha256": lines_sha256(lines)},
           "e2e": {"value": value, "unit": "genome-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def main():
    args = parse_args()
    if args.config in ("C3", "C4", "C5"):
        import bench_configs
        return bench_configs.main(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
