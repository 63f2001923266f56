"""ctypes wrapper around oracle/libkssd_oracle.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  Nothing under rabbitkssd_b200/ may (the product path has no CPU fallback).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Params(C.Structure):
    _fields_ = [("half_k", C.c_int), ("half_subk", C.c_int), ("drlevel", C.c_int),
                ("rev_add_move", C.c_int), ("half_outctx_len", C.c_int),
                ("dim_start", C.c_int), ("dim_end", C.c_int), ("kmer_size", C.c_uint),
                ("domask", C.c_uint64), ("tupmask", C.c_uint64),
                ("undomask0", C.c_uint64), ("undomask1", C.c_uint64)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "libkssd_oracle.so"])


def ref_binary():
    """Path of the unmodified reference binary (oracle/_ref/rabbit_kssd) or None."""
    p = os.path.join(_HERE, "_ref", "rabbit_kssd")
    return p if os.access(p, os.X_OK) else None


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libkssd_oracle.so")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(_HERE, "kssd_oracle.c")):
            build()
        L = C.CDLL(so)
        u8p, i8p = C.POINTER(C.c_uint8), C.POINTER(C.c_int8)
        u32p, u64p, i32p, f64p = (C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int32),
                                  C.POINTER(C.c_double))
        L.kssd_oracle_init_params.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Params)]
        L.kssd_oracle_init_params.restype = C.c_int
        L.kssd_oracle_shuffle_dim.argtypes = [C.c_int, C.c_int, C.c_int, i32p]
        L.kssd_oracle_shuffle_dim.restype = None
        L.kssd_oracle_scan_record.argtypes = [C.POINTER(Params), i32p, u8p, C.c_uint64, i8p, C.c_int, u64p, C.c_uint64]
        L.kssd_oracle_scan_record.restype = C.c_uint64
        L.kssd_oracle_finish_genome.argtypes = [u64p, C.c_uint64, C.c_int, C.c_int]
        L.kssd_oracle_finish_genome.restype = C.c_uint64
        L.kssd_oracle_index_build.argtypes = [u32p, u64p, C.c_uint32, u32p, u32p, u32p]
        L.kssd_oracle_index_build.restype = C.c_uint64
        L.kssd_oracle_count_row.argtypes = [u32p, C.c_uint64, u32p, u64p, C.c_uint64, u32p, C.c_uint32, i32p]
        L.kssd_oracle_count_row.restype = None
        L.kssd_oracle_distance.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, f64p]
        L.kssd_oracle_distance.restype = C.c_double
        L.kssd_oracle_max_threads.restype = C.c_int
        L.kssd_oracle_sketch_batch.argtypes = [C.POINTER(Params), i32p, u8p, i8p, u64p, u64p, C.c_uint32,
                                               C.c_int, C.c_int, C.c_int, u64p, C.c_uint64, u64p]
        L.kssd_oracle_sketch_batch.restype = None
        L.kssd_oracle_alldist.argtypes = [u32p, u64p, C.c_uint32, u32p, u64p, C.c_uint64, u32p,
                                          C.c_uint32, C.c_uint32, C.c_int, C.c_double, C.c_int,
                                          u32p, u32p, u32p, C.c_uint64]
        L.kssd_oracle_alldist.restype = C.c_uint64
        L.kssd_oracle_dist.argtypes = [u32p, u64p, C.c_uint32, C.c_uint32, u32p, C.c_uint32,
                                       u32p, u64p, C.c_uint64, u32p, C.c_int, C.c_double, C.c_int,
                                       u32p, u32p, u32p, C.c_uint64]
        L.kssd_oracle_dist.restype = C.c_uint64
        L.kssd_oracle_synth_genome.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_uint64, C.c_double, u8p]
        L.kssd_oracle_synth_genome.restype = None
        L.kssd_oracle_synth_fasta_files.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_uint64,
                                                    C.c_double, C.c_int]
        L.kssd_oracle_synth_fasta_files.restype = C.c_int
        _LIB = L
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def init_params(half_k, half_subk, drlevel, relax_guard=False):
    p = Params()
    rc = lib().kssd_oracle_init_params(half_k, half_subk, drlevel, int(relax_guard), C.byref(p))
    if rc != 0:
        raise ValueError("half_subk - drlevel must be >= 3 (reference aborts, common.cpp:37-39)")
    return p


def shuffle_dim(half_k, half_subk, drlevel):
    out = np.empty(1 << (4 * half_subk), dtype=np.int32)
    lib().kssd_oracle_shuffle_dim(half_k, half_subk, drlevel, _p(out, C.c_int32))
    return out


def max_threads():
    return lib().kssd_oracle_max_threads()


def sketch_batch(params, shuffled_dim, seq, rec_off, rec_first, qual=None, least_qual=0, least_num_kmer=1,
                 out_cap=None):
    """seq: uint8 array of concatenated records; rec_off[n_rec+1]; rec_first[n_genomes+1] (record index).
    Returns list of sorted unique uint64 arrays, one per genome."""
    seq = np.ascontiguousarray(seq, dtype=np.uint8)
    rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    rec_first = np.ascontiguousarray(rec_first, dtype=np.uint64)
    n_genomes = len(rec_first) - 1
    use64 = 1 if params.half_k - params.drlevel > 8 else 0
    if out_cap is None:
        glen = [int(rec_off[int(rec_first[g + 1])] - rec_off[int(rec_first[g])]) for g in range(n_genomes)]
        out_cap = max(glen) if glen else 1
        out_cap = max(out_cap, 1)
    out = np.empty(n_genomes * out_cap, dtype=np.uint64)
    out_n = np.zeros(n_genomes, dtype=np.uint64)
    q = None if qual is None else np.ascontiguousarray(qual).view(np.int8)
    lib().kssd_oracle_sketch_batch(C.byref(params), _p(shuffled_dim, C.c_int32), _p(seq, C.c_uint8), _p(q, C.c_int8),
                                   _p(rec_off, C.c_uint64), _p(rec_first, C.c_uint64), n_genomes,
                                   least_qual, least_num_kmer, use64, _p(out, C.c_uint64), out_cap,
                                   _p(out_n, C.c_uint64))
    res = []
    for g in range(n_genomes):
        n = int(out_n[g])
        assert n != 0xFFFFFFFFFFFFFFFF, "oracle out_cap overflow"
        res.append(out[g * out_cap: g * out_cap + n].copy())
    return res


def index_build(hashes, genome_off):
    hashes = np.ascontiguousarray(hashes, dtype=np.uint32)
    genome_off = np.ascontiguousarray(genome_off, dtype=np.uint64)
    n = len(genome_off) - 1
    total = int(genome_off[-1])
    dict_ = np.empty(max(total, 1), dtype=np.uint32)
    uh = np.empty(max(total, 1), dtype=np.uint32)
    uc = np.empty(max(total, 1), dtype=np.uint32)
    u = lib().kssd_oracle_index_build(_p(hashes, C.c_uint32), _p(genome_off, C.c_uint64), n,
                                      _p(dict_, C.c_uint32), _p(uh, C.c_uint32), _p(uc, C.c_uint32))
    uh, uc = uh[:u].copy(), uc[:u].copy()
    off = np.zeros(u + 1, dtype=np.uint64)
    np.cumsum(uc, out=off[1:])
    return dict_[:total].copy(), uh, uc, off


def distance(common, size0, size1, kmer_size, is_containment=0):
    j = C.c_double()
    d = lib().kssd_oracle_distance(int(common), int(size0), int(size1), int(kmer_size), int(is_containment), C.byref(j))
    return j.value, d


def alldist(hashes, genome_off, kmer_size, max_dist, is_containment=0, rows=None, want_hits=True):
    hashes = np.ascontiguousarray(hashes, dtype=np.uint32)
    genome_off = np.ascontiguousarray(genome_off, dtype=np.uint64)
    n = len(genome_off) - 1
    dict_, uh, uc, off = index_build(hashes, genome_off)
    r0, r1 = (0, n) if rows is None else rows
    cap = 1
    hr = hc = hm = None
    if want_hits:
        cap = lib().kssd_oracle_alldist(_p(hashes, C.c_uint32), _p(genome_off, C.c_uint64), n, _p(uh, C.c_uint32),
                                        _p(off, C.c_uint64), len(uh), _p(dict_, C.c_uint32), r0, r1, kmer_size,
                                        max_dist, is_containment, None, None, None, 0)
        hr = np.empty(max(cap, 1), np.uint32); hc = np.empty(max(cap, 1), np.uint32); hm = np.empty(max(cap, 1), np.uint32)
    tot = lib().kssd_oracle_alldist(_p(hashes, C.c_uint32), _p(genome_off, C.c_uint64), n, _p(uh, C.c_uint32),
                                    _p(off, C.c_uint64), len(uh), _p(dict_, C.c_uint32), r0, r1, kmer_size,
                                    max_dist, is_containment, _p(hr, C.c_uint32), _p(hc, C.c_uint32),
                                    _p(hm, C.c_uint32), cap)
    if not want_hits:
        return tot
    o = np.lexsort((hc[:tot], hr[:tot]))
    return hr[:tot][o], hc[:tot][o], hm[:tot][o]


def dist(q_hashes, q_off, r_hashes, r_off, kmer_size, max_dist, is_containment=0, rows=None, want_hits=True):
    q_hashes = np.ascontiguousarray(q_hashes, dtype=np.uint32)
    q_off = np.ascontiguousarray(q_off, dtype=np.uint64)
    r_hashes = np.ascontiguousarray(r_hashes, dtype=np.uint32)
    r_off = np.ascontiguousarray(r_off, dtype=np.uint64)
    nq, nr = len(q_off) - 1, len(r_off) - 1
    dict_, uh, uc, off = index_build(r_hashes, r_off)
    sizes = np.diff(r_off).astype(np.uint32)
    r0, r1 = (0, nq) if rows is None else rows
    args = [_p(q_hashes, C.c_uint32), _p(q_off, C.c_uint64), r0, r1, _p(sizes, C.c_uint32), nr, _p(uh, C.c_uint32),
            _p(off, C.c_uint64), len(uh), _p(dict_, C.c_uint32), kmer_size, max_dist, is_containment]
    if not want_hits:
        return lib().kssd_oracle_dist(*args, None, None, None, 0)
    cap = lib().kssd_oracle_dist(*args, None, None, None, 0)
    hr = np.empty(max(cap, 1), np.uint32); hc = np.empty(max(cap, 1), np.uint32); hm = np.empty(max(cap, 1), np.uint32)
    tot = lib().kssd_oracle_dist(*args, _p(hr, C.c_uint32), _p(hc, C.c_uint32), _p(hm, C.c_uint32), cap)
    o = np.lexsort((hc[:tot], hr[:tot]))
    return hr[:tot][o], hc[:tot][o], hm[:tot][o]


def synth_genome(seed, member, n_members, length, max_div=0.04):
    """ASCII bases of one member of a synthetic family: the bytes rabbitkssd_b200.synth.family() generates."""
    out = np.empty(length, dtype=np.uint8)
    lib().kssd_oracle_synth_genome(seed & 0xFFFFFFFFFFFFFFFF, member, n_members, length, max_div, _p(out, C.c_uint8))
    return out


def synth_fasta_files(directory, seed0, first, n, n_members, length, max_div=0.04, n_contigs=2):
    """Writes genomes [first, first+n) of the bench's family series as <directory>/g<i>.fna (OpenMP over files)."""
    got = lib().kssd_oracle_synth_fasta_files(directory.encode(), seed0 & 0xFFFFFFFFFFFFFFFF, first, n, n_members, length,
                                              max_div, n_contigs)
    if got != n:
        raise OSError(f"synth_fasta_files: wrote {got} of {n} files into {directory}")
