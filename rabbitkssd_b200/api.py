"""Host-side Python mirror of the hot-path interface, over the C-ABI (ctypes).  Used by tests and bench.py; the
C++ drop-in shims live in csrc/host/.  numpy arrays in, numpy arrays out; *_device variants take raw device
pointers (e.g. torch tensors' data_ptr()) and leave results in device memory."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


def _np_from_ptr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(C.addressof(ptr.contents))
    return np.frombuffer(buf, dtype=dtype, count=n).copy()


def shuffle_table(half_k, half_subk, drlevel):
    """The .shuf dimension-reduction table (host code of the library, src/shuffle.cpp:76-104): int32[16^half_subk]."""
    out = np.empty(1 << (4 * half_subk), dtype=np.int32)
    check(_lib.load().kssd_b200_shuffle_table(half_k, half_subk, drlevel, out.ctypes.data_as(_lib.i32p)))
    return out


class Context:
    def __init__(self, device=0):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        check(self.lib.kssd_b200_init(C.byref(self.h), device))
        self.params = None
        self.hash64 = False     # the parameters select 64-bit hashes (half_k - drlevel > 8, src/sketch.cpp:336)
        self.index64 = False    # the loaded index holds 64-bit hashes
        self.n_genomes = 0
        self._keep = []  # host arrays the library may read again until sketch_finish (overflow replay)

    def close(self):
        if self.h:
            self.lib.kssd_b200_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ----
    def set_stream(self, cuda_stream_ptr):
        check(self.lib.kssd_b200_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        check(self.lib.kssd_b200_sync(self.h))

    def set_profiling(self, on=True):
        check(self.lib.kssd_b200_set_profiling(self.h, int(on)))

    def set_relax_guard(self, on=True):
        """Accept half_subk - drlevel < 3 (L5K10), which the reference refuses (src/common.cpp:37-39)."""
        check(self.lib.kssd_b200_set_relax_guard(self.h, int(on)))

    def set_force_radix(self, on=True):
        """Tests: take the device-wide radix paths even where the shared-memory segment sorts apply."""
        check(self.lib.kssd_b200_set_force_radix(self.h, int(on)))

    def launch_count(self):
        return int(self.lib.kssd_b200_launch_count(self.h))

    def scan_kernel_time(self):
        ms, n, b = C.c_double(), C.c_uint64(), C.c_uint64()
        check(self.lib.kssd_b200_scan_kernel_time(self.h, C.byref(ms), C.byref(n), C.byref(b)))
        return ms.value, n.value, b.value

    def count_kernel_time(self):
        ms, n, p, a = C.c_double(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(self.lib.kssd_b200_count_kernel_time(self.h, C.byref(ms), C.byref(n), C.byref(p), C.byref(a)))
        return ms.value, n.value, p.value, a.value

    # ---- parameters ----
    def set_params(self, half_k, half_subk, drlevel, shuffled_dim):
        dim = np.ascontiguousarray(shuffled_dim, dtype=np.int32)
        if half_subk < 8 and len(dim) != 1 << (4 * half_subk):
            raise ValueError("shuffled_dim must have 16^half_subk entries")
        check(self.lib.kssd_b200_set_params(self.h, half_k, half_subk, drlevel, dim.ctypes.data_as(_lib.i32p)))
        self.params = (half_k, half_subk, drlevel)
        self.hash64 = half_k - drlevel > 8

    # ---- sketching ----
    def sketch_begin(self, n_genomes, has_qual=False, least_qual=0, least_num_kmer=1):
        check(self.lib.kssd_b200_sketch_begin(self.h, n_genomes, int(has_qual), least_qual, least_num_kmer))
        self.n_genomes = n_genomes
        self._keep = []

    def submit_host(self, seq, seg_off, seg_genome, qual=None, first_emit=0):
        """One pinned batch: copies the arrays into a library-owned pinned batch and submits it (async)."""
        seq = np.ascontiguousarray(seq, dtype=np.uint8)
        seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        seg_genome = np.ascontiguousarray(seg_genome, dtype=np.uint32)
        b = C.POINTER(_lib.Batch)()
        check(self.lib.kssd_b200_batch_acquire(self.h, C.byref(b), len(seq), len(seg_genome)))
        bt = b.contents
        C.memmove(bt.seq, seq.ctypes.data, len(seq))
        if qual is not None:
            q = np.ascontiguousarray(qual, dtype=np.uint8)
            C.memmove(bt.qual, q.ctypes.data, len(q))
        C.memmove(bt.seg_off, seg_off.ctypes.data, seg_off.nbytes)
        C.memmove(bt.seg_genome, seg_genome.ctypes.data, seg_genome.nbytes)
        bt.n_bytes = len(seq)
        bt.n_seg = len(seg_genome)
        bt.first_emit = first_emit
        check(self.lib.kssd_b200_batch_submit(self.h, b))

    def submit_raw_fasta(self, pieces):
        """One batch of RAW FASTA text, tokenised on the device.  pieces: list of (genome, text bytes / uint8 array, flags (0 line
        start, 1 inside a header line, 2 inside a sequence line), first) -- see kssd_batch.raw_fasta in include/kssd_b200.h.
        Every piece is padded with newlines to a multiple of 32."""
        sizes = [(len(t) + 31) // 32 * 32 for _, t, _, _ in pieces]
        total = sum(sizes)
        b = C.POINTER(_lib.Batch)()
        check(self.lib.kssd_b200_batch_acquire(self.h, C.byref(b), total, len(pieces)))
        bt = b.contents
        buf = np.full(total, 10, dtype=np.uint8)
        off = np.zeros(len(pieces) + 1, dtype=np.uint64)
        at = 0
        for i, ((g, t, fl, first), sz) in enumerate(zip(pieces, sizes)):
            t = np.frombuffer(bytes(t), dtype=np.uint8) if not isinstance(t, np.ndarray) else t
            buf[at:at + len(t)] = t
            off[i] = at
            bt.seg_genome[i] = g
            bt.seg_flags[i] = int(fl)
            bt.seg_first[i] = first
            at += sz
        off[len(pieces)] = at
        C.memmove(bt.seq, buf.ctypes.data, total)
        C.memmove(bt.seg_off, off.ctypes.data, off.nbytes)
        bt.n_bytes = total
        bt.n_seg = len(pieces)
        bt.first_emit = 0
        bt.raw_fasta = 1
        check(self.lib.kssd_b200_batch_submit(self.h, b))

    def submit_raw_fastq(self, pieces):
        """One batch of RAW four-line FASTQ text, tokenised on the device (kssd_batch.raw_fasta = KSSD_RAW_FASTQ).  pieces: list of
        (genome, text bytes holding whole records and ending with a newline); every piece is padded with 0x00 to a multiple of 32."""
        sizes = [(len(t) + 31) // 32 * 32 for _, t in pieces]
        total = sum(sizes)
        b = C.POINTER(_lib.Batch)()
        check(self.lib.kssd_b200_batch_acquire(self.h, C.byref(b), total, len(pieces)))
        bt = b.contents
        buf = np.zeros(total, dtype=np.uint8)
        off = np.zeros(len(pieces) + 1, dtype=np.uint64)
        at = 0
        for i, ((g, t), sz) in enumerate(zip(pieces, sizes)):
            t = np.frombuffer(bytes(t), dtype=np.uint8) if not isinstance(t, np.ndarray) else t
            buf[at:at + len(t)] = t
            off[i] = at
            bt.seg_genome[i] = g
            at += sz
        off[len(pieces)] = at
        C.memmove(bt.seq, buf.ctypes.data, total)
        C.memmove(bt.seg_off, off.ctypes.data, off.nbytes)
        bt.n_bytes = total
        bt.n_seg = len(pieces)
        bt.first_emit = 0
        bt.raw_fasta = 2
        check(self.lib.kssd_b200_batch_submit(self.h, b))

    def acquire_batch(self, min_capacity, min_segments):
        b = C.POINTER(_lib.Batch)()
        check(self.lib.kssd_b200_batch_acquire(self.h, C.byref(b), min_capacity, min_segments))
        return b

    def submit_batch(self, b):
        check(self.lib.kssd_b200_batch_submit(self.h, b))

    def scan_device(self, d_seq, n_bytes, d_seg_off, d_seg_genome, n_seg, d_qual=0, first_emit=0):
        check(self.lib.kssd_b200_sketch_scan_device(self.h, C.c_void_p(d_seq), C.c_void_p(d_qual), n_bytes,
                                                    C.c_void_p(d_seg_off), C.c_void_p(d_seg_genome), n_seg, first_emit))

    def scan_host(self, seq_ptr, n_bytes, seg_off, seg_genome, qual_ptr=0, chunk_bytes=0):
        """Stream a caller-owned (ideally pinned) host buffer through the double-buffered H2D pipeline."""
        seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        seg_genome = np.ascontiguousarray(seg_genome, dtype=np.uint32)
        self._keep.append((seg_off, seg_genome))  # must stay valid until sketch_finish (include/kssd_b200.h)
        check(self.lib.kssd_b200_sketch_scan_host(self.h, C.c_void_p(seq_ptr), C.c_void_p(qual_ptr), n_bytes,
                                                  C.c_void_p(seg_off.ctypes.data), C.c_void_p(seg_genome.ctypes.data),
                                                  len(seg_genome), chunk_bytes))

    def sketch_finish(self):
        """-> (hashes uint32[total], genome_off uint64[n_genomes+1]) on the host; per genome sorted unique."""
        if self.hash64:  # uint64 hashes
            hp, op = _lib.u64p(), _lib.u64p()
            check(self.lib.kssd_b200_sketch_finish64(self.h, C.byref(hp), C.byref(op)))
            off = _np_from_ptr(op, self.n_genomes + 1, np.uint64)
            return _np_from_ptr(hp, int(off[-1]), np.uint64), off
        hp, op = _lib.u32p(), _lib.u64p()
        check(self.lib.kssd_b200_sketch_finish(self.h, C.byref(hp), C.byref(op)))
        off = _np_from_ptr(op, self.n_genomes + 1, np.uint64)
        return _np_from_ptr(hp, int(off[-1]), np.uint32), off

    def sketch_finish_device(self):
        """-> (d_hashes ptr, d_genome_off ptr, total); library-owned device memory."""
        hp, op, tot = C.c_void_p(), C.c_void_p(), C.c_uint64()
        fn = self.lib.kssd_b200_sketch_finish_device64 if self.hash64 else self.lib.kssd_b200_sketch_finish_device
        check(fn(self.h, C.byref(hp), C.byref(op), C.byref(tot)))
        return hp.value or 0, op.value or 0, tot.value

    def sketch_concat_device(self, d_parts, stride_words, n_local, world):
        """Gather epilogue: `world` blocks [n_local sizes | hashes | padding] -> (d_hashes, d_genome_off, total, overflow)."""
        ph, po, tot, ov = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_int()
        check(self.lib.kssd_b200_sketch_concat_device(self.h, d_parts, stride_words, n_local, world, C.byref(ph), C.byref(po),
                                                      C.byref(tot), C.byref(ov)))
        return ph.value or 0, po.value or 0, tot.value, bool(ov.value)

    def peer_init(self, n_local, cap_hashes):
        """Allocates this rank's two peer buffers -> their IPC handles (128 bytes)."""
        buf = (C.c_uint8 * 128)()
        check(self.lib.kssd_b200_peer_init(self.h, n_local, cap_hashes, buf))
        return bytes(buf)

    def peer_connect(self, handles, world, rank):
        """handles: the 128-byte handle pairs of all ranks, rank order."""
        assert len(handles) == 128 * world
        buf = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
        check(self.lib.kssd_b200_peer_connect(self.h, buf, world, rank))

    def sketch_publish(self):
        check(self.lib.kssd_b200_sketch_publish(self.h))

    def sketch_gather_peers(self):
        """-> (d_hashes, d_genome_off, total, overflow) over all ranks; call after a cross-rank barrier on the stream."""
        ph, po, tot, ov = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_int()
        check(self.lib.kssd_b200_sketch_gather_peers(self.h, C.byref(ph), C.byref(po), C.byref(tot), C.byref(ov)))
        return ph.value or 0, po.value or 0, tot.value, bool(ov.value)

    def sketch(self, seq, genome_off, qual=None, least_qual=0, least_num_kmer=1, batch_bytes=None):
        """Sketch a packed batch (rabbitkssd_b200.synth.pack_batch layout) through pinned host batches.
        genome_off[n+1]: byte offsets; genome g owns [genome_off[g], genome_off[g+1]).  Large inputs are cut into
        batches of ~batch_bytes at genome boundaries."""
        genome_off = np.ascontiguousarray(genome_off, dtype=np.uint64)
        n = len(genome_off) - 1
        self.sketch_begin(n, qual is not None, least_qual, least_num_kmer)
        if batch_bytes is None:
            batch_bytes = 256 << 20
        g = 0
        while g < n:
            g1 = g + 1
            while g1 < n and int(genome_off[g1 + 1]) - int(genome_off[g]) <= batch_bytes:
                g1 += 1
            b0, b1 = int(genome_off[g]), int(genome_off[g1])
            so = (genome_off[g:g1 + 1] - genome_off[g]).astype(np.uint64)
            self.submit_host(seq[b0:b1], so, np.arange(g, g1, dtype=np.uint32), None if qual is None else qual[b0:b1])
            g = g1
        return self.sketch_finish()

    # ---- index ----
    def index_build(self, hashes, genome_off):
        """hashes: uint32, or uint64 for the 64-bit hash configuration (the dtype decides)."""
        hashes = np.ascontiguousarray(hashes)
        self.index64 = hashes.dtype == np.uint64
        if not self.index64:
            hashes = np.ascontiguousarray(hashes, dtype=np.uint32)
        genome_off = np.ascontiguousarray(genome_off, dtype=np.uint64)
        fn = self.lib.kssd_b200_index_build64 if self.index64 else self.lib.kssd_b200_index_build
        check(fn(self.h, C.c_void_p(hashes.ctypes.data), C.c_void_p(genome_off.ctypes.data), len(genome_off) - 1, 0))

    def index_build_device(self, d_hashes, d_genome_off, n_genomes, hash64=False):
        self.index64 = bool(hash64)
        fn = self.lib.kssd_b200_index_build64 if hash64 else self.lib.kssd_b200_index_build
        check(fn(self.h, C.c_void_p(d_hashes), C.c_void_p(d_genome_off), n_genomes, 1))

    def index_build_rows(self, hashes, genome_off, rows):
        """Index restricted to the alldist rows [rows[0], rows[1]) (multi-GPU row sharding)."""
        hashes = np.ascontiguousarray(hashes, dtype=np.uint32)
        genome_off = np.ascontiguousarray(genome_off, dtype=np.uint64)
        self.index64 = False
        check(self.lib.kssd_b200_index_build_rows(self.h, C.c_void_p(hashes.ctypes.data), C.c_void_p(genome_off.ctypes.data),
                                                  len(genome_off) - 1, 0, rows[0], rows[1]))

    def index_build_rows_device(self, d_hashes, d_genome_off, n_genomes, rows):
        self.index64 = False
        check(self.lib.kssd_b200_index_build_rows(self.h, C.c_void_p(d_hashes), C.c_void_p(d_genome_off), n_genomes, 1, rows[0], rows[1]))

    def index_build_range(self, d_hashes, d_genome_off, n_genomes, part, parts, d_block, block_words):
        """CSR of hash range `part` of `parts` packed into the caller's device block -> (overflow flag, words used)."""
        ov, used = C.c_int(), C.c_uint64()
        check(self.lib.kssd_b200_index_build_range(self.h, d_hashes, d_genome_off, n_genomes, part, parts, d_block, block_words,
                                                   C.byref(used), C.byref(ov)))
        return bool(ov.value), used.value

    def index_adopt_parts(self, d_blocks, block_words, parts, d_genome_off, n_genomes):
        """Exchanged pieces -> this context's index; -> overflow flag (some rank's piece did not fit)."""
        ov = C.c_int()
        check(self.lib.kssd_b200_index_adopt_parts(self.h, d_blocks, block_words, parts, d_genome_off, n_genomes, C.byref(ov)))
        return bool(ov.value)

    def index_export(self):
        """-> (uniq_hash u32[U] (u64 for a 64-bit index), uniq_count u32[U], dict u32[P]), ascending hash order"""
        U, P = C.c_uint64(), C.c_uint64()
        uh, uc, dc = (_lib.u64p() if self.index64 else _lib.u32p()), _lib.u32p(), _lib.u32p()
        fn = self.lib.kssd_b200_index_export64 if self.index64 else self.lib.kssd_b200_index_export
        check(fn(self.h, C.byref(U), C.byref(P), C.byref(uh), C.byref(uc), C.byref(dc)))
        return (_np_from_ptr(uh, U.value, np.uint64 if self.index64 else np.uint32), _np_from_ptr(uc, U.value, np.uint32),
                _np_from_ptr(dc, P.value, np.uint32))

    def index_load64(self, hashes, counts, postings, ref_sizes):
        """The content of the 64-bit .index / .dict files (formats.read_index_dict64), any hash order."""
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        counts = np.ascontiguousarray(counts, dtype=np.uint32)
        postings = np.ascontiguousarray(postings, dtype=np.uint32)
        ref_sizes = np.ascontiguousarray(ref_sizes, dtype=np.uint32)
        check(self.lib.kssd_b200_index_load64(self.h, len(hashes), hashes.ctypes.data_as(_lib.u64p), counts.ctypes.data_as(_lib.u32p),
                                              postings.ctypes.data_as(_lib.u32p), len(postings),
                                              ref_sizes.ctypes.data_as(_lib.u32p), len(ref_sizes)))
        self.index64 = True

    def index_load(self, counts, postings, ref_sizes):
        counts = np.ascontiguousarray(counts, dtype=np.uint32)
        postings = np.ascontiguousarray(postings, dtype=np.uint32)
        ref_sizes = np.ascontiguousarray(ref_sizes, dtype=np.uint32)
        check(self.lib.kssd_b200_index_load(self.h, len(counts), counts.ctypes.data_as(_lib.u32p),
                                            postings.ctypes.data_as(_lib.u32p), len(postings),
                                            ref_sizes.ctypes.data_as(_lib.u32p), len(ref_sizes)))
        self.index64 = False

    def index_device_views(self):
        U, P, N = C.c_uint64(), C.c_uint64(), C.c_uint32()
        a, b, c_, d = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self.lib.kssd_b200_index_device_views(self.h, C.byref(U), C.byref(P), C.byref(N), C.byref(a), C.byref(b),
                                                    C.byref(c_), C.byref(d)))
        return dict(n_uniq=U.value, total=P.value, n_ref=N.value, d_uniq_hash=a.value, d_uniq_off=b.value,
                    d_dict=c_.value, d_ref_sizes=d.value)

    def index_adopt_device(self, n_uniq, total, n_ref, d_uniq_hash, d_uniq_off, d_dict, d_ref_sizes):
        check(self.lib.kssd_b200_index_adopt_device(self.h, n_uniq, total, n_ref, C.c_void_p(d_uniq_hash),
                                                    C.c_void_p(d_uniq_off), C.c_void_p(d_dict), C.c_void_p(d_ref_sizes)))

    # ---- set algebra ----
    def set_union(self, hashes):
        """All hashes (uint32 or uint64, any order) -> the sorted distinct set (`rabbit_kssd union`)."""
        hashes = np.ascontiguousarray(hashes)
        dt = np.uint64 if hashes.dtype == np.uint64 else np.uint32
        hashes = np.ascontiguousarray(hashes, dtype=dt)
        out, n = C.c_void_p(), C.c_uint64()
        check(self.lib.kssd_b200_set_union(self.h, C.c_void_p(hashes.ctypes.data), len(hashes), hashes.itemsize, C.byref(out), C.byref(n)))
        if n.value == 0:
            return np.zeros(0, dt)
        return np.frombuffer((C.c_char * (n.value * hashes.itemsize)).from_address(out.value), dtype=dt).copy()

    def set_subtract(self, q_hashes, q_off, ref_set):
        """Query sketches (CSR) minus a sorted distinct reference set -> (hashes, genome_off) (`rabbit_kssd sub`)."""
        ref_set = np.ascontiguousarray(ref_set)
        dt = np.uint64 if ref_set.dtype == np.uint64 else np.uint32
        ref_set = np.ascontiguousarray(ref_set, dtype=dt)
        q_hashes = np.ascontiguousarray(q_hashes, dtype=dt)
        q_off = np.ascontiguousarray(q_off, dtype=np.uint64)
        nq = len(q_off) - 1
        out, off = C.c_void_p(), _lib.u64p()
        check(self.lib.kssd_b200_set_subtract(self.h, C.c_void_p(q_hashes.ctypes.data), C.c_void_p(q_off.ctypes.data), nq,
                                              C.c_void_p(ref_set.ctypes.data), len(ref_set), ref_set.itemsize, C.byref(out), C.byref(off)))
        goff = _np_from_ptr(off, nq + 1, np.uint64)
        n = int(goff[-1])
        h = np.frombuffer((C.c_char * (n * ref_set.itemsize)).from_address(out.value), dtype=dt).copy() if n else np.zeros(0, dt)
        return h, goff

    def set_subtract_noalloc(self, d_q_hashes, d_q_off, n_query, n_hashes, d_ref_set, n_ref, d_out_hashes, d_out_off):
        """set_subtract on caller-owned DEVICE buffers (raw addresses), scratch in the library's fixed __device__ array: nothing
        is allocated.  At most 4,194,304 query hashes (uint32) in 65,536 genomes; enqueued on the context's stream."""
        check(self.lib.kssd_b200_set_subtract_noalloc(self.h, C.c_void_p(d_q_hashes), C.c_void_p(d_q_off), int(n_query), int(n_hashes),
                                                      C.c_void_p(d_ref_set), int(n_ref), C.c_void_p(d_out_hashes), C.c_void_p(d_out_off)))

    # ---- distances ----
    def set_emit_zero_cells(self, on=True):
        check(self.lib.kssd_b200_set_emit_zero_cells(self.h, int(on)))

    def set_max_neighbor(self, n):
        """dist -N: the device reports per query row only the n smallest distances (+ ties); 0 = off."""
        check(self.lib.kssd_b200_set_max_neighbor(self.h, int(n)))

    def dist_rows(self, q_hashes, q_off, triangle, kmer_size, max_dist, containment=0, rows=None):
        """-> structured hits (row, col, common) that pass the device's conservative test (host refines)."""
        q_hashes = np.ascontiguousarray(q_hashes, dtype=np.uint64 if self.index64 else np.uint32)
        q_off = np.ascontiguousarray(q_off, dtype=np.uint64)
        nq = len(q_off) - 1
        r0, r1 = (0, nq) if rows is None else rows
        chunks = []

        def sink(user, hits, n):
            arr = np.frombuffer((C.c_char * (n * 12)).from_address(C.addressof(hits.contents)), dtype=np.uint32)
            chunks.append(arr.reshape(-1, 3).copy())
            return 0

        cb = _lib.SINK(sink)
        fn = self.lib.kssd_b200_dist_rows64 if self.index64 else self.lib.kssd_b200_dist_rows
        check(fn(self.h, C.c_void_p(q_hashes.ctypes.data), C.c_void_p(q_off.ctypes.data), nq, 0,
                 r0, r1, int(triangle), kmer_size, float(max_dist), int(containment), cb, None))
        if not chunks:
            return np.zeros((0, 3), dtype=np.uint32)
        return np.concatenate(chunks)

    def dist_rows_device(self, d_q_hashes, d_q_off, n_query, triangle, kmer_size, max_dist, containment=0, rows=None):
        r0, r1 = (0, n_query) if rows is None else rows
        hp, n = C.c_void_p(), C.c_uint64()
        fn = self.lib.kssd_b200_dist_rows_device64 if self.index64 else self.lib.kssd_b200_dist_rows_device
        check(fn(self.h, C.c_void_p(d_q_hashes), C.c_void_p(d_q_off), n_query, r0, r1,
                 int(triangle), kmer_size, float(max_dist), int(containment), C.byref(hp), C.byref(n)))
        return hp.value or 0, n.value


def refine_hits(hits, row_sizes, col_sizes, kmer_size, max_dist, triangle, containment=0):
    """Final keep/drop with the reference's exact double expressions (src/dist.cpp:216-232, :237-251; strict '<'
    for alldist :232, '<=' for dist :624).  numpy's log is the same libm log the reference calls.
    Returns (hits_kept, jorc, dist)."""
    if len(hits) == 0:
        return hits, np.zeros(0), np.zeros(0)
    c = hits[:, 2].astype(np.int64)
    s_row = np.asarray(row_sizes, dtype=np.int64)[hits[:, 0]]
    s_col = np.asarray(col_sizes, dtype=np.int64)[hits[:, 1]]
    empty = (s_row == 0) | (s_col == 0)
    with np.errstate(divide="ignore", invalid="ignore"):
        if not containment:
            v = np.where(empty, 0.0, c / (s_row + s_col - c).astype(np.float64))
            arg = (2 * v) / (1.0 + v)
        else:
            v = np.where(empty, 0.0, c / np.minimum(s_row, s_col).astype(np.float64))
            arg = v
        import math
        d = np.array([0.0 if x == 1.0 else 1.0 if x == 0.0 else -1.0 / kmer_size * math.log(a) for x, a in zip(v, arg)])
    keep = d < max_dist if triangle else d <= max_dist
    return hits[keep], v[keep], d[keep]
