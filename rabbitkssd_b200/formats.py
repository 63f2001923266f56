"""Readers/writers for the reference's frozen on-disk formats (harness side; the C++ host shims under
rabbitkssd_b200/csrc/host/ carry their own copies for the drop-in CLI).

.shuf    : {int id,k,subk,drlevel} + int32[16^subk]              (reference src/shuffle.cpp:8-23, :48-56)
.sketch  : sketchInfo_t {id,half_k,half_subk,drlevel,genomeNumber} + int nameLen[N] + int hashCnt[N]
           + per genome: name bytes, then u32 (or u64 when half_k-drlevel>8) hashes
                                                                 (reference src/sketch.cpp:1024-1068)
.index   : size_t hashSize; u64 totalIndex; u32 count[hashSize]  (32-bit path, src/sketch.cpp:1004-1008)
.dict    : u32 postings, ascending hash then ascending genome    (src/sketch.cpp:991-1000)
.alldist/.dist text lines                                        (src/dist.cpp:233, :642, :686)
"""
import struct

import numpy as np


def write_shuf(path, half_k, half_subk, drlevel, shuffled_dim):
    sid = (half_k << 8) + (half_subk << 4) + drlevel
    with open(path, "wb") as f:
        f.write(struct.pack("<4i", sid, half_k, half_subk, drlevel))
        f.write(np.ascontiguousarray(shuffled_dim, dtype=np.int32).tobytes())


def read_shuf(path):
    with open(path, "rb") as f:
        sid, k, s, l = struct.unpack("<4i", f.read(16))
        dim = np.fromfile(f, dtype=np.int32, count=1 << (4 * s))
    if len(dim) != 1 << (4 * s):
        raise IOError("read_shuf: truncated shuffle file " + path)
    return {"id": sid, "half_k": k, "half_subk": s, "drlevel": l}, dim


def write_sketch(path, half_k, half_subk, drlevel, names, hashes, genome_off):
    use64 = half_k - drlevel > 8
    n = len(names)
    sid = (half_k << 8) + (half_subk << 4) + drlevel
    off = np.asarray(genome_off, dtype=np.int64)
    with open(path, "wb") as f:
        f.write(struct.pack("<5i", sid, half_k, half_subk, drlevel, n))
        f.write(np.asarray([len(x.encode()) for x in names], dtype=np.int32).tobytes())
        f.write(np.diff(off).astype(np.int32).tobytes())
        h = np.ascontiguousarray(hashes, dtype=np.uint64 if use64 else np.uint32)
        for g in range(n):
            f.write(names[g].encode())
            f.write(h[off[g]:off[g + 1]].tobytes())


def read_sketch(path):
    """Returns (info dict, names, list of per-genome hash arrays in FILE order)."""
    with open(path, "rb") as f:
        sid, k, s, l, n = struct.unpack("<5i", f.read(20))
        use64 = k - l > 8
        name_len = np.fromfile(f, dtype=np.int32, count=n)
        cnt = np.fromfile(f, dtype=np.int32, count=n)
        names, sets = [], []
        for g in range(n):
            names.append(f.read(int(name_len[g])).decode())
            sets.append(np.fromfile(f, dtype=np.uint64 if use64 else np.uint32, count=int(cnt[g])))
    return {"id": sid, "half_k": k, "half_subk": s, "drlevel": l, "genomeNumber": n}, names, sets


def read_index_dict(index_path, dict_path):
    with open(index_path, "rb") as f:
        hash_size, total = struct.unpack("<QQ", f.read(16))
        counts = np.fromfile(f, dtype=np.uint32, count=hash_size)
    postings = np.fromfile(dict_path, dtype=np.uint32)
    if len(postings) != total or len(counts) != hash_size:
        raise IOError("read_index_dict: mismatched sizes")
    return counts, postings


def read_index_dict64(index_path, dict_path):
    """64-bit hash variant (half_k - drlevel > 8, src/sketch.cpp:904-969): .index = size_t n; u64 hash[n]; u32 count[n]
    in the iteration order of the reference's hash map, .dict = the posting lists in the same order.
    Returns (hashes u64[n], counts u32[n], postings u32[sum counts])."""
    with open(index_path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        hashes = np.fromfile(f, dtype=np.uint64, count=n)
        counts = np.fromfile(f, dtype=np.uint32, count=n)
    postings = np.fromfile(dict_path, dtype=np.uint32)
    if len(hashes) != n or len(counts) != n or int(counts.sum()) != len(postings):
        raise IOError("read_index_dict64: mismatched sizes")
    return hashes, counts, postings


def parse_dist_text(path):
    """Parse .alldist/.dist text into a set of canonical tuples
    (name_a, name_b, common, size0, size1, jorc_text, dist_text); header lines are skipped."""
    out = []
    with open(path, "r") as f:
        for line in f:
            line = line.rstrip("\n")
            if not line or line.startswith(" genome0"):
                continue
            a, b, trip, jorc, d = line.split("\t")
            c, s0, s1 = (int(x) for x in trip.split("|"))
            out.append((a, b, c, s0, s1, jorc, d))
    return out
