"""Deterministic synthetic inputs of the shapes SURVEY.md 8(d) names.

Everything is derived from a counter-based splitmix64 stream implemented in numpy uint64 arithmetic, so a
(seed, shape) pair gives the same bytes on every numpy version and machine: the golden fixtures under
tests/golden/ store only seeds + expected outputs and the inputs are regenerated at test time.
"""
import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def splitmix64(seed, n, offset=0):
    """n uint64 values of the splitmix64 sequence for `seed`, starting at counter `offset`."""
    with np.errstate(over="ignore"):
        z = (np.arange(offset + 1, offset + n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
             + np.uint64(seed & 0xFFFFFFFFFFFFFFFF))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def random_bases(seed, length):
    """uint8 ASCII array of i.i.d. uniform ACGT."""
    r = splitmix64(seed, (length + 31) // 32)
    codes = ((r[:, None] >> (np.arange(32, dtype=np.uint64) * np.uint64(2))) & np.uint64(3)).astype(np.uint8)
    return _ACGT[codes.reshape(-1)[:length]]


def mutate(seq, seed, rate):
    """i.i.d. substitutions at `rate`: a mutated base becomes one of the three OTHER bases."""
    n = len(seq)
    r = splitmix64(seed, n)
    thresh = np.uint64(int(rate * 2.0 ** 53))
    hit = (r >> np.uint64(11)) < thresh
    out = seq.copy()
    idx = np.nonzero(hit)[0]
    if len(idx):
        lut = np.full(256, 255, dtype=np.uint8)
        lut[_ACGT] = np.arange(4, dtype=np.uint8)
        old = lut[seq[idx] & 0xDF]
        shift = ((r[idx] & np.uint64(0x7FF)) % np.uint64(3)).astype(np.uint8) + 1
        ok = old != 255
        new = _ACGT[(old + shift) & 3]
        out[idx[ok]] = new[ok]
    return out


def sprinkle(seq, seed, n_runs=0, run_len=10, lowercase_frac=0.0):
    """Exercise BaseMap edge cases: runs of 'N' and lowercase stretches."""
    out = seq.copy()
    n = len(out)
    if n_runs and n > run_len:
        pos = (splitmix64(seed ^ 0xABCDEF, n_runs) % np.uint64(n - run_len)).astype(np.int64)
        for p in pos:
            out[p:p + run_len] = ord("N")
    if lowercase_frac > 0 and n > 100:
        k = max(1, int(n * lowercase_frac / 50))
        pos = (splitmix64(seed ^ 0x123457, k) % np.uint64(n - 50)).astype(np.int64)
        for p in pos:
            seg = out[p:p + 50]
            up = (seg >= 65) & (seg <= 90)
            seg[up] |= 0x20
    return out


def family(seed, n_members, length, max_div=0.04, n_contigs=1):
    """One 'bacterial family': base genome + (n_members-1) substitution variants with divergence spread
    evenly in (0, max_div].  Returns list of genomes; a genome is a list of uint8 contig arrays."""
    base = random_bases(seed, length)
    out = []
    for m in range(n_members):
        g = base if m == 0 else mutate(base, seed * 1000003 + m, max_div * m / max(1, n_members - 1))
        cuts = [length * c // n_contigs for c in range(n_contigs + 1)]
        out.append([g[cuts[c]:cuts[c + 1]] for c in range(n_contigs)])
    return out


def write_fasta(path, contigs, name="seq", width=80):
    with open(path, "wb") as f:
        for ci, c in enumerate(contigs):
            f.write(b">" + f"{name}_{ci} synthetic\n".encode())
            n = len(c)
            full = (n // width) * width
            if full:
                body = np.empty((n // width, width + 1), dtype=np.uint8)
                body[:, :width] = c[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(c[full:].tobytes() + b"\n")


def write_fastq(path, reads, quals, name="read"):
    with open(path, "wb") as f:
        for i, (r, q) in enumerate(zip(reads, quals)):
            f.write(b"@" + f"{name}_{i}\n".encode() + r.tobytes() + b"\n+\n" + q.tobytes() + b"\n")


def pack_batch(genomes, quals=None, sep=ord("\n"), align=16):
    """Concatenate parsed records into the flat batch layout the C-ABI takes: records of one genome are
    separated by one `sep` byte (any non-ACGT byte resets the valid-run counter exactly like a new record
    does, sketch.cpp:487-488), genomes start at `align`-byte multiples.
    Returns (seq uint8[n], qual uint8[n] or None, genome_off uint64[n_genomes+1])."""
    sizes = []
    for g in genomes:
        sizes.append(sum(len(c) for c in g) + max(0, len(g) - 1))
    offs = [0]
    for s in sizes:
        end = offs[-1] + s + 1  # at least one separator after every genome
        offs.append((end + align - 1) // align * align)
    total = offs[-1]
    seq = np.full(total, sep, dtype=np.uint8)
    qual = None if quals is None else np.zeros(total, dtype=np.uint8)
    for gi, g in enumerate(genomes):
        p = offs[gi]
        for ci, c in enumerate(g):
            seq[p:p + len(c)] = c
            if quals is not None:
                qual[p:p + len(c)] = quals[gi][ci]
            p += len(c) + 1
    return seq, qual, np.asarray(offs, dtype=np.uint64)


def oracle_records(genomes):
    """Same genomes as (seq, rec_off, rec_first) record tables for oracle.sketch_batch (no separators needed)."""
    rec_off = [0]
    rec_first = [0]
    parts = []
    for g in genomes:
        for c in g:
            parts.append(c)
            rec_off.append(rec_off[-1] + len(c))
        rec_first.append(len(rec_off) - 1)
    seq = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
    return seq, np.asarray(rec_off, np.uint64), np.asarray(rec_first, np.uint64)


def synth_sketches(seed, n_clusters, cluster_size, sketch_size, hash_bits, survival_step=0.04):
    """Directly synthesised sketches (SURVEY.md 8d, C2/C5 template): each cluster draws `sketch_size` root
    hashes uniform in [0, 2^hash_bits); member v keeps each root hash with probability 1 - survival_step*v
    (v = 0..cluster_size-1, clipped at 0.3) and tops up with private hashes to about `sketch_size`.
    Returns (hashes uint32 sorted unique per genome, genome_off uint64)."""
    chunks = []
    offs = [0]
    mask = np.uint64((1 << hash_bits) - 1)
    for c in range(n_clusters):
        root = np.unique((splitmix64(seed + 7919 * c, sketch_size) & mask).astype(np.uint32))
        for v in range(cluster_size):
            keep_p = max(0.3, 1.0 - survival_step * v)
            r = splitmix64(seed * 31 + c * 1009 + v, len(root) + sketch_size, offset=1 << 20)
            keep = (r[:len(root)] >> np.uint64(11)) < np.uint64(int(keep_p * 2.0 ** 53))
            kept = root[keep]
            n_priv = max(0, sketch_size - len(kept))
            priv = (r[len(root):len(root) + n_priv] & mask).astype(np.uint32)
            h = np.unique(np.concatenate([kept, priv]))
            chunks.append(h)
            offs.append(offs[-1] + len(h))
    hashes = np.concatenate(chunks) if chunks else np.zeros(0, np.uint32)
    return hashes, np.asarray(offs, dtype=np.uint64)


# ---- the same generator on a CUDA device (bench.py: genomes generated straight into HBM) ---------------------------
def _i64(v):
    v &= 0xFFFFFFFFFFFFFFFF
    return v - (1 << 64) if v >= (1 << 63) else v


def _splitmix64_torch(torch, seed, n, device):
    """splitmix64(seed, n) as int64 bit patterns (torch has no uint64 arithmetic: multiplies wrap, shifts are made logical)."""
    z = torch.arange(1, n + 1, dtype=torch.int64, device=device) * _i64(0x9E3779B97F4A7C15) + _i64(seed)
    z = (z ^ ((z >> 30) & ((1 << 34) - 1))) * _i64(0xBF58476D1CE4E5B9)
    z = (z ^ ((z >> 27) & ((1 << 37) - 1))) * _i64(0x94D049BB133111EB)
    return z ^ ((z >> 31) & ((1 << 33) - 1))


def family_device(torch, seed, n_members, length, max_div=0.04, device="cuda"):
    """family(seed, n_members, length, max_div) on the device: list of uint8 ASCII tensors (one per member, contigs not
    cut), byte-identical to the numpy generator."""
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=device)
    r = _splitmix64_torch(torch, seed, (length + 31) // 32, device)
    sh = torch.arange(32, dtype=torch.int64, device=device) * 2
    base = ((r[:, None] >> sh[None, :]) & 3).reshape(-1)[:length]
    out = []
    for m in range(n_members):
        codes = base
        if m:
            rate = max_div * m / max(1, n_members - 1)
            thresh = int(rate * 2.0 ** 53)
            q = _splitmix64_torch(torch, seed * 1000003 + m, length, device)
            hit = ((q >> 11) & ((1 << 53) - 1)) < thresh
            shift = (q & 0x7FF) % 3 + 1
            codes = torch.where(hit, (base + shift) & 3, base)
        out.append(lut[codes])
    return out
