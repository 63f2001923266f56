// Fused Kssd sketch scan for sm_100a.
//
// Replaces the per-base loop of the reference (src/sketch.cpp:487-532, same arithmetic in :196-231 and the FASTQ
// variants :273-307, :779-825).  Not a translation: the reference rolls two 64-bit tuples base by base and probes a
// hash map for EVERY k-mer.  Here the work is split into a streaming filter that touches every base with ~3
// integer instructions and an exact path that only sees the <1% of positions the filter lets through:
//
//   filter (every base)
//   * a lane owns 32 bases: one 256-bit load (LDG.E.256), the next tile is loaded into the same registers as soon
//     as the bytes are packed (register prefetch one round ahead, no L2 prefetch); lane 0 of the warp re-reads
//     the 32 bases before the tile as halo, so no state is carried between tiles (tile = 31 x 32 = 992 new
//     bytes per warp);
//   * 4 ASCII bytes -> 8 bits with one AND and one IMAD ((x & 0x06060606) * M gathers bits 1-2 of each byte:
//     A/a 0, C/c 1, T/t 2, G/g 3).  Validity (N, separators, quality) is NOT decided here;
//   * membership of the 2s-mer is tested WITHOUT deciding the canonical strand and for G consecutive k-mer end
//     positions at once: the shared-memory bitmap holds, for every retained 2s-mer and its reverse complement,
//     one bit per alignment g < G of the 2s-mer against the bases that all G windows have in common (up to 10
//     bases = 20 index bits).  One funnel shift, one shift+mask, one LDS, two shifts decide G positions;
//   * groups that pass are appended to a per-warp shared-memory queue (ballot + popc, no atomics) together with
//     the lane's packed words [P0,P1,A0,A1].
//   stage 1 (queued groups, 32 at a time, one per lane; no memory traffic)
//   * canonical dim_id of each of the group's G positions from the packed words, looked up in a second-level
//     shared-memory filter (2^18 bits indexed by a hash of the full dim_id); survivors go to a second queue.
//   stage 2 (queued positions, 32 at a time, one per lane; the exact path)
//   * re-reads the 2k raw bytes of the k-mer (three 128-bit loads + a select network), decodes them with the
//     reference's BaseMap semantics (common.h:27-37), checks validity of the whole window, then runs the
//     reference's 64-bit arithmetic: canonical min of tuple / reverse complement, dim_id, exact retained-subset
//     lookup, hash composition (sketch.cpp:508-524), quality check (sketch.cpp:280) for survivors only, owning
//     genome by binary search in the segment table, one warp-aggregated atomic append to the emit buffer.
#include <algorithm>
#include <cstring>

#include "common.cuh"

namespace kssd {

struct ScanArgs {
    const uint8_t *seq;
    const uint8_t *qual;
    uint32_t n_bytes;          // <= 2^31: launch_scan cuts larger buffers into pieces
    uint32_t first_emit;       // k-mers ending before this byte are context only
    const uint32_t *bitmap;    // membership bitmap followed by the second-level filter (copied to shared memory)
    const unsigned long long *table;  // exact retained-subset table: (dim_id << 32 | value), ~0 = empty
    unsigned long long *emit;  // (genome << 32 | hash); 64-bit hash configuration: the hash alone
    uint32_t *emit_gen;        // 64-bit hash configuration (half_k - drlevel > 8, src/sketch.cpp:336): genome of emit[i]; else null
    const uint64_t *seg_off;   // segment table of the buffer (coordinates: byte 0 of this launch = pos_base)
    const uint32_t *seg_genome;
    const uint64_t *seg_first; // optional, per segment: k-mers ending before this position (coordinates of seg_off) are context only
    uint32_t n_seg;
    uint64_t pos_base;
    unsigned long long *cursor;
    uint32_t *gcount;          // emitted keys per genome (sizes the per-genome segments of the shared-memory sort)
    uint32_t *emit_rank;       // rank of emit[i] among the keys of its genome (= position inside the genome's segment)
    uint32_t n_genomes;
    uint64_t emit_cap;
    uint32_t n_tiles;
    uint32_t n_interior;       // tiles [1, n_interior) lie completely inside the buffer (halo included)
    uint32_t bitmap_words;
    uint32_t addr_mask;        // ((1<<wbits)-1) << 2
    uint32_t mul29;            // 1u << 29
    int preshift;              // bit offset (mod 32) of the first index base of test 0 inside [P0,P1,A0,A1]
    int least_qual;
    ScanParams sp;
};

static constexpr int kScanThreads = 1024;
static constexpr int kScanWarps = kScanThreads / 32;
static constexpr int kLaneBases = 32;                          // bases owned by a lane
static constexpr int kTileLanes = 31;                          // lanes 1..31 own new bases, lane 0 is halo
static constexpr int kTileBytes = kTileLanes * kLaneBases;     // 992
static constexpr int kQueueCap = 64;                           // per-warp candidate / output queue entries
static constexpr int kWindowWords = 7;                         // 28 bytes >= 2k + G - 1 (k <= 12, G <= 4)
static constexpr uint32_t kMaxPiece = 1u << 30;                // bytes per launch (positions are 32-bit)
static constexpr int kL2Bits = 18;                             // second-level filter: 2^18 bits = 32 KB of shared memory
static constexpr int kL2Words = (1 << kL2Bits) / 32;

// ---- SWAR decode -----------------------------------------------------------------------------------------
// filter encoding: bits 1-2 of the ASCII byte (A/a 0, C/c 1, T/t 2, G/g 3); 4 bytes -> 8 bits in the top byte,
// first base in the LOWEST two bits.
__device__ __forceinline__ uint32_t pack4_filter(uint32_t x) {
    return (x & 0x06060606u) * ((1u << 23) | (1u << 17) | (1u << 11) | (1u << 5));
}
__device__ __forceinline__ uint32_t pack16_filter(uint4 v) {
    uint32_t p0 = pack4_filter(v.x), p1 = pack4_filter(v.y), p2 = pack4_filter(v.z), p3 = pack4_filter(v.w);
    uint32_t t01 = __byte_perm(p0, p1, 0x0073);  // byte0 = p0.top, byte1 = p1.top
    uint32_t t23 = __byte_perm(p2, p3, 0x7300);  // byte2 = p2.top, byte3 = p3.top
    return __byte_perm(t01, t23, 0x7610);
}

// exact decode of 4 ASCII bytes (reference encoding A/a 0, C/c 1, G/g 2, T/t 3 in bits [0,2) of each byte);
// err: non-zero byte <=> byte not in "ACGTacgt".  Bits b2,b1 of the four letters are the Gray code of 0..3; the
// other six bits are a function of them: b7 b6 b5 b3 = 0 1 x 0, b4 = b2 & ~b1 (only T), b0 = ~b4.
__device__ __forceinline__ void decode4(uint32_t x, uint32_t &code, uint32_t &err) {
    uint32_t s1 = x >> 1, s2 = x >> 2;
    code = (s1 ^ s2) & 0x03030303u;
    uint32_t z = s2 & ~s1 & 0x01010101u;
    uint32_t zz = z * 0x11u;
    uint32_t t = (x & 0xD9D9D9D9u) ^ 0x41414141u;
    err = t ^ zz;
}
// 4 codes (one per byte, first base in the lowest byte) -> 8 bits, first base most significant.
__device__ __forceinline__ uint32_t pack4_msb(uint32_t code) { return (code * 0x40100401u) >> 24; }
// per-byte non-zero flags of e -> 4 bits (bit i = byte i non-zero)
__device__ __forceinline__ uint32_t nz4(uint32_t e) {
    uint32_t f = (((e & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e) & 0x80808080u;
    return ((f >> 7) * 0x01020408u) >> 24 & 0xFu;
}

// one 256-bit global load (sm_100: LDG.E.256): a warp reads 1 KB contiguous with ONE instruction, so every
// 32-byte sector is requested once (two 128-bit loads per lane would touch each sector twice)
__device__ __forceinline__ void ldg256(const uint8_t *p, uint4 &v0, uint4 &v1) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v0.x), "=r"(v0.y), "=r"(v0.z), "=r"(v0.w), "=r"(v1.x), "=r"(v1.y), "=r"(v1.z), "=r"(v1.w)
                 : "l"(p));
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// 16 bytes at byte offset off (multiple of 16, may be negative); bytes outside [0, n_bytes) read as 0x00
__device__ __forceinline__ uint4 load_chunk(const uint8_t *base, int off, uint32_t n_bytes) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (off >= 0) {
        const uint32_t b = (uint32_t)off;
        if (b + 16 <= n_bytes) {
            v = __ldg(reinterpret_cast<const uint4 *>(base + b));
        } else if (b < n_bytes) {
            uint32_t w[4] = {0, 0, 0, 0};
            for (uint32_t j = 0; j < 16 && b + j < n_bytes; j++) w[j >> 2] |= (uint32_t)base[b + j] << (8 * (j & 3));
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    return v;
}

// 28 raw bytes starting at byte `start` (any alignment, may be negative): three aligned 128-bit loads and a
// select network instead of byte loads.
__device__ __forceinline__ void load_window(const uint8_t *base, int start, uint32_t n_bytes, uint32_t (&w)[kWindowWords]) {
    const int a0 = start & ~15;  // floor to 16 (two's complement)
    uint4 v0, v1, v2;
    if (a0 >= 0 && (uint32_t)a0 + 48 <= n_bytes) {
        const uint4 *p = reinterpret_cast<const uint4 *>(base + (uint32_t)a0);
        v0 = __ldg(p); v1 = __ldg(p + 1); v2 = __ldg(p + 2);
    } else {
        v0 = load_chunk(base, a0, n_bytes); v1 = load_chunk(base, a0 + 16, n_bytes); v2 = load_chunk(base, a0 + 32, n_bytes);
    }
    const uint32_t S[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
    const int o = start - a0;
    const bool q2 = o & 8, q1 = o & 4;
    const int r8 = (o & 3) * 8;
    uint32_t T[kWindowWords + 2];
#pragma unroll
    for (int i = 0; i < kWindowWords + 2; i++) T[i] = q2 ? S[i + 2] : S[i];
    uint32_t U[kWindowWords + 1];
#pragma unroll
    for (int i = 0; i < kWindowWords + 1; i++) U[i] = q1 ? T[i + 1] : T[i];
#pragma unroll
    for (int i = 0; i < kWindowWords; i++) w[i] = __funnelshift_r(U[i], U[i + 1], r8);
}

// reverse the order of the 32 two-bit fields of x
__device__ __forceinline__ uint64_t pairrev64(uint64_t x) {
    uint64_t r = __brevll(x);
    return ((r >> 1) & 0x5555555555555555ull) | ((r & 0x5555555555555555ull) << 1);
}

// reference arithmetic from the two packings of a k-mer (sketch.cpp:496-524):
//   tuple = bases first-most-significant, rvs = reverse complement; canonical min, dim_id of the inner 2s-mer
__device__ __forceinline__ void canonical_dim(const ScanParams &sp, uint64_t tuple, uint64_t rvs, uint64_t &uni, uint32_t &dim) {
    uni = tuple < rvs ? tuple : rvs;
    dim = (uint32_t)((uni & sp.domask) >> (2 * sp.hoc));
}

// Stage 2 (exact): one k-mer END position p per lane, everything re-derived from the raw bytes with the
// reference's semantics.  Returns true and the hash when the k-mer is retained.
__device__ __forceinline__ bool verify_kmer(const ScanArgs &a, uint32_t p, uint64_t &hash) {
    const ScanParams &sp = a.sp;
    const int L = sp.kmer_size;
    const int start = (int)p - (L - 1);
    uint32_t w[kWindowWords];
    load_window(a.seq, start, a.n_bytes, w);
    uint64_t C = 0;  // 2-bit codes of the 28 bytes, first byte most significant
    uint32_t E = 0;  // non-zero <=> some byte of the 28 is not a base
    uint32_t err[kWindowWords];
#pragma unroll
    for (int i = 0; i < kWindowWords; i++) {
        uint32_t code;
        decode4(w[i], code, err[i]);
        C = (C << 8) | pack4_msb(code);
        E |= err[i];
    }
    const uint64_t tuple = (C >> (2 * (4 * kWindowWords - L))) & sp.tupmask;
    const uint64_t rvs = (~pairrev64(tuple) >> (64 - 2 * L)) & sp.tupmask;
    uint64_t uni;
    uint32_t dim;
    canonical_dim(sp, tuple, rvs, uni, dim);
    uint32_t s = (dim * 0x9E3779B1u) >> sp.table_shift;
    unsigned long long e = __ldg(a.table + s);
    while (e != ~0ull && (uint32_t)(e >> 32) != dim) {
        s = (s + 1) & sp.table_mask;
        e = __ldg(a.table + s);
    }
    if (e == ~0ull) return false;  // not in the retained subset (sketch.cpp:509-511)
    const uint32_t lmask = (1u << L) - 1u;  // L <= 24
    if (E) {  // exact per-byte validity (sketch.cpp:494-505): every base of the window must be ACGTacgt
        uint32_t V = 0;
#pragma unroll
        for (int i = 0; i < kWindowWords; i++) V |= (~nz4(err[i]) & 0xFu) << (4 * i);
        if ((V & lmask) != lmask) return false;
    }
    if (p < a.first_emit) return false;  // context of a continued record
    if (a.qual) {  // valid only if (signed char)quality >= least_qual (sketch.cpp:280)
        uint32_t qw[kWindowWords];
        load_window(a.qual, start, a.n_bytes, qw);
        const uint32_t lq4 = (uint32_t)(a.least_qual & 0xFF) * 0x01010101u;
        uint32_t QV = 0;
#pragma unroll
        for (int i = 0; i < kWindowWords; i++) QV |= nz4(__vcmpges4(qw[i], lq4)) << (4 * i);
        if ((QV & lmask) != lmask) return false;
    }
    // dr_tuple, src/sketch.cpp:518-524: up to 4 * (half_k - drlevel) bits; the 32-bit configuration keeps the low word
    hash = (((uni & sp.undomask0) | ((uni & sp.undomask1) << (4 * sp.half_k - 4 * sp.hoc))) >> (4 * sp.drlevel)) | (uint32_t)e;
    return true;
}

// Stage 1 (still approximate, no memory traffic): one queued group per lane.  The group's bases come from the
// packed words saved at enqueue time ([P0,P1,A0,A1], filter encoding, garbage for non-bases); for each of its
// positions the canonical dim_id is looked up in the second-level shared-memory filter.  Returns a bit per
// position that still may be retained.
template <int G>
__device__ __forceinline__ uint32_t screen_group(const ScanArgs &a, const uint32_t *s_l2, uint32_t pos, uint4 sw) {
    const ScanParams &sp = a.sp;
    const int L = sp.kmer_size;
    const int e0 = (int)(pos % kLaneBases);
    int cnt = kLaneBases - e0;
    if (cnt > G) cnt = G;
    // 64 bits of [P0,P1,A0,A1,0,0] starting at the first base of the group's first k-mer
    const int b = 66 - 2 * L + 2 * e0;  // lane-local base i sits at bit 64 + 2i
    const bool w2 = b & 64, w1 = b & 32;
    const int sh = b & 31;
    const uint32_t t0 = w2 ? sw.z : sw.x, t1 = w2 ? sw.w : sw.y, t2 = w2 ? 0u : sw.z, t3 = w2 ? 0u : sw.w;
    const uint32_t x0 = w1 ? t1 : t0, x1 = w1 ? t2 : t1, x2 = w1 ? t3 : t2;
    uint32_t flo = __funnelshift_r(x0, x1, sh), fhi = __funnelshift_r(x1, x2, sh);
    // filter encoding (A0 C1 T2 G3) -> reference encoding (A0 C1 G2 T3)
    flo ^= (flo >> 1) & 0x55555555u;
    fhi ^= (fhi >> 1) & 0x55555555u;
    const uint64_t F = ((uint64_t)fhi << 32) | flo;  // base q of the window at bits [2q, 2q+2)
    const uint64_t Fr = pairrev64(F);                // base q at bits [62-2q, 64-2q)
    uint32_t pass = 0;
#pragma unroll
    for (int j = 0; j < G; j++) {
        const uint64_t tuple = (Fr >> (64 - 2 * (j + L))) & sp.tupmask;
        const uint64_t rvs = (~F >> (2 * j)) & sp.tupmask;
        uint64_t uni;
        uint32_t dim;
        canonical_dim(sp, tuple, rvs, uni, dim);
        const uint32_t h = (dim * 0x9E3779B1u) >> (32 - kL2Bits);
        const uint32_t word = s_l2[h >> 5];
        if (j < cnt && ((word >> (h & 31)) & 1u)) pass |= 1u << j;
    }
    return pass;
}

// G: k-mer end positions decided per bitmap test.  W0: word offset (0/1) of the first index base of test 0.
// A32: seq is 32-byte aligned, tiles are read with one 256-bit load per lane.
template <int G, int W0, bool A32>
__global__ void __launch_bounds__(kScanThreads, 1) kssd_scan_kernel(const __grid_constant__ ScanArgs a) {
    extern __shared__ uint4 s_mem4[];
    uint32_t *s_mem = reinterpret_cast<uint32_t *>(s_mem4);
    uint32_t *s_bitmap = s_mem;
    uint32_t *s_l2 = s_mem + a.bitmap_words;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(a.bitmap);  // bitmap followed by the second-level filter
        for (uint32_t i = threadIdx.x; i < (a.bitmap_words + kL2Words) / 4; i += blockDim.x) s_mem4[i] = __ldg(src + i);
    }
    __syncthreads();

    constexpr unsigned FULL = 0xFFFFFFFFu;
    constexpr int T = (kLaneBases + G - 1) / G;  // tests per lane
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    uint4 *cand_s = s_mem4 + (a.bitmap_words + kL2Words) / 4 + warp * kQueueCap;                      // [P0,P1,A0,A1] per candidate
    uint32_t *cand_pos = s_mem + a.bitmap_words + kL2Words + kScanWarps * kQueueCap * 4 + warp * kQueueCap;
    uint32_t *ver_pos = cand_pos + kScanWarps * kQueueCap;
    const uint32_t tile_stride = gridDim.x * kScanWarps;
    // round k: CTA c owns the 32 consecutive tiles [(k * grid + c) * 32, +32), one per warp -- all SMs stream one
    // contiguous window of HBM.  No L2 prefetch: a bulk prefetch per CTA window re-read 40% of the bytes from DRAM
    // (ncu, profiles/r01m_scan_prefetch_variants.txt) and one per warp tile cost more issue slots than it saved latency.
    // The loop variable is tile - 1 (tm1): "tm1 < n_interior - 1" is then the whole fast-path test (tile >= 1, tile and
    // its halo inside the buffer, and therefore also tile < n_tiles) -- tile 0 wraps to 0xFFFFFFFF and fails it.
    uint32_t tm1 = blockIdx.x * kScanWarps + (uint32_t)warp - 1u;
    if (tm1 + 1u >= a.n_tiles) return;
    const uint32_t n_int_m1 = a.n_interior ? a.n_interior - 1u : 0u;
    const uint32_t addr_mask = a.addr_mask;
    const uint32_t mul29 = a.mul29;  // 1 << 29, from the parameter block so that it is not folded back into a shift
    const int r = a.preshift;
    const char *smem_base = reinterpret_cast<const char *>(s_bitmap);
    uint32_t lt;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
    int n_cand = 0, n_ver = 0;
    // 32-bit shared-window addresses of this warp's queues (kept in registers; the enqueue round is hot)
    uint32_t cand_pos_sa = (uint32_t)__cvta_generic_to_shared(cand_pos);
    uint32_t cand_s_sa = (uint32_t)__cvta_generic_to_shared(cand_s);
    // byte offset of the lane's 32 bases from the start of tile tm1 (lane 0 = halo: the 32 bytes before tile tm1 + 1)
    uint32_t lane_off1 = (uint32_t)((lane - 1) * kLaneBases + kTileBytes);
    uint32_t keep = lane ? 0xFFFFFFFFu : 0u;  // halo lane: its positions belong to lane 31 of the previous tile
    // opaque to the optimiser AND to ptxas (which rematerialises anything it can re-derive from threadIdx in every
    // round: S2R + 5 integer ops): bounce the per-lane constants through the warp's still-empty queue
    asm volatile("" : "+r"(cand_pos_sa), "+r"(cand_s_sa));
    sts32(cand_pos_sa + lane * 4, lane_off1);
    sts32(cand_pos_sa + (32 + lane) * 4, keep);
    __syncwarp();
    asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(lane_off1) : "r"(cand_pos_sa + lane * 4) : "memory");
    asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(keep) : "r"(cand_pos_sa + (32 + lane) * 4) : "memory");
    __syncwarp();

    auto load_fast = [&](uint32_t t1, uint4 &v0, uint4 &v1) {  // tile t1 + 1, interior
        const uint8_t *p = a.seq + (t1 * kTileBytes + lane_off1);  // 32-bit offset: one IMAD + one 64-bit add
        if (A32) {
            ldg256(p, v0, v1);
        } else {
            v0 = __ldg(reinterpret_cast<const uint4 *>(p));
            v1 = __ldg(reinterpret_cast<const uint4 *>(p) + 1);
        }
    };
    auto load_edge = [&](uint32_t tl, uint4 &v0, uint4 &v1) {  // first / last tiles: bytes outside the buffer read as 0
        const int off = (int)(tl * kTileBytes) + (lane - 1) * kLaneBases;
        v0 = load_chunk(a.seq, off, a.n_bytes);
        v1 = load_chunk(a.seq, off + 16, a.n_bytes);
    };
    auto stage2 = [&](int base, int n) {  // exact path for queued positions [base, base+n), append survivors to the emit buffer
        uint32_t p = 0;
        uint64_t hash = 0;
        bool ok = false;
        if (lane < n) {
            p = ver_pos[base + lane];
            ok = verify_kmer(a, p, hash);
        }
        auto segment_of = [&](uint64_t gpos) {  // owning segment: largest s with seg_off[s] <= pos
            uint32_t lo = 0, hi = a.n_seg;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (__ldg(a.seg_off + mid) <= gpos) lo = mid; else hi = mid;
            }
            return lo;
        };
        uint32_t seg = 0;
        if (a.seg_first && ok) {  // raw-text batches: a piece may begin with context that an earlier piece already scanned
            seg = segment_of(a.pos_base + p);
            ok = a.pos_base + p >= __ldg(a.seg_first + seg);
        }
        const uint32_t m = __ballot_sync(FULL, ok);
        if (m) {
            unsigned long long at = 0;
            if (lane == 0) at = atomicAdd(a.cursor, (unsigned long long)__popc(m));
            at = __shfl_sync(FULL, at, 0) + __popc(m & lt);
            uint32_t genome = 0xFFFFFFFFu;  // lanes without a key to store
            if (ok && at < a.emit_cap) {
                if (!a.seg_first) seg = segment_of(a.pos_base + p);
                genome = __ldg(a.seg_genome + seg);
                if (a.emit_gen) {
                    a.emit[at] = hash;
                    a.emit_gen[at] = genome;
                } else {
                    a.emit[at] = ((unsigned long long)genome << 32) | (uint32_t)hash;
                }
                if (genome >= a.n_genomes) genome = 0xFFFFFFFFu;  // not a genome of this session: never dealt
            }
            // rank of the key inside its genome = running per-genome count; survivors of one warp almost always belong
            // to one genome, so one atomic per genome per warp (the counters are hot: all SMs stream the same genome)
            const uint32_t peers = __match_any_sync(FULL, genome);
            if (genome != 0xFFFFFFFFu) {
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(a.gcount + genome, (uint32_t)__popc(peers));
                base = __shfl_sync(peers, base, leader);
                a.emit_rank[at] = base + __popc(peers & lt);
            }
        }
    };
    auto stage1 = [&](int base, int n) {  // screen queued groups [base, base+n), queue surviving positions for stage 2
        uint32_t pos = 0, pass = 0;
        if (lane < n) {
            pos = cand_pos[base + lane];
            pass = screen_group<G>(a, s_l2, pos, cand_s[base + lane]);
        }
#pragma unroll
        for (int j = 0; j < G; j++) {
            const uint32_t m = __ballot_sync(FULL, (pass >> j) & 1u);
            if (m == 0) continue;
            if ((pass >> j) & 1u) ver_pos[n_ver + __popc(m & lt)] = pos + j;
            n_ver += __popc(m);
            __syncwarp();
            if (n_ver >= 32) {
                n_ver -= 32;
                stage2(n_ver, 32);
                __syncwarp();
            }
        }
    };

    uint4 cur0, cur1;
    if (tm1 < n_int_m1) load_fast(tm1, cur0, cur1); else load_edge(tm1 + 1u, cur0, cur1);

    while (true) {
        // ---- 32 bases -> 2 x 32 bits; the raw bytes are dead after this, so the SAME registers receive the next
        //      tile right away (software pipelining without a second register set or copies) ----
        const uint32_t A0 = pack16_filter(cur0), A1 = pack16_filter(cur1);
        const uint32_t this_tm1 = tm1;
        tm1 += tile_stride;
        bool more = true;
        if (tm1 < n_int_m1) {
            load_fast(tm1, cur0, cur1);
        } else {
            more = tm1 + 1u < a.n_tiles;
            if (more) load_edge(tm1 + 1u, cur0, cur1);
        }
        const uint32_t P0 = __shfl_up_sync(FULL, A0, 1), P1 = __shfl_up_sync(FULL, A1, 1);
        // S = [P0, P1, A0, A1, 0] shifted so that bit 0 is the first index base of test 0
        uint32_t S0, S1, S2;
        if (W0 == 0) {
            S0 = __funnelshift_r(P0, P1, r);
            S1 = __funnelshift_r(P1, A0, r);
            S2 = __funnelshift_r(A0, A1, r);
        } else {
            S0 = __funnelshift_r(P1, A0, r);
            S1 = __funnelshift_r(A0, A1, r);
            S2 = A1 >> r;
        }

        // ---- membership filter: T tests, bit t of hit <-> k-mer ends G*t .. G*t+G-1 of this lane ----
        uint32_t hit = 0;
#pragma unroll
        for (int t = T - 1; t >= 0; t--) {
            const int off = 2 * G * t;
            const int wd = off >> 5, sh = off & 31;
            const uint32_t lo = wd == 0 ? S0 : S1, hi = wd == 0 ? S1 : S2;
            const uint32_t x = sh == 0 ? lo : __funnelshift_r(lo, hi, sh);
            // x >> 3 as a multiply-high: runs on the FMA pipe, the ALU pipe is the one that is full
            uint32_t x3;
            asm("mul.hi.u32 %0, %1, %2;" : "=r"(x3) : "r"(x), "r"(mul29));
            const uint32_t word = *reinterpret_cast<const uint32_t *>(smem_base + (x3 & addr_mask));
            const uint32_t tt = __funnelshift_l(0u, word, x);  // word << (x & 31): tested bit -> MSB
            hit = __funnelshift_l(tt, hit, 1);                 // hit = hit << 1 | msb(tt)
        }
        hit &= keep;

        // ---- queue the groups that passed (ballot + popc, no atomics); screen 32 at a time ----
        uint32_t any = __ballot_sync(FULL, hit != 0);
        if (any) {
            const uint32_t lane_pos = this_tm1 * kTileBytes + lane_off1;
            const uint4 sw = make_uint4(P0, P1, A0, A1);
            do {
                const bool mine = hit != 0;
                if (mine) {
                    const int t = __ffs(hit) - 1;
                    hit &= hit - 1;
                    const uint32_t slot = (uint32_t)n_cand + __popc(any & lt);
                    sts32(cand_pos_sa + slot * 4, lane_pos + (uint32_t)(G * t));
                    sts128(cand_s_sa + slot * 16, sw);
                }
                n_cand += __popc(any);
                if (n_cand >= 32) {
                    __syncwarp();
                    n_cand -= 32;
                    stage1(n_cand, 32);
                    __syncwarp();
                }
                any = __ballot_sync(FULL, hit != 0);
            } while (any);
        }

        if (!more) break;
    }
    __syncwarp();
    if (n_cand) stage1(0, n_cand);
    __syncwarp();
    if (n_ver) stage2(0, n_ver);
}

// ---- host side ---------------------------------------------------------------------------------------------

static uint64_t revcomp_host(uint64_t t, int bits) {
    uint64_t r = 0;
    for (int i = 0; i < bits / 2; i++) {
        uint64_t b = (t >> (2 * i)) & 3;
        r = (r << 2) | (3 - b);
    }
    return r;
}

int build_tables(kssd_ctx *c, const int32_t *shuffled_dim) {
    ScanParams &sp = c->sp;
    const int m = 2 * sp.half_subk;  // bases in the 2s-mer
    // G consecutive windows share m - G + 1 bases; use up to 10 of them as bitmap index
    sp.group = std::max(1, std::min(4, m - 9));
    sp.nb = std::min(10, m - sp.group + 1);
    sp.wbits = 2 * sp.nb - 5;
    const uint64_t dim_size = 1ull << (4 * sp.half_subk);
    std::vector<uint32_t> keys, vals;
    for (uint64_t t = 0; t < dim_size; t++) {
        int32_t v = shuffled_dim[t];
        if (v < sp.dim_end && v >= 0) {  // dim_start = 0 (common.cpp:50)
            keys.push_back((uint32_t)t);
            vals.push_back((uint32_t)v);
        }
    }
    // exact table: open addressing at load <= 1/16, so that the first probe of a non-retained dim_id almost
    // always lands on an empty slot (one 64-bit load decides)
    uint64_t cap = 1024;
    while (cap < 16 * keys.size()) cap <<= 1;
    int lg = 0;
    while ((1ull << lg) < cap) lg++;
    sp.table_mask = (uint32_t)(cap - 1);
    sp.table_shift = 32 - lg;
    std::vector<unsigned long long> tab(cap, ~0ull);
    for (size_t i = 0; i < keys.size(); i++) {
        uint32_t slot = (keys[i] * 0x9E3779B1u) >> sp.table_shift;
        while (tab[slot] != ~0ull) slot = (slot + 1) & sp.table_mask;
        tab[slot] = ((unsigned long long)keys[i] << 32) | vals[i];
    }
    // bitmap: for every retained 2s-mer M and its reverse complement, and every alignment g < G of M against
    // the common span, one bit.  Index base i (i < nb) = base i + G-1-g of M, in the filter encoding
    // (code ^ code>>1: A 0, C 1, T 2, G 3), at index bits [2i, 2i+2); low 5 index bits select the bit of the word.
    const uint64_t n_bits = 1ull << (2 * sp.nb);
    const uint64_t n_words = std::max<uint64_t>(n_bits / 32, 4);
    std::vector<uint32_t> bm(n_words, 0);
    for (size_t i = 0; i < keys.size(); i++) {
        const uint64_t d = keys[i];
        const uint64_t both[2] = {d, revcomp_host(d, 4 * sp.half_subk)};
        for (int o = 0; o < 2; o++) {
            for (int g = 0; g < sp.group; g++) {
                uint32_t idx = 0;
                for (int b = 0; b < sp.nb; b++) {
                    const int mo = b + sp.group - 1 - g;                                   // base offset inside M (0 = first base)
                    const uint32_t code = (uint32_t)(both[o] >> (2 * (m - 1 - mo))) & 3u;  // reference encoding
                    idx |= (code ^ (code >> 1)) << (2 * b);
                }
                bm[idx >> 5] |= 1u << (31 - (idx & 31));
            }
        }
    }
    // second-level filter: one bit per retained (canonical) dim_id at a multiplicative hash of it
    bm.resize(n_words + kL2Words, 0);
    for (size_t i = 0; i < keys.size(); i++) {
        const uint32_t h = (keys[i] * 0x9E3779B1u) >> (32 - kL2Bits);
        bm[n_words + (h >> 5)] |= 1u << (h & 31);
    }
    cudaStream_t st = c->stream;
    KSSD_TRY(c->d_bitmap.reserve(bm.size() * 4, st));
    KSSD_TRY(c->d_tab_keys.reserve(cap * 8, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_bitmap.p, bm.data(), bm.size() * 4, cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_tab_keys.p, tab.data(), cap * 8, cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    return KSSD_OK;
}

template <int G, int W0, bool A32>
static int launch_one(kssd_ctx *c, const ScanArgs &a, size_t smem) {
    const uint64_t attr_bit = 1ull << (G * 4 + W0 * 2 + (A32 ? 1 : 0));
    if (!(c->attr_mask & attr_bit)) {
        KSSD_CUDA(cudaFuncSetAttribute(kssd_scan_kernel<G, W0, A32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
        c->attr_mask |= attr_bit;
    }
    uint64_t ctas = std::min<uint64_t>(((uint64_t)a.n_tiles + kScanWarps - 1) / kScanWarps, (uint64_t)c->sm_count);
    if (ctas == 0) return KSSD_OK;
    if (c->profiling) KSSD_TRY(c->t_scan.begin(c->stream));
    kssd_scan_kernel<G, W0, A32><<<(unsigned)ctas, kScanThreads, smem, c->stream>>>(a);
    KSSD_CUDA(cudaGetLastError());
    if (c->profiling) KSSD_TRY(c->t_scan.end(c->stream));
    c->launches++;
    c->t_scan.launches++;
    c->t_scan.units_a += a.n_bytes;
    return KSSD_OK;
}

template <int G>
static int launch_w0(kssd_ctx *c, const ScanArgs &a, size_t smem, int w0, bool a32) {
    if (a32) return w0 ? launch_one<G, 1, true>(c, a, smem) : launch_one<G, 0, true>(c, a, smem);
    return w0 ? launch_one<G, 1, false>(c, a, smem) : launch_one<G, 0, false>(c, a, smem);
}

int launch_scan(kssd_ctx *c, const uint8_t *d_seq, const uint8_t *d_qual, uint64_t n_bytes,
                const uint64_t *d_seg_off, const uint32_t *d_seg_genome, uint32_t n_seg, uint64_t first_emit,
                uint64_t pos_base, const uint64_t *d_seg_first) {
    if (n_bytes == 0 || n_seg == 0) return KSSD_OK;
    if ((reinterpret_cast<uintptr_t>(d_seq) & 15) || (d_qual && (reinterpret_cast<uintptr_t>(d_qual) & 15))) {
        set_error("sequence / quality device buffers must be 16-byte aligned");
        return KSSD_E_ARG;
    }
    const ScanParams &sp = c->sp;
    if (sp.kmer_size + sp.group - 1 > 4 * kWindowWords) {
        set_error("2*half_k + group - 1 = %d bytes exceeds the exact path's %d-byte window", sp.kmer_size + sp.group - 1, 4 * kWindowWords);
        return KSSD_E_UNSUPPORTED;
    }
    // test t of a lane covers k-mer ends G*t .. G*t+G-1 (lane-local); the bases common to their 2s-mer windows
    // start D = half_k + half_subk - G bases before base G*t.  [P0,P1,A0,A1] starts 32 bases before the lane.
    const int D = sp.half_k + sp.half_subk - sp.group;
    const int bit0 = 64 - 2 * D;
    if (bit0 < 0) { set_error("half_k + half_subk = %d too large for the 32-base context", sp.half_k + sp.half_subk); return KSSD_E_UNSUPPORTED; }
    // quality filter modes: compare is on a signed char (sketch.cpp:280)
    const bool use_qual = d_qual != nullptr && c->has_qual && c->least_qual > -128;
    if (use_qual && c->least_qual > 127) return KSSD_OK;  // no base can pass: nothing is emitted

    ScanArgs a;
    a.bitmap = c->d_bitmap.as<uint32_t>();
    a.table = c->d_tab_keys.as<unsigned long long>();
    a.emit = c->d_emit.as<unsigned long long>();
    a.emit_gen = c->hash64 ? c->d_emit_gen.as<uint32_t>() : nullptr;
    a.seg_off = d_seg_off;
    a.seg_genome = d_seg_genome;
    a.seg_first = d_seg_first;
    a.n_seg = n_seg;
    a.cursor = c->d_cursor.as<unsigned long long>();
    a.gcount = c->d_gcount.as<uint32_t>();
    a.emit_rank = c->d_emit_rank.as<uint32_t>();
    a.n_genomes = c->n_genomes;
    a.emit_cap = c->emit_cap;
    a.sp = sp;
    a.bitmap_words = (uint32_t)std::max<uint64_t>((1ull << (2 * sp.nb)) / 32, 4);
    a.addr_mask = ((1u << sp.wbits) - 1) << 2;
    a.mul29 = 1u << 29;
    a.preshift = bit0 & 31;
    a.least_qual = c->least_qual;
    const int w0 = bit0 >> 5;
    const size_t smem = (size_t)(a.bitmap_words + kL2Words) * 4 + (size_t)kScanWarps * kQueueCap * (sizeof(uint4) + 2 * sizeof(uint32_t));

    // positions inside a launch are 32-bit: cut the buffer into pieces; a piece re-reads `halo` bytes of context
    const uint64_t halo = 32;  // >= 2*half_k - 1, multiple of 16
    for (uint64_t b0 = 0; b0 < n_bytes; b0 += kMaxPiece) {
        const uint64_t b1 = std::min<uint64_t>(n_bytes, b0 + kMaxPiece);
        const uint64_t from = b0 ? b0 - halo : 0;
        a.seq = d_seq + from;
        a.pos_base = pos_base + from;
        const bool a32 = (reinterpret_cast<uintptr_t>(a.seq) & 31) == 0;
        a.qual = use_qual ? d_qual + from : nullptr;
        a.n_bytes = (uint32_t)(b1 - from);
        // k-mers ending inside the halo belong to the previous piece; the caller's first_emit applies to byte 0
        a.first_emit = (uint32_t)(b0 ? std::max<uint64_t>(halo, first_emit > from ? first_emit - from : 0)
                                     : std::min<uint64_t>(first_emit, 0xFFFFFFFFull));
        a.n_tiles = (a.n_bytes + kTileBytes - 1) / kTileBytes;
        a.n_interior = a.n_bytes / kTileBytes;
        int rc;
        switch (sp.group) {
        case 1: rc = launch_w0<1>(c, a, smem, w0, a32); break;
        case 2: rc = launch_w0<2>(c, a, smem, w0, a32); break;
        case 3: rc = launch_w0<3>(c, a, smem, w0, a32); break;
        case 4: rc = launch_w0<4>(c, a, smem, w0, a32); break;
        default: set_error("bad filter group %d", sp.group); return KSSD_E_ARG;
        }
        KSSD_TRY(rc);
    }
    return KSSD_OK;
}

}  // namespace kssd
