// Fused Kssd sketch scan for sm_100a.
//
// Replaces the per-base loop of the reference (src/sketch.cpp:487-532, same arithmetic in :196-231 and the FASTQ
// variants :273-307, :779-825).  Not a translation: the reference rolls two 64-bit tuples base by base and probes a
// hash map for EVERY k-mer.  Here the work is split into a streaming filter that touches every base with ~3
// integer instructions and an exact path that only sees the <1% of positions the filter lets through:
//
//   filter (every base)
//   * a lane owns 32 bases: two 128-bit coalesced loads; lane 0 of the warp re-reads the 32 bases before the
//     tile as halo, so no state is carried between tiles (tile = 31 x 32 = 992 new bytes per warp);
//   * 4 ASCII bytes -> 8 bits with one AND and one IMAD ((x & 0x06060606) * M gathers bits 1-2 of each byte:
//     A/a 0, C/c 1, T/t 2, G/g 3).  Validity (N, separators, quality) is NOT decided here;
//   * membership of the 2s-mer is tested WITHOUT deciding the canonical strand and for G consecutive k-mer end
//     positions at once: the shared-memory bitmap holds, for every retained 2s-mer and its reverse complement,
//     one bit per alignment g < G of the 2s-mer against the bases that all G windows have in common (up to 10
//     bases = 20 index bits).  One funnel shift, one shift+mask, one LDS, two shifts decide G positions;
//   * groups that pass are appended to a per-warp shared-memory queue (ballot + popc, no atomics).
//   exact path (queued groups, 32 at a time, one per lane)
//   * re-reads the 2k+G-1 raw bytes of the group (three 128-bit loads), decodes them with the reference's
//     BaseMap semantics (common.h:27-37), checks validity of each window, then runs the reference's 64-bit
//     arithmetic: canonical min of tuple / reverse complement, dim_id, exact retained-subset lookup, hash
//     composition (sketch.cpp:508-524), quality check (sketch.cpp:280) only for survivors, append to the emit buffer.
#include <algorithm>
#include <cstring>

#include "common.cuh"

namespace kssd {

struct ScanArgs {
    const uint8_t *seq;
    const uint8_t *qual;
    uint64_t n_bytes;
    const uint64_t *seg_off;
    const uint32_t *seg_genome;
    uint32_t n_seg;
    uint64_t first_emit;
    uint64_t pos_base;  // position of seq[0] in the coordinates of seg_off (chunked host submits)
    const uint32_t *bitmap;
    const uint32_t *tab_keys;
    const uint32_t *tab_vals;
    unsigned long long *emit;
    unsigned long long *cursor;
    uint64_t emit_cap;
    uint64_t n_tiles;
    uint32_t bitmap_words;
    uint32_t addr_mask;  // ((1<<wbits)-1) << 2
    int preshift;        // bit offset (mod 32) of the first index base of test 0 inside [P0,P1,A0,A1]
    int least_qual;
    ScanParams sp;
};

static constexpr int kScanThreads = 1024;
static constexpr int kLaneBases = 32;                  // bases owned by a lane
static constexpr int kTileLanes = 31;                  // lanes 1..31 own new bases, lane 0 is halo
static constexpr int kQueueCap = 64;                   // per-warp candidate queue entries
static constexpr int kWindowWords = 7;                 // 28 bytes >= 2k + G - 1 (k <= 12, G <= 4)

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

// 16-byte chunk c of the buffer; chunks outside [0, n_bytes) read as 0x00 (an invalid base)
__device__ __forceinline__ uint4 load_chunk(const uint8_t *base, long long c, uint64_t n_bytes) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (c >= 0) {
        uint64_t b = (uint64_t)c * 16;
        if (b + 16 <= n_bytes) {
            v = __ldg(reinterpret_cast<const uint4 *>(base + b));
        } else if (b < n_bytes) {
            uint32_t w[4] = {0, 0, 0, 0};
            for (int j = 0; j < 16 && b + j < n_bytes; j++) w[j >> 2] |= (uint32_t)base[b + j] << (8 * (j & 3));
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    return v;
}

// 28 raw bytes starting at byte `start` (any alignment, may be negative): three aligned 128-bit loads and a
// select network instead of byte loads.
__device__ __forceinline__ void load_window(const uint8_t *base, long long start, uint64_t n_bytes, uint32_t (&w)[kWindowWords]) {
    const long long c0 = start >> 4;  // floor
    const uint4 v0 = load_chunk(base, c0, n_bytes), v1 = load_chunk(base, c0 + 1, n_bytes), v2 = load_chunk(base, c0 + 2, n_bytes);
    const uint32_t S[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
    const int o = (int)(start - c0 * 16);
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

__device__ __forceinline__ uint64_t revcomp_bits(uint64_t t, int bits) {
    uint64_t r = __brevll(t);
    r = ((r >> 1) & 0x5555555555555555ull) | ((r & 0x5555555555555555ull) << 1);
    r = ~r;
    return r >> (64 - bits);
}

// Exact path for one queued group: k-mer end positions p0 .. p0+cnt-1 (buffer coordinates).
template <int G>
__device__ __noinline__ void exact_group(const ScanArgs &a, uint64_t p0) {
    const ScanParams &sp = a.sp;
    const int L = sp.kmer_size;
    int cnt = kLaneBases - (int)(p0 & (kLaneBases - 1));
    if (cnt > G) cnt = G;
    const long long start = (long long)p0 - (L - 1);
    uint32_t w[kWindowWords];
    load_window(a.seq, start, a.n_bytes, w);
    uint64_t C = 0;   // 2-bit codes of the 28 bytes, first byte most significant
    uint32_t V = 0;   // bit i = byte i is a valid base
#pragma unroll
    for (int i = 0; i < kWindowWords; i++) {
        uint32_t code, err;
        decode4(w[i], code, err);
        C = (C << 8) | pack4_msb(code);
        V |= (~nz4(err) & 0xFu) << (4 * i);
    }
    const uint32_t lmask = (L >= 32) ? 0xFFFFFFFFu : ((1u << L) - 1u);
    bool have_qual = false;
    uint32_t QV = 0;
    for (int j = 0; j < cnt; j++) {
        if (((V >> j) & lmask) != lmask) continue;  // an invalid base inside the window (sketch.cpp:494-505)
        const uint64_t pos = p0 + j;
        if (pos < a.first_emit) continue;            // context of a continued record
        const uint64_t tuple = (C >> (2 * (4 * kWindowWords - L - j))) & sp.tupmask;
        const uint64_t rvs = revcomp_bits(tuple, 2 * L);
        const uint64_t uni = tuple < rvs ? tuple : rvs;
        const uint32_t dim_id = (uint32_t)((uni & sp.domask) >> (2 * sp.hoc));
        uint32_t slot = (dim_id * 0x9E3779B1u) >> sp.table_shift;
        uint32_t pf = 0;
        bool found = false;
        while (true) {
            uint32_t k = __ldg(a.tab_keys + slot);
            if (k == dim_id) { pf = __ldg(a.tab_vals + slot); found = true; break; }
            if (k == 0xFFFFFFFFu) break;
            slot = (slot + 1) & sp.table_mask;
        }
        if (!found) continue;
        if (a.qual) {  // valid only if (signed char)quality >= least_qual (sketch.cpp:280), checked for survivors only
            if (!have_qual) {
                uint32_t qw[kWindowWords];
                load_window(a.qual, start, a.n_bytes, qw);
                const uint32_t lq4 = (uint32_t)(a.least_qual & 0xFF) * 0x01010101u;
#pragma unroll
                for (int i = 0; i < kWindowWords; i++) QV |= nz4(__vcmpges4(qw[i], lq4)) << (4 * i);
                have_qual = true;
            }
            if (((QV >> j) & lmask) != lmask) continue;
        }
        const uint64_t dr = (((uni & sp.undomask0) | ((uni & sp.undomask1) << (4 * sp.half_k - 4 * sp.hoc))) >> (4 * sp.drlevel)) | pf;
        // owning segment: largest s with seg_off[s] <= pos
        const uint64_t gpos = pos + a.pos_base;
        uint32_t lo = 0, hi = a.n_seg;
        while (hi - lo > 1) {
            uint32_t mid = (lo + hi) >> 1;
            if (a.seg_off[mid] <= gpos) lo = mid; else hi = mid;
        }
        const uint32_t g = a.seg_genome[lo];
        const unsigned long long idx = atomicAdd(a.cursor, 1ull);
        if (idx < a.emit_cap) a.emit[idx] = ((unsigned long long)g << 32) | (uint32_t)dr;
    }
}

// G: k-mer end positions decided per bitmap test.  W0: word offset (0/1) of the first index base of test 0.
template <int G, int W0>
__global__ void __launch_bounds__(kScanThreads, 1) kssd_scan_kernel(const __grid_constant__ ScanArgs a) {
    extern __shared__ uint32_t s_mem[];
    uint32_t *s_bitmap = s_mem;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(a.bitmap);
        uint4 *dst = reinterpret_cast<uint4 *>(s_bitmap);
        for (uint32_t i = threadIdx.x; i < a.bitmap_words / 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    constexpr unsigned FULL = 0xFFFFFFFFu;
    constexpr int T = (kLaneBases + G - 1) / G;  // tests per lane
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    unsigned long long *queue = reinterpret_cast<unsigned long long *>(s_mem + a.bitmap_words) + warp * kQueueCap;
    const uint64_t warps_total = (uint64_t)gridDim.x * (blockDim.x >> 5);
    // interleave warps of different CTAs over consecutive tiles so that all SMs stream the same HBM region
    uint64_t tile = (uint64_t)warp * gridDim.x + blockIdx.x;
    if (tile >= a.n_tiles) return;
    const uint32_t addr_mask = a.addr_mask;
    const int r = a.preshift;
    const char *smem_base = reinterpret_cast<const char *>(s_bitmap);
    const uint32_t lt = (1u << lane) - 1u;
    int qcount = 0;

    long long c = ((long long)tile * kTileLanes + lane - 1) * 2;  // first 16-byte chunk of this lane
    uint4 cur0 = load_chunk(a.seq, c, a.n_bytes), cur1 = load_chunk(a.seq, c + 1, a.n_bytes);

    while (true) {
        const uint64_t next_tile = tile + warps_total;
        const long long cn = ((long long)next_tile * kTileLanes + lane - 1) * 2;
        uint4 nxt0 = make_uint4(0, 0, 0, 0), nxt1 = nxt0;
        if (next_tile < a.n_tiles) {
            nxt0 = load_chunk(a.seq, cn, a.n_bytes);
            nxt1 = load_chunk(a.seq, cn + 1, a.n_bytes);
        }

        // ---- 32 bases -> 2 x 32 bits, previous lane's words as context ----
        const uint32_t A0 = pack16_filter(cur0), A1 = pack16_filter(cur1);
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
            const uint32_t word = *reinterpret_cast<const uint32_t *>(smem_base + ((x >> 3) & addr_mask));
            const uint32_t tt = __funnelshift_l(0u, word, x);  // word << (x & 31): tested bit -> MSB
            hit = __funnelshift_l(tt, hit, 1);                 // hit = hit << 1 | msb(tt)
        }
        if (lane == 0) hit = 0;  // halo lane: its positions belong to lane 31 of the previous tile

        // ---- queue the groups that passed; drain 32 at a time through the exact path ----
        const uint64_t lane_pos = (uint64_t)c * 16;
        while (true) {
            const uint32_t any = __ballot_sync(FULL, hit != 0);
            if (any == 0) break;
            if (hit) {
                const int t = __ffs(hit) - 1;
                hit &= hit - 1;
                queue[qcount + __popc(any & lt)] = lane_pos + (uint64_t)(G * t);
            }
            qcount += __popc(any);
            __syncwarp();
            if (qcount >= 32) {
                qcount -= 32;
                const uint64_t p0 = queue[qcount + lane];
                exact_group<G>(a, p0);
                __syncwarp();
            }
        }

        if (next_tile >= a.n_tiles) break;
        tile = next_tile;
        c = cn;
        cur0 = nxt0;
        cur1 = nxt1;
    }
    if (lane < qcount) exact_group<G>(a, queue[lane]);
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
    // exact table, open addressing
    uint64_t cap = 1024;
    while (cap < 2 * keys.size()) cap <<= 1;
    int lg = 0;
    while ((1ull << lg) < cap) lg++;
    sp.table_mask = (uint32_t)(cap - 1);
    sp.table_shift = 32 - lg;
    std::vector<uint32_t> tk(cap, 0xFFFFFFFFu), tv(cap, 0);
    for (size_t i = 0; i < keys.size(); i++) {
        uint32_t slot = (keys[i] * 0x9E3779B1u) >> sp.table_shift;
        while (tk[slot] != 0xFFFFFFFFu) slot = (slot + 1) & sp.table_mask;
        tk[slot] = keys[i];
        tv[slot] = vals[i];
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
                    const int mo = b + sp.group - 1 - g;                       // base offset inside M (0 = first base)
                    const uint32_t code = (uint32_t)(both[o] >> (2 * (m - 1 - mo))) & 3u;  // reference encoding
                    idx |= (code ^ (code >> 1)) << (2 * b);
                }
                bm[idx >> 5] |= 1u << (31 - (idx & 31));
            }
        }
    }
    cudaStream_t st = c->stream;
    KSSD_TRY(c->d_bitmap.reserve(n_words * 4, st));
    KSSD_TRY(c->d_tab_keys.reserve(cap * 4, st));
    KSSD_TRY(c->d_tab_vals.reserve(cap * 4, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_bitmap.p, bm.data(), n_words * 4, cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_tab_keys.p, tk.data(), cap * 4, cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_tab_vals.p, tv.data(), cap * 4, cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    return KSSD_OK;
}

template <int G, int W0>
static int launch_one(kssd_ctx *c, const ScanArgs &a, size_t smem) {
    static bool attr_done = false;  // one flag per template instantiation
    if (!attr_done) {
        KSSD_CUDA(cudaFuncSetAttribute(kssd_scan_kernel<G, W0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done = true;
    }
    uint64_t warps_per_cta = kScanThreads / 32;
    uint64_t ctas = std::min<uint64_t>((a.n_tiles + warps_per_cta - 1) / warps_per_cta, (uint64_t)c->sm_count);
    if (ctas == 0) return KSSD_OK;
    if (c->profiling) KSSD_TRY(c->t_scan.begin(c->stream));
    kssd_scan_kernel<G, W0><<<(unsigned)ctas, kScanThreads, smem, c->stream>>>(a);
    KSSD_CUDA(cudaGetLastError());
    if (c->profiling) KSSD_TRY(c->t_scan.end(c->stream));
    c->launches++;
    c->t_scan.launches++;
    c->t_scan.units_a += a.n_bytes;
    return KSSD_OK;
}

template <int G>
static int launch_w0(kssd_ctx *c, const ScanArgs &a, size_t smem, int w0) {
    return w0 ? launch_one<G, 1>(c, a, smem) : launch_one<G, 0>(c, a, smem);
}

int launch_scan(kssd_ctx *c, const uint8_t *d_seq, const uint8_t *d_qual, uint64_t n_bytes,
                const uint64_t *d_seg_off, const uint32_t *d_seg_genome, uint32_t n_seg, uint64_t first_emit,
                uint64_t pos_base) {
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
    ScanArgs a;
    a.seq = d_seq;
    a.n_bytes = n_bytes;
    a.seg_off = d_seg_off;
    a.seg_genome = d_seg_genome;
    a.n_seg = n_seg;
    a.first_emit = first_emit;
    a.pos_base = pos_base;
    a.bitmap = c->d_bitmap.as<uint32_t>();
    a.tab_keys = c->d_tab_keys.as<uint32_t>();
    a.tab_vals = c->d_tab_vals.as<uint32_t>();
    a.emit = c->d_emit.as<unsigned long long>();
    a.cursor = c->d_cursor.as<unsigned long long>();
    a.emit_cap = c->emit_cap;
    a.sp = sp;
    const uint64_t n_lane_chunks = (n_bytes + kLaneBases - 1) / kLaneBases;
    a.n_tiles = (n_lane_chunks + kTileLanes - 1) / kTileLanes;
    a.bitmap_words = (uint32_t)std::max<uint64_t>((1ull << (2 * sp.nb)) / 32, 4);
    a.addr_mask = ((1u << sp.wbits) - 1) << 2;
    // test t of a lane covers k-mer ends G*t .. G*t+G-1 (lane-local); the bases common to their 2s-mer windows
    // start D = half_k + half_subk - G bases before base G*t.  [P0,P1,A0,A1] starts 32 bases before the lane.
    const int D = sp.half_k + sp.half_subk - sp.group;
    const int bit0 = 64 - 2 * D;
    if (bit0 < 0) { set_error("half_k + half_subk = %d too large for the 32-base context", sp.half_k + sp.half_subk); return KSSD_E_UNSUPPORTED; }
    a.preshift = bit0 & 31;
    const int w0 = bit0 >> 5;
    size_t smem = (size_t)a.bitmap_words * 4 + (size_t)(kScanThreads / 32) * kQueueCap * sizeof(unsigned long long);
    // quality filter modes: compare is on a signed char (sketch.cpp:280)
    bool use_qual = d_qual != nullptr && c->has_qual && c->least_qual > -128;
    if (use_qual && c->least_qual > 127) return KSSD_OK;  // no base can pass: nothing is emitted
    a.qual = use_qual ? d_qual : nullptr;
    a.least_qual = c->least_qual;
    switch (sp.group) {
    case 1: return launch_w0<1>(c, a, smem, w0);
    case 2: return launch_w0<2>(c, a, smem, w0);
    case 3: return launch_w0<3>(c, a, smem, w0);
    case 4: return launch_w0<4>(c, a, smem, w0);
    default: set_error("bad filter group %d", sp.group); return KSSD_E_ARG;
    }
}

}  // namespace kssd
