// Fused Kssd sketch scan for sm_100a: ASCII load -> 2-bit pack + validity -> membership pre-filter in shared
// memory -> exact canonical k-mer / retained-subset lookup / hash emit for the (rare) survivors.
//
// Replaces the per-base loop of the reference (src/sketch.cpp:487-532, same arithmetic in :196-231 and the FASTQ
// variants :273-307, :779-825).  Not a translation: the reference rolls two 64-bit tuples base by base and
// probes a hash map for EVERY k-mer; here
//   * a warp owns a 512-byte tile: 32 lanes x one 128-bit coalesced load (16 bases per lane), lanes 0-1 re-read
//     the last 32 bases of the previous tile as halo so no state is carried between tiles;
//   * the 16 bases of a lane are decoded 4 at a time with SWAR arithmetic (2-bit code + exact BaseMap validity,
//     common.h:27-37) and packed into one 32-bit word; the halo comes from two warp shuffles;
//   * membership of the 2s-mer is tested WITHOUT deciding the canonical strand: the shared-memory bitmap holds
//     the retained set united with its reverse complement, indexed by 10 bases of the forward 2s-mer, so the
//     per-position cost is one funnel shift, one mask, one LDS and two shifts (no 64-bit arithmetic);
//   * only positions that pass the bitmap (<1%) run the reference's exact 64-bit arithmetic: canonical min of
//     tuple / reverse complement, dim_id, exact table lookup, hash composition, append to the emit buffer.
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
    int ysh;             // 2 + wbits
    int least_qual;
    ScanParams sp;
};

static constexpr int kTileLanes = 30;  // lanes 2..31 own new bases, lanes 0..1 are halo
static constexpr int kScanThreads = 1024;

// ---- SWAR decode of 4 ASCII bytes -------------------------------------------------------------------------
// code: A/a 0, C/c 1, G/g 2, T/t 3 in bits [0,2) of each byte;  err: non-zero byte <=> byte not in "ACGTacgt".
// Bits b2,b1 of the four letters are the Gray code of 0..3; the other six bits are a function of them:
// b7 b6 b5 b3 = 0 1 x 0, b4 = b2 & ~b1 (only T), b0 = ~b4.
__device__ __forceinline__ void decode4(uint32_t x, uint32_t &code, uint32_t &err) {
    uint32_t s1 = x >> 1, s2 = x >> 2;
    code = (s1 ^ s2) & 0x03030303u;
    uint32_t z = s2 & ~s1 & 0x01010101u;
    uint32_t zz = z * 0x11u;
    uint32_t t = (x & 0xD9D9D9D9u) ^ 0x41414141u;
    err = t ^ zz;
}

// 4 codes (one per byte, first base in the lowest byte) -> 8 bits in the TOP byte, first base most significant.
__device__ __forceinline__ uint32_t pack4_top(uint32_t code) { return code * 0x40100401u; }

// per-byte non-zero flags of e -> 4 bits (bit i = byte i non-zero)
__device__ __forceinline__ uint32_t nz4(uint32_t e) {
    uint32_t f = (((e & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e) & 0x80808080u;
    return ((f >> 7) * 0x01020408u) >> 24 & 0xFu;
}

__device__ __forceinline__ uint4 load_chunk(const uint8_t *base, long long c, uint64_t n_bytes) {
    uint4 v = make_uint4(0, 0, 0, 0);  // 0x00 bytes are invalid bases
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

__device__ __forceinline__ uint64_t revcomp_bits(uint64_t t, int bits) {
    uint64_t r = __brevll(t);
    r = ((r >> 1) & 0x5555555555555555ull) | ((r & 0x5555555555555555ull) << 1);
    r = ~r;
    return r >> (64 - bits);
}

// Exact path for one candidate position: the reference's arithmetic, sketch.cpp:508-524.
__device__ __noinline__ void exact_emit(const ScanArgs &a, uint32_t W, uint32_t W1, uint32_t W2, int i, uint64_t pos) {
    const ScanParams &sp = a.sp;
    int sh = 2 * (15 - i);
    uint64_t lo64 = ((uint64_t)W1 << 32) | W;
    uint64_t tuple = lo64 >> sh;
    if (sh) tuple |= (uint64_t)W2 << (64 - sh);
    tuple &= sp.tupmask;
    uint64_t rvs = revcomp_bits(tuple, 4 * sp.half_k);
    uint64_t uni = tuple < rvs ? tuple : rvs;
    uint32_t dim_id = (uint32_t)((uni & sp.domask) >> (2 * sp.hoc));
    uint32_t slot = (dim_id * 0x9E3779B1u) >> sp.table_shift;
    uint32_t pf;
    while (true) {
        uint32_t k = __ldg(a.tab_keys + slot);
        if (k == dim_id) { pf = __ldg(a.tab_vals + slot); break; }
        if (k == 0xFFFFFFFFu) return;
        slot = (slot + 1) & sp.table_mask;
    }
    if (pos < a.first_emit) return;
    uint64_t dr = (((uni & sp.undomask0) | ((uni & sp.undomask1) << (4 * sp.half_k - 4 * sp.hoc))) >> (4 * sp.drlevel)) | pf;
    // owning segment: largest s with seg_off[s] <= pos
    pos += a.pos_base;
    uint32_t lo = 0, hi = a.n_seg;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (a.seg_off[mid] <= pos) lo = mid; else hi = mid;
    }
    uint32_t g = a.seg_genome[lo];
    unsigned long long idx = atomicAdd(a.cursor, 1ull);
    if (idx < a.emit_cap) a.emit[idx] = ((unsigned long long)g << 32) | (uint32_t)dr;
}

template <int HOC, bool QUAL>
__global__ void __launch_bounds__(kScanThreads, 1) kssd_scan_kernel(const ScanArgs a) {
    extern __shared__ uint32_t s_bitmap[];
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(a.bitmap);
        uint4 *dst = reinterpret_cast<uint4 *>(s_bitmap);
        for (uint32_t i = threadIdx.x; i < a.bitmap_words / 4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const uint64_t warps_total = (uint64_t)gridDim.x * (blockDim.x >> 5);
    // interleave warps of different CTAs over consecutive tiles so that all SMs stream the same HBM region
    uint64_t tile = (uint64_t)(threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    const uint32_t addr_mask = a.addr_mask;
    const int ysh = a.ysh;
    const int L = a.sp.kmer_size;
    const char *smem_base = reinterpret_cast<const char *>(s_bitmap);
    uint32_t lq4 = 0;
    if (QUAL) lq4 = (uint32_t)(a.least_qual & 0xFF) * 0x01010101u;

    if (tile >= a.n_tiles) return;
    long long c = (long long)tile * kTileLanes - 2 + lane;
    uint4 cur = load_chunk(a.seq, c, a.n_bytes);
    uint4 curq = make_uint4(0, 0, 0, 0);
    if (QUAL) curq = load_chunk(a.qual, c, a.n_bytes);

    while (true) {
        uint64_t next_tile = tile + warps_total;
        long long cn = (long long)next_tile * kTileLanes - 2 + lane;
        uint4 nxt = make_uint4(0, 0, 0, 0), nxtq = make_uint4(0, 0, 0, 0);
        if (next_tile < a.n_tiles) {
            nxt = load_chunk(a.seq, cn, a.n_bytes);
            if (QUAL) nxtq = load_chunk(a.qual, cn, a.n_bytes);
        }

        // ---- decode 16 bases ----
        uint32_t c0, c1, c2, c3, e0, e1, e2, e3;
        decode4(cur.x, c0, e0);
        decode4(cur.y, c1, e1);
        decode4(cur.z, c2, e2);
        decode4(cur.w, c3, e3);
        if (QUAL) {  // valid only if (signed char)quality >= least_qual (sketch.cpp:280)
            e0 |= ~__vcmpges4(curq.x, lq4);
            e1 |= ~__vcmpges4(curq.y, lq4);
            e2 |= ~__vcmpges4(curq.z, lq4);
            e3 |= ~__vcmpges4(curq.w, lq4);
        }
        uint32_t p0 = pack4_top(c0), p1 = pack4_top(c1), p2 = pack4_top(c2), p3 = pack4_top(c3);
        uint32_t t01 = __byte_perm(p0, p1, 0x3700);
        uint32_t t23 = __byte_perm(p2, p3, 0x0037);
        uint32_t W = __byte_perm(t01, t23, 0x3254);
        uint32_t eall = e0 | e1 | e2 | e3;

        uint32_t W1 = __shfl_up_sync(FULL, W, 1);
        uint32_t W2 = __shfl_up_sync(FULL, W, 2);

        uint32_t okmask;
        if (__all_sync(FULL, eall == 0)) {
            okmask = lane >= 2 ? 0xFFFFu : 0u;
        } else {
            uint32_t V = (~(nz4(e0) | (nz4(e1) << 4) | (nz4(e2) << 8) | (nz4(e3) << 12))) & 0xFFFFu;
            uint32_t V1 = __shfl_up_sync(FULL, V, 1);
            uint32_t V2 = __shfl_up_sync(FULL, V, 2);
            // bit b <-> base (b - 32) relative to this lane's chunk
            uint64_t vm = ((uint64_t)V << 32) | ((uint64_t)V1 << 16) | V2;
            uint64_t r = vm;
            int len = 1;
            while (2 * len <= L) { r &= r << len; len *= 2; }
            if (len < L) r &= r << (L - len);
            okmask = lane >= 2 ? (uint32_t)(r >> 32) & 0xFFFFu : 0u;
        }

        // ---- membership pre-filter: 16 positions, bit i of hit <-> k-mer ending at base i ----
        uint32_t hit = 0;
#pragma unroll
        for (int i = 15; i >= 0; i--) {
            const int sh = 2 * (15 - i + HOC);
            uint32_t x;
            if (sh == 0) x = W;
            else if (sh < 32) x = __funnelshift_r(W, W1, sh);
            else if (sh == 32) x = W1;
            else x = __funnelshift_r(W1, W2, sh - 32);
            uint32_t word = *reinterpret_cast<const uint32_t *>(smem_base + (x & addr_mask));
            uint32_t y = x >> ysh;
            uint32_t t = __funnelshift_l(0u, word, y);  // word << (y & 31): tested bit -> MSB
            hit = __funnelshift_l(t, hit, 1);           // hit = hit << 1 | msb(t)
        }
        uint32_t cand = hit & okmask;
        if (cand) {
            uint64_t chunk_pos = (uint64_t)c * 16;
            do {
                int i = __ffs(cand) - 1;
                cand &= cand - 1;
                exact_emit(a, W, W1, W2, i, chunk_pos + i);
            } while (cand);
        }

        if (next_tile >= a.n_tiles) break;
        tile = next_tile;
        c = cn;
        cur = nxt;
        if (QUAL) curq = nxtq;
    }
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
    sp.nb = std::min(10, m - 1);
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
    // symmetrised bitmap over nb bases of the 2s-mer (skipping its last base)
    const uint64_t n_bits = 1ull << (2 * sp.nb);
    const uint64_t n_words = std::max<uint64_t>(n_bits / 32, 4);
    std::vector<uint32_t> bm(n_words, 0);
    const uint32_t idx_mask = (uint32_t)(n_bits - 1);
    const uint32_t wmask = (1u << sp.wbits) - 1;
    for (size_t i = 0; i < keys.size(); i++) {
        uint64_t d = keys[i];
        uint64_t both[2] = {d, revcomp_host(d, 4 * sp.half_subk)};
        for (int o = 0; o < 2; o++) {
            uint32_t I = (uint32_t)(both[o] >> 2) & idx_mask;
            uint32_t w = I & wmask, sel = I >> sp.wbits;
            bm[w] |= 1u << (31 - sel);
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

template <int HOC, bool QUAL>
static int launch_one(kssd_ctx *c, const ScanArgs &a, size_t smem) {
    static bool attr_done = false;  // one flag per template instantiation
    if (!attr_done) {
        KSSD_CUDA(cudaFuncSetAttribute(kssd_scan_kernel<HOC, QUAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done = true;
    }
    uint64_t warps_per_cta = kScanThreads / 32;
    uint64_t ctas = std::min<uint64_t>((a.n_tiles + warps_per_cta - 1) / warps_per_cta, (uint64_t)c->sm_count);
    if (ctas == 0) return KSSD_OK;
    if (c->profiling) KSSD_TRY(c->t_scan.begin(c->stream));
    kssd_scan_kernel<HOC, QUAL><<<(unsigned)ctas, kScanThreads, smem, c->stream>>>(a);
    KSSD_CUDA(cudaGetLastError());
    if (c->profiling) KSSD_TRY(c->t_scan.end(c->stream));
    c->launches++;
    c->t_scan.launches++;
    c->t_scan.units_a += a.n_bytes;
    return KSSD_OK;
}

template <bool QUAL>
static int launch_hoc(kssd_ctx *c, const ScanArgs &a, size_t smem) {
    switch (a.sp.hoc) {
    case 0: return launch_one<0, QUAL>(c, a, smem);
    case 1: return launch_one<1, QUAL>(c, a, smem);
    case 2: return launch_one<2, QUAL>(c, a, smem);
    case 3: return launch_one<3, QUAL>(c, a, smem);
    case 4: return launch_one<4, QUAL>(c, a, smem);
    case 5: return launch_one<5, QUAL>(c, a, smem);
    case 6: return launch_one<6, QUAL>(c, a, smem);
    case 7: return launch_one<7, QUAL>(c, a, smem);
    case 8: return launch_one<8, QUAL>(c, a, smem);
    default:
        set_error("half_k - half_subk = %d outside the supported range 0..8", a.sp.hoc);
        return KSSD_E_UNSUPPORTED;
    }
}

int launch_scan(kssd_ctx *c, const uint8_t *d_seq, const uint8_t *d_qual, uint64_t n_bytes,
                const uint64_t *d_seg_off, const uint32_t *d_seg_genome, uint32_t n_seg, uint64_t first_emit,
                uint64_t pos_base) {
    if (n_bytes == 0 || n_seg == 0) return KSSD_OK;
    if ((reinterpret_cast<uintptr_t>(d_seq) & 15) || (d_qual && (reinterpret_cast<uintptr_t>(d_qual) & 15))) {
        set_error("sequence / quality device buffers must be 16-byte aligned");
        return KSSD_E_ARG;
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
    a.sp = c->sp;
    uint64_t n_chunks = (n_bytes + 15) / 16;
    a.n_tiles = (n_chunks + kTileLanes - 1) / kTileLanes + 0;
    // the last tile must also cover chunk indices up to n_chunks-1: tile t covers [30t-2, 30t+30)
    a.bitmap_words = (uint32_t)std::max<uint64_t>((1ull << (2 * c->sp.nb)) / 32, 4);
    a.addr_mask = ((1u << c->sp.wbits) - 1) << 2;
    a.ysh = 2 + c->sp.wbits;
    size_t smem = (size_t)a.bitmap_words * 4;
    // quality filter modes: compare is on a signed char (sketch.cpp:280)
    bool use_qual = d_qual != nullptr && c->has_qual && c->least_qual > -128;
    if (use_qual && c->least_qual > 127) return KSSD_OK;  // no base can pass: nothing is emitted
    a.qual = use_qual ? d_qual : nullptr;
    a.least_qual = c->least_qual;
    return use_qual ? launch_hoc<true>(c, a, smem) : launch_hoc<false>(c, a, smem);
}

}  // namespace kssd
