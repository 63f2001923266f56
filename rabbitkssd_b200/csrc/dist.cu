// alldist / dist on the device: posting ranges per query row, shared-memory intersection counters, distance +
// threshold (+ top-N) epilogue, compacted (row, col, common) output.
//
// Replaces the per-row loops of index_tridist (src/dist.cpp:174-258) and index_dist (src/dist.cpp:560-692).  The
// reference keeps one dense int counter row per thread in DRAM, walks the posting list of every query hash into it
// and rescans all N cells per query with a libm log() per non-trivial cell.  Here
//
//  * alldist (triangle): NO per-hash lookups at all.  The index is walked once, list by list (row_work_kernel): the
//    posting at position p of a list belongs to genome g = dict[p], and what row g has to stream for that hash is
//    exactly the rest of the list, (p, end) -- ids ascend inside a list (src/sketch.cpp:979-985), so these are the
//    columns > g that src/dist.cpp:207 keeps.  Every row thus gets a work list of (begin, end) posting ranges; lists
//    of length one (a hash no other genome has) produce nothing.
//  * dist (rectangle): the lookup of a query hash (directory + binary search over the distinct hashes, L2 resident)
//    is fused into the counting kernel -- no per-hash ranges written to HBM and read back.
//  * dist_tile_kernel: a CTA owns TQ query rows x TC reference columns of 16-bit (32-bit when a sketch exceeds
//    65535 hashes) counters in shared memory.  The rows of a tile are processed CONCURRENTLY: the work of all rows
//    is cut into chunks of 32 posting ranges and the 16 warps take chunks round-robin, so a CTA has 16 independent
//    range -> postings -> shared-atomic chains in flight.  Short ranges are streamed one per lane (four loads in
//    flight), long ones by the whole warp (coalesced).  An increment that finds its counter at zero records the cell
//    in a shared-memory list ("first touch"), so the epilogue visits the touched cells only -- O(postings), not
//    O(cells): distance + threshold in fp64 for those, optional per-row top-N selection (bitonic sort of the
//    touched list), warp-aggregated append of the passing cells.
//  The device test is conservative (superset); the host recomputes with the reference's exact expressions.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "smem_sort.cuh"

namespace kssd {

static constexpr uint32_t kListCap = 4096;    // first-touch list entries per tile (u16 cell ids: a tile has < 65536 cells)
static constexpr uint32_t kLongList = 32;     // posting ranges longer than this are streamed by the whole warp
static constexpr int kDistThreads = 512;

// ---- alldist: per-row posting ranges from one walk over the index --------------------------------------------
// work[q_off[g] + i], i < fill[g - row_begin]: the ranges row g streams.  A row has at most |sketch(g)| ranges (one
// per hash), so the slots of row g start at q_off[g]; more than that means the query sketches are not the indexed
// ones (flags[0]).
// A warp takes 32 lists at a time and spreads their postings -- not the lists -- over its lanes: every posting but the last
// of a list is one emit, the emits of the 32 lists are numbered by a warp prefix sum and lane l takes emits l, l + 32, ...
// (the list of an emit is found by a 5-step search over the prefix sums held by the lanes).  All lanes thus run independent
// dict -> q_off -> atomic -> store chains, whatever the mix of singleton, short and thousand-member lists.
__global__ void __launch_bounds__(256) row_work_kernel(const uint64_t *__restrict__ uniq_off, uint64_t n_uniq, const uint32_t *__restrict__ dict,
                                                       const uint64_t *__restrict__ q_off, uint32_t row_begin, uint32_t row_end,
                                                       uint32_t *__restrict__ fill, uint2 *__restrict__ work, uint32_t *__restrict__ flags) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t u0 = warp * 32; u0 < n_uniq; u0 += n_warps * 32) {
        const uint64_t u = u0 + lane;
        uint32_t b = 0, e = 0;
        if (u < n_uniq) { b = (uint32_t)uniq_off[u]; e = (uint32_t)uniq_off[u + 1]; }
        const uint32_t mine = e > b ? e - b - 1 : 0;  // emits of this lane's list
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += y;
        }
        const uint32_t total = __shfl_sync(FULL, incl, 31);
        for (uint32_t k0 = 0; k0 < total; k0 += 32) {  // warp-uniform trip count
            const uint32_t k = k0 + lane;
            // list of emit k: the first lane j with incl[j] > k
            int lo = 0;
#pragma unroll
            for (int step = 16; step; step >>= 1) {
                const uint32_t v = __shfl_sync(FULL, incl, lo + step - 1);
                if (v <= k) lo += step;
            }
            const uint32_t jb = __shfl_sync(FULL, b, lo), je = __shfl_sync(FULL, e, lo), jincl = __shfl_sync(FULL, incl, lo);
            if (k < total) {
                const uint32_t p = je - 1 - (jincl - k);  // emit t of list j is posting b_j + t; t = k - excl_j, excl_j = incl_j - (e_j - b_j - 1)
                (void)jb;
                const uint32_t g = dict[p];
                if (g >= row_begin && g < row_end) {
                    const uint64_t o = q_off[g];
                    const uint32_t room = (uint32_t)(q_off[g + 1] - o);
                    const uint32_t slot = atomicAdd(fill + (g - row_begin), 1u);
                    if (slot < room) work[o + slot] = make_uint2(p + 1, je);
                    else atomicExch(flags, 1u);
                }
            }
        }
    }
}

// ---- count + epilogue ------------------------------------------------------------------------------------------
struct DistArgs {
    const uint64_t *q_off;
    const void *q_hashes;        // rectangle: the query hashes (u32, or u64 in the 64-bit hash configuration)
    const uint2 *work;           // triangle: posting ranges per row (row_work_kernel)
    const uint32_t *work_fill;
    const void *uniq_hash;       // rectangle: sorted distinct hashes of the index (u32 / u64)
    const uint64_t *uniq_off;
    const uint32_t *bucket;      // directory over uniq_hash (32-bit hashes)
    uint64_t n_uniq;
    int bucket_shift;
    uint32_t n_buckets;
    const uint32_t *dict;
    const uint32_t *ref_sizes;
    uint32_t row_begin, row_end;
    uint32_t n_ref;
    uint32_t tq, tc;             // tile rows, tile cols
    uint32_t n_col_tiles;
    const uint32_t *tile_rows;   // triangle, one column tile: first row of every tile ([n_tiles] = row_end); a tile then holds only the
                                 // columns right of its first row, so tiles further down the triangle take more rows.  Else null
    int triangle;
    int containment;
    int zero_pass;               // cells with common == 0 (or an empty sketch) pass the threshold
    uint32_t top_n;              // > 0: per row and column tile only the top_n smallest distances (+ ties) are reported
    uint32_t list_word_off;      // word offset of the first-touch list behind the counters in shared memory
    uint32_t sort_word_off;      // word offset of the top-N sort keys (8-byte aligned)
    double neg_inv_k;            // -1.0 / kmer_size
    double dist_limit;           // conservative upper bound on the distance
    kssd_hit *hits;
    unsigned long long *hit_cursor;
    uint64_t hit_cap;
    unsigned long long *posting_total;
};

// increment a counter; true when this was the first increment of the cell
template <typename CNT> __device__ __forceinline__ bool bump(CNT *cnt, uint32_t idx);
template <> __device__ __forceinline__ bool bump<uint32_t>(uint32_t *cnt, uint32_t idx) { return atomicAdd(cnt + idx, 1u) == 0; }
template <> __device__ __forceinline__ bool bump<uint16_t>(uint16_t *cnt, uint32_t idx) {
    // no native 16-bit shared atomics: add 1 or 1<<16 to the containing word (counts never reach 65536)
    const uint32_t old = atomicAdd(reinterpret_cast<uint32_t *>(cnt) + (idx >> 1), (idx & 1) ? 0x10000u : 1u);
    return ((idx & 1) ? (old >> 16) : (old & 0xFFFFu)) == 0;
}

// posting range of query hash h in the index; false when the hash is not indexed
template <typename H> __device__ __forceinline__ bool find_list(const DistArgs &a, H h, uint32_t &b, uint32_t &en);
template <> __device__ __forceinline__ bool find_list<uint32_t>(const DistArgs &a, uint32_t h, uint32_t &b, uint32_t &en) {
    const uint32_t bk = h >> a.bucket_shift;
    if (bk >= a.n_buckets) return false;  // hashes above the largest indexed one have no bucket
    const uint32_t *uh = reinterpret_cast<const uint32_t *>(a.uniq_hash);
    uint32_t lo = a.bucket[bk];
    const uint32_t top = a.bucket[bk + 1];
    uint32_t hi = top;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (uh[mid] < h) lo = mid + 1; else hi = mid;
    }
    if (lo >= top || uh[lo] != h) return false;
    b = (uint32_t)a.uniq_off[lo];
    en = (uint32_t)a.uniq_off[lo + 1];
    return true;
}
template <> __device__ __forceinline__ bool find_list<uint64_t>(const DistArgs &a, uint64_t h, uint32_t &b, uint32_t &en) {
    const uint64_t *uh = reinterpret_cast<const uint64_t *>(a.uniq_hash);
    uint64_t lo = 0, hi = a.n_uniq;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (uh[mid] < h) lo = mid + 1; else hi = mid;
    }
    if (lo >= a.n_uniq || uh[lo] != h) return false;
    b = (uint32_t)a.uniq_off[lo];
    en = (uint32_t)a.uniq_off[lo + 1];
    return true;
}

template <typename CNT, typename H, bool TRI>
__global__ void __launch_bounds__(kDistThreads, 2) dist_tile_kernel(const __grid_constant__ DistArgs a) {
    extern __shared__ uint4 s_raw[];
    __shared__ uint32_t list_n;
    __shared__ uint32_t s_chunk[33];   // first chunk of every row of the tile; [32] = chunks of the tile
    __shared__ uint32_t s_rown[32];    // posting ranges (triangle) / query hashes (rectangle) of every row
    __shared__ uint64_t s_rowe[32];    // q_off of every row
    __shared__ unsigned long long s_post;
    __shared__ uint32_t s_row_first[32];
    CNT *cnt = reinterpret_cast<CNT *>(s_raw);
    uint16_t *list = reinterpret_cast<uint16_t *>(reinterpret_cast<uint32_t *>(s_raw) + a.list_word_off);
    constexpr unsigned FULL = 0xFFFFFFFFu;

    uint32_t r0, r1, c0, c1, tc;
    if (TRI && a.tile_rows) {
        r0 = a.tile_rows[blockIdx.x];
        r1 = a.tile_rows[blockIdx.x + 1];
        c0 = min((r0 + 1) & ~7u, a.n_ref);  // rows stay 16-byte aligned
        c1 = a.n_ref;
        tc = max((c1 - c0 + 7) & ~7u, 8u);
    } else {
        const uint32_t row_tile = blockIdx.x / a.n_col_tiles;
        const uint32_t col_tile = blockIdx.x % a.n_col_tiles;
        r0 = a.row_begin + row_tile * a.tq;
        r1 = min(r0 + a.tq, a.row_end);
        c0 = col_tile * a.tc;
        c1 = min(c0 + a.tc, a.n_ref);
        tc = a.tc;
    }
    if (TRI && c1 <= r0 + 1) return;  // whole tile is on or below the diagonal
    const uint32_t n_rows = r1 - r0;
    const bool clip = a.n_col_tiles > 1;
    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;

    {   // zero the counters
        const uint32_t n16 = (uint32_t)(((uint64_t)n_rows * tc * sizeof(CNT) + 15) / 16);
        for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) s_raw[i] = make_uint4(0, 0, 0, 0);
    }
    if (warp == 0) {  // row table: lane r owns row r of the tile
        uint32_t n = 0;
        uint64_t e0 = 0;
        if ((uint32_t)lane < n_rows) {
            e0 = a.q_off[r0 + lane];
            const uint32_t size = (uint32_t)(a.q_off[r0 + lane + 1] - e0);
            n = TRI ? min(a.work_fill[r0 + lane - a.row_begin], size) : size;
        }
        const uint32_t chunks = (n + 31) >> 5;
        uint32_t x = chunks;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, x, o);
            if (lane >= o) x += y;
        }
        s_chunk[lane] = x - chunks;
        s_rown[lane] = n;
        s_rowe[lane] = e0;
        if (lane == 31) { s_chunk[32] = x; list_n = 0; s_post = 0; }
    }
    __syncthreads();

    // ---- accumulate: warps take chunks of 32 posting ranges round-robin over ALL rows of the tile ----
    const uint32_t my_cs = (uint32_t)lane < n_rows ? s_chunk[lane] : 0xFFFFFFFFu;
    const uint32_t total_chunks = s_chunk[32];
    auto fetch = [&](uint32_t c, uint32_t &b, uint32_t &en, uint32_t &r) {  // c is warp-uniform
        b = en = 0;
        r = 0;
        if (c >= total_chunks) return;
        r = 31 - __clz(__ballot_sync(FULL, my_cs <= c));  // last row whose first chunk is <= c
        const uint32_t idx = (c - s_chunk[r]) * 32 + lane;
        if (idx < s_rown[r]) {
            const uint64_t e = s_rowe[r] + idx;
            if (TRI) {
                const uint2 w = a.work[e];
                b = w.x;
                en = w.y;
            } else {
                if (!find_list<H>(a, reinterpret_cast<const H *>(a.q_hashes)[e], b, en)) b = en = 0;
            }
        }
    };
    auto touch = [&](uint32_t r, uint32_t id) {
        const uint32_t cell = r * tc + (id - c0);
        if (bump<CNT>(cnt, cell)) {
            const uint32_t s = atomicAdd(&list_n, 1u);
            if (s < kListCap) list[s] = (uint16_t)cell;
        }
    };
    unsigned long long posted = 0;
    uint32_t c = warp, nb, nen, nr;
    fetch(c, nb, nen, nr);
    while (c < total_chunks) {
        uint32_t b = nb, en = nen;
        const uint32_t r = nr;
        c += n_warps;
        fetch(c, nb, nen, nr);  // the next chunk's ranges are in flight while this one is streamed
        if (clip && b < en) {   // restrict to ids in [c0, c1)
            uint32_t l = b, h = en;
            while (l < h) { const uint32_t m = (l + h) >> 1; if (a.dict[m] < c0) l = m + 1; else h = m; }
            b = l;
            h = en;
            while (l < h) { const uint32_t m = (l + h) >> 1; if (a.dict[m] < c1) l = m + 1; else h = m; }
            en = l;
        }
        const uint32_t len = en - b;
        posted += len;
        // long ranges: the whole warp strides over one range at a time (coalesced)
        uint32_t longs = __ballot_sync(FULL, len > kLongList);
        while (longs) {
            const int src = __ffs(longs) - 1;
            longs &= longs - 1;
            const uint32_t bb = __shfl_sync(FULL, b, src), ee = __shfl_sync(FULL, en, src);
            for (uint32_t p = bb + lane; p < ee; p += 32) touch(r, a.dict[p]);
        }
        // short ranges: one lane per range, four independent posting loads in flight
        if (len <= kLongList)
            for (uint32_t p = b; p < en; p += 4) {
                uint32_t id[4];
#pragma unroll
                for (int q = 0; q < 4; q++) id[q] = p + q < en ? a.dict[p + q] : 0xFFFFFFFFu;
#pragma unroll
                for (int q = 0; q < 4; q++)
                    if (id[q] != 0xFFFFFFFFu) touch(r, id[q]);
            }
    }
    for (int o = 16; o; o >>= 1) posted += __shfl_xor_sync(FULL, posted, o);
    if (lane == 0 && posted) atomicAdd(&s_post, posted);
    __syncthreads();
    if (threadIdx.x == 0 && s_post) atomicAdd(a.posting_total, s_post);

    // ---- epilogue: distance + threshold, compacted output ----
    auto distance = [&](uint32_t row, uint32_t j, uint32_t cm) -> double {  // device-side value; the host recomputes exactly
        const uint32_t size_q = (uint32_t)(a.q_off[row + 1] - a.q_off[row]);
        const uint32_t size_r = a.ref_sizes[j];
        double v;  // jaccard or containment
        if (!a.containment) v = (double)cm / (double)(int)(size_q + size_r - cm);
        else v = (double)cm / (double)min(size_q, size_r);
        if (v == 1.0) return 0.0;
        return a.containment ? a.neg_inv_k * log(v) : a.neg_inv_k * log((2 * v) / (1.0 + v));
    };
    auto append = [&](bool pass, uint32_t row, uint32_t j, uint32_t cm) {  // warp-aggregated; all 32 lanes call it
        const uint32_t bal = __ballot_sync(FULL, pass);
        if (bal) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(a.hit_cursor, (unsigned long long)__popc(bal));
            base = __shfl_sync(FULL, base, 0);
            if (pass) {
                const unsigned long long at = base + __popc(bal & ((1u << lane) - 1));
                if (at < a.hit_cap) { kssd_hit h; h.row = row; h.col = j; h.common = cm; a.hits[at] = h; }
            }
        }
    };
    const uint32_t n_touch = list_n;
    if (!a.zero_pass && n_touch <= kListCap) {
        const uint32_t n_pad = (n_touch + 31) & ~31u;
        if (a.top_n == 0 || n_touch <= a.top_n) {  // every touched cell that passes the threshold
            for (uint32_t t = threadIdx.x; t < n_pad; t += blockDim.x) {
                bool pass = false;
                uint32_t row = 0, j = 0, cm = 0;
                if (t < n_touch) {
                    const uint32_t ci = list[t];
                    const uint32_t r = ci / tc;
                    row = r0 + r;
                    j = c0 + (ci - r * tc);
                    cm = cnt[ci];
                    pass = distance(row, j, cm) <= a.dist_limit;
                }
                append(pass, row, j, cm);
            }
            return;
        }
        // top-N (src/dist.cpp:625-640 keeps the maxNeighbor smallest distances of a row): sort the passing touched cells
        // by (row, distance); a cell is reported when fewer than top_n cells of its row precede it or when its
        // distance ties with the top_n-th (within 2^-31 relative + one quantum: the host's exact heap decides).
        // key = row (6 bits) | distance bits >> 21 (42 bits, monotone for d >= 0) | cell (16 bits)
        unsigned long long *skey = reinterpret_cast<unsigned long long *>(reinterpret_cast<uint32_t *>(s_raw) + a.sort_word_off);
        for (uint32_t t = threadIdx.x; t < n_touch; t += blockDim.x) {
            const uint32_t ci = list[t];
            const uint32_t r = ci / tc;
            const double d = distance(r0 + r, c0 + (ci - r * tc), cnt[ci]);
            unsigned long long key = ~0ull;
            if (d <= a.dist_limit) {
                const unsigned long long dk = ((unsigned long long)__double_as_longlong(d) >> 21) & ((1ull << 42) - 1);
                key = ((unsigned long long)r << 58) | (dk << 16) | ci;
            }
            skey[t] = key;
        }
        if (threadIdx.x < 32) s_row_first[threadIdx.x] = 0;
        __syncthreads();
        bitonic_sort_smem<unsigned long long>(skey, n_touch);
        for (uint32_t t = threadIdx.x; t < n_touch; t += blockDim.x) {
            const unsigned long long k = skey[t];
            if (k != ~0ull && (t == 0 || (skey[t - 1] >> 58) != (k >> 58))) s_row_first[(uint32_t)(k >> 58)] = t;
        }
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < n_pad; t += blockDim.x) {
            bool pass = false;
            uint32_t row = 0, j = 0, cm = 0;
            if (t < n_touch) {
                const unsigned long long k = skey[t];
                if (k != ~0ull) {
                    const uint32_t r = (uint32_t)(k >> 58), ci = (uint32_t)(k & 0xFFFFu);
                    const uint32_t first = s_row_first[r];
                    pass = t - first < a.top_n;
                    if (!pass) {
                        const unsigned long long kth = skey[first + a.top_n - 1];  // same row: it has more than top_n entries
                        pass = ((k >> 16) & ((1ull << 42) - 1)) <= ((kth >> 16) & ((1ull << 42) - 1)) + 1;
                    }
                    row = r0 + r;
                    j = c0 + (ci - r * tc);
                    cm = cnt[ci];
                }
            }
            append(pass, row, j, cm);
        }
        return;
    }
    if (!a.zero_pass) {
        // the first-touch list overflowed (a tile with more than kListCap non-zero cells): read the counters 16 bytes at a
        // time, collect the non-zero cells of a chunk in the list, evaluate those.  A chunk has at most kListCap cells.
        constexpr uint32_t V = 16 / sizeof(CNT);
        __syncthreads();
        if (threadIdx.x == 0) list_n = 0;
        __syncthreads();
        const uint32_t n_vec = (n_rows * tc) / V;  // tc is a multiple of 8
        for (uint32_t vb = 0; vb < n_vec; vb += blockDim.x) {
            const uint32_t v = vb + threadIdx.x;
            if (v < n_vec) {
                const uint4 q = s_raw[v];
                if (q.x | q.y | q.z | q.w) {
                    const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                    for (uint32_t k = 0; k < V; k++) {
                        const uint32_t cval = sizeof(CNT) == 2 ? (w4[k >> 1] >> (16 * (k & 1))) & 0xFFFFu : w4[k & 3];
                        if (cval) list[atomicAdd(&list_n, 1u)] = (uint16_t)(v * V + k);
                    }
                }
            }
            __syncthreads();
            const uint32_t n_list = list_n;
            for (uint32_t t = threadIdx.x; t < ((n_list + 31) & ~31u); t += blockDim.x) {
                bool pass = false;
                uint32_t row = 0, j = 0, cm = 0;
                if (t < n_list) {
                    const uint32_t ci = list[t];
                    const uint32_t r = ci / tc;
                    row = r0 + r;
                    j = c0 + (ci - r * tc);
                    cm = cnt[ci];
                    pass = distance(row, j, cm) <= a.dist_limit;
                }
                append(pass, row, j, cm);
            }
            __syncthreads();
            if (threadIdx.x == 0) list_n = 0;
            __syncthreads();
        }
        return;
    }
    for (uint32_t r = 0; r < n_rows; r++) {  // every zero cell passes too: dense output
        const uint32_t row = r0 + r;
        const CNT *row_cnt = cnt + (uint64_t)r * tc;
        uint32_t jstart = c0;
        if (TRI && row + 1 > jstart) jstart = row + 1;
        const uint32_t span = c1 > jstart ? c1 - jstart : 0;
        const uint32_t span_pad = (span + 31) & ~31u;
        for (uint32_t t = threadIdx.x; t < span_pad; t += blockDim.x) {
            bool pass = false;
            uint32_t j = jstart + t, cm = 0;
            if (t < span) {
                cm = row_cnt[j - c0];
                pass = cm == 0 ? true : distance(row, j, cm) <= a.dist_limit;
            }
            append(pass, row, j, cm);
        }
    }
}

template <typename CNT, typename H, bool TRI>
static int launch_tiles(kssd_ctx *c, const DistArgs &a, uint32_t n_tiles, size_t smem, uint64_t attr_bit) {
    if (!(c->attr_mask & attr_bit)) {
        KSSD_CUDA(cudaFuncSetAttribute(dist_tile_kernel<CNT, H, TRI>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        c->attr_mask |= attr_bit;
    }
    dist_tile_kernel<CNT, H, TRI><<<n_tiles, kDistThreads, smem, c->stream>>>(a);
    KSSD_CUDA(cudaGetLastError());
    return KSSD_OK;
}

int dist_rows_device(kssd_ctx *c, const void *d_q_hashes, const uint64_t *d_q_off, uint32_t n_query, uint32_t row_begin,
                     uint32_t row_end, int triangle, int kmer_size, double max_dist, int containment, uint64_t *n_hits) {
    *n_hits = 0;
    if (!c->have_index) { set_error("dist: no index loaded (index_build / index_load first)"); return KSSD_E_ARG; }
    if (row_end > n_query || row_begin > row_end) { set_error("dist: bad row range"); return KSSD_E_ARG; }
    if (kmer_size <= 0) { set_error("dist: kmer_size must be positive"); return KSSD_E_ARG; }
    if (c->n_post >= (1ull << 32)) { set_error("dist: more than 2^32 postings is not supported yet"); return KSSD_E_UNSUPPORTED; }
    if (triangle && n_query != c->n_ref) {
        set_error("dist: triangle (alldist) needs the indexed sketches as queries (%u queries, %u indexed)", n_query, c->n_ref);
        return KSSD_E_ARG;
    }
    if (c->index_row_end != 0xFFFFFFFFu && (!triangle || row_begin < c->index_row_begin || row_end > c->index_row_end)) {
        set_error("dist: the index was built for alldist rows [%u, %u) only (index_build_rows)", c->index_row_begin, c->index_row_end);
        return KSSD_E_ARG;
    }
    if (row_begin == row_end || c->n_ref == 0) return KSSD_OK;
    cudaStream_t st = c->stream;
    const uint32_t rows = row_end - row_begin;

    KSSD_TRY(c->h_scalar.reserve(64));
    uint64_t *hs = c->h_scalar.as<uint64_t>();
    KSSD_TRY(c->d_hit_cursor.reserve(64, st));
    unsigned long long *d_cursor = c->d_hit_cursor.as<unsigned long long>();  // [0] hits, [1] postings streamed, [2] flags
    unsigned long long *d_post_total = d_cursor + 1;
    uint32_t *d_flags = reinterpret_cast<uint32_t *>(d_cursor + 2);
    KSSD_CUDA(cudaMemsetAsync(d_cursor, 0, 24, st));
    if (c->profiling) KSSD_TRY(c->t_count.begin(st));
    if (triangle) {
        // posting ranges per row; slots of row g start at q_off[g] (<= list_slots in total: queries are the indexed sketches)
        KSSD_TRY(c->d_q_list_begin.reserve(std::max<uint64_t>(c->list_slots, 1) * sizeof(uint2), st));
        KSSD_TRY(c->d_q_list_end.reserve((uint64_t)rows * 4, st));
        KSSD_CUDA(cudaMemsetAsync(c->d_q_list_end.p, 0, (uint64_t)rows * 4, st));
        if (c->n_uniq) {
            const unsigned grid = (unsigned)std::min<uint64_t>((c->n_uniq + 255) / 256, (uint64_t)c->sm_count * 16);
            row_work_kernel<<<grid, 256, 0, st>>>(c->d_uniq_off.as<uint64_t>(), c->n_uniq, c->d_dict.as<uint32_t>(), d_q_off, row_begin, row_end,
                                                  c->d_q_list_end.as<uint32_t>(), c->d_q_list_begin.as<uint2>(), d_flags);
            KSSD_CUDA(cudaGetLastError());
            c->launches++;
        }
    }

    // tile geometry
    const bool wide = c->max_ref_size > 65535;
    const uint32_t cnt_bytes = wide ? 4 : 2;
    const uint32_t top_n = c->max_neighbor > 0 && c->max_neighbor < kListCap ? (uint32_t)c->max_neighbor : 0;
    // counters + 8 KB list (+ 32 KB of sort keys for top-N) <= ~104 KB: two CTAs per SM
    const uint32_t smem_budget = top_n ? 64 * 1024 : 100 * 1024;  // 2 x (100 + 8) KB fit the 227 KB of an SM; 50,000 references (config 3) fit one column tile
    const uint32_t cap = smem_budget / cnt_bytes;
    uint32_t tc = std::min<uint32_t>(c->n_ref, cap);
    tc = (tc + 7) & ~7u;  // keep rows 16-byte aligned
    if (tc > cap) tc = cap & ~7u;
    const uint32_t n_col_tiles = (c->n_ref + tc - 1) / tc;
    uint32_t tq = std::max<uint32_t>(1, cap / tc);
    tq = std::min<uint32_t>(tq, 32);
    // enough CTAs to fill the machine a few times over
    const uint32_t want_tiles = (uint32_t)c->sm_count * 4;
    const uint32_t tq_fill = std::max<uint32_t>(1, rows / want_tiles);
    tq = std::min(tq, tq_fill);
    uint32_t n_row_tiles = (rows + tq - 1) / tq;
    // triangle with one column tile: pack the rows by the columns they can still see (src/dist.cpp:207, j > i)
    const uint32_t *d_tile_rows = nullptr;
    size_t cnt_smem = ((size_t)tq * tc * cnt_bytes + 15) & ~(size_t)15;
    if (triangle && n_col_tiles == 1 && rows > 1) {
        // rows per tile further down the triangle: measured (profiles/r02i, r02j) -- taller tiles mean fewer CTAs and a longer
        // tail, and the kernel is bound by shared-memory wavefronts, not by per-tile latency
        uint32_t q_max = tq;
        if (const char *e = getenv("KSSD_B200_TILE_QMAX")) q_max = std::max(1, atoi(e));
        q_max = std::min<uint32_t>(32, std::max<uint32_t>(q_max, 1));
        const uint64_t key[4] = {((uint64_t)row_begin << 32) | row_end, c->n_ref, cap, q_max};
        if (memcmp(key, c->tile_key, sizeof(key)) != 0 || c->d_tile_rows.p == nullptr) {
            std::vector<uint32_t> starts;
            for (uint32_t r = row_begin; r < row_end;) {
                starts.push_back(r);
                const uint32_t cc0 = std::min((r + 1) & ~7u, c->n_ref);
                const uint32_t w = std::max<uint32_t>((c->n_ref - cc0 + 7) & ~7u, 8);
                r += std::min<uint32_t>(std::max<uint32_t>(cap / w, 1), q_max);
            }
            starts.push_back(row_end);
            KSSD_TRY(c->h_tile_rows.reserve(starts.size() * 4));
            memcpy(c->h_tile_rows.p, starts.data(), starts.size() * 4);
            KSSD_TRY(c->d_tile_rows.reserve(starts.size() * 4, st));
            KSSD_CUDA(cudaMemcpyAsync(c->d_tile_rows.p, c->h_tile_rows.p, starts.size() * 4, cudaMemcpyHostToDevice, st));
            memcpy(c->tile_key, key, sizeof(key));
            c->n_tile_rows = (uint32_t)starts.size() - 1;
        }
        d_tile_rows = c->d_tile_rows.as<uint32_t>();
        n_row_tiles = c->n_tile_rows;
        cnt_smem = ((size_t)cap * cnt_bytes + 15) & ~(size_t)15;  // a tile uses at most `cap` counters
    }
    const size_t smem = cnt_smem + kListCap * 2 + (top_n ? kListCap * 8 : 0);

    DistArgs a;
    a.q_off = d_q_off;
    a.q_hashes = d_q_hashes;
    a.work = c->d_q_list_begin.as<uint2>();
    a.work_fill = c->d_q_list_end.as<uint32_t>();
    a.uniq_hash = c->index_hash64 ? c->d_uniq_hash64.p : c->d_uniq_hash.p;
    a.uniq_off = c->d_uniq_off.as<uint64_t>();
    a.bucket = c->d_bucket.as<uint32_t>();
    a.n_uniq = c->n_uniq;
    a.bucket_shift = c->bucket_shift;
    a.n_buckets = c->n_buckets;
    a.dict = c->d_dict.as<uint32_t>();
    a.ref_sizes = c->d_ref_sizes.as<uint32_t>();
    a.row_begin = row_begin;
    a.row_end = row_end;
    a.n_ref = c->n_ref;
    a.tq = tq;
    a.tc = tc;
    a.n_col_tiles = n_col_tiles;
    a.tile_rows = d_tile_rows;
    a.triangle = triangle;
    a.containment = containment;
    // distance of a zero cell is exactly 1.0 (dist.cpp:227-228): strict '<' for alldist, '<=' for dist
    a.zero_pass = c->emit_zero_cells && (triangle ? (1.0 < max_dist) : (1.0 <= max_dist));
    a.top_n = top_n;
    a.neg_inv_k = -1.0 / kmer_size;
    a.dist_limit = max_dist + std::fabs(max_dist) * 1e-9 + 1e-12;  // +inf stays +inf: every non-zero cell is reported
    a.hit_cursor = d_cursor;
    a.posting_total = d_post_total;
    a.list_word_off = (uint32_t)(cnt_smem / 4);
    a.sort_word_off = (uint32_t)((cnt_smem + kListCap * 2) / 4);

    uint64_t hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    if (hit_cap == 0) {
        const uint64_t guess = a.zero_pass ? (uint64_t)rows * c->n_ref : std::max<uint64_t>(1 << 20, (uint64_t)rows * 64);
        KSSD_TRY(c->d_hits.reserve(guess * sizeof(kssd_hit), st));
        hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    }
    const uint32_t n_tiles = n_row_tiles * n_col_tiles;
    for (int attempt = 0; attempt < 2; attempt++) {
        a.hits = c->d_hits.as<kssd_hit>();
        a.hit_cap = hit_cap;
        if (attempt) {
            KSSD_CUDA(cudaMemsetAsync(d_cursor, 0, 16, st));
            if (c->profiling) KSSD_TRY(c->t_count.begin(st));
        }
        int rc;
        if (triangle) {
            rc = wide ? launch_tiles<uint32_t, uint32_t, true>(c, a, n_tiles, smem, 1ull << 42) : launch_tiles<uint16_t, uint32_t, true>(c, a, n_tiles, smem, 1ull << 43);
        } else if (c->index_hash64) {
            rc = wide ? launch_tiles<uint32_t, uint64_t, false>(c, a, n_tiles, smem, 1ull << 44) : launch_tiles<uint16_t, uint64_t, false>(c, a, n_tiles, smem, 1ull << 45);
        } else {
            rc = wide ? launch_tiles<uint32_t, uint32_t, false>(c, a, n_tiles, smem, 1ull << 46) : launch_tiles<uint16_t, uint32_t, false>(c, a, n_tiles, smem, 1ull << 47);
        }
        KSSD_TRY(rc);
        if (c->profiling) KSSD_TRY(c->t_count.end(st));
        c->launches++;
        c->t_count.launches++;
        KSSD_CUDA(cudaMemcpyAsync(hs, d_cursor, 24, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        if (*reinterpret_cast<const uint32_t *>(hs + 2)) {
            set_error("dist: triangle (alldist) needs the indexed sketches as queries: a row has more posting ranges than hashes");
            return KSSD_E_ARG;
        }
        if (hs[0] <= hit_cap) {
            c->t_count.units_a += hs[1];  // posting entries streamed
            c->t_count.units_b += hs[1];  // = shared-memory atomic increments (one per posting)
            *n_hits = hs[0];
            return KSSD_OK;
        }
        // output buffer too small: grow to the exact need and redo the tile pass (the work lists stay valid)
        KSSD_TRY(c->d_hits.reserve(hs[0] * sizeof(kssd_hit), st));
        hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    }
    set_error("dist: hit buffer overflow after regrow");
    return KSSD_E_OVERFLOW;
}

}  // namespace kssd
