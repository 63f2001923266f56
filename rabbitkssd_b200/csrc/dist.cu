// alldist / dist on the device: posting-list lookup, shared-memory intersection counters, distance + threshold
// epilogue, compacted (row, col, common) output.
//
// Replaces the per-row loops of index_tridist (src/dist.cpp:174-258) and index_dist (src/dist.cpp:560-692).  The
// reference keeps one dense int counter row per thread in DRAM and rescans all N cells of it per query with a
// libm log() per non-trivial cell.  Here a CTA owns a tile of TQ query rows x TC reference columns whose
// counters live in shared memory (16-bit when no sketch exceeds 65535 hashes); posting lists are streamed from
// the CSR index (L2 resident for every BASELINE config) and accumulated with shared-memory atomics; the
// epilogue scans the tile from shared memory, rejects zero cells with one compare, evaluates the distance for
// the few non-zero cells and appends passing cells with warp-aggregated atomics.  The device test is
// conservative (superset); the host shim recomputes with the reference's exact expressions.
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace kssd {

// ---- step 1: per query hash, the posting range it has to stream --------------------------------------------
// grid = (rows, chunks): the row is known from blockIdx.x, so nothing is searched but the hash itself (a coarse
// directory narrows it to a handful of distinct hashes) and, for alldist, the first posting id > row.
__global__ void __launch_bounds__(128) list_lookup_kernel(const uint32_t *__restrict__ q_hashes, const uint64_t *__restrict__ q_off,
                                                          uint32_t row_begin, const uint32_t *__restrict__ uniq_hash,
                                                          const uint64_t *__restrict__ uniq_off, const uint32_t *__restrict__ bucket,
                                                          int bucket_shift, uint32_t n_buckets, const uint32_t *__restrict__ dict,
                                                          int triangle, uint32_t *__restrict__ lb, uint32_t *__restrict__ le,
                                                          unsigned long long *__restrict__ posting_total) {
    const uint32_t row = row_begin + blockIdx.x;
    const uint64_t e_base = q_off[row_begin];
    const uint64_t e0 = q_off[row], e1 = q_off[row + 1];
    unsigned long long s = 0;
    for (uint64_t e = e0 + (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; e < e1; e += (uint64_t)gridDim.y * blockDim.x) {
        const uint32_t h = q_hashes[e];
        const uint32_t bk = h >> bucket_shift;
        uint32_t b = 0, en = 0;
        if (bk < n_buckets) {  // hashes above the largest indexed one have no bucket
            uint32_t lo = bucket[bk], hi = bucket[bk + 1];
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (uniq_hash[mid] < h) lo = mid + 1; else hi = mid;
            }
            if (lo < bucket[bk + 1] && uniq_hash[lo] == h) {
                b = (uint32_t)uniq_off[lo];
                en = (uint32_t)uniq_off[lo + 1];
                if (triangle) {  // first posting id > row (ids ascending inside a list)
                    if (en - b <= 8) {
                        uint32_t skip = 0;
                        for (uint32_t p = b; p < en; p++) skip += dict[p] <= row;  // independent loads
                        b += skip;
                    } else {
                        uint32_t l2 = b, h2 = en;
                        while (l2 < h2) {
                            const uint32_t mid = (l2 + h2) >> 1;
                            if (dict[mid] <= row) l2 = mid + 1; else h2 = mid;
                        }
                        b = l2;
                    }
                }
            }
        }
        lb[e - e_base] = b;
        le[e - e_base] = en;
        s += en - b;
    }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(posting_total, s);
}

// ---- step 2: count + epilogue -------------------------------------------------------------------------------
static constexpr uint32_t kListCap = 4096;  // cells per epilogue chunk = 512 threads x 8 (16-bit) cells

struct DistArgs {
    const uint64_t *q_off;
    const uint32_t *lb, *le;
    const uint32_t *dict;
    const uint32_t *ref_sizes;
    uint64_t e_base;        // q_off[row_begin]
    uint32_t row_begin, row_end;
    uint32_t n_ref;
    uint32_t tq, tc;        // tile rows, tile cols
    uint32_t n_col_tiles;
    int triangle;
    int containment;
    int zero_pass;          // cells with common == 0 (or an empty sketch) pass the threshold
    uint32_t list_word_off; // word offset of the non-zero-cell list behind the counters in shared memory
    double neg_inv_k;       // -1.0 / kmer_size
    double dist_limit;      // conservative upper bound on the distance
    kssd_hit *hits;
    unsigned long long *hit_cursor;
    uint64_t hit_cap;
};

template <typename CNT> __device__ __forceinline__ void bump(CNT *cnt, uint32_t idx);
template <> __device__ __forceinline__ void bump<uint32_t>(uint32_t *cnt, uint32_t idx) { atomicAdd(cnt + idx, 1u); }
template <> __device__ __forceinline__ void bump<uint16_t>(uint16_t *cnt, uint32_t idx) {
    // no native 16-bit shared atomics: add 1 or 1<<16 to the containing word (counts never reach 65536)
    atomicAdd(reinterpret_cast<uint32_t *>(cnt) + (idx >> 1), (idx & 1) ? 0x10000u : 1u);
}

template <typename CNT>
__global__ void __launch_bounds__(512) dist_tile_kernel(const DistArgs a) {
    extern __shared__ uint4 s_raw[];
    CNT *cnt = reinterpret_cast<CNT *>(s_raw);
    const uint32_t row_tile = blockIdx.x / a.n_col_tiles;
    const uint32_t col_tile = blockIdx.x % a.n_col_tiles;
    const uint32_t r0 = a.row_begin + row_tile * a.tq;
    const uint32_t r1 = min(r0 + a.tq, a.row_end);
    const uint32_t c0 = col_tile * a.tc;
    const uint32_t c1 = min(c0 + a.tc, a.n_ref);
    if (a.triangle && c1 <= r0 + 1) return;  // whole tile is on or below the diagonal
    const uint32_t n_rows = r1 - r0;
    const uint32_t tc = a.tc;
    const bool clip = a.n_col_tiles > 1;

    // zero counters
    {
        const uint32_t n16 = (uint32_t)(((uint64_t)n_rows * tc * sizeof(CNT) + 15) / 16);
        for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) s_raw[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    // accumulate: warps take 32 query hashes (= 32 posting lists) at a time
    for (uint32_t r = 0; r < n_rows; r++) {
        const uint64_t e0 = a.q_off[r0 + r] - a.e_base, e1 = a.q_off[r0 + r + 1] - a.e_base;
        CNT *row_cnt = cnt + (uint64_t)r * tc;
        uint32_t nxt_b = 0, nxt_en = 0;  // ranges of the NEXT batch are fetched while the current one is streamed
        if (e0 + (uint64_t)warp * 32 + lane < e1) { nxt_b = a.lb[e0 + (uint64_t)warp * 32 + lane]; nxt_en = a.le[e0 + (uint64_t)warp * 32 + lane]; }
        for (uint64_t eb = e0 + (uint64_t)warp * 32; eb < e1; eb += (uint64_t)n_warps * 32) {
            uint32_t b = nxt_b, en = nxt_en;
            {
                const uint64_t e2 = eb + (uint64_t)n_warps * 32 + lane;
                nxt_b = nxt_en = 0;
                if (e2 < e1) { nxt_b = a.lb[e2]; nxt_en = a.le[e2]; }
            }
            if (clip && b < en) {  // restrict to ids in [c0, c1)
                uint32_t l = b, h = en;
                while (l < h) { uint32_t m = (l + h) >> 1; if (a.dict[m] < c0) l = m + 1; else h = m; }
                b = l;
                h = en;
                while (l < h) { uint32_t m = (l + h) >> 1; if (a.dict[m] < c1) l = m + 1; else h = m; }
                en = l;
            }
            uint32_t len = en - b;
            // long lists: the whole warp strides over one list at a time (coalesced)
            uint32_t longs = __ballot_sync(0xFFFFFFFFu, len > 16);
            while (longs) {
                int src = __ffs(longs) - 1;
                longs &= longs - 1;
                uint32_t bb = __shfl_sync(0xFFFFFFFFu, b, src), ee = __shfl_sync(0xFFFFFFFFu, en, src);
                for (uint32_t p = bb + lane; p < ee; p += 32) bump<CNT>(row_cnt, a.dict[p] - c0);
            }
            // short lists: one lane per list, four independent posting loads in flight
            if (len <= 16)
                for (uint32_t p = b; p < en; p += 4) {
                    uint32_t id[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) id[q] = p + q < en ? a.dict[p + q] : 0xFFFFFFFFu;
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        if (id[q] != 0xFFFFFFFFu) bump<CNT>(row_cnt, id[q] - c0);
                }
        }
    }
    __syncthreads();

    // epilogue: scan the tile, distance + threshold, compacted output
    auto evaluate = [&](uint32_t row, uint32_t j, uint32_t c) -> bool {  // conservative device-side threshold test
        const uint32_t size_q = (uint32_t)(a.q_off[row + 1] - a.q_off[row]);
        const uint32_t size_r = a.ref_sizes[j];
        double v;  // jaccard or containment
        if (!a.containment) v = (double)c / (double)(int)(size_q + size_r - c);
        else v = (double)c / (double)min(size_q, size_r);
        double d;
        if (v == 1.0) d = 0.0;
        else if (!a.containment) d = a.neg_inv_k * log((2 * v) / (1.0 + v));
        else d = a.neg_inv_k * log(v);
        return d <= a.dist_limit;
    };
    auto append = [&](bool pass, uint32_t row, uint32_t j, uint32_t c) {  // warp-aggregated; all 32 lanes call it
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, pass);
        if (bal) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(a.hit_cursor, (unsigned long long)__popc(bal));
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (pass) {
                const unsigned long long at = base + __popc(bal & ((1u << lane) - 1));
                if (at < a.hit_cap) { kssd_hit h; h.row = row; h.col = j; h.common = c; a.hits[at] = h; }
            }
        }
    };
    if (!a.zero_pass) {
        // Almost every cell is zero: read the counters 16 bytes at a time, collect the non-zero cells of a chunk in a
        // shared-memory list (a cell is non-zero only if a posting with a legal column id was counted into it), then
        // evaluate the listed cells.  A chunk has at most kListCap cells, so the list cannot overflow.
        constexpr uint32_t V = 16 / sizeof(CNT);
        uint16_t *list = reinterpret_cast<uint16_t *>(reinterpret_cast<uint32_t *>(s_raw) + a.list_word_off);  // a tile has < 65536 cells
        __shared__ uint32_t list_n;
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
                uint32_t row = 0, j = 0, c = 0;
                if (t < n_list) {
                    const uint32_t ci = list[t];
                    const uint32_t r = ci / tc;
                    row = r0 + r;
                    j = c0 + (ci - r * tc);
                    c = cnt[ci];
                    pass = evaluate(row, j, c);
                }
                append(pass, row, j, c);
            }
            __syncthreads();
            if (threadIdx.x == 0) list_n = 0;
            // the next chunk's atomics come after its own __syncthreads-free region: make the reset visible first
            __syncthreads();
        }
        return;
    }
    for (uint32_t r = 0; r < n_rows; r++) {  // every zero cell passes too: dense output
        const uint32_t row = r0 + r;
        const CNT *row_cnt = cnt + (uint64_t)r * tc;
        uint32_t jstart = c0;
        if (a.triangle && row + 1 > jstart) jstart = row + 1;
        const uint32_t span = c1 > jstart ? c1 - jstart : 0;
        const uint32_t span_pad = (span + 31) & ~31u;
        for (uint32_t t = threadIdx.x; t < span_pad; t += blockDim.x) {
            bool pass = false;
            uint32_t j = jstart + t, c = 0;
            if (t < span) {
                c = row_cnt[j - c0];
                pass = c == 0 ? true : evaluate(row, j, c);
            }
            append(pass, row, j, c);
        }
    }
}

int dist_rows_device(kssd_ctx *c, const uint32_t *d_q_hashes, const uint64_t *d_q_off, uint32_t n_query, uint32_t row_begin,
                     uint32_t row_end, int triangle, int kmer_size, double max_dist, int containment, uint64_t *n_hits) {
    *n_hits = 0;
    if (!c->have_index) { set_error("dist: no index loaded (index_build / index_load first)"); return KSSD_E_ARG; }
    if (row_end > n_query || row_begin > row_end) { set_error("dist: bad row range"); return KSSD_E_ARG; }
    if (kmer_size <= 0) { set_error("dist: kmer_size must be positive"); return KSSD_E_ARG; }
    if (c->n_post >= (1ull << 32)) { set_error("dist: more than 2^32 postings is not supported yet"); return KSSD_E_UNSUPPORTED; }
    if (row_begin == row_end || c->n_ref == 0) return KSSD_OK;
    cudaStream_t st = c->stream;

    // element range of the row block (two scalars from the device offsets)
    KSSD_TRY(c->h_scalar.reserve(64));
    uint64_t *hs = c->h_scalar.as<uint64_t>();
    KSSD_CUDA(cudaMemcpyAsync(hs, d_q_off + row_begin, 8, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaMemcpyAsync(hs + 1, d_q_off + row_end, 8, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    const uint64_t e0 = hs[0], e1 = hs[1], n_el = e1 - e0;

    KSSD_TRY(c->d_q_list_begin.reserve(std::max<uint64_t>(n_el, 1) * 4, st));
    KSSD_TRY(c->d_q_list_end.reserve(std::max<uint64_t>(n_el, 1) * 4, st));
    KSSD_TRY(c->d_hit_cursor.reserve(64, st));
    unsigned long long *d_cursor = c->d_hit_cursor.as<unsigned long long>();
    unsigned long long *d_post_total = d_cursor + 1;
    KSSD_CUDA(cudaMemsetAsync(d_cursor, 0, 16, st));
    if (n_el) {
        const uint32_t n_rows_blk = row_end - row_begin;
        const uint64_t avg_len = n_el / n_rows_blk + 1;
        const uint32_t chunks = (uint32_t)std::min<uint64_t>(64, (avg_len + 255) / 256);  // ~2 hashes per thread
        list_lookup_kernel<<<dim3(n_rows_blk, chunks), 128, 0, st>>>(
            d_q_hashes, d_q_off, row_begin, c->d_uniq_hash.as<uint32_t>(), c->d_uniq_off.as<uint64_t>(),
            c->d_bucket.as<uint32_t>(), c->bucket_shift, c->n_buckets, c->d_dict.as<uint32_t>(), triangle, c->d_q_list_begin.as<uint32_t>(), c->d_q_list_end.as<uint32_t>(), d_post_total);
        KSSD_CUDA(cudaGetLastError());
        c->launches++;
    }

    // tile geometry
    const bool wide = c->max_ref_size > 65535;
    const uint32_t cnt_bytes = wide ? 4 : 2;
    const uint32_t smem_budget = 96 * 1024;  // + 8 KB list = 104 KB: two CTAs per SM, one streams postings while the other scans
    const uint32_t cap = smem_budget / cnt_bytes;
    uint32_t tc = std::min<uint32_t>(c->n_ref, cap);
    tc = (tc + 7) & ~7u;  // keep rows 16-byte aligned
    if (tc > cap) tc = cap & ~7u;
    const uint32_t n_col_tiles = (c->n_ref + tc - 1) / tc;
    const uint32_t rows = row_end - row_begin;
    uint32_t tq = std::max<uint32_t>(1, cap / tc);
    tq = std::min<uint32_t>(tq, 32);
    // enough CTAs to fill the machine a few times over
    uint32_t want_tiles = (uint32_t)c->sm_count * 4;
    uint32_t tq_fill = std::max<uint32_t>(1, rows / want_tiles);
    tq = std::min(tq, tq_fill);
    const uint32_t n_row_tiles = (rows + tq - 1) / tq;
    const size_t cnt_smem = ((size_t)tq * tc * cnt_bytes + 15) & ~(size_t)15;
    const size_t smem = cnt_smem + kListCap * 2;

    DistArgs a;
    a.q_off = d_q_off;
    a.lb = c->d_q_list_begin.as<uint32_t>();
    a.le = c->d_q_list_end.as<uint32_t>();
    a.dict = c->d_dict.as<uint32_t>();
    a.ref_sizes = c->d_ref_sizes.as<uint32_t>();
    a.e_base = e0;
    a.row_begin = row_begin;
    a.row_end = row_end;
    a.n_ref = c->n_ref;
    a.tq = tq;
    a.tc = tc;
    a.n_col_tiles = n_col_tiles;
    a.triangle = triangle;
    a.containment = containment;
    // distance of a zero cell is exactly 1.0 (dist.cpp:227-228): strict '<' for alldist, '<=' for dist
    a.zero_pass = c->emit_zero_cells && (triangle ? (1.0 < max_dist) : (1.0 <= max_dist));
    a.neg_inv_k = -1.0 / kmer_size;
    a.dist_limit = max_dist + std::fabs(max_dist) * 1e-9 + 1e-12;
    a.hit_cursor = d_cursor;
    a.list_word_off = (uint32_t)(cnt_smem / 4);

    const uint64_t attr_bit = 1ull << (42);
    if (!(c->attr_mask & attr_bit)) {
        KSSD_CUDA(cudaFuncSetAttribute(dist_tile_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
        KSSD_CUDA(cudaFuncSetAttribute(dist_tile_kernel<uint32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
        c->attr_mask |= attr_bit;
    }
    uint64_t hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    if (hit_cap == 0) {
        uint64_t guess = a.zero_pass ? (uint64_t)rows * c->n_ref : std::max<uint64_t>(1 << 20, (uint64_t)rows * 64);
        KSSD_TRY(c->d_hits.reserve(guess * sizeof(kssd_hit), st));
        hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    }
    for (int attempt = 0; attempt < 2; attempt++) {
        a.hits = c->d_hits.as<kssd_hit>();
        a.hit_cap = hit_cap;
        KSSD_CUDA(cudaMemsetAsync(d_cursor, 0, 8, st));
        if (c->profiling) KSSD_TRY(c->t_count.begin(st));
        if (wide) dist_tile_kernel<uint32_t><<<n_row_tiles * n_col_tiles, 512, smem, st>>>(a);
        else dist_tile_kernel<uint16_t><<<n_row_tiles * n_col_tiles, 512, smem, st>>>(a);
        KSSD_CUDA(cudaGetLastError());
        if (c->profiling) KSSD_TRY(c->t_count.end(st));
        c->launches++;
        c->t_count.launches++;
        KSSD_CUDA(cudaMemcpyAsync(hs, d_cursor, 16, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        if (hs[0] <= hit_cap) {
            c->t_count.units_a += hs[1];  // posting entries streamed
            c->t_count.units_b += hs[1];  // = shared-memory atomic increments (one per posting)
            *n_hits = hs[0];
            return KSSD_OK;
        }
        // output buffer too small: grow to the exact need and redo the tile pass
        KSSD_TRY(c->d_hits.reserve(hs[0] * sizeof(kssd_hit), st));
        hit_cap = c->d_hits.cap / sizeof(kssd_hit);
    }
    set_error("dist: hit buffer overflow after regrow");
    return KSSD_E_OVERFLOW;
}

}  // namespace kssd
