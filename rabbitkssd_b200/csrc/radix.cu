// Hand-written device-wide LSD radix sort of 64-bit keys (8-bit digits, stable), plus the unique / count-filter
// compaction and the split into per-genome segments.  Together they replace the reference's per-genome
// unordered_set / unordered_map containers (src/sketch.cpp:526-550, :832-849): keys are (genome << 32 | hash), so
// one sort groups by genome and orders each genome's hashes ("segmented sort + unique per genome").
//
// Only the digit positions that can differ are sorted (hash bits and genome-id bits), so a 28-bit hash with
// 10,000 genomes costs 6 passes instead of 8.
#include "common.cuh"

namespace kssd {

static constexpr int kRadixThreads = 256;
static constexpr int kRadixItems = 16;  // keys per thread
static constexpr int kRadixTile = kRadixThreads * kRadixItems;
static constexpr int kWarps = kRadixThreads / 32;

// Per-CTA digit histogram -> counts[digit * n_tiles + tile]
__global__ void __launch_bounds__(kRadixThreads) radix_hist_kernel(const uint64_t *__restrict__ keys, uint64_t n, int shift,
                                                                   uint32_t *__restrict__ counts, uint32_t n_tiles) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * kRadixTile;
#pragma unroll 4
    for (int r = 0; r < kRadixItems; r++) {
        uint64_t i = base + (uint64_t)r * kRadixThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255], 1u);
    }
    __syncthreads();
    counts[(uint64_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// One CTA per digit: exclusive scan over tiles of counts[digit][*]; digit totals -> totals[digit]
__global__ void __launch_bounds__(256) radix_scan_tiles_kernel(uint32_t *counts, uint32_t n_tiles, uint32_t *totals) {
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t carry;
    uint32_t *row = counts + (uint64_t)blockIdx.x * n_tiles;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n_tiles; base += 256) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < n_tiles ? row[i] : 0;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sums[w] = x;
        __syncthreads();
        uint32_t wp = 0;
        for (int k = 0; k < w; k++) wp += warp_sums[k];
        uint32_t excl = carry + wp + x - v;
        if (i < n_tiles) row[i] = excl;
        __syncthreads();
        if (threadIdx.x == 255) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

// totals[256] -> exclusive digit bases (single warp-sized CTA of 256 threads)
__global__ void radix_scan_digits_kernel(uint32_t *totals, uint64_t *digit_base) {
    __shared__ uint64_t s[256];
    s[threadIdx.x] = totals[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t acc = 0;
        for (int d = 0; d < 256; d++) { uint64_t v = s[d]; s[d] = acc; acc += v; }
    }
    __syncthreads();
    digit_base[threadIdx.x] = s[threadIdx.x];
}

// Stable scatter.  Warp w owns keys [w*512, (w+1)*512) of the tile and walks them 32 at a time in order;
// __match_any_sync groups equal digits inside a 32-key row, a per-warp shared counter carries the running count.
// Keys are first placed at their tile-local sorted position in shared memory and then written out in that order, so
// that consecutive threads write consecutive addresses of the same digit run (coalesced) instead of 8-byte
// scattered stores.
__global__ void __launch_bounds__(kRadixThreads) radix_scatter_kernel(const uint64_t *__restrict__ keys, uint64_t *__restrict__ out,
                                                                      uint64_t n, int shift, const uint32_t *__restrict__ counts,
                                                                      const uint64_t *__restrict__ digit_base, uint32_t n_tiles) {
    __shared__ uint64_t skeys[kRadixTile];
    __shared__ uint32_t wcount[kWarps][256];
    __shared__ uint32_t tstart[256];   // tile-local start of each digit run
    __shared__ uint64_t gbase[256];    // global position of the first key of this tile's digit run
    __shared__ uint32_t wsum[kWarps];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kWarps * 256; i += kRadixThreads) (&wcount[0][0])[i] = 0;
    __syncthreads();
    const uint64_t tbase = (uint64_t)blockIdx.x * kRadixTile;
    const uint64_t wbase = tbase + (uint64_t)w * (32 * kRadixItems);
    uint64_t key[kRadixItems];
    uint32_t rank[kRadixItems];
    const uint32_t lt = (1u << lane) - 1;
#pragma unroll
    for (int r = 0; r < kRadixItems; r++) {
        uint64_t i = wbase + r * 32 + lane;
        bool live = i < n;
        key[r] = live ? keys[i] : ~0ull;
        uint32_t d = (uint32_t)(key[r] >> shift) & 255;
        uint32_t peers = __match_any_sync(0xFFFFFFFFu, live ? d : 256u + 0u);
        uint32_t before = __popc(peers & lt);
        uint32_t old = 0;
        int leader = __ffs(peers) - 1;
        if (live && lane == leader) {
            old = wcount[w][d];
            wcount[w][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xFFFFFFFFu, old, leader);
        rank[r] = old + before;
        __syncwarp();
    }
    __syncthreads();
    // per digit: exclusive prefix over warps, tile total; then exclusive scan of the totals over the 256 digits
    {
        const int d = threadIdx.x;
        uint32_t acc = 0;
#pragma unroll
        for (int k = 0; k < kWarps; k++) { uint32_t v = wcount[k][d]; wcount[k][d] = acc; acc += v; }
        uint32_t x = acc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) wsum[w] = x;
        __syncthreads();
        uint32_t wp = 0;
        for (int k = 0; k < w; k++) wp += wsum[k];
        tstart[d] = wp + x - acc;
        gbase[d] = digit_base[d] + counts[(uint64_t)d * n_tiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kRadixItems; r++) {
        uint64_t i = wbase + r * 32 + lane;
        if (i < n) {
            uint32_t d = (uint32_t)(key[r] >> shift) & 255;
            skeys[tstart[d] + wcount[w][d] + rank[r]] = key[r];
        }
    }
    __syncthreads();
    const uint32_t n_tile = (uint32_t)min((uint64_t)kRadixTile, n - tbase);
    for (uint32_t i = threadIdx.x; i < n_tile; i += kRadixThreads) {
        const uint64_t k = skeys[i];
        const uint32_t d = (uint32_t)(k >> shift) & 255;
        out[gbase[d] + (i - tstart[d])] = k;
    }
}

// NOTE the histogram kernel strides keys by thread while the scatter kernel assigns contiguous 512-key blocks to
// warps; both cover exactly the same tile [blockIdx*4096, +4096), which is all the histogram depends on.

int radix_sort_u64(kssd_ctx *c, uint64_t *d_keys, uint64_t *d_tmp, uint64_t n, int lo_begin, int lo_end, int hi_begin,
                   int hi_end, uint64_t **d_sorted) {
    *d_sorted = d_keys;
    if (n <= 1) return KSSD_OK;
    if (n >= (1ull << 32) * 16) { set_error("radix sort: too many keys"); return KSSD_E_ARG; }
    uint32_t n_tiles = (uint32_t)((n + kRadixTile - 1) / kRadixTile);
    KSSD_TRY(c->d_sort_counts.reserve(((uint64_t)256 * n_tiles + 256) * 4 + 256 * 8, c->stream));
    uint32_t *counts = c->d_sort_counts.as<uint32_t>();
    uint32_t *totals = counts + (uint64_t)256 * n_tiles;
    uint64_t *digit_base = reinterpret_cast<uint64_t *>(totals + 256);
    uint64_t *src = d_keys, *dst = d_tmp;
    int ranges[2][2] = {{lo_begin, lo_end}, {hi_begin, hi_end}};
    for (int rg = 0; rg < 2; rg++) {
        for (int shift = ranges[rg][0]; shift < ranges[rg][1]; shift += 8) {
            // a final partial digit still sorts 8 bits; higher bits inside it are constant or belong to the
            // next range, which is harmless for an LSD sort as long as later passes cover them again
            radix_hist_kernel<<<n_tiles, kRadixThreads, 0, c->stream>>>(src, n, shift, counts, n_tiles);
            radix_scan_tiles_kernel<<<256, 256, 0, c->stream>>>(counts, n_tiles, totals);
            radix_scan_digits_kernel<<<1, 256, 0, c->stream>>>(totals, digit_base);
            radix_scatter_kernel<<<n_tiles, kRadixThreads, 0, c->stream>>>(src, dst, n, shift, counts, digit_base, n_tiles);
            KSSD_CUDA(cudaGetLastError());
            c->launches += 4;
            uint64_t *t = src; src = dst; dst = t;
        }
    }
    *d_sorted = src;
    return KSSD_OK;
}

// ---- unique / count filter ---------------------------------------------------------------------------------
// keep[i] = key[i] starts a run && the run has >= min_count members (sorted => run members are contiguous).
__device__ __forceinline__ bool keep_key(const uint64_t *k, uint64_t i, uint64_t n, int min_count) {
    uint64_t v = k[i];
    if (i > 0 && k[i - 1] == v) return false;
    if (min_count <= 1) return true;
    uint64_t j = i + (uint64_t)(min_count - 1);
    return j < n && k[j] == v;
}

__global__ void __launch_bounds__(256) unique_count_kernel(const uint64_t *__restrict__ k, uint64_t n, int min_count,
                                                           uint32_t *__restrict__ tile_counts) {
    __shared__ uint32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * 4096;
    uint32_t mine = 0;
    for (int r = 0; r < 16; r++) {
        uint64_t i = base + r * 256 + threadIdx.x;
        if (i < n && keep_key(k, i, n, min_count)) mine++;
    }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&total, mine);
    __syncthreads();
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

// single CTA: exclusive scan of tile_counts (64-bit result), total -> *n_out
__global__ void __launch_bounds__(1024) scan_tiles64_kernel(const uint32_t *tile_counts, uint32_t n_tiles, uint64_t *tile_base,
                                                            uint64_t *n_out) {
    __shared__ uint64_t warp_sums[32];
    __shared__ uint64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint64_t v = i < n_tiles ? tile_counts[i] : 0;
        uint64_t x = v;
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sums[w] = x;
        __syncthreads();
        uint64_t wp = 0;
        for (int k2 = 0; k2 < w; k2++) wp += warp_sums[k2];
        uint64_t excl = carry + wp + x - v;
        if (i < n_tiles) tile_base[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_out = carry;
}

__global__ void __launch_bounds__(256) unique_write_kernel(const uint64_t *__restrict__ k, uint64_t n, int min_count,
                                                           const uint64_t *__restrict__ tile_base, uint64_t *__restrict__ out) {
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t running;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint64_t base = (uint64_t)blockIdx.x * 4096;
    uint64_t obase = tile_base[blockIdx.x];
    for (int r = 0; r < 16; r++) {  // rows in order => output stays sorted
        uint64_t i = base + r * 256 + threadIdx.x;
        bool kp = i < n && keep_key(k, i, n, min_count);
        uint32_t bal = __ballot_sync(0xFFFFFFFFu, kp);
        if (lane == 0) warp_sums[w] = __popc(bal);
        __syncthreads();
        uint32_t wp = 0, tot = 0;
        for (int q = 0; q < 8; q++) { uint32_t s = warp_sums[q]; if (q < w) wp += s; tot += s; }
        if (kp) out[obase + running + wp + __popc(bal & ((1u << lane) - 1))] = k[i];
        __syncthreads();
        if (threadIdx.x == 0) running += tot;
        __syncthreads();
    }
}

int unique_keys(kssd_ctx *c, const uint64_t *d_sorted, uint64_t n, int min_count, uint64_t *d_out_keys, uint64_t *n_out) {
    *n_out = 0;
    if (n == 0) return KSSD_OK;
    uint32_t n_tiles = (uint32_t)((n + 4095) / 4096);
    KSSD_TRY(c->d_sort_counts.reserve((uint64_t)n_tiles * 12 + 64, c->stream));
    uint64_t *tile_base = c->d_sort_counts.as<uint64_t>();
    uint64_t *d_n = tile_base + n_tiles;
    uint32_t *tile_counts = reinterpret_cast<uint32_t *>(d_n + 1);
    unique_count_kernel<<<n_tiles, 256, 0, c->stream>>>(d_sorted, n, min_count, tile_counts);
    scan_tiles64_kernel<<<1, 1024, 0, c->stream>>>(tile_counts, n_tiles, tile_base, d_n);
    unique_write_kernel<<<n_tiles, 256, 0, c->stream>>>(d_sorted, n, min_count, tile_base, d_out_keys);
    KSSD_CUDA(cudaGetLastError());
    c->launches += 3;
    KSSD_TRY(c->h_scalar.reserve(64));
    KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, d_n, 8, cudaMemcpyDeviceToHost, c->stream));
    KSSD_CUDA(cudaStreamSynchronize(c->stream));
    *n_out = *c->h_scalar.as<uint64_t>();
    return KSSD_OK;
}

// keys sorted unique (genome<<32|hash) -> hashes[n], genome_off[n_genomes+1] (lower bound of genome g)
__global__ void split_hash_kernel(const uint64_t *__restrict__ k, uint64_t n, uint32_t *__restrict__ hashes) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) hashes[i] = (uint32_t)k[i];
}
__global__ void genome_off_kernel(const uint64_t *__restrict__ k, uint64_t n, uint32_t n_genomes, uint64_t *__restrict__ off) {
    uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g > n_genomes) return;
    uint64_t target = g << 32, lo = 0, hi = n;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (k[mid] < target) lo = mid + 1; else hi = mid;
    }
    off[g] = lo;
}

int split_genomes(kssd_ctx *c, const uint64_t *d_keys, uint64_t n, uint32_t n_genomes, uint32_t *d_hashes, uint64_t *d_genome_off) {
    if (n) split_hash_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_keys, n, d_hashes);
    genome_off_kernel<<<(n_genomes + 1 + 255) / 256, 256, 0, c->stream>>>(d_keys, n, n_genomes, d_genome_off);
    KSSD_CUDA(cudaGetLastError());
    c->launches += n ? 2 : 1;
    return KSSD_OK;
}

}  // namespace kssd
