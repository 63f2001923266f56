// Inverted-index build on the device: hash -> ascending genome ids, kept as CSR over the DISTINCT hashes.
// Replaces transSketches' 32-bit path (src/sketch.cpp:970-1017), which allocates 2^(4(k-l)) std::vectors and
// sweeps them serially (a fixed 14-19 s for L3K10).  Here: expand (hash,genome) pairs, one
// stable radix sort on the hash bits (input is genome-major, so ids stay ascending inside a posting list,
// matching the reference's push_back order sketch.cpp:979-985), then head detection + scan for the CSR.
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace kssd {

__global__ void scan_tiles64_kernel(const uint32_t *tile_counts, uint32_t n_tiles, uint64_t *tile_base, uint64_t *n_out);

// one thread per element: genome by binary search in genome_off
__global__ void __launch_bounds__(256) make_pairs_kernel(const uint32_t *__restrict__ hashes, const uint64_t *__restrict__ genome_off,
                                                         uint32_t n_genomes, uint64_t total, uint64_t *__restrict__ pairs) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    uint32_t lo = 0, hi = n_genomes;  // largest g with genome_off[g] <= i
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (genome_off[mid] <= i) lo = mid; else hi = mid;
    }
    pairs[i] = ((uint64_t)hashes[i] << 32) | lo;
}

__global__ void __launch_bounds__(256) ref_sizes_kernel(const uint64_t *__restrict__ genome_off, uint32_t n, uint32_t *__restrict__ sizes,
                                                        uint32_t *__restrict__ max_size) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = 0;
    if (g < n) { s = (uint32_t)(genome_off[g + 1] - genome_off[g]); sizes[g] = s; }
    for (int o = 16; o; o >>= 1) s = max(s, __shfl_xor_sync(0xFFFFFFFFu, s, o));
    if ((threadIdx.x & 31) == 0 && s) atomicMax(max_size, s);
}

__device__ __forceinline__ bool is_head(const uint64_t *p, uint64_t i) { return i == 0 || (p[i] >> 32) != (p[i - 1] >> 32); }

__global__ void __launch_bounds__(256) head_count_kernel(const uint64_t *__restrict__ p, uint64_t n, uint32_t *__restrict__ tile_counts,
                                                         uint32_t *__restrict__ dict) {
    __shared__ uint32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * 4096;
    uint32_t mine = 0;
    for (int r = 0; r < 16; r++) {
        uint64_t i = base + r * 256 + threadIdx.x;
        if (i < n) {
            dict[i] = (uint32_t)p[i];
            if (is_head(p, i)) mine++;
        }
    }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&total, mine);
    __syncthreads();
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(256) head_write_kernel(const uint64_t *__restrict__ p, uint64_t n, const uint64_t *__restrict__ tile_base,
                                                         uint32_t *__restrict__ uniq_hash, uint64_t *__restrict__ uniq_off) {
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t running;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint64_t base = (uint64_t)blockIdx.x * 4096;
    uint64_t obase = tile_base[blockIdx.x];
    for (int r = 0; r < 16; r++) {
        uint64_t i = base + r * 256 + threadIdx.x;
        bool kp = i < n && is_head(p, i);
        uint32_t bal = __ballot_sync(0xFFFFFFFFu, kp);
        if (lane == 0) warp_sums[w] = __popc(bal);
        __syncthreads();
        uint32_t wp = 0, tot = 0;
        for (int q = 0; q < 8; q++) { uint32_t s = warp_sums[q]; if (q < w) wp += s; tot += s; }
        if (kp) {
            uint64_t u = obase + running + wp + __popc(bal & ((1u << lane) - 1));
            uniq_hash[u] = (uint32_t)(p[i] >> 32);
            uniq_off[u] = i;
        }
        __syncthreads();
        if (threadIdx.x == 0) running += tot;
        __syncthreads();
    }
}

__global__ void set_u64_kernel(uint64_t *p, const uint64_t *idx, uint64_t v) { p[*idx] = v; }

// Coarse directory over the sorted distinct hashes: bucket[b] = first distinct hash u with (hash >> shift) >= b, so a
// query hash is found with ~2 binary-search steps inside its bucket (list_lookup_kernel).  Built by one pass over
// the distinct hashes: entry u fills the buckets between its predecessor's and its own.
static constexpr int kBucketBits = 20;
__global__ void __launch_bounds__(256) bucket_fill_kernel(const uint32_t *__restrict__ uniq_hash, uint32_t n_uniq, int shift,
                                                          uint32_t n_buckets, uint32_t *__restrict__ bucket) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u > n_uniq) return;
    const uint32_t lo = u == 0 ? 0 : (uniq_hash[u - 1] >> shift) + 1;
    const uint32_t hi = u == n_uniq ? n_buckets : min(uniq_hash[u] >> shift, n_buckets);
    for (uint32_t b = lo; b <= hi; b++) bucket[b] = u;
}

// Few distinct hashes (small inputs: the smoke test, config 1): the kernel above would let one thread fill up to
// n_buckets / n_uniq entries serially; here every bucket finds its own entry by binary search instead.
__global__ void __launch_bounds__(256) bucket_search_kernel(const uint32_t *__restrict__ uniq_hash, uint32_t n_uniq, int shift, uint32_t n_buckets,
                                                            uint32_t *__restrict__ bucket) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > n_buckets) return;
    uint32_t lo = 0, hi = n_uniq;  // first u with (uniq_hash[u] >> shift) >= b
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((uniq_hash[mid] >> shift) < b) lo = mid + 1; else hi = mid;
    }
    bucket[b] = lo;
}

int index_build_buckets(kssd_ctx *c, int64_t known_max_hash) {
    cudaStream_t st = c->stream;
    if (c->n_uniq >= (1ull << 32)) { set_error("index: more than 2^32 distinct hashes"); return KSSD_E_UNSUPPORTED; }
    uint32_t mx = known_max_hash >= 0 ? (uint32_t)known_max_hash : 0;
    if (c->n_uniq && known_max_hash < 0) {
        KSSD_TRY(c->h_scalar.reserve(64));
        KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, c->d_uniq_hash.as<uint32_t>() + (c->n_uniq - 1), 4, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        mx = *c->h_scalar.as<uint32_t>();
    }
    int bits = 0;
    while (bits < 32 && (mx >> bits)) bits++;
    c->bucket_shift = bits > kBucketBits ? bits - kBucketBits : 0;
    c->n_buckets = (mx >> c->bucket_shift) + 1;  // <= 2^kBucketBits
    KSSD_TRY(c->d_bucket.reserve(((uint64_t)c->n_buckets + 1) * 4, st));
    if (c->n_uniq * 4 < c->n_buckets)
        bucket_search_kernel<<<(c->n_buckets + 1 + 255) / 256, 256, 0, st>>>(c->d_uniq_hash.as<uint32_t>(), (uint32_t)c->n_uniq, c->bucket_shift, c->n_buckets,
                                                                            c->d_bucket.as<uint32_t>());
    else
        bucket_fill_kernel<<<(unsigned)((c->n_uniq + 1 + 255) / 256), 256, 0, st>>>(c->d_uniq_hash.as<uint32_t>(), (uint32_t)c->n_uniq, c->bucket_shift,
                                                                                   c->n_buckets, c->d_bucket.as<uint32_t>());
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

// ---- fast path: hash buckets sorted in shared memory ------------------------------------------------------------
// (hash, genome) pairs are dealt into 2^B buckets by the top hash bits (histogram + cursors: two passes of global
// atomics over the L2-resident bucket table), and every bucket (~1 k pairs) is sorted by ONE CTA in shared memory
// on the full (hash << 32 | genome) key, which also gives the ascending genome ids inside a posting list
// (the reference's push_back order, sketch.cpp:979-985).  ~50 B of HBM traffic per posting instead of the
// 4 radix passes x 48 B of the device-wide sort below (kept for bucket overflow, i.e. heavily skewed hashes).
static constexpr uint32_t kBucketCapSmall = 3072, kBucketCapLarge = 8192;  // pairs per bucket: 24 KB / 64 KB of shared memory; the hash density falls
                                                                           // linearly (see range_begin), so the first buckets hold 2 x the mean of ~750-1000
#ifndef KSSD_SUB_BUCKETS
#define KSSD_SUB_BUCKETS 512  // measured: 512 and 256 are 7 % faster than 1024 for the index build (profiles/r01y_index_variants.txt)
#endif
#ifndef KSSD_SORT_THREADS
#define KSSD_SORT_THREADS 256
#endif
#ifndef KSSD_BUCKET_LOG
#define KSSD_BUCKET_LOG 10
#endif
static constexpr uint32_t kSubBuckets = KSSD_SUB_BUCKETS;                               // counting-sort buckets inside a bucket

// Both passes can be restricted (a) to the hashes in [h_lo, h_last] (multi-GPU: every rank builds the CSR of one hash
// range, index_build_range; buckets are then relative to h_lo) and (b) to the postings a block of alldist rows needs
// (index_build_rows): genomes >= g_begin only, and only hashes whose bit is set in `filter`, a hashed bitmap of the hashes
// that occur in the rows of the block (false positives only add lists that no row of the block streams).
struct PostingFilter {
    const uint32_t *bits;  // null: no filter
    uint32_t mask;         // bitmap size in bits - 1
    uint32_t g_begin;      // postings of genomes below this are dropped
};
__device__ __forceinline__ uint32_t filter_slot(uint32_t h, uint32_t mask) { return (h * 0x9E3779B1u) >> 3 & mask; }
__device__ __forceinline__ bool filter_pass(const PostingFilter &f, uint32_t h) {
    if (!f.bits) return true;
    const uint32_t s = filter_slot(h, f.mask);
    return (__ldg(f.bits + (s >> 5)) >> (s & 31)) & 1u;
}

// rows [row_begin, row_end) -> filter bits
__global__ void __launch_bounds__(256) filter_set_kernel(const uint32_t *__restrict__ hashes, const uint64_t *__restrict__ genome_off, uint32_t row_begin,
                                                         uint32_t row_end, uint32_t mask, uint32_t *__restrict__ bits) {
    const uint64_t e0 = genome_off[row_begin], e1 = genome_off[row_end];
    for (uint64_t e = e0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < e1; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t s = filter_slot(hashes[e], mask);
        atomicOr(bits + (s >> 5), 1u << (s & 31));
    }
}

// One thread per hash (the sketches are genome-major, so "genomes >= g_begin" is "elements >= genome_off[g_begin]").
// Counts the pairs per bucket and draws each pair's arrival rank inside its bucket (0xFFFFFFFF: the pair is not indexed).
__global__ void __launch_bounds__(256) bucket_hist_kernel(const uint32_t *__restrict__ hashes, const uint64_t *__restrict__ genome_off, uint64_t total,
                                                          int shift, uint32_t last_bucket, uint32_t h_lo, uint32_t h_last, const PostingFilter f,
                                                          uint32_t *__restrict__ bcount, uint32_t *__restrict__ rank) {
    const uint64_t i = (f.g_begin ? genome_off[f.g_begin] : 0) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const uint32_t h = hashes[i];
    uint32_t r = 0xFFFFFFFFu;
    if (h >= h_lo && h <= h_last && filter_pass(f, h)) r = atomicAdd(bcount + min((h - h_lo) >> shift, last_bucket), 1u);  // arrival order
    rank[i] = r;
}

// the slot inside the bucket was drawn by the histogram pass
__global__ void __launch_bounds__(128) bucket_deal_kernel(const uint32_t *__restrict__ hashes, const uint64_t *__restrict__ genome_off, int shift,
                                                          uint32_t last_bucket, uint32_t h_lo, uint32_t g_begin, const uint64_t *__restrict__ bstart,
                                                          const uint32_t *__restrict__ rank, uint64_t *__restrict__ pairs) {
    const uint32_t g = blockIdx.x;
    if (g < g_begin) return;
    const uint64_t e1 = genome_off[g + 1];
    for (uint64_t e = genome_off[g] + (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; e < e1; e += (uint64_t)gridDim.y * blockDim.x) {
        const uint32_t r = rank[e];
        if (r == 0xFFFFFFFFu) continue;
        const uint32_t h = hashes[e];
        const uint32_t b = min((h - h_lo) >> shift, last_bucket);
        pairs[bstart[b] + r] = ((uint64_t)h << 32) | g;
    }
}

// One CTA per bucket.  Writes the postings of the bucket (dict), its distinct hashes + list starts (compacted at the
// head of the bucket's range of tmp_hash / tmp_rel), their number, and folds the largest hash into *max_hash.
__global__ void __launch_bounds__(KSSD_SORT_THREADS) bucket_sort_kernel(const uint64_t *__restrict__ pairs, const uint64_t *__restrict__ bstart, int shift,
                                                          uint32_t h_lo, uint32_t cap, uint32_t *__restrict__ dict, uint32_t *__restrict__ tmp_hash,
                                                          uint32_t *__restrict__ tmp_rel, uint32_t *__restrict__ hcount,
                                                          uint32_t *__restrict__ flags /* [0] overflow, [1] max hash */) {
    extern __shared__ uint64_t s_pairs[];
    __shared__ uint32_t warp_sums[KSSD_SORT_THREADS / 32];
    __shared__ uint32_t running;
    const uint32_t b = blockIdx.x;
    const uint64_t b0 = bstart[b];
    const uint64_t n64 = bstart[b + 1] - b0;
    if (n64 == 0 || n64 > cap) {
        if (threadIdx.x == 0) {
            hcount[b] = 0;
            if (n64) atomicExch(flags, 1u);
        }
        return;
    }
    const uint32_t n = (uint32_t)n64;
    if (threadIdx.x == 0) running = 0;
    {   // sub-bucket = the hash bits below the bucket prefix (clamped: the last bucket also holds out-of-range hashes)
        constexpr int kSubLog = KSSD_SUB_BUCKETS == 2048 ? 11 : KSSD_SUB_BUCKETS == 1024 ? 10 : KSSD_SUB_BUCKETS == 512 ? 9 : 8;
        const int sub_shift = shift > kSubLog ? shift - kSubLog : 0;
        const uint32_t first = (uint32_t)(((uint64_t)b << shift) >> sub_shift);
        count_sort_smem<uint64_t>(pairs + b0, s_pairs, reinterpret_cast<uint32_t *>(s_pairs + cap), n, kSubBuckets,
                                  [=](uint64_t k) { return min((((uint32_t)(k >> 32) - h_lo) >> sub_shift) - first, kSubBuckets - 1); });
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t c0 = 0; c0 < n; c0 += blockDim.x) {
        const uint32_t i = c0 + threadIdx.x;
        bool head = false;
        uint64_t v = 0;
        if (i < n) {
            v = s_pairs[i];
            dict[b0 + i] = (uint32_t)v;
            head = i == 0 || (uint32_t)(s_pairs[i - 1] >> 32) != (uint32_t)(v >> 32);
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, head);
        if (lane == 0) warp_sums[w] = __popc(bal);
        __syncthreads();
        uint32_t wp = 0, tot = 0;
        for (int q = 0; q < KSSD_SORT_THREADS / 32; q++) { const uint32_t x = warp_sums[q]; if (q < w) wp += x; tot += x; }
        if (head) {
            const uint64_t at = b0 + running + wp + __popc(bal & ((1u << lane) - 1));
            tmp_hash[at] = (uint32_t)(v >> 32);
            tmp_rel[at] = i;
        }
        __syncthreads();
        if (threadIdx.x == 0) running += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        hcount[b] = running;
        atomicMax(flags + 1, (uint32_t)(s_pairs[n - 1] >> 32));
    }
}

__global__ void __launch_bounds__(128) bucket_pack_kernel(const uint64_t *__restrict__ bstart, const uint64_t *__restrict__ hstart,
                                                          const uint32_t *__restrict__ tmp_hash, const uint32_t *__restrict__ tmp_rel,
                                                          uint32_t n_buckets, uint32_t *__restrict__ uniq_hash,
                                                          uint64_t *__restrict__ uniq_off) {
    const uint32_t b = blockIdx.x;
    const uint64_t b0 = bstart[b], u0 = hstart[b];
    const uint32_t n = (uint32_t)(hstart[b + 1] - u0);
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) {
        uniq_hash[u0 + j] = tmp_hash[b0 + j];
        uniq_off[u0 + j] = b0 + tmp_rel[b0 + j];
    }
    if (b == n_buckets - 1 && threadIdx.x == 0) uniq_off[hstart[n_buckets]] = bstart[n_buckets];
}

__global__ void segment_scan_kernel(const uint32_t *counts, uint32_t n, uint64_t *start, uint32_t *max_out);  // segsort.cu

// Returns KSSD_OK with *done = false when a bucket overflowed (caller falls back to the radix path).
// [h_lo, h_last]: the hash range to index (0 .. 0xFFFFFFFF = everything); range_bits = bits of (h_last - h_lo) the
// buckets are cut from; expect = postings expected in the range (sizes the buckets); *n_post = postings found.
static int index_build_bucketed(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total,
                                uint32_t h_lo, uint32_t h_last, int range_bits, uint64_t expect, uint32_t *d_max_size, bool *done,
                                uint64_t *n_uniq, uint64_t *n_post, uint32_t *max_hash, PostingFilter filter = PostingFilter{nullptr, 0, 0}) {
    cudaStream_t st = c->stream;
    *done = false;
    int B = 0;
    while (B < 18 && (expect >> (B + KSSD_BUCKET_LOG)) > 0) B++;  // ~1 k pairs per bucket on average
    if (B > range_bits) B = range_bits;
    const int shift = range_bits - B;
    const uint32_t nb = 1u << B;
    // scratch: [bstart nb+1][hstart nb+1] u64, then [bcount nb][bcur nb][hcount nb][flags 4] u32
    KSSD_TRY(c->d_sort_counts.reserve(((uint64_t)nb + 1) * 16 + (uint64_t)nb * 12 + 16, st));
    uint64_t *bstart = c->d_sort_counts.as<uint64_t>();
    uint64_t *hstart = bstart + nb + 1;
    uint32_t *bcount = reinterpret_cast<uint32_t *>(hstart + nb + 1);
    uint32_t *bcur = bcount + nb, *hcount = bcur + nb, *flags = hcount + nb;
    KSSD_CUDA(cudaMemsetAsync(bcount, 0, (uint64_t)nb * 12 + 16, st));
    uint32_t *tmp_hash = c->d_pairs_tmp.as<uint32_t>();
    uint32_t *tmp_rel = tmp_hash + total;
    const uint64_t avg_len = total / n_genomes + 1;
    const uint32_t chunks = (uint32_t)std::min<uint64_t>(64, (avg_len + 255) / 256);
    // (a restricted build launches threads for all hashes; those past the end after the g_begin offset exit at once)
    bucket_hist_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_hashes, d_genome_off, total, shift, nb - 1, h_lo, h_last, filter, bcount,
                                                                       tmp_hash /* rank, dead before the sort */);
    segment_scan_kernel<<<1, 1024, 0, st>>>(bcount, nb, bstart, nullptr);
    bucket_deal_kernel<<<dim3(n_genomes, chunks), 128, 0, st>>>(d_hashes, d_genome_off, shift, nb - 1, h_lo, filter.g_begin, bstart, tmp_hash,
                                                                c->d_pairs.as<uint64_t>());
    KSSD_CUDA(cudaGetLastError());
    c->launches += 3;
    const uint64_t attr_bit = 1ull << (40);
    if (!(c->attr_mask & attr_bit)) {
        KSSD_CUDA(cudaFuncSetAttribute(bucket_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kBucketCapLarge * 8 + kSubBuckets * 4)));
        c->attr_mask |= attr_bit;
    }
    KSSD_TRY(c->h_scalar.reserve(64));
    uint64_t *hs = c->h_scalar.as<uint64_t>();
    const uint32_t *fl = reinterpret_cast<const uint32_t *>(hs + 1);
    // small shared-memory footprint first (more CTAs per SM); one retry with the large one when a bucket overflowed
    for (uint32_t cap : {kBucketCapSmall, kBucketCapLarge}) {
        KSSD_CUDA(cudaMemsetAsync(flags, 0, 16, st));
        bucket_sort_kernel<<<nb, KSSD_SORT_THREADS, (size_t)cap * 8 + kSubBuckets * 4, st>>>(c->d_pairs.as<uint64_t>(), bstart, shift, h_lo, cap, c->d_dict.as<uint32_t>(),
                                                                               tmp_hash, tmp_rel, hcount, flags);
        segment_scan_kernel<<<1, 1024, 0, st>>>(hcount, nb, hstart, nullptr);
        bucket_pack_kernel<<<nb, 128, 0, st>>>(bstart, hstart, tmp_hash, tmp_rel, nb, c->d_uniq_hash.as<uint32_t>(), c->d_uniq_off.as<uint64_t>());
        KSSD_CUDA(cudaGetLastError());
        c->launches += 3;
        KSSD_CUDA(cudaMemcpyAsync(hs, hstart + nb, 8, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaMemcpyAsync(hs + 1, flags, 8, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaMemcpyAsync(hs + 2, d_max_size, 4, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaMemcpyAsync(hs + 3, bstart + nb, 8, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        c->max_ref_size = *reinterpret_cast<const uint32_t *>(hs + 2);
        if (!fl[0]) break;
    }
    if (fl[0]) return KSSD_OK;  // a bucket does not fit shared memory: caller takes the radix path
    *n_uniq = hs[0];
    *n_post = hs[3];
    *max_hash = fl[1];
    *done = true;
    return KSSD_OK;
}

static int index_scratch(kssd_ctx *c, uint32_t n_genomes, uint64_t total, const uint64_t *d_genome_off, uint32_t **d_max) {
    cudaStream_t st = c->stream;
    KSSD_TRY(c->d_ref_sizes.reserve((uint64_t)(n_genomes + 2) * 4, st));
    *d_max = c->d_ref_sizes.as<uint32_t>() + n_genomes;
    KSSD_CUDA(cudaMemsetAsync(*d_max, 0, 4, st));
    if (n_genomes) {
        ref_sizes_kernel<<<(n_genomes + 255) / 256, 256, 0, st>>>(d_genome_off, n_genomes, c->d_ref_sizes.as<uint32_t>(), *d_max);
        c->launches++;
    }
    KSSD_TRY(c->d_pairs.reserve(std::max<uint64_t>(total, 1) * 8, st));
    KSSD_TRY(c->d_pairs_tmp.reserve(std::max<uint64_t>(total, 1) * 8, st));
    KSSD_TRY(c->d_dict.reserve(std::max<uint64_t>(total, 1) * 4, st));
    KSSD_TRY(c->d_uniq_hash.reserve(std::max<uint64_t>(total, 1) * 4, st));
    KSSD_TRY(c->d_uniq_off.reserve((total + 1) * 8, st));
    return KSSD_OK;
}

// ---- multi-GPU: one hash range per rank, the pieces are exchanged and adopted ---------------------------------
// Hash range of part p of `parts`: the leading hash bits are the outer bases of the CANONICAL k-mer (the smaller of
// the two strands, src/sketch.cpp:508), so their density falls linearly, F(x) = 2x - x^2; the boundaries
// x_p = 1 - sqrt(1 - p/parts) give every part the same expected number of postings.
static uint32_t range_begin(int hash_bits, uint32_t part, uint32_t parts) {
    if (part == 0) return 0;
    const double x = 1.0 - sqrt(1.0 - (double)part / (double)parts);
    const double span = hash_bits >= 32 ? 4294967296.0 : (double)(1ull << hash_bits);
    const double v = floor(x * span);
    return v >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)v;
}

// block layout (u32 words): [0] distinct hashes U, [1] postings P, [2] max hash, [3] reserved, then uniq_hash[U],
// list starts relative to the piece [U + 1], postings [P].
__global__ void __launch_bounds__(256) piece_pack_kernel(const uint32_t *__restrict__ uniq_hash, const uint64_t *__restrict__ uniq_off,
                                                         const uint32_t *__restrict__ dict, uint32_t U, uint32_t P, uint32_t max_hash,
                                                         uint32_t *__restrict__ block) {
    const uint64_t n = 4ull + U + (U + 1ull) + P;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t v;
        if (i < 4) v = i == 0 ? U : i == 1 ? P : i == 2 ? max_hash : 0u;
        else if (i < 4ull + U) v = uniq_hash[i - 4];
        else if (i < 5ull + 2ull * U) v = (uint32_t)uniq_off[i - 4 - U];
        else v = dict[i - 5 - 2ull * U];
        block[i] = v;
    }
}

int index_build_range(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total, uint32_t part,
                      uint32_t parts, uint32_t *d_block, uint64_t block_words, uint64_t *used_words, int *overflow) {
    cudaStream_t st = c->stream;
    *overflow = 0;
    *used_words = 0;
    if (block_words < 8) { set_error("index_build_range: block too small"); return KSSD_E_ARG; }
    if (total >= (1ull << 32) || n_genomes == 0) {
        *overflow = 1;
        KSSD_CUDA(cudaMemsetAsync(d_block, 0xFF, 16, st));
        return KSSD_OK;
    }
    uint32_t *d_max;
    KSSD_TRY(index_scratch(c, n_genomes, total, d_genome_off, &d_max));
    int hash_bits = c->have_params ? 4 * (c->sp.half_k - c->sp.drlevel) : 32;
    if (hash_bits > 32) hash_bits = 32;
    const uint32_t h_lo = range_begin(hash_bits, part, parts);
    const uint32_t h_last = part + 1 == parts ? 0xFFFFFFFFu : range_begin(hash_bits, part + 1, parts) - 1u;
    uint64_t U = 0, P = 0;
    uint32_t mh = 0;
    bool done = false;
    if (total && h_last >= h_lo) {
        const uint64_t width = part + 1 == parts ? (1ull << hash_bits) - std::min<uint64_t>(h_lo, 1ull << hash_bits) : (uint64_t)h_last - h_lo + 1;
        int range_bits = 1;
        while (range_bits < 32 && (1ull << range_bits) < width) range_bits++;
        KSSD_TRY(index_build_bucketed(c, d_hashes, d_genome_off, n_genomes, total, h_lo, h_last, range_bits, total / parts + 1, d_max, &done, &U,
                                      &P, &mh));
        if (!done) *overflow = 1;  // heavily skewed hashes: callers build the whole index on every rank instead
    }
    if (5ull + 2 * U + P > block_words) *overflow = 1;
    if (*overflow) {  // header 0xFFFFFFFF...: every rank sees it after the exchange (index_adopt_parts reports it)
        KSSD_CUDA(cudaMemsetAsync(d_block, 0xFF, 16, st));
        return KSSD_OK;
    }
    *used_words = 5ull + 2 * U + P;
    if (U == 0) KSSD_CUDA(cudaMemsetAsync(c->d_uniq_off.p, 0, 8, st));
    piece_pack_kernel<<<(unsigned)std::min<uint64_t>((5ull + 2 * U + P + 255) / 256, 148 * 16), 256, 0, st>>>(
        c->d_uniq_hash.as<uint32_t>(), c->d_uniq_off.as<uint64_t>(), c->d_dict.as<uint32_t>(), (uint32_t)U, (uint32_t)P, mh, d_block);
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

// pieces (parts blocks of block_words u32, hash ranges ascending) -> the ctx's CSR: one CTA row per piece
__global__ void __launch_bounds__(256) piece_adopt_kernel(const uint32_t *__restrict__ blocks, uint64_t block_words,
                                                          const uint64_t *__restrict__ base /* [2 * parts]: distinct-hash base, posting base */,
                                                          uint32_t *__restrict__ uniq_hash, uint64_t *__restrict__ uniq_off, uint32_t *__restrict__ dict) {
    const uint32_t r = blockIdx.y;
    const uint32_t *blk = blocks + (uint64_t)r * block_words;
    const uint32_t U = blk[0], P = blk[1];
    const uint64_t ub = base[2 * r], pb = base[2 * r + 1];
    const uint64_t n = (uint64_t)U + P;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (i < U) {
            uniq_hash[ub + i] = blk[4 + i];
            uniq_off[ub + i] = pb + blk[4 + U + i];
        } else {
            dict[pb + (i - U)] = blk[5 + 2ull * U + (i - U)];
        }
    }
}

int index_adopt_parts(kssd_ctx *c, const uint32_t *d_blocks, uint64_t block_words, uint32_t parts, const uint64_t *d_genome_off, uint32_t n_genomes,
                      int *overflow) {
    cudaStream_t st = c->stream;
    *overflow = 0;
    c->have_index = false;
    c->index_hash64 = false;
    if (parts == 0 || parts > 64) { set_error("index_adopt_parts: 1..64 parts"); return KSSD_E_ARG; }
    KSSD_TRY(c->h_scalar.reserve(64 * 16 + 64));
    uint32_t *hd = c->h_scalar.as<uint32_t>();  // 4 header words per part
    KSSD_CUDA(cudaMemcpy2DAsync(hd, 16, d_blocks, block_words * 4, 16, parts, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    uint64_t base[128];
    uint64_t U = 0, P = 0;
    uint32_t mh = 0;
    for (uint32_t r = 0; r < parts; r++) {
        const uint64_t u = hd[4 * r], p = hd[4 * r + 1];
        if (5ull + 2 * u + p > block_words) { *overflow = 1; return KSSD_OK; }  // the piece's rank reported overflow (header left as it was) or a bad block
        base[2 * r] = U;
        base[2 * r + 1] = P;
        U += u;
        P += p;
        if (u) mh = std::max(mh, hd[4 * r + 2]);
    }
    uint32_t *d_max;
    KSSD_TRY(index_scratch(c, n_genomes, P, d_genome_off, &d_max));
    KSSD_TRY(c->d_sort_counts.reserve(128 * 8, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_sort_counts.p, base, (size_t)parts * 16, cudaMemcpyHostToDevice, st));
    if (U + P) {
        piece_adopt_kernel<<<dim3(148 * 2, parts), 256, 0, st>>>(d_blocks, block_words, c->d_sort_counts.as<uint64_t>(), c->d_uniq_hash.as<uint32_t>(),
                                                                 c->d_uniq_off.as<uint64_t>(), c->d_dict.as<uint32_t>());
        c->launches++;
    }
    KSSD_CUDA(cudaMemcpyAsync(c->d_uniq_off.as<uint64_t>() + U, &P, 8, cudaMemcpyHostToDevice, st));  // P lives on this frame: synchronise below
    KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, d_max, 4, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    c->max_ref_size = *c->h_scalar.as<uint32_t>();
    c->n_ref = n_genomes;
    c->n_post = P;
    c->list_slots = P;
    c->index_row_begin = 0;
    c->index_row_end = 0xFFFFFFFFu;
    c->n_uniq = U;
    KSSD_TRY(index_build_buckets(c, U ? (int64_t)mh : -1));
    c->have_index = true;
    return KSSD_OK;
}

int index_build_device(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total,
                       int hash_bits_override, uint32_t row_begin, uint32_t row_end) {
    cudaStream_t st = c->stream;
    c->have_index = false;
    c->index_hash64 = false;
    c->n_ref = n_genomes;
    c->n_post = total;
    c->list_slots = total;
    const bool restricted = row_end != 0xFFFFFFFFu && !(row_begin == 0 && row_end >= n_genomes);
    c->index_row_begin = restricted ? row_begin : 0;
    c->index_row_end = restricted ? row_end : 0xFFFFFFFFu;
    uint32_t *d_max;
    KSSD_TRY(index_scratch(c, n_genomes, total, d_genome_off, &d_max));
    uint64_t n_uniq = 0;
    int64_t known_max = -1;
    bool done = false;
    int hash_bits = c->have_params ? 4 * (c->sp.half_k - c->sp.drlevel) : 32;
    if (hash_bits_override) hash_bits = hash_bits_override;  // 64-bit hash configuration: the "hashes" are dictionary ranks
    if (hash_bits > 32) hash_bits = 32;
    if (restricted) {
        // alldist rows [row_begin, row_end) only (multi-GPU row sharding): a row streams the lists of ITS hashes,
        // and of those only the ids above itself (src/dist.cpp:207) -- index just those postings.  Work ~ 1/ranks of the full
        // build, which every rank would otherwise repeat.
        if (row_begin >= row_end || row_end > n_genomes) { set_error("index_build_rows: bad row range"); return KSSD_E_ARG; }
        if (c->force_radix || total >= (1ull << 32)) { set_error("index_build_rows: needs the bucketed build (< 2^32 postings)"); return KSSD_E_UNSUPPORTED; }
        PostingFilter f;
        f.mask = (1u << 27) - 1;  // 16 MB hashed bitmap: L2 resident; ~1 % false positives at 1.5 M hashes
        f.g_begin = row_begin;
        KSSD_TRY(c->d_filter.reserve(((uint64_t)f.mask + 1) / 8, st));
        f.bits = c->d_filter.as<uint32_t>();
        KSSD_CUDA(cudaMemsetAsync(c->d_filter.p, 0, ((uint64_t)f.mask + 1) / 8, st));
        const uint64_t own = total / n_genomes * (row_end - row_begin) + 1;  // ~ hashes in the rows of the block
        filter_set_kernel<<<(unsigned)std::min<uint64_t>((own + 255) / 256, (uint64_t)c->sm_count * 8), 256, 0, st>>>(d_hashes, d_genome_off, row_begin, row_end,
                                                                                                              f.mask, c->d_filter.as<uint32_t>());
        KSSD_CUDA(cudaGetLastError());
        c->launches++;
        uint32_t mh = 0;
        uint64_t n_post = 0;
        if (total) KSSD_TRY(index_build_bucketed(c, d_hashes, d_genome_off, n_genomes, total, 0u, 0xFFFFFFFFu, hash_bits, 2 * own + 1024, d_max, &done, &n_uniq, &n_post, &mh, f));
        if (!done && total) {  // a crowded bucket (the estimate of the passing postings was far off): callers build the full index
            set_error("index_build_rows: bucket overflow");
            return KSSD_E_OVERFLOW;
        }
        known_max = mh;
        c->n_post = n_post;
    } else if (total && n_genomes && !c->force_radix && total < (1ull << 32)) {
        uint32_t mh = 0;
        uint64_t n_post = 0;
        KSSD_TRY(index_build_bucketed(c, d_hashes, d_genome_off, n_genomes, total, 0u, 0xFFFFFFFFu, hash_bits, total, d_max, &done, &n_uniq, &n_post, &mh));
        if (done) known_max = mh;
    }
    if (done) {
        // CSR is in place
    } else if (total) {
        make_pairs_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_hashes, d_genome_off, n_genomes, total,
                                                                          c->d_pairs.as<uint64_t>());
        c->launches++;
        uint64_t *sorted = nullptr;
        KSSD_TRY(radix_sort_u64(c, c->d_pairs.as<uint64_t>(), c->d_pairs_tmp.as<uint64_t>(), total, 0, 0, 32, 32 + hash_bits, &sorted));
        uint32_t n_tiles = (uint32_t)((total + 4095) / 4096);
        KSSD_TRY(c->d_sort_counts.reserve((uint64_t)n_tiles * 12 + 64, st));
        uint64_t *tile_base = c->d_sort_counts.as<uint64_t>();
        uint64_t *d_n = tile_base + n_tiles;
        uint32_t *tile_counts = reinterpret_cast<uint32_t *>(d_n + 1);
        head_count_kernel<<<n_tiles, 256, 0, st>>>(sorted, total, tile_counts, c->d_dict.as<uint32_t>());
        scan_tiles64_kernel<<<1, 1024, 0, st>>>(tile_counts, n_tiles, tile_base, d_n);
        head_write_kernel<<<n_tiles, 256, 0, st>>>(sorted, total, tile_base, c->d_uniq_hash.as<uint32_t>(), c->d_uniq_off.as<uint64_t>());
        set_u64_kernel<<<1, 1, 0, st>>>(c->d_uniq_off.as<uint64_t>(), d_n, total);
        KSSD_CUDA(cudaGetLastError());
        c->launches += 4;
        KSSD_TRY(c->h_scalar.reserve(64));
        KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, d_n, 8, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.as<uint64_t>() + 1, d_max, 4, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        n_uniq = c->h_scalar.as<uint64_t>()[0];
        c->max_ref_size = *reinterpret_cast<uint32_t *>(c->h_scalar.as<uint64_t>() + 1);
    } else {
        KSSD_CUDA(cudaMemsetAsync(c->d_uniq_off.p, 0, 8, st));
        c->max_ref_size = 0;
    }
    c->n_uniq = n_uniq;
    if (restricted) {  // alldist rows only: the work lists come from a walk over the lists, no hash is ever looked up
        c->n_buckets = 0;
        c->bucket_shift = 0;
    } else {
        KSSD_TRY(index_build_buckets(c, known_max));
    }
    c->have_index = true;
    return KSSD_OK;
}

}  // namespace kssd
