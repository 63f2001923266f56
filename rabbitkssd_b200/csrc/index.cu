// Inverted-index build on the device: hash -> ascending genome ids, kept as CSR over the DISTINCT hashes.
// Replaces transSketches' 32-bit path (src/sketch.cpp:970-1017), which allocates 2^(4(k-l)) std::vectors and
// sweeps them serially (a fixed 14-19 s for L3K10).  Here: expand (hash,genome) pairs, one
// stable radix sort on the hash bits (input is genome-major, so ids stay ascending inside a posting list,
// matching the reference's push_back order sketch.cpp:979-985), then head detection + scan for the CSR.
#include <algorithm>

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

// Coarse directory over the sorted distinct hashes: a query hash is then found with ~5 binary-search steps inside
// its bucket instead of ~21 over the whole array (list_lookup_kernel).
static constexpr int kBucketBits = 16;
__global__ void __launch_bounds__(256) bucket_start_kernel(const uint32_t *__restrict__ uniq_hash, uint64_t n_uniq, int shift,
                                                           uint32_t *__restrict__ bucket) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > (1u << kBucketBits)) return;
    uint64_t lo = 0, hi = n_uniq;  // first u with (uniq_hash[u] >> shift) >= b
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if ((uniq_hash[mid] >> shift) < b) lo = mid + 1; else hi = mid;
    }
    bucket[b] = (uint32_t)lo;
}

int index_build_buckets(kssd_ctx *c) {
    cudaStream_t st = c->stream;
    KSSD_TRY(c->d_bucket.reserve(((1u << kBucketBits) + 1) * 4, st));
    uint32_t mx = 0;
    if (c->n_uniq) {
        KSSD_TRY(c->h_scalar.reserve(64));
        KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, c->d_uniq_hash.as<uint32_t>() + (c->n_uniq - 1), 4, cudaMemcpyDeviceToHost, st));
        KSSD_CUDA(cudaStreamSynchronize(st));
        mx = *c->h_scalar.as<uint32_t>();
    }
    int bits = 0;
    while (bits < 32 && (mx >> bits)) bits++;
    c->bucket_shift = bits > kBucketBits ? bits - kBucketBits : 0;
    bucket_start_kernel<<<((1u << kBucketBits) + 1 + 255) / 256, 256, 0, st>>>(c->d_uniq_hash.as<uint32_t>(), c->n_uniq, c->bucket_shift,
                                                                              c->d_bucket.as<uint32_t>());
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

int index_build_device(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total) {
    cudaStream_t st = c->stream;
    c->have_index = false;
    c->n_ref = n_genomes;
    c->n_post = total;
    KSSD_TRY(c->d_ref_sizes.reserve((uint64_t)(n_genomes + 2) * 4, st));
    uint32_t *d_max = c->d_ref_sizes.as<uint32_t>() + n_genomes;
    KSSD_CUDA(cudaMemsetAsync(d_max, 0, 4, st));
    if (n_genomes) {
        ref_sizes_kernel<<<(n_genomes + 255) / 256, 256, 0, st>>>(d_genome_off, n_genomes, c->d_ref_sizes.as<uint32_t>(), d_max);
        c->launches++;
    }
    KSSD_TRY(c->d_pairs.reserve(std::max<uint64_t>(total, 1) * 8, st));
    KSSD_TRY(c->d_pairs_tmp.reserve(std::max<uint64_t>(total, 1) * 8, st));
    KSSD_TRY(c->d_dict.reserve(std::max<uint64_t>(total, 1) * 4, st));
    KSSD_TRY(c->d_uniq_hash.reserve(std::max<uint64_t>(total, 1) * 4, st));
    KSSD_TRY(c->d_uniq_off.reserve((total + 1) * 8, st));
    uint64_t n_uniq = 0;
    if (total) {
        make_pairs_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_hashes, d_genome_off, n_genomes, total,
                                                                          c->d_pairs.as<uint64_t>());
        c->launches++;
        int hash_bits = c->have_params ? 4 * (c->sp.half_k - c->sp.drlevel) : 32;
        if (hash_bits > 32) hash_bits = 32;
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
    KSSD_TRY(index_build_buckets(c));
    c->have_index = true;
    return KSSD_OK;
}

}  // namespace kssd
