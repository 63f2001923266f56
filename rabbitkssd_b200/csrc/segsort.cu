// Segment sort in shared memory: the fast path of "per-genome sort + unique" (and of the index build's buckets).
//
// Replaces the reference's per-genome unordered_set / unordered_map containers (src/sketch.cpp:526-550, :832-849,
// count filter :704-706).  The scan kernel counts the keys it emits per genome, so the (genome, hash) keys can be
// dealt into per-genome segments with one pass (the scan kernel also records each key's rank inside its genome) and
// every segment is then sorted by ONE CTA entirely in shared memory (counting-bucket sort on the leading hash bits;
// bitonic network as fallback), deduplicated / count-filtered there and written back.  HBM traffic: 8 B + 4 B per key
// for the deal, 4 B + 4 B for the sort, 4 B + 4 B for the final packing -- against 6 radix passes x 48 B per key
// for the device-wide sort it replaces (radix.cu, still used when a segment does not fit shared memory).
#include <cstring>

#include "common.cuh"
#include "smem_sort.cuh"

namespace kssd {

static constexpr uint32_t kSegSortMaxKeys = 56 * 1024;  // u32 keys: 224 KB of shared memory

// Counting-bucket sort of n keys from global memory into shared memory: hashes are uniform, so NB >= n/4 buckets on
// the leading key bits hold a handful of keys each -- one shared-memory atomic to count, one to place, then a
// per-thread insertion sort of each tiny bucket, instead of the
// ~log^2(n)/2 compare-exchange rounds of the bitonic network (which remains the fallback when some bucket holds
// more than 32 keys, e.g. one hash shared by a thousand genomes).
// bucket(key) must be non-decreasing in key and < NB.  NB: power of two >= 32.  cnt: NB counters.
// On return dst[0, n) is sorted.
template <typename K, typename F>
__device__ __forceinline__ void count_sort_smem(const K *src, K *dst, uint32_t *cnt, uint32_t n, uint32_t NB, F bucket) {
    __shared__ uint32_t cs_warp[32];
    __shared__ uint32_t cs_max;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (uint32_t i = threadIdx.x; i < NB; i += blockDim.x) cnt[i] = 0;
    if (threadIdx.x == 0) cs_max = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) atomicAdd(cnt + bucket(src[i]), 1u);
    __syncthreads();
    {   // exclusive scan of the counters, in place: thread t owns `per` consecutive counters
        const uint32_t per = NB > blockDim.x ? NB / blockDim.x : 1;  // both powers of two
        const uint32_t first = threadIdx.x * per;
        uint32_t sum = 0, mx = 0;
        if (first < NB)
            for (uint32_t q = 0; q < per; q++) { const uint32_t v = cnt[first + q]; sum += v; mx = max(mx, v); }
        uint32_t x = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
        if (lane == 31) cs_warp[w] = x;
        if (lane == 0 && mx) atomicMax(&cs_max, mx);
        __syncthreads();
        uint32_t wp = 0;
        for (int q = 0; q < w; q++) wp += cs_warp[q];
        uint32_t run = wp + x - sum;
        if (first < NB)
            for (uint32_t q = 0; q < per; q++) { const uint32_t v = cnt[first + q]; cnt[first + q] = run; run += v; }
        __syncthreads();
    }
    if (cs_max > 32) {  // crowded bucket: sort the whole segment with the network instead
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
        __syncthreads();
        bitonic_sort_smem<K>(dst, n);
        return;
    }
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const K k = src[i];
        dst[atomicAdd(cnt + bucket(k), 1u)] = k;
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < NB; b += blockDim.x) {  // cnt[b] is now the END of bucket b: insertion sort inside
        const uint32_t beg = b ? cnt[b - 1] : 0, end = cnt[b];
        for (uint32_t i = beg + 1; i < end; i++) {
            const K k = dst[i];
            uint32_t j = i;
            while (j > beg && dst[j - 1] > k) { dst[j] = dst[j - 1]; j--; }
            dst[j] = k;
        }
    }
    __syncthreads();
}

// counts[n] -> start[n+1] (exclusive, 64-bit), *max_out = largest count.  Single CTA; a thread owns 16 consecutive
// counts per round, so 16 k segments need one round (two barriers).
__global__ void __launch_bounds__(1024) segment_scan_kernel(const uint32_t *__restrict__ counts, uint32_t n, uint64_t *__restrict__ start,
                                                            uint32_t *__restrict__ max_out) {
    constexpr int kItems = 16;
    __shared__ uint64_t warp_sums[32];
    __shared__ uint32_t warp_max[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t mx = 0;
    uint64_t carry = 0;  // kept identically by every thread
    for (uint32_t base = 0; base < n; base += 1024 * kItems) {
        const uint32_t first = base + threadIdx.x * kItems;
        uint32_t v[kItems];
        uint64_t sum = 0;
#pragma unroll
        for (int q = 0; q < kItems; q++) {
            v[q] = first + q < n ? counts[first + q] : 0;
            sum += v[q];
            mx = max(mx, v[q]);
        }
        uint64_t x = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sums[w] = x;
        __syncthreads();
        uint64_t wp = 0, tot = 0;
        for (int q = 0; q < 32; q++) { const uint64_t t = warp_sums[q]; if (q < w) wp += t; tot += t; }
        uint64_t run = carry + wp + x - sum;
#pragma unroll
        for (int q = 0; q < kItems; q++) {
            if (first + q < n) start[first + q] = run;
            run += v[q];
        }
        carry += tot;
        __syncthreads();
    }
    for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    if (lane == 0) warp_max[w] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        start[n] = carry;
        uint32_t m = 0;
        for (int q = 0; q < 32; q++) m = max(m, warp_max[q]);
        if (max_out) *max_out = m;
    }
}

// Deal the emitted (genome << 32 | hash) keys into per-genome segments: the scan kernel recorded the rank of every key
// inside its genome, so this is a plain permutation (no atomics).
__global__ void __launch_bounds__(256) genome_deal_kernel(const uint64_t *__restrict__ emit, const uint32_t *__restrict__ emit_rank, uint64_t n,
                                                          uint32_t n_genomes, const uint64_t *__restrict__ seg_start, uint32_t *__restrict__ seg) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t key = emit[i];
    const uint32_t g = (uint32_t)(key >> 32);
    if (g < n_genomes) seg[seg_start[g] + emit_rank[i]] = (uint32_t)key;
}

// One CTA per genome: sort the segment in shared memory, keep run heads whose run has >= min_count members
// (min_count = 1: plain set, sketch.cpp:526-529; FASTQ -n: sketch.cpp:704-706), write them back to the head of the
// segment and the number kept to ucount[genome].
__global__ void __launch_bounds__(1024) genome_sort_unique_kernel(uint32_t *__restrict__ seg, const uint64_t *__restrict__ seg_start,
                                                                  int min_count, uint32_t n_buckets /* 0: bitonic only */, int bucket_shift,
                                                                  uint32_t key_cap, uint32_t *__restrict__ ucount) {
    extern __shared__ uint32_t s_keys[];
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t running;
    const uint32_t g = blockIdx.x;
    const uint64_t b0 = seg_start[g];
    const uint32_t n = (uint32_t)(seg_start[g + 1] - b0);
    uint32_t *base = seg + b0;
    if (n == 0) {
        if (threadIdx.x == 0) ucount[g] = 0;
        return;
    }
    if (threadIdx.x == 0) running = 0;
    if (n_buckets) {
        const uint32_t last = n_buckets - 1;
        count_sort_smem<uint32_t>(base, s_keys, s_keys + key_cap, n, n_buckets, [=](uint32_t h) { return min(h >> bucket_shift, last); });
    } else {
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) s_keys[i] = base[i];
        __syncthreads();
        bitonic_sort_smem<uint32_t>(s_keys, n);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (uint32_t c0 = 0; c0 < n; c0 += blockDim.x) {
        const uint32_t i = c0 + threadIdx.x;
        bool kp = false;
        uint32_t v = 0;
        if (i < n) {
            v = s_keys[i];
            kp = i == 0 || s_keys[i - 1] != v;
            if (kp && min_count > 1) {
                const uint32_t j = i + (uint32_t)(min_count - 1);
                kp = j < n && s_keys[j] == v;
            }
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, kp);
        if (lane == 0) warp_sums[w] = __popc(bal);
        __syncthreads();
        uint32_t wp = 0, tot = 0;
        for (int q = 0; q < n_warps; q++) { const uint32_t x = warp_sums[q]; if (q < w) wp += x; tot += x; }
        if (kp) base[running + wp + __popc(bal & ((1u << lane) - 1))] = v;
        __syncthreads();
        if (threadIdx.x == 0) running += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) ucount[g] = running;
}

// Pack the kept heads of every segment into the final CSR (hashes, genome_off).
__global__ void __launch_bounds__(128) genome_pack_kernel(const uint32_t *__restrict__ seg, const uint64_t *__restrict__ seg_start,
                                                          const uint64_t *__restrict__ genome_off, uint32_t *__restrict__ hashes) {
    const uint32_t g = blockIdx.x;
    const uint64_t o0 = genome_off[g];
    const uint32_t n = (uint32_t)(genome_off[g + 1] - o0);
    const uint32_t *src = seg + seg_start[g];
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) hashes[o0 + i] = src[i];
}

// gcount -> seg_start (+ max count in d_cursor word 2); runs on the stream, no sync.
int genome_segments(kssd_ctx *c) {
    cudaStream_t st = c->stream;
    KSSD_TRY(c->d_seg_start.reserve((uint64_t)(c->n_genomes + 1) * 8, st));
    uint32_t *d_max = reinterpret_cast<uint32_t *>(c->d_cursor.as<unsigned long long>() + 2);
    segment_scan_kernel<<<1, 1024, 0, st>>>(c->d_gcount.as<uint32_t>(), c->n_genomes, c->d_seg_start.as<uint64_t>(), d_max);
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

int genome_sort_unique(kssd_ctx *c, const uint64_t *d_emit, uint64_t n_emit, uint32_t max_count, int min_count, uint64_t *n_keep) {
    cudaStream_t st = c->stream;
    const uint32_t ng = c->n_genomes;
    *n_keep = 0;
    KSSD_TRY(c->d_ucount.reserve((uint64_t)std::max<uint32_t>(ng, 1) * 4, st));
    KSSD_TRY(c->d_sort_tmp.reserve(std::max<uint64_t>(n_emit, 1) * 4, st));
    KSSD_TRY(c->d_hashes.reserve(std::max<uint64_t>(n_emit, 1) * 4, st));
    KSSD_TRY(c->d_genome_off.reserve((uint64_t)(ng + 1) * 8, st));
    if (ng == 0) return KSSD_OK;
    uint32_t *seg = c->d_sort_tmp.as<uint32_t>();
    uint64_t *seg_start = c->d_seg_start.as<uint64_t>();
    if (n_emit) {
        genome_deal_kernel<<<(unsigned)((n_emit + 255) / 256), 256, 0, st>>>(d_emit, c->d_emit_rank.as<uint32_t>(), n_emit, ng, seg_start, seg);
        c->launches++;
    }
    const uint64_t attr_bit = 1ull << (41);
    if (!(c->attr_mask & attr_bit)) {
        KSSD_CUDA(cudaFuncSetAttribute(genome_sort_unique_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kSegSortMaxKeys * 4)));
        c->attr_mask |= attr_bit;
    }
    const int threads = max_count <= 2048 ? 256 : max_count <= 8192 ? 512 : 1024;
    const uint32_t key_cap = std::max<uint32_t>(max_count, 1);
    // counting-bucket sort while keys + counters fit (<= 32 k keys); the largest segments use the network alone
    uint32_t nb = 0;
    int shift = 0;
    if (max_count <= 32768) {
        nb = 32;
        while (nb < max_count && nb < 8192) nb <<= 1;
        int lg = 0;
        while ((1u << lg) < nb) lg++;
        const int hash_bits = c->hash64 ? c->rank_bits : std::min(32, 4 * (c->sp.half_k - c->sp.drlevel));
        shift = std::max(0, hash_bits - lg);
    }
    const size_t smem = ((size_t)key_cap + nb) * 4;
    genome_sort_unique_kernel<<<ng, threads, smem, st>>>(seg, seg_start, min_count, nb, shift, key_cap, c->d_ucount.as<uint32_t>());
    uint64_t *goff = c->d_genome_off.as<uint64_t>();
    segment_scan_kernel<<<1, 1024, 0, st>>>(c->d_ucount.as<uint32_t>(), ng, goff, nullptr);
    genome_pack_kernel<<<ng, 128, 0, st>>>(seg, seg_start, goff, c->d_hashes.as<uint32_t>());
    KSSD_CUDA(cudaGetLastError());
    c->launches += 3;
    KSSD_TRY(c->h_scalar.reserve(64));
    KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, goff + ng, 8, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    *n_keep = *c->h_scalar.as<uint64_t>();
    return KSSD_OK;
}

// ---- multi-GPU gather: concatenate the sketch shards of all ranks -----------------------------------------------
// Block of rank r = [n_local sizes][hashes, genome-major][padding], `stride` u32 words.  The blocks are either the
// pieces of one NCCL all-gather (contiguous in local memory) or -- the path bench.py uses -- the ranks' OWN
// buffers read in place over NVLink through CUDA-IPC peer mappings (PartTable holds one pointer per rank): gather
// and concatenation are then ONE pass of peer loads, no collective moves the data.  No host round trip either:
// sizes -> counts (one array), segment_scan_kernel -> offsets, one CTA per genome copies its hashes.
struct PartTable { const uint32_t *p[kMaxPeers]; };

// Peer barrier without a collective: after its shard is complete, rank r stores its publish count into slot r of the flag
// array of EVERY rank (remote stores over NVLink, behind a system-scope fence); a reader spins on its OWN flag array until
// all slots have reached the step it is about to read.  flags == nullptr: the caller has synchronised the ranks already.
__device__ __forceinline__ void peer_wait(const volatile uint32_t *flags, uint32_t world, uint32_t step, unsigned long long *result) {
    if (flags && threadIdx.x == 0) {
        const long long t0 = clock64();
        for (uint32_t r = 0; r < world; r++)
            while ((int32_t)(flags[r] - step) < 0) {
                if (clock64() - t0 > (20ll << 30)) {  // ~10 s: a peer died; the host reports it instead of hanging
                    atomicMax(result + 1, 2ull);
                    r = world;
                    break;
                }
                __nanosleep(200);
            }
        __threadfence_system();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) concat_counts_kernel(const PartTable parts, uint32_t n_local, uint32_t n_total, uint32_t *__restrict__ counts,
                                                            const uint32_t *flags, uint32_t world, uint32_t step, unsigned long long *result) {
    peer_wait(flags, world, step, result);
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n_total) counts[g] = __ldcv(parts.p[g / n_local] + g % n_local);
}

// The hashes of rank r's genomes are one contiguous run both in its block and in the concatenated array (genome-major,
// rank order): blockIdx.y = rank, the CTAs of a rank stride over its run with coalesced loads from the peer's memory.
__global__ void __launch_bounds__(256) concat_copy_kernel(const PartTable parts, uint64_t stride, uint32_t n_local, uint32_t world,
                                                          const uint64_t *__restrict__ off, uint32_t *__restrict__ out,
                                                          unsigned long long *__restrict__ result /* [0] total, [1] overflow */) {
    const uint32_t r = blockIdx.y;
    const uint64_t r0 = off[(uint64_t)r * n_local], r1 = off[(uint64_t)(r + 1) * n_local];
    if (r1 - r0 > stride - n_local) {  // the shard of rank r did not fit its block: nothing of it is copied
        if (threadIdx.x == 0 && blockIdx.x == 0) atomicMax(result + 1, 1ull);
        return;
    }
    const uint32_t *src = parts.p[r] + n_local;
    const uint64_t n = r1 - r0, step = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * step < n; i += 4 * step) {  // four independent loads in flight per thread (NVLink latency)
        const uint32_t a = __ldcv(src + i), b = __ldcv(src + i + step), c2 = __ldcv(src + i + 2 * step), d = __ldcv(src + i + 3 * step);
        out[r0 + i] = a;
        out[r0 + i + step] = b;
        out[r0 + i + 2 * step] = c2;
        out[r0 + i + 3 * step] = d;
    }
    for (; i < n; i += step) out[r0 + i] = __ldcv(src + i);
    if (r == world - 1 && blockIdx.x == 0 && threadIdx.x == 0) result[0] = r1;
}

static int sketch_concat_table(kssd_ctx *c, const PartTable &parts, uint64_t stride, uint32_t n_local, uint32_t world, uint64_t *total, int *overflow,
                               const uint32_t *d_flags = nullptr, uint32_t step = 0) {
    cudaStream_t st = c->stream;
    const uint32_t n_total = n_local * world;
    const uint64_t result_word = ((uint64_t)n_total + 1) & ~1ull;  // [counts n_total][pad][result: total, overflow (u64 each)]
    KSSD_TRY(c->d_cat_counts.reserve(result_word * 4 + 16, st));
    KSSD_TRY(c->d_cat_off.reserve(((uint64_t)n_total + 1) * 8, st));
    KSSD_TRY(c->d_cat_hashes.reserve(std::max<uint64_t>((stride - n_local) * world, 1) * 4, st));  // upper bound: every block full
    unsigned long long *result = reinterpret_cast<unsigned long long *>(c->d_cat_counts.as<uint32_t>() + result_word);
    KSSD_CUDA(cudaMemsetAsync(result, 0, 16, st));
    concat_counts_kernel<<<(n_total + 255) / 256, 256, 0, st>>>(parts, n_local, n_total, c->d_cat_counts.as<uint32_t>(), d_flags, world, step, result);
    segment_scan_kernel<<<1, 1024, 0, st>>>(c->d_cat_counts.as<uint32_t>(), n_total, c->d_cat_off.as<uint64_t>(), nullptr);
    concat_copy_kernel<<<dim3((unsigned)std::max(1, c->sm_count * 4 / (int)world), world), 256, 0, st>>>(parts, stride, n_local, world, c->d_cat_off.as<uint64_t>(),
                                                                                                         c->d_cat_hashes.as<uint32_t>(), result);
    KSSD_CUDA(cudaGetLastError());
    c->launches += 3;
    KSSD_TRY(c->h_scalar.reserve(64));
    KSSD_CUDA(cudaMemcpyAsync(c->h_scalar.p, result, 16, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    *total = c->h_scalar.as<uint64_t>()[0];
    if (c->h_scalar.as<uint64_t>()[1] == 2) {
        set_error("sketch_gather_peers: a peer did not publish its shard within ~10 s (peer barrier timed out)");
        return KSSD_E_CUDA;
    }
    *overflow = c->h_scalar.as<uint64_t>()[1] != 0;
    c->cat_total = *total;
    c->cat_n = *overflow ? 0 : n_total;
    return KSSD_OK;
}

int sketch_concat(kssd_ctx *c, const uint32_t *d_parts, uint64_t stride, uint32_t n_local, uint32_t world, uint64_t *total, int *overflow) {
    if (world > kMaxPeers) { set_error("sketch_concat: at most %d ranks", kMaxPeers); return KSSD_E_ARG; }
    PartTable t;
    for (uint32_t r = 0; r < kMaxPeers; r++) t.p[r] = r < world ? d_parts + (uint64_t)r * stride : nullptr;
    return sketch_concat_table(c, t, stride, n_local, world, total, overflow);
}

// ---- peer buffers: every rank publishes its shard in its own memory, the others read it over NVLink ---------------
struct FlagTable { uint32_t *p[kMaxPeers]; };  // slot `rank` of every rank's flag array

__global__ void __launch_bounds__(256) publish_kernel(const uint32_t *__restrict__ hashes, const uint64_t *__restrict__ genome_off, uint32_t n_local,
                                                      uint64_t n_copy, uint32_t *__restrict__ block, uint32_t *done, const FlagTable flags, uint32_t world,
                                                      uint32_t step) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_local + n_copy; i += (uint64_t)gridDim.x * blockDim.x)
        block[i] = i < n_local ? (uint32_t)(genome_off[i + 1] - genome_off[i]) : hashes[i - n_local];
    // the last CTA to finish tells every rank that this shard is complete
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(done, 1u) == gridDim.x - 1) {
        *done = 0;
        __threadfence_system();
        for (uint32_t r = 0; r < world; r++) *reinterpret_cast<volatile uint32_t *>(flags.p[r]) = step;
    }
}

int peer_init(kssd_ctx *c, uint32_t n_local, uint64_t cap_hashes, uint8_t *handles_out) {
    PeerState &ps = c->peers;
    if (ps.own[0]) { set_error("peer_init: already initialised"); return KSSD_E_ARG; }
    ps.n_local = n_local;
    ps.stride = (uint64_t)n_local + cap_hashes;
    for (int b = 0; b < 2; b++) {
        // [sizes | hashes | 16 spare bytes | flags: one u32 per rank | done counter]; plain cudaMalloc: the whole allocation is what an IPC handle maps
        KSSD_CUDA(cudaMalloc(&ps.own[b], ps.stride * 4 + 16 + (kMaxPeers + 4) * 4));
        KSSD_CUDA(cudaMemset(reinterpret_cast<uint8_t *>(ps.own[b]) + ps.stride * 4, 0, 16 + (kMaxPeers + 4) * 4));
        cudaIpcMemHandle_t h;
        KSSD_CUDA(cudaIpcGetMemHandle(&h, ps.own[b]));
        static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
        memcpy(handles_out + 64 * b, &h, 64);
    }
    return KSSD_OK;
}

int peer_connect(kssd_ctx *c, const uint8_t *handles, uint32_t world, uint32_t rank) {
    PeerState &ps = c->peers;
    if (!ps.own[0] || world == 0 || world > kMaxPeers || rank >= world) { set_error("peer_connect: peer_init first; 1..%d ranks", kMaxPeers); return KSSD_E_ARG; }
    for (uint32_t r = 0; r < world; r++)
        for (int b = 0; b < 2; b++) {
            if (r == rank) { ps.peer[b][r] = ps.own[b]; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, handles + ((size_t)r * 2 + b) * 64, 64);
            void *p = nullptr;
            KSSD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
            ps.peer[b][r] = static_cast<uint32_t *>(p);
        }
    ps.world = world;
    ps.rank = rank;
    return KSSD_OK;
}

void peer_release(kssd_ctx *c) {
    PeerState &ps = c->peers;
    for (uint32_t r = 0; r < ps.world; r++)
        for (int b = 0; b < 2; b++)
            if (r != ps.rank && ps.peer[b][r]) cudaIpcCloseMemHandle(ps.peer[b][r]);
    for (int b = 0; b < 2; b++)
        if (ps.own[b]) cudaFree(ps.own[b]);
    ps = PeerState();
}

// the ctx's finished sketch (d_hashes, d_genome_off of sketch_finish) -> its own peer buffer; stream-ordered, no sync
int sketch_publish(kssd_ctx *c) {
    PeerState &ps = c->peers;
    if (!ps.world) { set_error("sketch_publish: peer_connect first"); return KSSD_E_ARG; }
    if (c->n_genomes != ps.n_local) { set_error("sketch_publish: %u genomes in the session, peer buffers hold %u", c->n_genomes, ps.n_local); return KSSD_E_ARG; }
    if (c->hash64) { set_error("sketch_publish: the peer gather carries 32-bit hashes only"); return KSSD_E_UNSUPPORTED; }
    ps.flip ^= 1;  // double buffer: the peers may still be reading the previous step's block
    ps.step++;
    const uint64_t n_copy = std::min<uint64_t>(c->sketch_total, ps.stride - ps.n_local);  // a larger shard shows as overflow at the readers
    const uint64_t n = ps.n_local + n_copy;
    FlagTable ft;
    for (uint32_t r = 0; r < kMaxPeers; r++) ft.p[r] = r < ps.world ? ps.peer[0][r] + ps.stride + 4 + ps.rank : nullptr;  // flags live behind buffer 0
    publish_kernel<<<(unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 8), 256, 0, c->stream>>>(c->d_hashes.as<uint32_t>(), c->d_genome_off.as<uint64_t>(),
                                                                                                ps.n_local, n_copy, ps.own[ps.flip],
                                                                                                ps.own[0] + ps.stride + 4 + kMaxPeers, ft, ps.world, ps.step);
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

// after a cross-rank barrier on the stream (every rank's publish has completed): pull + concatenate all shards
int sketch_gather_peers(kssd_ctx *c, uint64_t *total, int *overflow) {
    PeerState &ps = c->peers;
    if (!ps.world) { set_error("sketch_gather_peers: peer_connect first"); return KSSD_E_ARG; }
    PartTable t;
    for (uint32_t r = 0; r < kMaxPeers; r++) t.p[r] = r < ps.world ? ps.peer[ps.flip][r] : nullptr;
    // the concat kernels wait in place until every rank's flag says that its current shard is complete: no collective
    return sketch_concat_table(c, t, ps.stride, ps.n_local, ps.world, total, overflow, ps.own[0] + ps.stride + 4, ps.step);
}

}  // namespace kssd
