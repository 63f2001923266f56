// Set algebra on sketches: the arithmetic of `rabbit_kssd union` and `rabbit_kssd sub`.
//
// The reference keeps a bitmap over the whole hash space (2^32 or 2^(4(k-l)) bits, one copy per consumer thread):
// union = OR all hashes of all genomes into it and list the set bits (src/subCommand.cpp:307-543); sub = build the
// bitmap of the reference sketches, then drop from every query genome the hashes whose bit is set (:545-794).
// Here: union = one device-wide radix sort + unique of all hashes (hash64.cu: dict_build), sub = one binary search
// per query hash in that sorted set, compaction per genome in shared memory order (the surviving hashes keep their
// order inside a genome, as the reference's filter loop does).
#include <algorithm>

#include "common.cuh"

namespace kssd {

__global__ void __launch_bounds__(256) widen_kernel(const uint32_t *__restrict__ in, uint64_t n, uint64_t *__restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = in[i];
}
__global__ void __launch_bounds__(256) narrow_kernel(const uint64_t *__restrict__ in, uint64_t n, uint32_t *__restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = (uint32_t)in[i];
}

template <typename H> __device__ __forceinline__ bool in_sorted_set(const H *__restrict__ set, uint64_t n, H v) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(set + mid) < v) lo = mid + 1; else hi = mid;
    }
    return lo < n && __ldg(set + lo) == v;
}

// One CTA per query genome: the hashes that are NOT in the reference set, order kept, written to the head of the genome's
// own range of `tmp`; their number to kept[g].
template <typename H>
__global__ void __launch_bounds__(256) subtract_kernel(const H *__restrict__ q, const uint64_t *__restrict__ q_off, const H *__restrict__ ref, uint64_t n_ref,
                                                       H *__restrict__ tmp, uint32_t *__restrict__ kept) {
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t running;
    const uint32_t g = blockIdx.x;
    const uint64_t b0 = q_off[g];
    const uint32_t n = (uint32_t)(q_off[g + 1] - b0);
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t c0 = 0; c0 < n; c0 += blockDim.x) {
        const uint32_t i = c0 + threadIdx.x;
        bool kp = false;
        H v = 0;
        if (i < n) {
            v = q[b0 + i];
            kp = !in_sorted_set<H>(ref, n_ref, v);
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, kp);
        if (lane == 0) warp_sums[w] = __popc(bal);
        __syncthreads();
        uint32_t wp = 0, tot = 0;
        for (int k = 0; k < 8; k++) { const uint32_t x = warp_sums[k]; if (k < w) wp += x; tot += x; }
        if (kp) tmp[b0 + running + wp + __popc(bal & ((1u << lane) - 1))] = v;
        __syncthreads();
        if (threadIdx.x == 0) running += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) kept[g] = running;
}

template <typename H>
__global__ void __launch_bounds__(128) subtract_pack_kernel(const H *__restrict__ tmp, const uint64_t *__restrict__ q_off, const uint64_t *__restrict__ out_off,
                                                            H *__restrict__ out) {
    const uint32_t g = blockIdx.x;
    const uint64_t o0 = out_off[g];
    const uint32_t n = (uint32_t)(out_off[g + 1] - o0);
    const H *src = tmp + q_off[g];
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) out[o0 + i] = src[i];
}

__global__ void segment_scan_kernel(const uint32_t *counts, uint32_t n, uint64_t *start, uint32_t *max_out);  // segsort.cu

static unsigned setop_grid(const kssd_ctx *c, uint64_t n) { return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)c->sm_count * 16)); }

// hashes (host, hash_bytes = 4 | 8) -> sorted distinct set in c->h_setop (host), *n_out entries of the same width
int set_union_host(kssd_ctx *c, const void *hashes, uint64_t n, int hash_bytes, uint64_t *n_out) {
    cudaStream_t st = c->stream;
    *n_out = 0;
    KSSD_TRY(c->d_setop_in.reserve(std::max<uint64_t>(n, 1) * 8, st));
    uint64_t *d_vals = c->d_setop_in.as<uint64_t>();
    if (n && hash_bytes == 4) {  // staged at the tail of the buffer, widened in place from the front
        uint32_t *d32 = reinterpret_cast<uint32_t *>(d_vals) + n;
        KSSD_CUDA(cudaMemcpyAsync(d32, hashes, n * 4, cudaMemcpyHostToDevice, st));
        KSSD_TRY(c->d_setop_tmp.reserve(n * 8, st));
        widen_kernel<<<setop_grid(c, n), 256, 0, st>>>(d32, n, c->d_setop_tmp.as<uint64_t>());
        KSSD_CUDA(cudaGetLastError());
        c->launches++;
        d_vals = c->d_setop_tmp.as<uint64_t>();
    } else if (n) {
        KSSD_CUDA(cudaMemcpyAsync(d_vals, hashes, n * 8, cudaMemcpyHostToDevice, st));
    }
    uint64_t n_set = 0;
    KSSD_TRY(dict_build(c, d_vals, n, hash_bytes == 4 ? 32 : 64, c->d_setop_out, &n_set));
    KSSD_TRY(c->h_setop.reserve(std::max<uint64_t>(n_set, 1) * 8));
    if (n_set) {
        if (hash_bytes == 4) {
            KSSD_TRY(c->d_setop_tmp.reserve(n_set * 4, st));
            narrow_kernel<<<setop_grid(c, n_set), 256, 0, st>>>(c->d_setop_out.as<uint64_t>(), n_set, c->d_setop_tmp.as<uint32_t>());
            KSSD_CUDA(cudaGetLastError());
            c->launches++;
            KSSD_CUDA(cudaMemcpyAsync(c->h_setop.p, c->d_setop_tmp.p, n_set * 4, cudaMemcpyDeviceToHost, st));
        } else {
            KSSD_CUDA(cudaMemcpyAsync(c->h_setop.p, c->d_setop_out.p, n_set * 8, cudaMemcpyDeviceToHost, st));
        }
    }
    KSSD_CUDA(cudaStreamSynchronize(st));
    *n_out = n_set;
    return KSSD_OK;
}

template <typename H>
static int set_subtract_t(kssd_ctx *c, const H *q, const uint64_t *q_off, uint32_t n_q, const H *ref, uint64_t n_ref) {
    cudaStream_t st = c->stream;
    const uint64_t total = q_off[n_q];
    KSSD_TRY(c->d_setop_in.reserve(std::max<uint64_t>(total, 1) * sizeof(H), st));
    KSSD_TRY(c->d_setop_tmp.reserve(std::max<uint64_t>(total, 1) * sizeof(H), st));
    KSSD_TRY(c->d_setop_out.reserve(std::max<uint64_t>(total, 1) * sizeof(H), st));
    KSSD_TRY(c->d_h64_a.reserve(std::max<uint64_t>(n_ref, 1) * sizeof(H), st));
    KSSD_TRY(c->d_q_off.reserve((uint64_t)(n_q + 1) * 8, st));
    KSSD_TRY(c->d_seg_start.reserve((uint64_t)(n_q + 1) * 8, st));
    KSSD_TRY(c->d_ucount.reserve((uint64_t)std::max<uint32_t>(n_q, 1) * 4, st));
    if (total) KSSD_CUDA(cudaMemcpyAsync(c->d_setop_in.p, q, total * sizeof(H), cudaMemcpyHostToDevice, st));
    if (n_ref) KSSD_CUDA(cudaMemcpyAsync(c->d_h64_a.p, ref, n_ref * sizeof(H), cudaMemcpyHostToDevice, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_q_off.p, q_off, (uint64_t)(n_q + 1) * 8, cudaMemcpyHostToDevice, st));
    uint64_t *out_off = c->d_seg_start.as<uint64_t>();
    if (n_q) {
        subtract_kernel<H><<<n_q, 256, 0, st>>>(c->d_setop_in.as<H>(), c->d_q_off.as<uint64_t>(), c->d_h64_a.as<H>(), n_ref, c->d_setop_tmp.as<H>(),
                                              c->d_ucount.as<uint32_t>());
        segment_scan_kernel<<<1, 1024, 0, st>>>(c->d_ucount.as<uint32_t>(), n_q, out_off, nullptr);
        subtract_pack_kernel<H><<<n_q, 128, 0, st>>>(c->d_setop_tmp.as<H>(), c->d_q_off.as<uint64_t>(), out_off, c->d_setop_out.as<H>());
        KSSD_CUDA(cudaGetLastError());
        c->launches += 3;
    } else {
        KSSD_CUDA(cudaMemsetAsync(out_off, 0, 8, st));
    }
    KSSD_TRY(c->h_setop.reserve(std::max<uint64_t>(total, 1) * sizeof(H)));
    KSSD_TRY(c->h_genome_off.reserve((uint64_t)(n_q + 1) * 8));
    KSSD_CUDA(cudaMemcpyAsync(c->h_genome_off.p, out_off, (uint64_t)(n_q + 1) * 8, cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    const uint64_t n_keep = c->h_genome_off.as<uint64_t>()[n_q];
    if (n_keep) KSSD_CUDA(cudaMemcpyAsync(c->h_setop.p, c->d_setop_out.p, n_keep * sizeof(H), cudaMemcpyDeviceToHost, st));
    KSSD_CUDA(cudaStreamSynchronize(st));
    return KSSD_OK;
}

int set_subtract_host(kssd_ctx *c, const void *q, const uint64_t *q_off, uint32_t n_q, const void *ref, uint64_t n_ref, int hash_bytes) {
    return hash_bytes == 4 ? set_subtract_t<uint32_t>(c, static_cast<const uint32_t *>(q), q_off, n_q, static_cast<const uint32_t *>(ref), n_ref)
                           : set_subtract_t<uint64_t>(c, static_cast<const uint64_t *>(q), q_off, n_q, static_cast<const uint64_t *>(ref), n_ref);
}

}  // namespace kssd
