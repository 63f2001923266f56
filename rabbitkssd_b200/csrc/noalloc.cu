// kssd_b200_set_subtract_noalloc: the arithmetic of `rabbit_kssd sub` (setops.cu) for callers that cannot allocate --
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The allocating entry point (kssd_b200_set_subtract) stages host buffers into grow-only device buffers.  This one takes
// DEVICE pointers for every input and output and keeps its two pieces of scratch -- the per-genome compaction area and the
// per-genome survivor counts -- in ONE file-scope __device__ array of fixed size (below): every GPU that loads the library
// holds these 16.25 MiB until the process exits.  It calls no cudaMalloc / cudaFree / allocating library routine.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at zero:
// the compaction area is written before it is read (each genome's CTA fills [q_off[g], q_off[g] + kept[g]) and the pack
// kernel reads exactly that range), and every count kept[g], g < n_query, is stored by its CTA before the prefix sum runs.
// Calls that overlap on the GPU (two streams, two contexts on the same device) RACE on the array: serialise them.
#include <mutex>

#include "common.cuh"

namespace kssd {

static constexpr uint64_t kNoAllocMaxHashes = 4ull << 20;   // query hashes per call (uint32): 16 MiB of compaction area
static constexpr uint32_t kNoAllocMaxGenomes = 64u << 10;   // query genomes per call: 256 KiB of counts
static constexpr size_t kNoAllocTmpBytes = kNoAllocMaxHashes * sizeof(uint32_t);
static constexpr size_t kNoAllocBytes = kNoAllocTmpBytes + (size_t)kNoAllocMaxGenomes * sizeof(uint32_t);

// In HOST code this name is a stand-in for the symbol, not a device address: kernels get the pointer that
// cudaGetSymbolAddress returns (noalloc_scratch below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// the array's device address on the CURRENT device, looked up once per device
static int noalloc_scratch(int device, unsigned char **out) {
    static std::mutex mu;
    static unsigned char *addr[64] = {};
    if (device < 0 || device >= 64) { set_error("set_subtract_noalloc: device %d out of range", device); return KSSD_E_ARG; }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // the caller made `device` current (check_ctx)
        if (e != cudaSuccess || !p) {
            set_error("cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", device, cudaGetErrorString(e));
            return KSSD_E_CUDA;
        }
        addr[device] = static_cast<unsigned char *>(p);
    }
    *out = addr[device];
    return KSSD_OK;
}

}  // namespace kssd

using namespace kssd;

extern "C" int kssd_b200_set_subtract_noalloc(kssd_ctx *c, const uint32_t *d_q_hashes, const uint64_t *d_q_off, uint32_t n_query,
                                              uint64_t n_hashes, const uint32_t *d_ref_set, uint64_t n_ref, uint32_t *d_out_hashes,
                                              uint64_t *d_out_off) {
    KSSD_TRY(check_ctx(c));
    if (!d_q_off || !d_out_off || (n_hashes && (!d_q_hashes || !d_out_hashes)) || (n_ref && !d_ref_set)) {
        set_error("set_subtract_noalloc: bad arguments");
        return KSSD_E_ARG;
    }
    if (n_hashes > kNoAllocMaxHashes || n_query > kNoAllocMaxGenomes) {
        set_error("set_subtract_noalloc: %llu query hashes in %u genomes exceed the fixed scratch (at most %llu hashes, %u genomes per call); "
                  "use kssd_b200_set_subtract",
                  (unsigned long long)n_hashes, n_query, (unsigned long long)kNoAllocMaxHashes, kNoAllocMaxGenomes);
        return KSSD_E_ARG;
    }
    unsigned char *scratch = nullptr;
    KSSD_TRY(noalloc_scratch(c->device, &scratch));
    uint32_t *tmp = reinterpret_cast<uint32_t *>(scratch);
    uint32_t *kept = reinterpret_cast<uint32_t *>(scratch + kNoAllocTmpBytes);
    cudaStream_t st = c->stream;
    if (n_query == 0) {
        KSSD_CUDA(cudaMemsetAsync(d_out_off, 0, sizeof(uint64_t), st));
        return KSSD_OK;
    }
    subtract_kernel<uint32_t><<<n_query, 256, 0, st>>>(d_q_hashes, d_q_off, d_ref_set, n_ref, tmp, kept);
    segment_scan_kernel<<<1, 1024, 0, st>>>(kept, n_query, d_out_off, nullptr);
    subtract_pack_kernel<uint32_t><<<n_query, 128, 0, st>>>(tmp, d_q_off, d_out_off, d_out_hashes);
    KSSD_CUDA(cudaGetLastError());
    c->launches += 3;
    return KSSD_OK;
}
