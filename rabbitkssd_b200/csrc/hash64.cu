// 64-bit hash configuration (half_k - drlevel > 8: src/sketch.cpp:336 use64, :904-969 transSketches, src/dist.cpp:35-82,
// :180-192, :442-487, :566-576).
//
// The reference switches every container to 64-bit keys and to hash maps (unordered_set<uint64_t> per genome,
// unordered_map<uint64_t, vector<uint32_t>> as the index).  Here the 64-bit hashes are DICTIONARY-ENCODED instead: the
// sorted distinct hashes of a sketch session (or of the sketches an index is built from) form a dictionary, every hash
// is replaced by its u32 rank in it, and the whole 32-bit machinery -- per-genome shared-memory sort + unique (+ count
// filter), bucketed index build, work lists, counters -- runs unchanged on the ranks: set semantics, posting lists and
// intersection counts depend on hash EQUALITY only, and rank order = hash order keeps every sorted output sorted.
// Cost over the 32-bit path: one device-wide radix sort of the hashes (8-bit digits over 4 * (half_k - drlevel) bits) and
// one binary search per hash.
#include <algorithm>

#include "common.cuh"

namespace kssd {

__device__ __forceinline__ uint32_t lower_bound64(const uint64_t *__restrict__ dict, uint64_t n, uint64_t v) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(dict + mid) < v) lo = mid + 1; else hi = mid;
    }
    return (uint32_t)lo;
}

__global__ void __launch_bounds__(256) rank64_kernel(const uint64_t *__restrict__ vals, uint64_t n, const uint64_t *__restrict__ dict,
                                                     uint64_t n_dict, uint32_t *__restrict__ ranks) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        ranks[i] = lower_bound64(dict, n_dict, vals[i]);
}

// emit[i]: hash -> (genome << 32 | rank of the hash): the key layout of the 32-bit configuration
__global__ void __launch_bounds__(256) encode64_kernel(unsigned long long *__restrict__ emit, const uint32_t *__restrict__ gen, uint64_t n,
                                                       const uint64_t *__restrict__ dict, uint64_t n_dict) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        emit[i] = ((unsigned long long)gen[i] << 32) | lower_bound64(dict, n_dict, emit[i]);
}

__global__ void __launch_bounds__(256) gather64_kernel(const uint32_t *__restrict__ idx, uint64_t n, const uint64_t *__restrict__ dict,
                                                       uint64_t *__restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = dict[idx[i]];
}

static unsigned grid_for(const kssd_ctx *c, uint64_t n) { return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)c->sm_count * 16)); }

// sorted distinct values of d_vals[0, n) -> dict (the input is left untouched)
int dict_build(kssd_ctx *c, const uint64_t *d_vals, uint64_t n, int hash_bits, DevBuf &dict, uint64_t *n_dict) {
    cudaStream_t st = c->stream;
    *n_dict = 0;
    KSSD_TRY(dict.reserve(std::max<uint64_t>(n, 1) * 8, st));
    if (n == 0) return KSSD_OK;
    KSSD_TRY(c->d_h64_a.reserve(n * 8, st));
    KSSD_TRY(c->d_h64_b.reserve(n * 8, st));
    KSSD_CUDA(cudaMemcpyAsync(c->d_h64_a.p, d_vals, n * 8, cudaMemcpyDeviceToDevice, st));
    uint64_t *sorted = nullptr;
    KSSD_TRY(radix_sort_u64(c, c->d_h64_a.as<uint64_t>(), c->d_h64_b.as<uint64_t>(), n, 0, std::min(64, std::max(8, hash_bits)), 0, 0, &sorted));
    return unique_keys(c, sorted, n, 1, dict.as<uint64_t>(), n_dict);
}

int dict_ranks(kssd_ctx *c, const uint64_t *d_vals, uint64_t n, const uint64_t *d_dict, uint64_t n_dict, uint32_t *d_ranks) {
    if (n == 0) return KSSD_OK;
    rank64_kernel<<<grid_for(c, n), 256, 0, c->stream>>>(d_vals, n, d_dict, n_dict, d_ranks);
    KSSD_CUDA(cudaGetLastError());
    c->launches++;
    return KSSD_OK;
}

static int bits_for(uint64_t n) {  // ranks are < n
    int b = 1;
    while (b < 32 && (1ull << b) < n) b++;
    return b;
}

// End of a sketch session in the 64-bit configuration: d_emit holds n_emit bare hashes, d_emit_gen their genomes.
int session_encode64(kssd_ctx *c, uint64_t n_emit, uint64_t *n_dict) {
    const int hash_bits = std::min(64, 4 * (c->sp.half_k - c->sp.drlevel));
    KSSD_TRY(dict_build(c, c->d_emit.as<uint64_t>(), n_emit, hash_bits, c->d_dict64, n_dict));
    if (*n_dict >= (1ull << 32)) { set_error("sketch: more than 2^32 distinct 64-bit hashes in one session"); return KSSD_E_UNSUPPORTED; }
    c->rank_bits = bits_for(*n_dict);
    if (n_emit) {
        encode64_kernel<<<grid_for(c, n_emit), 256, 0, c->stream>>>(c->d_emit.as<unsigned long long>(), c->d_emit_gen.as<uint32_t>(), n_emit,
                                                                   c->d_dict64.as<uint64_t>(), *n_dict);
        KSSD_CUDA(cudaGetLastError());
        c->launches++;
    }
    return KSSD_OK;
}

int session_decode64(kssd_ctx *c, uint64_t n_keep) {
    KSSD_TRY(c->d_hashes64.reserve(std::max<uint64_t>(n_keep, 1) * 8, c->stream));
    if (n_keep) {
        gather64_kernel<<<grid_for(c, n_keep), 256, 0, c->stream>>>(c->d_hashes.as<uint32_t>(), n_keep, c->d_dict64.as<uint64_t>(),
                                                                   c->d_hashes64.as<uint64_t>());
        KSSD_CUDA(cudaGetLastError());
        c->launches++;
    }
    return KSSD_OK;
}

// transSketches, 64-bit path (src/sketch.cpp:904-969): hash -> ascending genome ids.  Dictionary = the sorted distinct
// hashes (d_uniq_hash64); the CSR (d_uniq_off, d_dict) comes from the 32-bit build over the ranks, whose distinct "hashes"
// are then exactly 0 .. n_dict-1, so list u of the CSR belongs to dictionary entry u.
int index_build_device64(kssd_ctx *c, const uint64_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total) {
    c->have_index = false;
    c->index_hash64 = false;
    if (total >= (1ull << 32)) { set_error("index: more than 2^32 postings is not supported"); return KSSD_E_UNSUPPORTED; }
    const int hash_bits = c->have_params && c->hash64 ? std::min(64, 4 * (c->sp.half_k - c->sp.drlevel)) : 64;
    uint64_t n_dict = 0;
    KSSD_TRY(dict_build(c, d_hashes, total, hash_bits, c->d_uniq_hash64, &n_dict));
    KSSD_TRY(c->d_ranks.reserve(std::max<uint64_t>(total, 1) * 4, c->stream));
    KSSD_TRY(dict_ranks(c, d_hashes, total, c->d_uniq_hash64.as<uint64_t>(), n_dict, c->d_ranks.as<uint32_t>()));
    KSSD_TRY(index_build_device(c, c->d_ranks.as<uint32_t>(), d_genome_off, n_genomes, total, bits_for(n_dict)));
    if (c->n_uniq != n_dict) {
        c->have_index = false;
        set_error("index (64-bit hashes): %llu posting lists for %llu distinct hashes", (unsigned long long)c->n_uniq, (unsigned long long)n_dict);
        return KSSD_E_CUDA;
    }
    c->index_hash64 = true;
    return KSSD_OK;
}

}  // namespace kssd
