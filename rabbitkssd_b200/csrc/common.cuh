// Internal declarations shared by the translation units of libkssd_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/kssd_b200.h"

namespace kssd {

void set_error(const char *fmt, ...);

#define KSSD_CUDA(call)                                                                              \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) {                                                                    \
            kssd::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));  \
            return KSSD_E_CUDA;                                                                      \
        }                                                                                            \
    } while (0)

#define KSSD_TRY(call)               \
    do {                             \
        int r__ = (call);            \
        if (r__ != KSSD_OK) return r__; \
    } while (0)

// Device buffer that only grows (HBM is plentiful: 180 GB per GPU; reallocation is the cost to avoid).
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes, cudaStream_t st, bool keep = false);
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct HostBuf {  // pinned
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// Scan-kernel parameters derived from (half_k, half_subk, drlevel); reference: src/common.cpp:35-78.
struct ScanParams {
    int half_k, half_subk, drlevel;
    int hoc;          // half_outctx_len = half_k - half_subk
    int kmer_size;    // 2*half_k
    int group;        // G: consecutive k-mer end positions decided by one bitmap test (1..4)
    int nb;           // bases common to the G windows that index the membership bitmap (<= 10)
    int wbits;        // bitmap word-index bits = 2*nb - 5
    uint64_t tupmask, domask, undomask0, undomask1;
    uint32_t table_mask;  // exact table capacity - 1
    int table_shift;      // 32 - log2(capacity)
    int dim_end;
};

// One scan submitted in the current sketch session whose keys are not yet known to have fit the emit buffer.  When the
// buffer overflows (repeat-rich input), the unverified scans are undone and run again into a grown buffer
// (capi_sketch.inc: verify_scans); the reference never fails on repeats (src/sketch.cpp:526-529).
struct ReplayItem {
    int kind;            // 0: device buffers (a staged batch or the caller's resident buffer), 1: the caller's host buffers (scan_host)
    int stage;           // staging buffer the scan read (-1: none)
    const uint8_t *seq, *qual;
    uint64_t n_bytes;
    const uint64_t *seg_off;
    const uint32_t *seg_genome;
    uint32_t n_seg;
    uint64_t first_emit, chunk_bytes;
    const uint64_t *seg_first;  // raw-text batches: per-segment first new position (device), else null
};

struct TimedLaunches {
    std::vector<cudaEvent_t> ev;  // pairs
    std::vector<cudaEvent_t> pool;
    uint64_t launches = 0, units_a = 0, units_b = 0;
    int begin(cudaStream_t st);
    int end(cudaStream_t st);
    int collect(double *ms);  // syncs the recorded events, recycles them
};

}  // namespace kssd

struct kssd_batch_priv {
    cudaEvent_t h2d_done;
    bool in_flight = false;
    bool held = false;  // acquired and not yet submitted
    kssd::HostBuf seq, qual, seg_off, seg_genome, seg_flags, seg_first;
};

namespace kssd {
static constexpr int kMaxPeers = 64;
// multi-GPU gather over peer memory: two buffers per rank (double buffered), mapped into every rank with CUDA IPC
struct PeerState {
    uint32_t world = 0, rank = 0, n_local = 0;
    uint64_t stride = 0;  // u32 words per block: n_local sizes + capacity in hashes
    int flip = 0;
    uint32_t step = 0;    // publishes so far: the value this rank writes into its flag at every peer (peer barrier, segsort.cu)
    uint32_t *own[2] = {nullptr, nullptr};
    uint32_t *peer[2][kMaxPeers] = {};
};
}  // namespace kssd

struct kssd_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaStream_t copy_stream = nullptr;
    uint64_t launches = 0;
    bool profiling = false;
    uint64_t attr_mask = 0;    // kernels whose dynamic shared-memory limit was raised ON THIS DEVICE (cudaFuncSetAttribute is per device)
    bool force_radix = false;  // tests: take the device-wide radix paths even when the shared-memory fast paths apply

    // parameters + tables
    bool have_params = false;
    bool relax_guard = false;  // accept half_subk - drlevel < 3 (L5K10 of BASELINE config 4: the reference refuses it, common.cpp:37-39)
    kssd::ScanParams sp;
    kssd::DevBuf d_bitmap, d_tab_keys;  // membership bitmap, exact (dim_id<<32|value) table

    // sketch session
    bool in_session = false;
    uint32_t n_genomes = 0;
    int has_qual = 0, least_qual = 0, least_num_kmer = 1;
    uint64_t bases_submitted = 0;
    std::vector<kssd::ReplayItem> replay;  // scans not yet verified against the emit capacity, in submission order
    uint64_t verified_cursor = 0;          // emit cursor after the last verified scan
    bool replaying = false;
    bool raw_text_seen = false;            // the session took raw FASTA text (fasta.cu)
    kssd::HostBuf h_snap;                  // pinned: emit cursor after the last scan of staging buffer 0 / 1
    kssd::DevBuf d_emit;      // u64 keys (genome<<32 | hash)
    kssd::DevBuf d_emit_rank; // u32 rank of each key inside its genome
    kssd::DevBuf d_emit_gen;  // 64-bit hash configuration: u32 genome of each key (d_emit then holds the bare hashes)
    uint64_t emit_cap = 0;
    kssd::DevBuf d_cursor;    // u64 cursor (+ spare words)
    kssd::DevBuf d_gcount;    // u32 emitted keys per genome (written by the scan kernel)
    kssd::DevBuf d_seg_start, d_ucount;  // per-genome segment starts (u64), unique counts
    kssd::DevBuf d_stage_seq[2], d_stage_qual[2], d_stage_segoff[2], d_stage_seggen[2];
    // raw FASTA text batches (fasta.cu): the raw bytes + piece table of a batch, the compacted piece table; scratch
    kssd::DevBuf d_stage_raw[2], d_stage_rawseg[2], d_stage_cseg[2], d_fa_scratch;
    cudaEvent_t stage_free[2] = {nullptr, nullptr};
    int stage_next = 0;
    static constexpr int kBatches = 4;  // pinned batches: a host may fill two kinds (parsed records, raw text) while two are in flight
    kssd_batch batches[kBatches];
    kssd_batch_priv bpriv[kBatches];
    bool batches_init = false;
    int batch_next = 0;
    kssd::DevBuf d_sort_tmp, d_sort_counts, d_keys_unique;
    kssd::DevBuf d_hashes, d_genome_off;
    kssd::PeerState peers;
    uint64_t cat_total = 0;  // hashes in d_cat_hashes (valid while cat_n != 0)
    uint32_t cat_n = 0;      // genomes in d_cat_off
    kssd::DevBuf d_cat_hashes, d_cat_off, d_cat_counts;  // sketch shards of all ranks, concatenated (multi-GPU gather epilogue)
    kssd::HostBuf h_hashes, h_genome_off, h_scalar;
    uint64_t sketch_total = 0;

    // index (CSR over distinct hashes)
    bool have_index = false;
    // 64-bit hash configuration (half_k - drlevel > 8, src/sketch.cpp:336), hash64.cu: hashes are dictionary-encoded -- the
    // sorted distinct hashes of a session / of an index form a dictionary and everything downstream (per-genome sort +
    // unique, index build, counters) runs on the u32 RANKS in it, i.e. through the 32-bit kernels
    bool hash64 = false;       // the parameters select 64-bit hashes: sketch sessions produce u64 hashes
    int rank_bits = 0;         // bits of the ranks in the current session's dictionary (sizes the sort buckets)
    bool index_hash64 = false; // the loaded index is over 64-bit hashes: d_uniq_hash64 holds its sorted distinct hashes
    kssd::DevBuf d_h64_a, d_h64_b;   // scratch: u64 keys being sorted
    kssd::DevBuf d_dict64;           // session dictionary: sorted distinct u64 hashes of all genomes of the session
    kssd::DevBuf d_hashes64;         // finished sketch, decoded: u64 hashes (CSR with d_genome_off)
    kssd::DevBuf d_uniq_hash64;      // index dictionary: sorted distinct u64 hashes; d_uniq_off / d_dict as in the 32-bit index
    kssd::DevBuf d_ranks;            // index build: u32 rank of every input hash in the index dictionary
    kssd::HostBuf h_hashes64, h_uniq_hash64;
    uint64_t n_uniq = 0, n_post = 0;
    uint64_t list_slots = 0;   // hashes of all indexed sketches: bounds the alldist work lists (a row has <= |sketch| ranges)
    uint32_t index_row_begin = 0, index_row_end = 0xFFFFFFFFu;  // index restricted to these alldist rows (index_build_rows)
    kssd::DevBuf d_filter;     // index_build_rows: hashed bitmap of the hashes of the row block
    uint32_t n_ref = 0;
    uint32_t max_ref_size = 0;
    kssd::DevBuf d_uniq_hash, d_uniq_off, d_dict, d_ref_sizes, d_pairs, d_pairs_tmp;
    kssd::DevBuf d_bucket;    // bucket_start[b] = first distinct hash with (hash >> bucket_shift) >= b; n_buckets + 1 entries
    int bucket_shift = 0;
    uint32_t n_buckets = 0;
    kssd::HostBuf h_uniq_hash, h_uniq_count, h_dict;

    // distances
    bool emit_zero_cells = true;  // report cells with common == 0 when they pass the threshold (kssd_b200_set_emit_zero_cells)
    uint64_t max_neighbor = 0;    // > 0: dist reports only the top-N cells of a row (kssd_b200_set_max_neighbor)
    kssd::DevBuf d_q_hashes, d_q_off, d_hits, d_hit_cursor, d_rowwork;
    kssd::DevBuf d_q_list_begin;  // alldist: posting ranges per row (uint2 begin, end), slots of row g at q_off[g]
    kssd::DevBuf d_q_list_end;    // alldist: ranges filled per row
    kssd::HostBuf h_hits;
    kssd::DevBuf d_tile_rows;     // alldist: first row of every tile of the packed triangle (dist.cu)
    kssd::HostBuf h_tile_rows;
    uint64_t tile_key[4] = {0, 0, 0, 0};
    uint32_t n_tile_rows = 0;

    // set algebra (setops.cu)
    kssd::DevBuf d_setop_in, d_setop_tmp, d_setop_out;
    kssd::HostBuf h_setop;

    kssd::TimedLaunches t_scan, t_count;
};

namespace kssd {
// ---- tokenisers (fasta.cu, fastq.cu): compacted output of a tile ----
// The kept bytes of a tile are contiguous in the output: stage them in shared memory at their tile-local offsets, then the
// CTA writes the range with 16-byte stores (a thread storing its own bytes one by one reached 200 GB/s, profiles/r02z3).
// The staging buffer is shifted by the output address modulo 16, so that aligned 16-byte groups of the buffer are aligned
// 16-byte groups of global memory; the partial groups at both ends are written byte by byte (their other bytes belong to the
// neighbouring tiles).
__device__ __forceinline__ void store_tile_bytes(const uint8_t *s_buf, uint32_t shift, uint32_t cnt, uint8_t *out) {
    uint8_t *base = out - shift;  // 16-byte aligned; buffer byte i <-> base[i], the data is [shift, end)
    const uint32_t end = shift + cnt;
    const uint32_t v0 = (shift + 15) / 16, v1 = end / 16;  // whole 16-byte groups [v0, v1)
    if (v1 > v0) {
        for (uint32_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) reinterpret_cast<uint4 *>(base)[v] = reinterpret_cast<const uint4 *>(s_buf)[v];
        for (uint32_t i = shift + threadIdx.x; i < v0 * 16; i += blockDim.x) base[i] = s_buf[i];
        for (uint32_t i = v1 * 16 + threadIdx.x; i < end; i += blockDim.x) base[i] = s_buf[i];
    } else {
        for (uint32_t i = shift + threadIdx.x; i < end; i += blockDim.x) base[i] = s_buf[i];
    }
}

// sketch.cu
int launch_scan(kssd_ctx *c, const uint8_t *d_seq, const uint8_t *d_qual, uint64_t n_bytes,
                const uint64_t *d_seg_off, const uint32_t *d_seg_genome, uint32_t n_seg, uint64_t first_emit,
                uint64_t pos_base = 0, const uint64_t *d_seg_first = nullptr);
// fasta.cu
int fasta_tokenise(kssd_ctx *c, const uint8_t *d_raw, uint64_t n_raw, const uint64_t *d_seg_off, const uint8_t *d_seg_flags, const uint32_t *d_seg_first,
                   uint32_t n_seg, uint8_t *d_out, uint64_t *d_cseg_off, uint64_t *d_cseg_first, uint32_t *d_anomaly, uint64_t *out_bytes);
// fastq.cu
int fastq_tokenise(kssd_ctx *c, const uint8_t *d_raw, uint64_t n_raw, const uint64_t *d_seg_off, uint32_t n_seg, uint8_t *d_out_seq,
                   uint8_t *d_out_qual, uint64_t *d_cseg_off, uint32_t *d_anomaly, uint64_t *out_bytes);
int build_tables(kssd_ctx *c, const int32_t *shuffled_dim);
// radix.cu
int radix_sort_u64(kssd_ctx *c, uint64_t *d_keys, uint64_t *d_tmp, uint64_t n, int lo_bits_begin, int lo_bits_end,
                   int hi_bits_begin, int hi_bits_end, uint64_t **d_sorted);
int unique_keys(kssd_ctx *c, const uint64_t *d_sorted, uint64_t n, int min_count, uint64_t *d_out_keys,
                uint64_t *n_out);
int genome_sort_unique(kssd_ctx *c, const uint64_t *d_emit, uint64_t n_emit, uint32_t max_count, int min_count,
                       uint64_t *n_keep);   // segsort.cu: shared-memory per-genome sort + unique (fast path)
int sketch_concat(kssd_ctx *c, const uint32_t *d_parts, uint64_t stride, uint32_t n_local, uint32_t world, uint64_t *total,
                  int *overflow);  // segsort.cu: multi-GPU gather epilogue
int peer_init(kssd_ctx *c, uint32_t n_local, uint64_t cap_hashes, uint8_t *handles_out);
int peer_connect(kssd_ctx *c, const uint8_t *handles, uint32_t world, uint32_t rank);
void peer_release(kssd_ctx *c);
int sketch_publish(kssd_ctx *c);
int sketch_gather_peers(kssd_ctx *c, uint64_t *total, int *overflow);
int genome_segments(kssd_ctx *c);           // segsort.cu: gcount -> seg_start + max count (no sync)
int split_genomes(kssd_ctx *c, const uint64_t *d_keys, uint64_t n, uint32_t n_genomes, uint32_t *d_hashes,
                  uint64_t *d_genome_off);
// index.cu
int index_build_device(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes,
                       uint64_t total, int hash_bits_override = 0, uint32_t row_begin = 0, uint32_t row_end = 0xFFFFFFFFu);
// hash64.cu: dictionary encoding of 64-bit hashes
int dict_build(kssd_ctx *c, const uint64_t *d_vals, uint64_t n, int hash_bits, DevBuf &dict, uint64_t *n_dict);
int dict_ranks(kssd_ctx *c, const uint64_t *d_vals, uint64_t n, const uint64_t *d_dict, uint64_t n_dict, uint32_t *d_ranks);
int session_encode64(kssd_ctx *c, uint64_t n_emit, uint64_t *n_dict);   // d_emit: hash -> (genome << 32 | rank)
int session_decode64(kssd_ctx *c, uint64_t n_keep);                     // d_hashes (ranks) -> d_hashes64
int index_build_device64(kssd_ctx *c, const uint64_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total);

int index_build_range(kssd_ctx *c, const uint32_t *d_hashes, const uint64_t *d_genome_off, uint32_t n_genomes, uint64_t total, uint32_t part,
                      uint32_t parts, uint32_t *d_block, uint64_t block_words, uint64_t *used_words, int *overflow);
int index_adopt_parts(kssd_ctx *c, const uint32_t *d_blocks, uint64_t block_words, uint32_t parts, const uint64_t *d_genome_off, uint32_t n_genomes,
                      int *overflow);
int index_build_buckets(kssd_ctx *c, int64_t known_max_hash = -1);  // after d_uniq_hash / n_uniq are in place
// setops.cu
int set_union_host(kssd_ctx *c, const void *hashes, uint64_t n, int hash_bytes, uint64_t *n_out);
int set_subtract_host(kssd_ctx *c, const void *q, const uint64_t *q_off, uint32_t n_q, const void *ref, uint64_t n_ref, int hash_bytes);
// dist.cu
int dist_rows_device(kssd_ctx *c, const void *d_q_hashes, const uint64_t *d_q_off, uint32_t n_query,
                     uint32_t row_begin, uint32_t row_end, int triangle, int kmer_size, double max_dist,
                     int containment, uint64_t *n_hits);
}  // namespace kssd
