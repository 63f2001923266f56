/*
 * kssd_b200.h -- C-ABI of libkssd_b200.so: the B200 (sm_100a) implementation of RabbitKSSD's two
 * data-parallel hot paths.  Plain pointers and sizes only; no C++/torch types.
 *
 * The reference has no FFI layer: its seam is five free C++ functions.  The host shims under
 * rabbitkssd_b200/csrc/host/ keep those five signatures and call the entry points below; each entry point
 * cites the reference code it replaces (paths relative to the RabbitKSSD tree).
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error (KSSD_E_*); text via kssd_b200_last_error().  The
 *    library never calls exit(); there is NO CPU fallback: without a usable CUDA device init fails.
 *  - one kssd_ctx per GPU (one process or one host thread per GPU); a ctx is not thread-safe.
 *  - pointers are HOST pointers unless the parameter is named d_* or the call ends in _device.
 *  - hashes are uint32 when half_k - drlevel <= 8 and uint64 otherwise (sketch.cpp:336 use64); set_params decides,
 *    and the entry points that carry hashes come in two widths (the *64 ones for the 64-bit configuration).
 *    Calling the wrong width returns KSSD_E_ARG.
 */
#ifndef KSSD_B200_H
#define KSSD_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KSSD_OK 0
#define KSSD_E_CUDA (-1)        /* CUDA runtime error (text has the cudaError string) */
#define KSSD_E_ARG (-2)         /* bad argument / call order */
#define KSSD_E_UNSUPPORTED (-3) /* shape the library does not implement (half_k > 12, >= 2^32 postings); the text says which */
#define KSSD_E_PARAM (-4)       /* half_subk - drlevel < 3: the reference aborts here (common.cpp:37-39) */
#define KSSD_E_SYNTAX (-6)      /* raw FASTA text that kseq would not read as plain FASTA (sketch_finish): parse it on the host */
#define KSSD_E_OVERFLOW (-5)    /* an output buffer stayed too small after regrowing; text says which and the size needed */

typedef struct kssd_ctx kssd_ctx;

const char *kssd_b200_last_error(void);
int kssd_b200_version(void);

/* Host only, no device needed: the dimension-reduction table `rabbit_kssd shuffle -k -s -l` writes into a .shuf file
 * (src/shuffle.cpp:48-54, :76-104; glibc srand/rand Fisher-Yates).  out: 16^half_subk ints.  The table is an input of
 * the GPU path (kssd_b200_set_params); generating it stays on the CPU as in the reference. */
int kssd_b200_shuffle_table(int half_k, int half_subk, int drlevel, int32_t *out);

/* Number of usable CUDA devices (0 and KSSD_E_CUDA when there is none): a host with one context per GPU (dist_gpu.cpp)
 * sizes its pool with it. */
int kssd_b200_device_count(int *n);
/* Create a context on CUDA device `device`.  Fails (KSSD_E_CUDA) when no device is usable. */
int kssd_b200_init(kssd_ctx **ctx, int device);
void kssd_b200_destroy(kssd_ctx *ctx);
/* Run all work of this ctx on the caller's cudaStream_t (e.g. a torch.cuda.Stream); NULL = library-owned
 * non-blocking stream.  The legacy default stream has handle 0 and therefore cannot be passed: callers that mix
 * their own kernels with the library's must create an explicit stream. */
int kssd_b200_set_stream(kssd_ctx *ctx, void *cuda_stream);
int kssd_b200_sync(kssd_ctx *ctx);

/* ---- parameters + shuffle table --------------------------------------------------------------------
 * replaces initParameter (src/common.cpp:35-78) and the retained-subset map build (src/sketch.cpp:338-345,
 * :616-623).  shuffled_dim = the 16^half_subk ints of the .shuf file (src/shuffle.cpp:8-23), borrowed for
 * the call.  Uploads the masks, a symmetrised membership bitmap and the exact retained-subset table. */
int kssd_b200_set_params(kssd_ctx *ctx, int half_k, int half_subk, int drlevel, const int32_t *shuffled_dim);
/* on = 1: set_params accepts half_subk - drlevel < 3 (still > 0), which the reference refuses (common.cpp:37-39,
 * shuffle.cpp:30-33) -- BASELINE config 4 names L5K10 (half_subk 7, drlevel 5).  There is no reference output for such a
 * configuration; the arithmetic is the same.  Default 0: KSSD_E_PARAM with the reference's message. */
int kssd_b200_set_relax_guard(kssd_ctx *ctx, int on);

/* ---- sketching --------------------------------------------------------------------------------------
 * replaces the per-base scan + per-genome set of sketchFastaFile / sketchFastqFile and the consumer tasks
 * (src/sketch.cpp:173-314, :487-550, :777-849).  Input is PARSED sequence (no FASTA/FASTQ syntax): a flat
 * byte buffer in which the records of one genome are separated by at least one non-ACGT byte (a separator
 * resets the valid-run counter exactly as a new record does, sketch.cpp:487-488), cut into segments; segment
 * i = bytes [seg_off[i], seg_off[i+1]) and belongs to genome seg_genome[i] (index in [0, n_genomes)).
 * The first byte after a genome's last record must also be a separator (or the end of the batch).
 * A record continued from the previous batch is prefixed by its last 2*half_k-1 bytes and first_emit is set
 * to that count: k-mers ending before first_emit are context only. */
#define KSSD_SEG_IN_HEADER 1
#define KSSD_SEG_MID_LINE 2
#define KSSD_RAW_FASTQ 2
typedef struct kssd_batch {
    uint8_t *seq;         /* pinned, library owned, capacity bytes */
    uint8_t *qual;        /* pinned, same layout; used only when the session was begun with has_qual */
    uint64_t capacity;    /* bytes available in seq / qual */
    uint64_t n_bytes;     /* filled by caller */
    uint64_t *seg_off;    /* pinned, seg_capacity+1 entries; filled by caller, seg_off[0] == 0 */
    uint32_t *seg_genome; /* pinned, seg_capacity entries; filled by caller */
    uint32_t seg_capacity;
    uint32_t n_seg;       /* filled by caller */
    uint64_t first_emit;  /* filled by caller (0 when nothing is continued) */
    void *priv;
    /* Raw FASTA text, tokenised ON THE DEVICE (replaces kseq's record loop, src/kseq.h:175-215, and RabbitFX's formatter,
     * RabbitFX/io/Formater.cpp:87-129, for plain uncompressed FASTA): raw_fasta = 1 and seq holds PIECES of file text instead
     * of parsed records.  Piece i = bytes [seg_off[i], seg_off[i+1]) of genome seg_genome[i]; seg_off[i] is a multiple of 32
     * (pad the end of a piece with '\n'; a piece must not end with a '\r' whose '\n' is in the next one).  A piece is a whole
     * file, or a slice of a large one that begins at a line start (seg_flags[i] = 0), inside a header line (KSSD_SEG_IN_HEADER)
     * or inside a sequence line (KSSD_SEG_MID_LINE); a slice that continues a record is prefixed with the last
     * >= 2*half_k - 1 sequence bytes before it (plus a '\n' when the slice itself begins at a line start) and seg_first[i] =
     * length of that prefix: k-mers ending inside it were already found by the previous slice.  Header lines become one
     * separator byte.  Text kseq would not
     * take as plain FASTA (a line starting with '+' after a header / sequence line) makes sketch_finish return
     * KSSD_E_SYNTAX.  Not for sessions with quality.
     * raw_fasta = KSSD_RAW_FASTQ: raw four-line FASTQ text, tokenised on the device (replaces kseq's FASTQ branch,
     * src/kseq.h:201-213, and the per-read loops src/sketch.cpp:273-307, :779-825 keep their arithmetic in the scan); sessions
     * with quality only.  A piece is a WHOLE NUMBER OF RECORDS ('@' line, sequence line, '+' line, quality line), ends with
     * '\n' and is padded with 0x00 to a multiple of 32; seg_flags / seg_first are not read.  Reads never share k-mers, so
     * slices of a large file need no context.  Text that kseq would parse differently from "line number modulo four"
     * (multi-line records, sequence and quality of different lengths, a record line that does not begin with '@', stray 0x00)
     * makes sketch_finish return KSSD_E_SYNTAX. */
    int raw_fasta;        /* filled by caller (0, 1 or KSSD_RAW_FASTQ); batch_acquire resets it to 0 */
    uint8_t *seg_flags;   /* pinned, seg_capacity entries */
    uint32_t *seg_first;  /* pinned, seg_capacity entries */
} kssd_batch;

/* Start a sketch session for n_genomes genomes.  least_qual / least_num_kmer are the -Q / -n options
 * (raw ASCII compare quality[i] >= least_qual, sketch.cpp:280; keep hashes seen >= least_num_kmer times,
 * sketch.cpp:704-706); has_qual = 0 for FASTA. */
int kssd_b200_sketch_begin(kssd_ctx *ctx, uint32_t n_genomes, int has_qual, int least_qual, int least_num_kmer);
/* Double-buffered pinned staging: acquire blocks until one of the ctx's batches is free. */
int kssd_b200_batch_acquire(kssd_ctx *ctx, kssd_batch **batch, uint64_t min_capacity, uint32_t min_segments);
/* Async H2D + scan kernel; the batch may be re-acquired once the copy has drained. */
int kssd_b200_batch_submit(kssd_ctx *ctx, kssd_batch *batch);
/* Same scan on a buffer ALREADY resident in device memory (bench `value`, multi-GPU pipelines). d_seq must be
 * 16-byte aligned; d_seg_off/d_seg_genome are device arrays.  All buffers passed to sketch_scan_device / _scan_host must
 * stay valid and unchanged until sketch_finish returns: when a repeat-rich input emits more keys than the emit buffer holds
 * (it is sized for 8 x the expected bases/16^drlevel), the library grows the buffer and runs the affected scans again
 * instead of failing -- the reference's unordered_set never fails on repeats (src/sketch.cpp:526-529). */
int kssd_b200_sketch_scan_device(kssd_ctx *ctx, const uint8_t *d_seq, const uint8_t *d_qual, uint64_t n_bytes,
                                 const uint64_t *d_seg_off, const uint32_t *d_seg_genome, uint32_t n_seg,
                                 uint64_t first_emit);
/* Same scan straight from the CALLER's host buffers (best pinned with cudaHostAlloc/cudaHostRegister: the copy
 * is then asynchronous): the buffer is streamed H2D in chunk_bytes pieces (0 = 256 MiB) through the ctx's two
 * staging buffers, copy of chunk i+1 overlapping the scan of chunk i; chunk borders carry a 32-byte context
 * prefix so that no k-mer is lost or emitted twice.  seg_off / seg_genome as above, over the whole buffer. */
int kssd_b200_sketch_scan_host(kssd_ctx *ctx, const uint8_t *seq, const uint8_t *qual, uint64_t n_bytes,
                               const uint64_t *seg_off, const uint32_t *seg_genome, uint32_t n_seg,
                               uint64_t chunk_bytes);
/* Per-genome sort + unique (+ count filter): hashes[genome_off[g] .. genome_off[g+1]) ascending, unique.
 * Host arrays are library owned and valid until the next sketch_begin / destroy. */
int kssd_b200_sketch_finish(kssd_ctx *ctx, const uint32_t **hashes, const uint64_t **genome_off);
/* Same, results stay in device memory (library owned). */
int kssd_b200_sketch_finish_device(kssd_ctx *ctx, const uint32_t **d_hashes, const uint64_t **d_genome_off,
                                   uint64_t *total);
/* 64-bit hash configuration (half_k - drlevel > 8: sketch.cpp:336, hashSet64 of sketch_t, sketch.h:22): the same
 * session calls, hashes come out as uint64. */
int kssd_b200_sketch_finish64(kssd_ctx *ctx, const uint64_t **hashes, const uint64_t **genome_off);
int kssd_b200_sketch_finish_device64(kssd_ctx *ctx, const uint64_t **d_hashes, const uint64_t **d_genome_off,
                                     uint64_t *total);
/* Multi-GPU gather epilogue (the sketch shards of all ranks are all-gathered before the index is
 * built; the reference has one address space, its counterpart is the `sketches` vector filled under omp critical,
 * src/sketch.cpp:455-560).  d_parts (device) holds `world` blocks of stride_words u32 each, block r =
 * [n_local per-genome sketch sizes][the hashes of rank r's genomes, genome-major][padding] -- what one all-gather
 * of equally sized send buffers delivers.  Writes the concatenated CSR (world * n_local genomes, rank order) into
 * library-owned device memory, valid until the next call.  *overflow != 0: some rank's shard did not fit its block
 * (identical on every rank, so all ranks can fall back together); the outputs are then invalid. */
int kssd_b200_sketch_concat_device(kssd_ctx *ctx, const uint32_t *d_parts, uint64_t stride_words, uint32_t n_local,
                                   uint32_t world, const uint32_t **d_hashes, const uint64_t **d_genome_off,
                                   uint64_t *total, int *overflow);

/* Gather over PEER MEMORY (NVLink, CUDA IPC) instead of a collective: every rank keeps its shard in two buffers of
 * its own (double buffered), mapped into all ranks of the node; gathering is then the concat kernels reading the
 * peers' buffers in place -- one pass of peer loads, no NCCL transfer, no staging copy.
 *   peer_init      allocates the two buffers ([n_local sizes | cap_hashes hashes] each) and returns their IPC
 *                  handles (2 x 64 bytes) for the caller to exchange (any transport; bench.py uses one all-gather);
 *   peer_connect   handles of all ranks in rank order (world x 128 bytes) -> peer mappings;
 *   sketch_publish after sketch_finish(_device): this rank's sketch -> its current buffer (asynchronous); the last CTA of
 *                  the copy then stores this rank's publish count into its flag slot at EVERY rank (remote stores behind a
 *                  system-scope fence) -- the barrier is part of the kernel, there is no collective in the step;
 *   sketch_gather_peers   the concat kernels first wait (on the device) until every rank's flag has reached the current step,
 *                  then read the peers' shards in place: concatenated CSR as kssd_b200_sketch_concat_device returns it.
 *                  Every rank must publish once per step.  A peer that does not publish within ~10 s: KSSD_E_CUDA.
 * A buffer is rewritten two steps later; by then every reader has passed the barrier of the step in between. */
int kssd_b200_peer_init(kssd_ctx *ctx, uint32_t n_local, uint64_t cap_hashes, uint8_t *handles_out);
int kssd_b200_peer_connect(kssd_ctx *ctx, const uint8_t *handles, uint32_t world, uint32_t rank);
int kssd_b200_sketch_publish(kssd_ctx *ctx);
int kssd_b200_sketch_gather_peers(kssd_ctx *ctx, const uint32_t **d_hashes, const uint64_t **d_genome_off,
                                  uint64_t *total, int *overflow);

/* ---- inverted index -----------------------------------------------------------------------------------
 * replaces transSketches, 32-bit path (src/sketch.cpp:970-1017): hash -> ascending genome ids.  The index is
 * kept in device memory as CSR over the DISTINCT hashes.  index_export gives the host the content of the
 * two files: dict = the .dict payload; (uniq_hash, uniq_count) = the non-zero entries of the .index count
 * array (the host shim scatters them into the dense 2^(4(k-l)) array the file format wants). */
#define KSSD_PTR_HOST 0
#define KSSD_PTR_DEVICE 1
int kssd_b200_index_build(kssd_ctx *ctx, const uint32_t *hashes, const uint64_t *genome_off, uint32_t n_genomes,
                          int ptr_kind);
/* Multi-GPU alldist (the triangle is sharded by query-row blocks; every rank needs "the index"): the index
 * RESTRICTED to what rows [row_begin, row_end) of the triangle stream -- only the posting lists of hashes that occur in
 * those rows' sketches (selected through a hashed bitmap: a false positive only adds a list no row of the block walks)
 * and of these only the ids >= row_begin (src/dist.cpp:207 looks at columns > row).  Every rank then sorts ~1/ranks of
 * the postings instead of repeating the whole build.  kssd_b200_dist_rows* accepts only triangle = 1 with rows inside
 * the block afterwards; index_export returns the restricted lists.  KSSD_E_OVERFLOW: call kssd_b200_index_build. */
int kssd_b200_index_build_rows(kssd_ctx *ctx, const uint32_t *hashes, const uint64_t *genome_off, uint32_t n_genomes,
                               int ptr_kind, uint32_t row_begin, uint32_t row_end);
int kssd_b200_index_export(kssd_ctx *ctx, uint64_t *n_uniq, uint64_t *total, const uint32_t **uniq_hash,
                           const uint32_t **uniq_count, const uint32_t **dict);
/* 64-bit hash configuration: replaces transSketches' hash-map path (src/sketch.cpp:904-969).  index_export64 returns the
 * content of the sparse .index file (size_t n; uint64 hash[n]; uint32 count[n], sketch.cpp:951-958) and of .dict, lists
 * in ASCENDING hash order (the reference writes its hash map's iteration order; readers rebuild a map, dist.cpp:45-78). */
int kssd_b200_index_build64(kssd_ctx *ctx, const uint64_t *hashes, const uint64_t *genome_off, uint32_t n_genomes,
                            int ptr_kind);
int kssd_b200_index_export64(kssd_ctx *ctx, uint64_t *n_uniq, uint64_t *total, const uint64_t **uniq_hash,
                             const uint32_t **uniq_count, const uint32_t **dict);
/* replaces the 64-bit .index/.dict load of index_tridist / index_dist (src/dist.cpp:35-82, :442-487): n_hash (hash,
 * count) entries in ANY order, dict = the posting lists in the same order. */
int kssd_b200_index_load64(kssd_ctx *ctx, uint64_t n_hash, const uint64_t *hash, const uint32_t *counts,
                           const uint32_t *dict, uint64_t total, const uint32_t *ref_sizes, uint32_t n_ref);
/* replaces the .index/.dict load + prefix sum of index_tridist / index_dist (src/dist.cpp:83-130, :488-523):
 * counts = the dense count array of the .index file (hash_size entries), dict = the .dict payload. */
int kssd_b200_index_load(kssd_ctx *ctx, uint64_t hash_size, const uint32_t *counts, const uint32_t *dict,
                         uint64_t total, const uint32_t *ref_sizes, uint32_t n_ref);
/* Multi-GPU index build (north_star: "the reference index replicated by NCCL over NVLink"): rank
 * `part` of `parts` builds the CSR of ONE hash range of transSketches' table (src/sketch.cpp:970-1017) from the
 * sketches of ALL genomes (device pointers, as index_build) and packs it into d_block, block_words u32 of caller
 * memory laid out for one all-gather of equally sized blocks: [distinct hashes U, postings P, max hash, 0,
 * uniq_hash[U], list starts[U+1], postings[P]].  Ranges are cut where the hash density (leading bits = outer
 * bases of the canonical k-mer) predicts equal posting counts.  *overflow != 0: the piece does not fit the block
 * (or the shared-memory bucket sort does not apply); the block header then says so to every rank.
 * *used_words = 5 + 2U + P: only that prefix of the block needs to travel (exchange the sizes first and gather
 * blocks of the largest one; kssd_b200_index_adopt_parts takes that smaller stride). */
int kssd_b200_index_build_range(kssd_ctx *ctx, const uint32_t *d_hashes, const uint64_t *d_genome_off,
                                uint32_t n_genomes, uint32_t part, uint32_t parts, uint32_t *d_block,
                                uint64_t block_words, uint64_t *used_words, int *overflow);
/* Adopts the exchanged pieces (`parts` blocks of block_words u32 in device memory, hash ranges ascending) as this
 * context's index: concatenated CSR, lookup directory, reference sizes from d_genome_off.  *overflow != 0: some
 * rank reported overflow -- identical on every rank; call kssd_b200_index_build on the full input instead. */
int kssd_b200_index_adopt_parts(kssd_ctx *ctx, const uint32_t *d_blocks, uint64_t block_words, uint32_t parts,
                                const uint64_t *d_genome_off, uint32_t n_genomes, int *overflow);
/* Device views of the CSR index (for replication to other GPUs with one NCCL broadcast). */
int kssd_b200_index_device_views(kssd_ctx *ctx, uint64_t *n_uniq, uint64_t *total, uint32_t *n_ref,
                                 const uint32_t **d_uniq_hash, const uint64_t **d_uniq_off,
                                 const uint32_t **d_dict, const uint32_t **d_ref_sizes);
/* Adopt a replicated index: device arrays are COPIED into ctx-owned memory. */
int kssd_b200_index_adopt_device(kssd_ctx *ctx, uint64_t n_uniq, uint64_t total, uint32_t n_ref,
                                 const uint32_t *d_uniq_hash, const uint64_t *d_uniq_off, const uint32_t *d_dict,
                                 const uint32_t *d_ref_sizes);

/* ---- distances ----------------------------------------------------------------------------------------
 * replaces the per-row posting accumulation + row epilogue of index_tridist (src/dist.cpp:174-258) and
 * index_dist (src/dist.cpp:560-692).  Query rows [row_begin,row_end) of the given query sketches are run
 * against the loaded index (query-row sharding across GPUs).  A hit is a cell whose distance
 * passes max_dist under a CONSERVATIVE device-side test; the caller recomputes Jaccard/containment and the
 * distance with the reference's exact double expressions (dist.cpp:216-232, :237-251) for the final
 * keep/drop, which makes the pair set and the printed numbers identical by construction.
 *   triangle = 1: alldist, only cols > row (dist.cpp:207), strict  `<`  (dist.cpp:232);
 *   triangle = 0: dist, all cols, inclusive `<=` (dist.cpp:624). */
typedef struct kssd_hit { uint32_t row, col, common; } kssd_hit;
/* A cell with common == 0 has distance exactly 1.0 (dist.cpp:227-228), so with -D >= 1 (the CLI default) EVERY cell
 * passes and the hit list is the dense rows x cols matrix.  on = 0 makes the device report only cells with
 * common > 0; the caller then knows that every unreported cell has common == 0, Jaccard 0 and distance 1
 * (dist_gpu.cpp fills them in, in reference order, where the threshold lets them pass).  Default: on = 1. */
int kssd_b200_set_emit_zero_cells(kssd_ctx *ctx, int on);
/* dist -N (src/dist.cpp:599, :625-640, :662-676: a bounded max-heap keeps the maxNeighbor smallest distances of a query
 * row).  max_neighbor > 0 makes the device select per query row: only the max_neighbor cells with the smallest distances
 * (plus everything tied with the last of them, to 2^-31 relative) among those that pass max_dist are reported -- per
 * column tile when the references do not fit one tile, so the result is a SUPERSET of the row's top-N which the caller's
 * exact heap (dist_gpu.cpp) finishes.  Cells with common == 0 are never candidates on the device: the caller fills them
 * in when a row has fewer than max_neighbor reported cells and the threshold lets distance 1.0 pass.  0 = off (default). */
int kssd_b200_set_max_neighbor(kssd_ctx *ctx, uint64_t max_neighbor);
typedef int (*kssd_sink)(void *user, const kssd_hit *hits, size_t n);
/* Query sketches to device memory ONCE, for callers that walk the rows in several dist_rows calls (a host thread per GPU
 * streaming row blocks to its result files, dist_gpu.cpp): returns library-owned device arrays to pass with
 * KSSD_PTR_DEVICE, valid until the next queries_upload / host-pointer dist_rows / index_build from host pointers.
 * hash_bytes = 4 (uint32 hashes) or 8 (64-bit configuration). */
int kssd_b200_queries_upload(kssd_ctx *ctx, const void *q_hashes, const uint64_t *q_off, uint32_t n_query, int hash_bytes,
                             const void **d_q_hashes, const uint64_t **d_q_off);
int kssd_b200_dist_rows(kssd_ctx *ctx, const uint32_t *q_hashes, const uint64_t *q_off, uint32_t n_query,
                        int ptr_kind, uint32_t row_begin, uint32_t row_end, int triangle, int kmer_size,
                        double max_dist, int containment, kssd_sink sink, void *user);
/* Same, hits stay in device memory (library owned until the next call); *n_hits = number written. */
int kssd_b200_dist_rows_device(kssd_ctx *ctx, const uint32_t *d_q_hashes, const uint64_t *d_q_off,
                               uint32_t n_query, uint32_t row_begin, uint32_t row_end, int triangle,
                               int kmer_size, double max_dist, int containment, const kssd_hit **d_hits,
                               uint64_t *n_hits);

/* 64-bit hash configuration (src/dist.cpp:180-192, :566-576): query hashes are uint64, the index was built / loaded by
 * the *64 entry points.  Everything else as above. */
int kssd_b200_dist_rows64(kssd_ctx *ctx, const uint64_t *q_hashes, const uint64_t *q_off, uint32_t n_query,
                          int ptr_kind, uint32_t row_begin, uint32_t row_end, int triangle, int kmer_size,
                          double max_dist, int containment, kssd_sink sink, void *user);
int kssd_b200_dist_rows_device64(kssd_ctx *ctx, const uint64_t *d_q_hashes, const uint64_t *d_q_off,
                                 uint32_t n_query, uint32_t row_begin, uint32_t row_end, int triangle,
                                 int kmer_size, double max_dist, int containment, const kssd_hit **d_hits,
                                 uint64_t *n_hits);

/* ---- set algebra on sketches ----------------------------------------------------------
 * The arithmetic of `rabbit_kssd union` (src/subCommand.cpp:307-543: OR every hash of every genome into a bitmap over the
 * hash space, list the set bits) and `rabbit_kssd sub` (:545-794: drop from every query genome the hashes present in the
 * reference sketches).  hash_bytes = 4 (uint32) or 8 (64-bit configuration); host pointers; results are library owned
 * until the next call.
 *   set_union     all hashes, any order, duplicates allowed -> the sorted distinct set (the order the bitmap scan gives);
 *   set_subtract  query sketches (CSR) minus a SORTED DISTINCT reference set (e.g. set_union's output): CSR of the
 *                 surviving hashes, their order inside a genome kept. */
int kssd_b200_set_union(kssd_ctx *ctx, const void *hashes, uint64_t n, int hash_bytes, const void **set, uint64_t *n_set);
int kssd_b200_set_subtract(kssd_ctx *ctx, const void *q_hashes, const uint64_t *q_off, uint32_t n_query, const void *ref_set,
                           uint64_t n_ref, int hash_bytes, const void **hashes, const uint64_t **genome_off);

/* set_subtract without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic
 * as kssd_b200_set_subtract for 32-bit hashes, but every input and output is a DEVICE pointer supplied by the caller
 * (d_out_hashes: room for n_hashes entries, d_out_off: n_query + 1 entries; n_hashes = q_off[n_query]) and the scratch is
 * a fixed 16.25 MiB __device__ array inside the library (held by every GPU that loads it, until exit).
 * Limits: n_hashes <= 4,194,304 and n_query <= 65,536 per call, larger sizes return KSSD_E_ARG.  The work is enqueued on
 * the context's stream (kssd_b200_sync waits for it); no cudaMalloc / cudaFree / allocating routine is called.  All calls
 * on a device share the array: calls that overlap on the GPU race -- serialise them (one stream, or sync between calls). */
int kssd_b200_set_subtract_noalloc(kssd_ctx *ctx, const uint32_t *d_q_hashes, const uint64_t *d_q_off, uint32_t n_query,
                                   uint64_t n_hashes, const uint32_t *d_ref_set, uint64_t n_ref, uint32_t *d_out_hashes,
                                   uint64_t *d_out_off);

/* ---- introspection for bench.py / tests ---------------------------------------------------------------- */
/* Bracket the scan / count kernels with CUDA events on the ctx stream (off by default). */
int kssd_b200_set_profiling(kssd_ctx *ctx, int on);
/* Tests: on = 1 takes the device-wide radix-sort paths even where the shared-memory segment sorts apply (per-genome
 * sort + unique, index build), so that both implementations are checked against the oracle. */
int kssd_b200_set_force_radix(kssd_ctx *ctx, int on);
/* Number of kernels this ctx has launched since init (bench.py's gpu_launches). */
uint64_t kssd_b200_launch_count(kssd_ctx *ctx);
/* Time (ms, CUDA events on the ctx stream) and launch count of the scan kernel since the last call. */
int kssd_b200_scan_kernel_time(kssd_ctx *ctx, double *ms, uint64_t *launches, uint64_t *bytes);
/* Posting-accumulate kernel: time, launches, posting entries streamed, smem atomic increments. */
int kssd_b200_count_kernel_time(kssd_ctx *ctx, double *ms, uint64_t *launches, uint64_t *postings,
                                uint64_t *atomics);

#ifdef __cplusplus
}
#endif
#endif /* KSSD_B200_H */
