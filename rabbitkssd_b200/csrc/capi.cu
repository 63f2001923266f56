// extern "C" entry points of libkssd_b200.so (declared in include/kssd_b200.h) + small runtime helpers.
#include <algorithm>
#include <cstdarg>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace kssd {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int DevBuf::reserve(size_t bytes, cudaStream_t st, bool keep) {
    if (bytes <= cap) return KSSD_OK;
    size_t ncap = std::max(bytes, cap + cap / 2);
    ncap = (ncap + 255) & ~(size_t)255;
    void *np = nullptr;
    KSSD_CUDA(cudaMalloc(&np, ncap));
    if (p) {
        KSSD_CUDA(cudaStreamSynchronize(st));
        if (keep) KSSD_CUDA(cudaMemcpy(np, p, cap, cudaMemcpyDeviceToDevice));
        KSSD_CUDA(cudaFree(p));
    }
    p = np;
    cap = ncap;
    return KSSD_OK;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}
int HostBuf::reserve(size_t bytes) {
    if (bytes <= cap) return KSSD_OK;
    size_t ncap = std::max(bytes, cap + cap / 2);
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    KSSD_CUDA(cudaMallocHost(&p, ncap));
    cap = ncap;
    return KSSD_OK;
}
void HostBuf::release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
}

int TimedLaunches::begin(cudaStream_t st) {
    cudaEvent_t e[2];
    for (int i = 0; i < 2; i++) {
        if (!pool.empty()) { e[i] = pool.back(); pool.pop_back(); }
        else KSSD_CUDA(cudaEventCreate(&e[i]));
    }
    ev.push_back(e[0]);
    ev.push_back(e[1]);
    KSSD_CUDA(cudaEventRecord(e[0], st));
    return KSSD_OK;
}
int TimedLaunches::end(cudaStream_t st) {
    KSSD_CUDA(cudaEventRecord(ev.back(), st));
    return KSSD_OK;
}
int TimedLaunches::collect(double *ms) {
    double total = 0;
    for (size_t i = 0; i + 1 < ev.size(); i += 2) {
        KSSD_CUDA(cudaEventSynchronize(ev[i + 1]));
        float t = 0;
        KSSD_CUDA(cudaEventElapsedTime(&t, ev[i], ev[i + 1]));
        total += t;
        pool.push_back(ev[i]);
        pool.push_back(ev[i + 1]);
    }
    ev.clear();
    *ms = total;
    return KSSD_OK;
}

static int check_ctx(kssd_ctx *c) {
    if (!c) { set_error("null context"); return KSSD_E_ARG; }
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) { set_error("cudaSetDevice(%d): %s", c->device, cudaGetErrorString(e)); return KSSD_E_CUDA; }
    return KSSD_OK;
}

}  // namespace kssd

using namespace kssd;

extern "C" {

const char *kssd_b200_last_error(void) { return g_err; }
int kssd_b200_version(void) { return 100; }

// Host only: the shuffle table is an INPUT of the GPU path and its generation stays on the CPU, as in the reference
// (src/shuffle.cpp:76-104 shuffleN/shuffle, :48-54 write_shuffle_dim_file): identity permutation of 16^half_subk ints,
// Fisher-Yates with glibc srand(23), then again with srand(id), id = (half_k << 8) + (half_subk << 4) + drlevel,
// j = rand() % (i + 1).  Same glibc => the same table as `rabbit_kssd shuffle` writes.
int kssd_b200_shuffle_table(int half_k, int half_subk, int drlevel, int32_t *out) {
    if (!out || half_subk < 1 || half_subk > 7 || half_k < half_subk || drlevel < 0 || drlevel >= half_subk) {
        set_error("shuffle_table: need 1 <= half_subk <= 7, half_k >= half_subk, 0 <= drlevel < half_subk");
        return KSSD_E_ARG;
    }
    const int n = 1 << (4 * half_subk);
    for (int i = 0; i < n; i++) out[i] = i;
    const unsigned seeds[2] = {23u, (unsigned)((half_k << 8) + (half_subk << 4) + drlevel)};
    for (unsigned seed : seeds) {
        srand(seed);
        for (int i = n - 1; i > 0; i--) std::swap(out[i], out[rand() % (i + 1)]);
    }
    return KSSD_OK;
}

int kssd_b200_device_count(int *n) {
    if (!n) { set_error("device_count: null out pointer"); return KSSD_E_ARG; }
    *n = 0;
    int k = 0;
    cudaError_t e = cudaGetDeviceCount(&k);
    if (e != cudaSuccess || k == 0) {
        set_error("no usable CUDA device (%s); libkssd_b200 has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
        return KSSD_E_CUDA;
    }
    *n = k;
    return KSSD_OK;
}

int kssd_b200_init(kssd_ctx **out, int device) {
    if (!out) { set_error("null out pointer"); return KSSD_E_ARG; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no usable CUDA device (%s); libkssd_b200 has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
        return KSSD_E_CUDA;
    }
    if (device < 0 || device >= n) { set_error("device %d out of range (0..%d)", device, n - 1); return KSSD_E_ARG; }
    KSSD_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    KSSD_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
        return KSSD_E_CUDA;
    }
    kssd_ctx *c = new kssd_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    KSSD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    KSSD_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    c->own_stream = true;
    memset(c->batches, 0, sizeof(c->batches));
    *out = c;
    return KSSD_OK;
}

void kssd_b200_destroy(kssd_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    peer_release(c);
    DevBuf *dev[] = {&c->d_bitmap, &c->d_tab_keys, &c->d_emit, &c->d_cursor, &c->d_stage_seq[0], &c->d_stage_seq[1],
                     &c->d_stage_qual[0], &c->d_stage_qual[1], &c->d_stage_segoff[0], &c->d_stage_segoff[1], &c->d_stage_seggen[0],
                     &c->d_stage_seggen[1], &c->d_sort_tmp, &c->d_sort_counts, &c->d_keys_unique, &c->d_hashes, &c->d_genome_off,
                     &c->d_uniq_hash, &c->d_uniq_off, &c->d_dict, &c->d_ref_sizes, &c->d_pairs, &c->d_pairs_tmp, &c->d_bucket, &c->d_q_hashes,
                     &c->d_q_off, &c->d_q_list_begin, &c->d_q_list_end, &c->d_hits, &c->d_hit_cursor, &c->d_rowwork, &c->d_gcount, &c->d_seg_start, &c->d_ucount,
                     &c->d_emit_rank, &c->d_cat_hashes, &c->d_cat_off, &c->d_cat_counts, &c->d_emit_gen, &c->d_h64_a, &c->d_h64_b, &c->d_dict64,
                     &c->d_hashes64, &c->d_uniq_hash64, &c->d_ranks};
    for (DevBuf *b : dev) b->release();
    HostBuf *host[] = {&c->h_hashes, &c->h_genome_off, &c->h_scalar, &c->h_uniq_hash, &c->h_uniq_count, &c->h_dict, &c->h_hits, &c->h_snap, &c->h_hashes64, &c->h_uniq_hash64};
    for (HostBuf *b : host) b->release();
    for (int i = 0; i < kssd_ctx::kBatches; i++) {
        c->bpriv[i].seq.release();
        c->bpriv[i].qual.release();
        c->bpriv[i].seg_off.release();
        c->bpriv[i].seg_genome.release();
        c->bpriv[i].seg_flags.release();
        c->bpriv[i].seg_first.release();
        if (c->batches_init) cudaEventDestroy(c->bpriv[i].h2d_done);
    }
    for (int i = 0; i < 2; i++) {
        if (c->stage_free[i]) cudaEventDestroy(c->stage_free[i]);
        c->d_stage_raw[i].release();
        c->d_stage_rawseg[i].release();
        c->d_stage_cseg[i].release();
    }
    c->d_fa_scratch.release();
    c->d_tile_rows.release();
    c->d_setop_in.release();
    c->d_setop_tmp.release();
    c->d_setop_out.release();
    c->h_setop.release();
    c->h_tile_rows.release();
    c->d_filter.release();
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    delete c;
}

int kssd_b200_set_stream(kssd_ctx *c, void *cuda_stream) {
    KSSD_TRY(check_ctx(c));
    KSSD_CUDA(cudaStreamSynchronize(c->stream));
    if (cuda_stream == nullptr) {
        if (!c->own_stream) {
            KSSD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
            c->own_stream = true;
        }
        return KSSD_OK;
    }
    if (c->own_stream) cudaStreamDestroy(c->stream);
    c->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    c->own_stream = false;
    return KSSD_OK;
}

int kssd_b200_sync(kssd_ctx *c) {
    KSSD_TRY(check_ctx(c));
    KSSD_CUDA(cudaStreamSynchronize(c->stream));
    return KSSD_OK;
}

int kssd_b200_set_profiling(kssd_ctx *c, int on) {
    KSSD_TRY(check_ctx(c));
    c->profiling = on != 0;
    return KSSD_OK;
}

int kssd_b200_set_relax_guard(kssd_ctx *c, int on) {
    KSSD_TRY(check_ctx(c));
    c->relax_guard = on != 0;
    return KSSD_OK;
}

int kssd_b200_set_force_radix(kssd_ctx *c, int on) {
    KSSD_TRY(check_ctx(c));
    c->force_radix = on != 0;
    return KSSD_OK;
}

int kssd_b200_set_params(kssd_ctx *c, int half_k, int half_subk, int drlevel, const int32_t *shuffled_dim) {
    KSSD_TRY(check_ctx(c));
    if (!shuffled_dim) { set_error("set_params: null shuffle table"); return KSSD_E_ARG; }
    if (half_subk - drlevel < 3 && !c->relax_guard) {  // common.cpp:37-39
        set_error("the half_subk - drlevel should at least 3; current half_subk and drlevel are: %d and %d", half_subk, drlevel);
        return KSSD_E_PARAM;
    }
    if (half_k < half_subk || half_subk < 1 || half_subk >= 8 || drlevel < 0 || half_k > 16 || drlevel >= half_subk) {  // shuffle.cpp:26-33
        set_error("set_params: need half_subk <= half_k <= 16, 1 <= half_subk < 8 (got k=%d s=%d l=%d)", half_k, half_subk, drlevel);
        return KSSD_E_ARG;
    }
    c->hash64 = half_k - drlevel > 8;  // sketch.cpp:336 use64: hashes of up to 4 * (half_k - drlevel) bits (hash64.cu)
    ScanParams &sp = c->sp;
    sp.half_k = half_k;
    sp.half_subk = half_subk;
    sp.drlevel = drlevel;
    sp.hoc = half_k - half_subk;
    sp.kmer_size = 2 * half_k;
    sp.dim_end = 1 << 4 * (half_subk - drlevel);
    const int comp_bittl = 64 - 4 * half_k;
    sp.tupmask = 0xffffffffffffffffULL >> comp_bittl;
    sp.domask = (sp.tupmask >> (4 * sp.hoc)) << (2 * sp.hoc);
    const uint64_t undomask = (sp.tupmask ^ sp.domask) & sp.tupmask;
    sp.undomask1 = undomask & (sp.tupmask >> ((half_k + half_subk) * 2));
    sp.undomask0 = undomask ^ sp.undomask1;
    c->have_params = false;
    KSSD_TRY(build_tables(c, shuffled_dim));
    c->have_params = true;
    return KSSD_OK;
}

uint64_t kssd_b200_launch_count(kssd_ctx *c) { return c ? c->launches : 0; }

int kssd_b200_scan_kernel_time(kssd_ctx *c, double *ms, uint64_t *launches, uint64_t *bytes) {
    KSSD_TRY(check_ctx(c));
    double t = 0;
    KSSD_TRY(c->t_scan.collect(&t));
    if (ms) *ms = t;
    if (launches) *launches = c->t_scan.launches;
    if (bytes) *bytes = c->t_scan.units_a;
    c->t_scan.launches = c->t_scan.units_a = c->t_scan.units_b = 0;
    return KSSD_OK;
}

int kssd_b200_count_kernel_time(kssd_ctx *c, double *ms, uint64_t *launches, uint64_t *postings, uint64_t *atomics) {
    KSSD_TRY(check_ctx(c));
    double t = 0;
    KSSD_TRY(c->t_count.collect(&t));
    if (ms) *ms = t;
    if (launches) *launches = c->t_count.launches;
    if (postings) *postings = c->t_count.units_a;
    if (atomics) *atomics = c->t_count.units_b;
    c->t_count.launches = c->t_count.units_a = c->t_count.units_b = 0;
    return KSSD_OK;
}

}  // extern "C"

#include "capi_sketch.inc"
#include "capi_dist.inc"
