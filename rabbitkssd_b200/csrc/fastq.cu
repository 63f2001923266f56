// FASTQ tokenisation on the device: raw text of four-line FASTQ records -> the sequence bytes and the
// quality bytes, position by position, as the scan kernel reads them.
//
// Replaces, for plain (uncompressed) four-line FASTQ, what kseq does on the host (src/kseq.h:175-215: '@' line = name, the next
// line the sequence, a line starting with '+', then quality lines until as many quality bytes as bases were read; '\r' in
// front of '\n' dropped, :143) and RabbitFX's FASTQ formatter for the big-file path (RabbitFX/io/Formater.cpp).  A batch
// holds PIECES of raw text; a piece is a whole number of records (the host cuts at every fourth line end), ends with '\n' and is
// padded with 0x00 to a multiple of 32 bytes.  The role of a byte is its line number modulo four, so line state is a plain prefix
// sum of newline counts over the whole batch (every piece holds a multiple of four lines; this is checked).  Output: the
// sequence lines and the quality lines, each line followed by ONE separator ('\n': a non-base byte resets the scan's valid-run
// counter exactly as a new read does, src/sketch.cpp:273-283), compacted into two parallel buffers -- byte i of the one is the
// base whose quality is byte i of the other.
//
// Anything that kseq would parse differently from "line number modulo four" raises the anomaly flag (the session then ends
// with KSSD_E_SYNTAX and the host falls back to its parser): a record line not starting with '@', a third line not starting
// with '+', a sequence line starting with '@', '>' or '+', sequence and quality of different lengths (multi-line records), a
// piece that is not a whole number of records, 0x00 bytes that are not padding.
#include "common.cuh"

namespace kssd {

static constexpr int kFqChunk = 32;             // bytes per thread
static constexpr int kFqThreads = 256;          // chunks per tile

struct FqArgs {
    const uint8_t *raw;          // n_chunks * 32 bytes
    uint32_t n_chunks;
    uint32_t n_tiles;
    const uint64_t *seg_off;     // raw offsets of the pieces, multiples of 32; [n_seg] = n_chunks * 32
    uint32_t n_seg;
    uint32_t *tile_nl;           // newlines per tile
    const uint64_t *tile_nl_base;// exclusive scan of tile_nl, [n_tiles] = total
    uint32_t *keep_s, *keep_q;   // per chunk: bit i = byte i goes to the sequence / quality output
    uint16_t *pre_s, *pre_q;     // kept bytes of the tile before the chunk
    uint32_t *tile_cnt_s, *tile_cnt_q;
    const uint64_t *tile_base_s, *tile_base_q;  // exclusive scans, [n_tiles] = totals
    uint8_t *out_s, *out_q;
    uint64_t *cseg_off;          // [n_seg + 1] offsets of the pieces in the outputs
    uint32_t *anomaly;
};

struct FqChunk {
    uint32_t w[8];   // the 32 bytes
    uint32_t prev;   // the byte before the chunk ('\n' before the batch)
    uint32_t next;   // the byte after the chunk (0 after the batch)
    bool seg_start;  // the chunk is the first of its piece
    bool next_start; // the next chunk begins a piece (or the batch ends)
};

__device__ __forceinline__ void fq_load_bytes(const FqArgs &a, uint32_t c, uint32_t (&w)[8]) {
    const uint4 *p = reinterpret_cast<const uint4 *>(a.raw + (uint64_t)c * kFqChunk);
    const uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
    w[0] = v0.x; w[1] = v0.y; w[2] = v0.z; w[3] = v0.w;
    w[4] = v1.x; w[5] = v1.y; w[6] = v1.z; w[7] = v1.w;
}
__device__ __forceinline__ uint32_t fq_byte(const uint32_t (&w)[8], int i) { return (w[i >> 2] >> (8 * (i & 3))) & 0xFFu; }

__device__ __forceinline__ uint32_t fq_count_newlines(const uint32_t (&w)[8]) {
    uint32_t n = 0;
#pragma unroll
    for (int i = 0; i < kFqChunk; i++) n += fq_byte(w, i) == '\n';
    return n;
}

// exclusive prefix sum of v over the CTA; *total (optional, thread 0 writes it) = the sum
__device__ __forceinline__ uint32_t fq_block_exclusive_sum(uint32_t v, uint32_t *s_warp /* [blockDim/32] */, uint32_t *total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o) x += y;
    }
    __syncthreads();  // s_warp may still be read from a previous call
    if (lane == 31) s_warp[w] = x;
    __syncthreads();
    uint32_t before = 0, tot = 0;
    for (int q = 0; q < n_warps; q++) { const uint32_t t = s_warp[q]; if (q < w) before += t; tot += t; }
    if (total && threadIdx.x == 0) *total = tot;
    return before + x - v;
}

__global__ void __launch_bounds__(kFqThreads) fq_count_kernel(const FqArgs a) {
    __shared__ uint32_t s_warp[kFqThreads / 32];
    const uint32_t c = blockIdx.x * kFqThreads + threadIdx.x;
    uint32_t n = 0;
    if (c < a.n_chunks) {
        uint32_t w[8];
        fq_load_bytes(a, c, w);
        n = fq_count_newlines(w);
    }
    fq_block_exclusive_sum(n, s_warp, a.tile_nl + blockIdx.x);
}

__device__ __forceinline__ void fq_load(const FqArgs &a, uint32_t c, FqChunk &k) {
    fq_load_bytes(a, c, k.w);
    const uint64_t pos = (uint64_t)c * kFqChunk;
    uint32_t lo = 0, hi = a.n_seg;  // piece that contains the chunk: largest s with seg_off[s] <= pos
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(a.seg_off + mid) <= pos) lo = mid; else hi = mid;
    }
    k.seg_start = __ldg(a.seg_off + lo) == pos;
    k.next_start = __ldg(a.seg_off + lo + 1) == pos + kFqChunk;
    k.prev = c ? (uint32_t)__ldg(a.raw + pos - 1) : (uint32_t)'\n';
    k.next = c + 1 < a.n_chunks ? (uint32_t)__ldg(a.raw + pos + kFqChunk) : 0u;
}

// Roles of the 32 bytes of a chunk whose first byte lies on line `line` (batch-wide count of '\n' before it): the masks of
// the bytes that go to the sequence and to the quality output, and whether the text breaks a rule (header comment).
__device__ __forceinline__ bool fq_classify(const FqChunk &k, uint32_t line, uint32_t &keep_s, uint32_t &keep_q) {
    bool bad = k.seg_start && (line & 3u) != 0;  // the pieces before this one do not hold a multiple of four lines
    keep_s = keep_q = 0;
#pragma unroll
    for (int i = 0; i < kFqChunk; i++) {
        const uint32_t b = fq_byte(k.w, i);
        const uint32_t before = i ? fq_byte(k.w, i ? i - 1 : 0) : k.prev;
        const uint32_t after = i + 1 < kFqChunk ? fq_byte(k.w, i + 1 < kFqChunk ? i + 1 : 0) : k.next;
        const bool line_start = before == '\n' || (i == 0 && k.seg_start);
        const uint32_t role = line & 3u;
        if (b == 0) {  // padding: only after the last record of a piece, up to the start of the next piece
            bad |= role != 0;
            bad |= after != 0 && !(i == kFqChunk - 1 && k.next_start);
        } else {
            bad |= before == 0 && !(i == 0 && k.seg_start);  // text after padding inside a piece
            if (line_start) {
                bad |= role == 0 && b != '@';
                bad |= role == 2 && b != '+';
                bad |= role == 1 && (b == '@' || b == '>' || b == '+');
            }
            const bool take = b == '\n' || !(b == '\r' && after == '\n');  // the line end stays as the separator
            if (role == 1 && take) keep_s |= 1u << i;
            if (role == 3 && take) keep_q |= 1u << i;
            if (b == '\n') line++;
        }
    }
    return bad;
}

__global__ void __launch_bounds__(kFqThreads) fq_classify_kernel(const FqArgs a) {
    __shared__ uint32_t s_warp[kFqThreads / 32];
    const uint32_t c = blockIdx.x * kFqThreads + threadIdx.x;
    const bool live = c < a.n_chunks;
    FqChunk k;
    uint32_t nl = 0;
    if (live) {
        fq_load(a, c, k);
        nl = fq_count_newlines(k.w);
    }
    const uint32_t line = (uint32_t)a.tile_nl_base[blockIdx.x] + fq_block_exclusive_sum(nl, s_warp, nullptr);  // modulo 2^32: only modulo 4 matters
    uint32_t ks = 0, kq = 0;
    if (live) {
        if (fq_classify(k, line, ks, kq)) atomicExch(a.anomaly, 1u);
        a.keep_s[c] = ks;
        a.keep_q[c] = kq;
    }
    const uint32_t ns = __popc(ks), nq = __popc(kq);
    const uint32_t ps = fq_block_exclusive_sum(ns, s_warp, a.tile_cnt_s + blockIdx.x);
    const uint32_t pq = fq_block_exclusive_sum(nq, s_warp, a.tile_cnt_q + blockIdx.x);
    if (live) {
        a.pre_s[c] = (uint16_t)ps;
        a.pre_q[c] = (uint16_t)pq;
    }
}

// kept bytes go through shared memory and leave with 16-byte stores (store_tile_bytes, common.cuh)
__global__ void __launch_bounds__(kFqThreads) fq_scatter_kernel(const FqArgs a) {
    __shared__ __align__(16) uint8_t s_buf[kFqChunk * kFqThreads + 16];
    const uint32_t c = blockIdx.x * kFqThreads + threadIdx.x;
    const bool live = c < a.n_chunks;
    uint32_t ks = 0, kq = 0, ps = 0, pq = 0;
    uint32_t w[8];
    if (live) {
        ks = a.keep_s[c];
        kq = a.keep_q[c];
        ps = a.pre_s[c];
        pq = a.pre_q[c];
        if (ks | kq) fq_load_bytes(a, c, w);
    }
    const uint64_t s_base = a.tile_base_s[blockIdx.x], q_base = a.tile_base_q[blockIdx.x];
    const uint32_t s_cnt = a.tile_cnt_s[blockIdx.x], q_cnt = a.tile_cnt_q[blockIdx.x];
    // sequence bytes
    uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(a.out_s + s_base) & 15u);
    uint8_t *o = s_buf + shift + ps;
    for (uint32_t m = ks; m; m &= m - 1) {
        const int i = __ffs(m) - 1;
        *o++ = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
    }
    __syncthreads();
    store_tile_bytes(s_buf, shift, s_cnt, a.out_s + s_base);
    __syncthreads();
    // quality bytes + the per-record length check
    shift = (uint32_t)(reinterpret_cast<uintptr_t>(a.out_q + q_base) & 15u);
    o = s_buf + shift + pq;
    bool bad = false;
    for (uint32_t m = kq; m; m &= m - 1) {
        const int i = __ffs(m) - 1;
        const uint32_t b = (w[i >> 2] >> (8 * (i & 3))) & 0xFFu;
        *o++ = (uint8_t)b;
        if (b == '\n') {  // end of a record: as many quality bytes as sequence bytes (both counts include the line's separator)
            const uint32_t below = (1u << i) - 1u;
            bad |= s_base + ps + __popc(ks & below) != q_base + pq + __popc(kq & below) + 1;
        }
    }
    if (bad) atomicExch(a.anomaly, 1u);
    __syncthreads();
    store_tile_bytes(s_buf, shift, q_cnt, a.out_q + q_base);
}

// piece table of the outputs for the scan kernel + the batch-wide checks
__global__ void __launch_bounds__(256) fq_segments_kernel(const FqArgs a) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s > a.n_seg) return;
    if (s == a.n_seg) {
        a.cseg_off[s] = a.tile_base_s[a.n_tiles];
        if ((a.tile_nl_base[a.n_tiles] & 3ull) || a.tile_base_s[a.n_tiles] != a.tile_base_q[a.n_tiles]) atomicExch(a.anomaly, 1u);
        return;
    }
    const uint32_t c = (uint32_t)(a.seg_off[s] / kFqChunk);  // a piece starts at a chunk
    a.cseg_off[s] = a.tile_base_s[c / kFqThreads] + a.pre_s[c];
}

__global__ void segment_scan_kernel(const uint32_t *counts, uint32_t n, uint64_t *start, uint32_t *max_out);  // segsort.cu

// d_raw[0, n_raw) (n_raw a multiple of 32) + piece offsets (device) -> d_out_seq / d_out_qual (capacity n_raw rounded up to
// 16, 16-byte aligned), d_cseg_off[n_seg + 1].  Stream-ordered, no host synchronisation: *out_bytes is an upper bound of
// the compacted size; the tail of d_out_seq holds separators.
int fastq_tokenise(kssd_ctx *c, const uint8_t *d_raw, uint64_t n_raw, const uint64_t *d_seg_off, uint32_t n_seg, uint8_t *d_out_seq,
                   uint8_t *d_out_qual, uint64_t *d_cseg_off, uint32_t *d_anomaly, uint64_t *out_bytes) {
    cudaStream_t st = c->stream;
    if (n_raw % kFqChunk || n_raw >= (1ull << 32) * kFqChunk) { set_error("fastq: raw batch must be a multiple of 32 bytes"); return KSSD_E_ARG; }
    FqArgs a;
    a.raw = d_raw;
    a.n_chunks = (uint32_t)(n_raw / kFqChunk);
    a.n_tiles = (a.n_chunks + kFqThreads - 1) / kFqThreads;
    a.seg_off = d_seg_off;
    a.n_seg = n_seg;
    // scratch: 3 x u64[n_tiles+1] scans | 3 x u32[n_tiles] counts | 2 x u32[n_chunks] masks | 2 x u16[n_chunks] offsets
    const uint64_t nt = a.n_tiles, nc = a.n_chunks;
    KSSD_TRY(c->d_fa_scratch.reserve(3 * (nt + 1) * 8 + 3 * nt * 4 + 2 * nc * 4 + 2 * nc * 2 + 64, st));
    uint64_t *scans = c->d_fa_scratch.as<uint64_t>();
    uint64_t *nl_base = scans, *base_s = scans + (nt + 1), *base_q = scans + 2 * (nt + 1);
    a.tile_nl_base = nl_base;
    a.tile_base_s = base_s;
    a.tile_base_q = base_q;
    a.tile_nl = reinterpret_cast<uint32_t *>(scans + 3 * (nt + 1));
    a.tile_cnt_s = a.tile_nl + nt;
    a.tile_cnt_q = a.tile_cnt_s + nt;
    a.keep_s = a.tile_cnt_q + nt;
    a.keep_q = a.keep_s + nc;
    a.pre_s = reinterpret_cast<uint16_t *>(a.keep_q + nc);
    a.pre_q = a.pre_s + nc;
    a.out_s = d_out_seq;
    a.out_q = d_out_qual;
    a.cseg_off = d_cseg_off;
    a.anomaly = d_anomaly;
    const uint64_t cap = (n_raw + 15) & ~15ull;
    *out_bytes = cap;
    KSSD_CUDA(cudaMemsetAsync(d_out_seq, '\n', cap, st));
    if (a.n_chunks == 0 || n_seg == 0) return KSSD_OK;
    fq_count_kernel<<<a.n_tiles, kFqThreads, 0, st>>>(a);
    segment_scan_kernel<<<1, 1024, 0, st>>>(a.tile_nl, a.n_tiles, nl_base, nullptr);
    fq_classify_kernel<<<a.n_tiles, kFqThreads, 0, st>>>(a);
    segment_scan_kernel<<<1, 1024, 0, st>>>(a.tile_cnt_s, a.n_tiles, base_s, nullptr);
    segment_scan_kernel<<<1, 1024, 0, st>>>(a.tile_cnt_q, a.n_tiles, base_q, nullptr);
    fq_scatter_kernel<<<a.n_tiles, kFqThreads, 0, st>>>(a);
    fq_segments_kernel<<<(n_seg + 1 + 255) / 256, 256, 0, st>>>(a);
    KSSD_CUDA(cudaGetLastError());
    c->launches += 7;
    return KSSD_OK;
}

}  // namespace kssd
