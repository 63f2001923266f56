// FASTA tokenisation on the device: raw file text -> the parsed layout the scan kernel reads.
//
// Replaces, for plain (uncompressed) FASTA, what kseq does on the host (src/kseq.h:175-215: a record starts at a line that
// begins with '>' -- or '@' --, the rest of that line is name + comment, the following lines up to the next such line are
// the sequence with the line ends removed, a '\r' in front of a '\n' is dropped, :143) and what RabbitFX's formatter does
// for the big-file path (RabbitFX/io/Formater.cpp:87-129).  The host only reads file bytes into pinned memory; a batch holds
// PIECES of raw text (a piece = a whole file or a slice of a large one that starts at a line start, inside a header line or
// inside a sequence line), each piece padded with '\n' to a multiple of 32 bytes.  Output per piece: its sequence bytes with every header
// line replaced by ONE separator byte (any non-ACGT byte resets the scan's valid-run counter exactly as a new record does,
// src/sketch.cpp:487-488) and one more separator after the piece.
//
// Line state ("is this byte inside a header line") is a prefix property of the text: a chunk of 32 bytes either contains a
// line start (then the state after the chunk is decided inside it) or inherits.  Three levels resolve it: per-tile
// aggregates (fa_tile_state_kernel), one single-CTA scan over the tiles (fa_scan_state_kernel), and the per-chunk scan
// inside a tile when the bytes are classified (fa_classify_kernel, which also counts the kept bytes); kept bytes are then
// scattered to their compacted positions (fa_scatter_kernel).
#include "common.cuh"

namespace kssd {

static constexpr int kFaChunk = 32;             // bytes per thread
static constexpr int kFaThreads = 256;          // chunks per tile
static constexpr int kFaTile = kFaChunk * kFaThreads;

struct FaArgs {
    const uint8_t *raw;        // n_chunks * 32 bytes
    uint32_t n_chunks;
    uint32_t n_tiles;
    const uint64_t *seg_off;   // raw offsets of the pieces, multiples of 32; [n_seg] = n_chunks * 32
    const uint8_t *seg_flags;  // bit 0: the piece starts inside a header line; bit 1: inside a sequence line; else at a line start
    const uint32_t *seg_first; // raw offset inside the piece from which k-mers are new (bytes before it are context)
    uint32_t n_seg;
    uint8_t *tile_state;       // per tile: 0 inherit, 1 sequence line, 2 header line (state after the tile)
    uint8_t *tile_in;          // state before the tile (1 / 2)
    uint32_t *chunk_keep;      // per chunk: bit i = byte i is kept
    uint16_t *chunk_pre;       // kept bytes of the tile before the chunk
    uint32_t *tile_count;
    const uint64_t *tile_base; // exclusive scan of tile_count, [n_tiles] = total
    uint8_t *out;              // compacted text (pre-filled with separators)
    uint64_t *cseg_off;        // [n_seg + 1] compacted offsets of the pieces
    uint64_t *cseg_first;      // [n_seg] compacted position of seg_first
    uint32_t *anomaly;         // set when the text is not plain FASTA (a sequence line starting with '+': kseq switches to FASTQ there)
};

__device__ __forceinline__ bool fa_is_header_char(uint32_t b) { return b == '>' || b == '@'; }

// piece that contains chunk c: largest s with seg_off[s] <= c * 32
__device__ __forceinline__ uint32_t fa_segment_of(const FaArgs &a, uint32_t c) {
    const uint64_t pos = (uint64_t)c * kFaChunk;
    uint32_t lo = 0, hi = a.n_seg;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(a.seg_off + mid) <= pos) lo = mid; else hi = mid;
    }
    return lo;
}

struct FaChunk {
    uint32_t w[8];       // the 32 bytes
    uint32_t next;       // the byte after the chunk ('\n' at the end of the piece)
    uint32_t seg;
    bool seg_start;      // first chunk of its piece
    bool forced_hdr;     // seg_start: state of the first line
    bool mid_line;       // the piece starts inside a sequence line: its first byte is not a line start
};

__device__ __forceinline__ void fa_load(const FaArgs &a, uint32_t c, FaChunk &k) {
    const uint4 *p = reinterpret_cast<const uint4 *>(a.raw + (uint64_t)c * kFaChunk);
    const uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
    k.w[0] = v0.x; k.w[1] = v0.y; k.w[2] = v0.z; k.w[3] = v0.w;
    k.w[4] = v1.x; k.w[5] = v1.y; k.w[6] = v1.z; k.w[7] = v1.w;
    k.seg = fa_segment_of(a, c);
    const uint64_t s0 = __ldg(a.seg_off + k.seg), s1 = __ldg(a.seg_off + k.seg + 1);
    const uint64_t pos = (uint64_t)c * kFaChunk;
    k.seg_start = pos == s0;
    k.next = pos + kFaChunk < s1 ? (uint32_t)__ldg(a.raw + pos + kFaChunk) : (uint32_t)'\n';
    const uint32_t fl = __ldg(a.seg_flags + k.seg);
    k.mid_line = (fl & 2u) != 0;
    k.forced_hdr = k.seg_start && ((fl & 1u) || (!k.mid_line && fa_is_header_char(k.w[0] & 0xFFu)));
}

__device__ __forceinline__ uint32_t fa_byte(const FaChunk &k, int i) { return i < kFaChunk ? (k.w[i >> 2] >> (8 * (i & 3))) & 0xFFu : k.next; }

// state after the chunk: 0 inherit (no line start inside), 1 sequence line, 2 header line
__device__ __forceinline__ uint32_t fa_chunk_state(const FaChunk &k) {
    uint32_t st = k.seg_start ? (k.forced_hdr ? 2u : 1u) : 0u;
#pragma unroll
    for (int i = 0; i < kFaChunk; i++)
        if (fa_byte(k, i) == '\n') st = fa_is_header_char(fa_byte(k, i + 1)) ? 2u : 1u;
    return st;
}

// inclusive "last non-zero value" scan over the CTA; returns the EXCLUSIVE result for this thread (0: nothing before it)
__device__ __forceinline__ uint32_t fa_block_exclusive_last(uint32_t v, uint32_t *s_warp /* [blockDim/32] */, uint32_t *tile_total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o && x == 0) x = y;
    }
    if (lane == 31) s_warp[w] = x;
    uint32_t excl = __shfl_up_sync(0xFFFFFFFFu, x, 1);
    if (lane == 0) excl = 0;
    __syncthreads();
    uint32_t before = 0;
    for (int q = 0; q < w; q++) { const uint32_t t = s_warp[q]; if (t) before = t; }
    if (tile_total && threadIdx.x == 0) {
        uint32_t t = 0;
        for (int q = 0; q < n_warps; q++) { const uint32_t u = s_warp[q]; if (u) t = u; }
        *tile_total = t;
    }
    __syncthreads();
    return excl ? excl : before;
}

__global__ void __launch_bounds__(kFaThreads) fa_tile_state_kernel(const FaArgs a) {
    __shared__ uint32_t s_warp[kFaThreads / 32];
    __shared__ uint32_t s_total;
    const uint32_t c = blockIdx.x * kFaThreads + threadIdx.x;
    uint32_t st = 0;
    if (c < a.n_chunks) {
        FaChunk k;
        fa_load(a, c, k);
        st = fa_chunk_state(k);
    }
    fa_block_exclusive_last(st, s_warp, &s_total);
    if (threadIdx.x == 0) a.tile_state[blockIdx.x] = (uint8_t)s_total;
}

// single CTA: tile_in[t] = last non-zero tile_state before t (1 when there is none: a piece always starts with a forced state)
__global__ void __launch_bounds__(1024) fa_scan_state_kernel(const FaArgs a) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_total;
    uint32_t carry = 1;
    for (uint32_t base = 0; base < a.n_tiles; base += 1024) {
        const uint32_t t = base + threadIdx.x;
        const uint32_t v = t < a.n_tiles ? a.tile_state[t] : 0;
        const uint32_t ex = fa_block_exclusive_last(v, s_warp, &s_total);
        if (t < a.n_tiles) a.tile_in[t] = (uint8_t)(ex ? ex : carry);
        if (s_total) carry = s_total;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kFaThreads) fa_classify_kernel(const FaArgs a) {
    __shared__ uint32_t s_warp[kFaThreads / 32];
    __shared__ uint32_t s_cnt[kFaThreads / 32];
    const uint32_t c = blockIdx.x * kFaThreads + threadIdx.x;
    const bool live = c < a.n_chunks;
    FaChunk k;
    uint32_t st_out = 0;
    if (live) {
        fa_load(a, c, k);
        st_out = fa_chunk_state(k);
    }
    uint32_t st = fa_block_exclusive_last(st_out, s_warp, nullptr);
    if (st == 0) st = a.tile_in[blockIdx.x];
    uint32_t keep = 0;
    if (live) {
        bool hdr = k.seg_start ? k.forced_hdr : st == 2;
        bool bad = k.seg_start && !k.forced_hdr && !k.mid_line && (k.w[0] & 0xFFu) == '+';
#pragma unroll
        for (int i = 0; i < kFaChunk; i++) {
            const uint32_t b = fa_byte(k, i);
            if (b == '\n') {
                if (hdr) keep |= 1u << i;  // the header line becomes one separator byte
                const uint32_t nb = fa_byte(k, i + 1);
                hdr = fa_is_header_char(nb);
                bad |= nb == '+';
            } else if (!hdr && !(b == '\r' && fa_byte(k, i + 1) == '\n')) {
                keep |= 1u << i;
            }
        }
        if (bad) atomicExch(a.anomaly, 1u);
        a.chunk_keep[c] = keep;
    }
    // kept bytes of the tile before this chunk
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t n = __popc(keep);
    uint32_t x = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_cnt[w] = x;
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (int q = 0; q < kFaThreads / 32; q++) { const uint32_t t = s_cnt[q]; if (q < w) before += t; total += t; }
    if (live) a.chunk_pre[c] = (uint16_t)(before + x - n);
    if (threadIdx.x == 0) a.tile_count[blockIdx.x] = total;
}

// The kept bytes of a tile, with the one separator that follows every piece that ends inside it, are contiguous in the
// output: staged in shared memory (pre-filled with separators), written with 16-byte stores (store_tile_bytes).
__global__ void __launch_bounds__(kFaThreads) fa_scatter_kernel(const FaArgs a) {
    __shared__ __align__(16) uint8_t s_buf[kFaTile + kFaThreads + 16];  // at most one piece boundary per chunk
    __shared__ uint32_t s_seg[2];
    const uint32_t c = blockIdx.x * kFaThreads + threadIdx.x;
    const bool live = c < a.n_chunks;
    const uint32_t seg = live ? fa_segment_of(a, c) : 0;
    const uint32_t n_live = min((uint32_t)kFaThreads, a.n_chunks - blockIdx.x * kFaThreads);
    if (threadIdx.x == 0) s_seg[0] = seg;
    if (threadIdx.x == n_live - 1) s_seg[1] = seg;
    for (uint32_t v = threadIdx.x; v < sizeof(s_buf) / 16; v += blockDim.x) reinterpret_cast<uint4 *>(s_buf)[v] = make_uint4(0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au);
    __syncthreads();
    // one separator follows every piece: pieces before this one shift the output by their number
    uint8_t *out = a.out + a.tile_base[blockIdx.x] + s_seg[0];
    const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(out) & 15u);
    if (live) {
        uint32_t keep = a.chunk_keep[c];
        if (keep) {
            const uint4 *p = reinterpret_cast<const uint4 *>(a.raw + (uint64_t)c * kFaChunk);
            const uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
            const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            uint8_t *o = s_buf + shift + a.chunk_pre[c] + (seg - s_seg[0]);
            while (keep) {
                const int i = __ffs(keep) - 1;
                keep &= keep - 1;
                *o++ = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
            }
        }
    }
    __syncthreads();
    store_tile_bytes(s_buf, shift, a.tile_count[blockIdx.x] + (s_seg[1] - s_seg[0]), out);
}

// compacted piece table for the scan kernel
__global__ void __launch_bounds__(256) fa_segments_kernel(const FaArgs a) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s > a.n_seg) return;
    if (s == a.n_seg) {
        a.cseg_off[s] = a.tile_base[a.n_tiles] + a.n_seg;
        return;
    }
    auto compacted = [&](uint64_t raw_pos) {  // compacted position of a raw position inside piece s
        const uint32_t c = (uint32_t)(raw_pos / kFaChunk), r = (uint32_t)(raw_pos % kFaChunk);
        return a.tile_base[c / kFaThreads] + a.chunk_pre[c] + __popc(a.chunk_keep[c] & ((1u << r) - 1u)) + s;
    };
    const uint64_t s0 = a.seg_off[s], s1 = a.seg_off[s + 1];
    a.cseg_off[s] = compacted(s0);
    const uint64_t f = s0 + a.seg_first[s];
    a.cseg_first[s] = f < s1 ? compacted(f) : compacted(s1 - kFaChunk) + __popc(a.chunk_keep[(s1 - kFaChunk) / kFaChunk]);
}

__global__ void segment_scan_kernel(const uint32_t *counts, uint32_t n, uint64_t *start, uint32_t *max_out);  // segsort.cu

// d_raw[0, n_raw) (n_raw a multiple of 32) + piece table (device) -> d_out (capacity n_raw + n_seg, 16-byte aligned),
// d_cseg_off[n_seg + 1], d_cseg_first[n_seg].  Stream-ordered, no host synchronisation: *out_bytes = n_raw + n_seg rounded
// up to 16 is an upper bound of the compacted size; the tail of d_out holds separators.
int fasta_tokenise(kssd_ctx *c, const uint8_t *d_raw, uint64_t n_raw, const uint64_t *d_seg_off, const uint8_t *d_seg_flags, const uint32_t *d_seg_first,
                   uint32_t n_seg, uint8_t *d_out, uint64_t *d_cseg_off, uint64_t *d_cseg_first, uint32_t *d_anomaly, uint64_t *out_bytes) {
    cudaStream_t st = c->stream;
    if (n_raw % kFaChunk || n_raw >= (1ull << 32) * kFaChunk) { set_error("fasta: raw batch must be a multiple of 32 bytes"); return KSSD_E_ARG; }
    FaArgs a;
    a.raw = d_raw;
    a.n_chunks = (uint32_t)(n_raw / kFaChunk);
    a.n_tiles = (a.n_chunks + kFaThreads - 1) / kFaThreads;
    a.seg_off = d_seg_off;
    a.seg_flags = d_seg_flags;
    a.seg_first = d_seg_first;
    a.n_seg = n_seg;
    // scratch: tile_base u64[n_tiles+1] | tile_count u32[n_tiles] | chunk_keep u32[n_chunks] | chunk_pre u16[n_chunks] | tile_state, tile_in u8[n_tiles]
    const uint64_t nt = a.n_tiles, nc = a.n_chunks;
    KSSD_TRY(c->d_fa_scratch.reserve((nt + 1) * 8 + nt * 4 + nc * 4 + ((nc * 2 + 3) & ~3ull) + 2 * nt + 64, st));
    uint8_t *base = c->d_fa_scratch.as<uint8_t>();
    uint64_t *tile_base = reinterpret_cast<uint64_t *>(base);
    a.tile_base = tile_base;
    a.tile_count = reinterpret_cast<uint32_t *>(tile_base + nt + 1);
    a.chunk_keep = a.tile_count + nt;
    a.chunk_pre = reinterpret_cast<uint16_t *>(a.chunk_keep + nc);
    a.tile_state = reinterpret_cast<uint8_t *>(a.chunk_pre) + ((nc * 2 + 3) & ~3ull);
    a.tile_in = a.tile_state + nt;
    a.out = d_out;
    a.cseg_off = d_cseg_off;
    a.cseg_first = d_cseg_first;
    a.anomaly = d_anomaly;
    const uint64_t cap = (n_raw + n_seg + 15) & ~15ull;
    *out_bytes = cap;
    KSSD_CUDA(cudaMemsetAsync(d_out, '\n', cap, st));
    if (a.n_chunks == 0 || n_seg == 0) return KSSD_OK;
    fa_tile_state_kernel<<<a.n_tiles, kFaThreads, 0, st>>>(a);
    fa_scan_state_kernel<<<1, 1024, 0, st>>>(a);
    fa_classify_kernel<<<a.n_tiles, kFaThreads, 0, st>>>(a);
    segment_scan_kernel<<<1, 1024, 0, st>>>(a.tile_count, a.n_tiles, tile_base, nullptr);
    fa_scatter_kernel<<<a.n_tiles, kFaThreads, 0, st>>>(a);
    fa_segments_kernel<<<(n_seg + 1 + 255) / 256, 256, 0, st>>>(a);
    KSSD_CUDA(cudaGetLastError());
    c->launches += 6;
    return KSSD_OK;
}

}  // namespace kssd
