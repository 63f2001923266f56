// RawSlicer: the host half of the raw-FASTA path (kssd_batch.raw_fasta, include/kssd_b200.h; device half: csrc/fasta.cu).
// Kept free of reference headers so that tests/test_raw_slicer.py can compile it alone.
#pragma once
#ifndef _GNU_SOURCE
#define _GNU_SOURCE  // memrchr
#endif
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstring>

#include "kssd_b200.h"

namespace kssd_host {

inline bool is_hdr_char(unsigned char c) { return c == '>' || c == '@'; }

struct RawSlicer {
    static constexpr int kCtx = 31;  // >= 2 * half_k - 1 for half_k <= 16
    bool hdr = false;     // the line that holds the next byte is a header line
    bool at_ls = true;    // the next byte is the first of a line
    unsigned char ctx[kCtx];
    int n_ctx = 0;

    // prefix + flags of the piece that starts at the current position (file offset > 0)
    int next_prefix(unsigned char *out, size_t &len) const {
        len = 0;
        if (hdr && !at_ls) return KSSD_SEG_IN_HEADER;
        if (n_ctx == 0) return at_ls ? 0 : KSSD_SEG_MID_LINE;
        memcpy(out, ctx, n_ctx);
        len = n_ctx;
        if (at_ls) out[len++] = '\n';
        return KSSD_SEG_MID_LINE;
    }

    void push(unsigned char c) {
        if (n_ctx == kCtx) memmove(ctx, ctx + 1, --n_ctx);
        ctx[n_ctx++] = c;
    }

    // advance over the piece p[0, n) that was just read; `next` = the byte after it (or '\n' at the end of the file).
    // Only the tail matters: find, walking back line by line with memrchr (long lines cost nothing), a position from which
    // a forward replay gives the state at the end -- a header line (nothing before it matters), a line start with > kCtx
    // bytes after it, or the piece start with the state this object already holds.
    void advance(const unsigned char *p, size_t n, unsigned char next) {
        size_t from = 0;
        bool fresh = false;        // replay from `from` with an empty context instead of the current state
        bool fresh_at_ls = false;  // ... `from` is a line start (else: inside a sequence line, or at the '\n' that ends a header line)
        bool fresh_hdr = false;
        size_t i = n, kept = 0;
        while (true) {
            const unsigned char *nl = i ? (const unsigned char *)memrchr(p, '\n', i) : nullptr;
            if (!nl) {  // p[0, i) belongs to the line the piece started in (or, at_ls, to a line that starts at p[0])
                const bool first_is_hdr = at_ls ? (n > 0 && is_hdr_char(p[0])) : hdr;
                if (first_is_hdr && i < n) { from = i; fresh = true; fresh_hdr = true; }                     // p[i] is the '\n' that ends it
                else if (!first_is_hdr && kept + i > (size_t)kCtx + 1 && i > (size_t)kCtx + 2) { from = i - (kCtx + 2); fresh = true; }
                break;  // else: replay the whole piece from the current state
            }
            const size_t ls = (size_t)(nl - p) + 1;
            if (ls < i && is_hdr_char(p[ls])) { from = ls; fresh = true; fresh_at_ls = true; break; }
            kept += i - ls;
            if (kept > (size_t)kCtx + 1) { from = ls; fresh = true; fresh_at_ls = true; break; }
            i = ls - 1;
        }
        if (fresh) {
            n_ctx = 0;
            at_ls = fresh_at_ls;
            hdr = fresh_hdr;
        }
        for (size_t j = from; j < n; j++) {
            const unsigned char c = p[j];
            if (at_ls) {
                hdr = is_hdr_char(c);
                at_ls = false;
                if (hdr) n_ctx = 0;
            }
            if (c == '\n') at_ls = true;
            else if (!hdr && !(c == '\r' && (j + 1 < n ? p[j + 1] : next) == '\n')) push(c);
        }
    }
};

// Raw four-line FASTQ text (fastq.cu takes pieces of WHOLE records): how many of the len bytes at p are whole records, i.e. the
// offset just behind the last '\n' whose number is a multiple of four -- one counting pass and at most four memrchr calls.
// 0: not even one whole record.
inline size_t fastq_whole_records(const unsigned char *p, size_t len) {
    size_t lines = 0, i = 0;
    for (; i + 8 <= len; i += 8) {  // eight bytes at a time: 0x80 in every byte that equals '\n', exactly
        uint64_t x;
        memcpy(&x, p + i, 8);
        x ^= 0x0A0A0A0A0A0A0A0Aull;
        const uint64_t z = ~(((x & 0x7F7F7F7F7F7F7F7Full) + 0x7F7F7F7F7F7F7F7Full) | x | 0x7F7F7F7F7F7F7F7Full);
        lines += (size_t)__builtin_popcountll(z);
    }
    for (; i < len; i++) lines += p[i] == '\n';
    const unsigned char *e = len ? static_cast<const unsigned char *>(memrchr(p, '\n', len)) : nullptr;
    for (size_t skip = lines % 4; e && skip; skip--) e = e > p ? static_cast<const unsigned char *>(memrchr(p, '\n', (size_t)(e - p))) : nullptr;
    return e ? (size_t)(e - p) + 1 : 0;
}

}  // namespace kssd_host
