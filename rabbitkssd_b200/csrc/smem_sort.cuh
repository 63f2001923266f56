// Shared-memory sorting primitives used by the per-genome sort (segsort.cu), the index build (index.cu) and the
// top-N selection of the distance epilogue (dist.cu).
#pragma once
#include <stdint.h>

namespace kssd {

// Ascending-only bitonic network over s[0, n): every compare-exchange puts the smaller key at the lower index, so
// the (virtual) +inf keys at indices >= n never move and need no storage.  All threads of the CTA must call it
// with the same n; on return the result is visible to every thread.
template <typename K>
__device__ __forceinline__ void bitonic_sort_smem(K *s, uint32_t n) {
    if (n < 2) return;
    const uint32_t np = 1u << (32 - __clz(n - 1));  // next power of two >= n
    const uint32_t half = np >> 1;
    for (uint32_t k = 2; k <= np; k <<= 1) {
        const uint32_t hk = k >> 1;
        const int lg = 31 - __clz(hk);
        for (uint32_t i = threadIdx.x; i < half; i += blockDim.x) {  // flip: i-th pair of its k-block is (off, k-1-off)
            const uint32_t blk = i >> lg, off = i & (hk - 1);
            const uint32_t a = blk * k + off, b = blk * k + (k - 1 - off);
            if (b < n) {
                const K x = s[a], y = s[b];
                if (x > y) { s[a] = y; s[b] = x; }
            }
        }
        __syncthreads();
        for (uint32_t j = hk >> 1; j >= 1; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < half; i += blockDim.x) {
                const uint32_t a = ((i & ~(j - 1)) << 1) | (i & (j - 1)), b = a + j;
                if (b < n) {
                    const K x = s[a], y = s[b];
                    if (x > y) { s[a] = y; s[b] = x; }
                }
            }
            __syncthreads();
        }
    }
}

}  // namespace kssd
