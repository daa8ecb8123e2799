// Staging of a block's SH rows through shared memory (preprocess_fwd / preprocess_bwd, 1 < M <= 16).
// The rows of the 256 Gaussians of a block form one contiguous span of global memory.  The span is brought in with
// cp.async (no register staging, so every word of the span is in flight at once): 16-byte copies of the whole span
// when the shared-memory rows are unpadded and the span is 16-byte aligned, else 4-byte copies with consecutive
// lanes on consecutive words (coalesced for any row size), rows of culled Gaussians skipped.  Rows in shared memory
// have an odd stride, so per-thread row access is conflict-free.  row = word / row_words is one mul.hi.
#pragma once
#include <cstdint>

namespace wgs {

constexpr int SH_STAGE_MAX_M = 16;
constexpr int SH_STAGE_THREADS = 256;
__host__ __device__ inline int sh_row_stride(int row_words) { return row_words | 1; }
inline size_t sh_stage_bytes(int M) { return (size_t)SH_STAGE_THREADS * sh_row_stride(3 * M) * sizeof(float); }

// exact floor(w / rw) for w < 2^16, rw <= 48:  umulhi(w, ceil(2^32 / rw))
__device__ __forceinline__ uint32_t sh_div_magic(uint32_t rw) { return 0xFFFFFFFFu / rw + 1u; }

__device__ __forceinline__ void sh_stage_load(float* stage, const uint8_t* vis, const float* __restrict__ src,
                                              uint32_t nw, uint32_t rw, uint32_t rs) {
    const uint32_t pad = rs - rw;
    if (pad == 0 && ((uintptr_t)src & 15u) == 0 && (nw & 3u) == 0) {
        for (uint32_t w = 4 * threadIdx.x; w < nw; w += 4 * SH_STAGE_THREADS) {
            const unsigned s = (unsigned)__cvta_generic_to_shared(stage + w);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(src + w) : "memory");
        }
    } else {
        const uint32_t magic = sh_div_magic(rw);
#pragma unroll 4
        for (uint32_t w = threadIdx.x; w < nw; w += SH_STAGE_THREADS) {
            const uint32_t row = __umulhi(w, magic);
            if (vis[row]) {
                const unsigned s = (unsigned)__cvta_generic_to_shared(stage + w + row * pad);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(src + w) : "memory");
            }
        }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// dst[w] = stage (0 for culled rows), or dst[w] += stage for the visible rows when accumulating
template <int U>
__device__ __forceinline__ void sh_stage_store(const float* stage, const uint8_t* vis, float* __restrict__ dst,
                                               uint32_t nw, uint32_t rw, uint32_t rs, bool accumulate) {
    const uint32_t pad = rs - rw, magic = sh_div_magic(rw);
    for (uint32_t w0 = threadIdx.x; w0 < nw; w0 += SH_STAGE_THREADS * U) {
        float v[U];
        bool pin[U], pv[U];
#pragma unroll
        for (int j = 0; j < U; j++) {
            const uint32_t w = w0 + SH_STAGE_THREADS * j;
            const uint32_t row = __umulhi(w, magic);
            pin[j] = w < nw;
            pv[j] = pin[j] && vis[row];
            v[j] = pv[j] ? stage[w + row * pad] : 0.f;
        }
        if (accumulate) {
#pragma unroll
            for (int j = 0; j < U; j++)
                if (pv[j]) v[j] += dst[w0 + SH_STAGE_THREADS * j];
#pragma unroll
            for (int j = 0; j < U; j++)
                if (pv[j]) dst[w0 + SH_STAGE_THREADS * j] = v[j];
        } else {
#pragma unroll
            for (int j = 0; j < U; j++)
                if (pin[j]) dst[w0 + SH_STAGE_THREADS * j] = v[j];
        }
    }
}

}  // namespace wgs
