// Staging of SH rows through shared memory (preprocess_fwd / preprocess_bwd, 1 < M <= 16), warp-local.
// The rows of the 32 Gaussians of a warp form one contiguous span of global memory.  The warp brings in the rows of
// its visible Gaussians one row per step -- lanes on consecutive words, so every request is a contiguous 128-B (+
// remainder) piece whatever the row size -- with cp.async (no register staging: all rows are in flight at once),
// each lane then reads ITS row from shared memory (odd row stride: conflict-free) and, in the backward pass,
// leaves its dL_dsh there; the warp writes the rows back the same way.  Only the warp touches its rows, so the
// steps are ordered by __syncwarp alone and the warps of a block run their phases independently.
#pragma once
#include <cstdint>

namespace wgs {

constexpr int SH_STAGE_MAX_M = 16;        // rows of at most 48 words: two requests per row
__host__ __device__ inline int sh_row_stride(int row_words) { return row_words | 1; }
inline size_t sh_stage_bytes(int M, int threads) { return (size_t)threads * sh_row_stride(3 * M) * sizeof(float); }

__device__ __forceinline__ void sh_cp_async4(float* smem_dst, const float* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gsrc) : "memory");
}

// wstage / wsrc: the warp's first row in shared / global memory; vmask: ballot of the lanes' visibility
__device__ __forceinline__ void sh_stage_load(float* wstage, const float* __restrict__ wsrc, unsigned vmask,
                                              uint32_t rw, uint32_t rs) {
    const uint32_t lane = threadIdx.x & 31;
    const bool second = lane + 32 < rw, first = lane < rw;
    while (vmask) {
        const uint32_t i = __ffs(vmask) - 1;
        vmask &= vmask - 1;
        float* d = wstage + i * rs + lane;
        const float* s = wsrc + i * rw + lane;
        if (first) sh_cp_async4(d, s);
        if (second) sh_cp_async4(d + 32, s + 32);
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
}

// rows [0, nrows) of the warp: wdst = staged row (0 for culled rows), or += for the visible rows when accumulating
__device__ __forceinline__ void sh_stage_store(const float* wstage, float* __restrict__ wdst, unsigned vmask,
                                               uint32_t nrows, uint32_t rw, uint32_t rs, int accumulate) {
    const uint32_t lane = threadIdx.x & 31;
    const bool second = lane + 32 < rw, first = lane < rw;
    __syncwarp();
    if (accumulate) {
        while (vmask) {
            const uint32_t i = __ffs(vmask) - 1;
            vmask &= vmask - 1;
            const float* s = wstage + i * rs + lane;
            float* d = wdst + i * rw + lane;
            if (accumulate == 2) {   // several backwards add to this buffer concurrently (WgsBackwardArgs::accumulate)
                if (first) atomicAdd(d, s[0]);
                if (second) atomicAdd(d + 32, s[32]);
            } else {
                if (first) d[0] += s[0];
                if (second) d[32] += s[32];
            }
        }
    } else {
#pragma unroll 4
        for (uint32_t i = 0; i < nrows; i++) {
            const bool v = (vmask >> i) & 1u;
            const float* s = wstage + i * rs + lane;
            float* d = wdst + i * rw + lane;
            if (first) d[0] = v ? s[0] : 0.f;
            if (second) d[32] = v ? s[32] : 0.f;
        }
    }
}

}  // namespace wgs
