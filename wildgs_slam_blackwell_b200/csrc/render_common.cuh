// Shared pieces of the forward and backward compositing kernels: cp.async helpers and the
// shared-memory stage one batch of gathered Gaussian records lives in.
#pragma once
#include "common.cuh"

namespace wgs {

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

constexpr int BATCH = 256;

struct __align__(16) FwdStage {
    float4 rec0[BATCH];  // x, y, conic.x, conic.y
    float4 rec1[BATCH];  // conic.z, opacity, depth, ln(255*opacity)
    float4 col[BATCH];   // r, g, b, -
    uint32_t id[BATCH];
};


// Gather batch entry `slot` (Gaussian `id`) into a stage: 3 x 16-byte cp.async (LDGSTS).
__device__ __forceinline__ void stage_gather(FwdStage& st, int slot, uint32_t id, const GeomView& g) {
    st.id[slot] = id;
    cp_async16(&st.rec0[slot], g.rec + 2 * (size_t)id);
    cp_async16(&st.rec1[slot], g.rec + 2 * (size_t)id + 1);
    cp_async16(&st.col[slot], g.col + id);
}

}  // namespace wgs
