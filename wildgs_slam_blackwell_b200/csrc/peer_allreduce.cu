// Gradient-bucket allreduce over NVLink peer memory (the one exchange step of the
// mapping window).  Every rank's bucket lives in a symmetric allocation that all peers have mapped
// (and, where the fabric offers it, in an NVSwitch multicast mapping).  One kernel per rank:
//   block-level start barrier  (flags in peer memory; proves every peer's backward has finished)
//   two-shot reduction of this rank's 1/world slice:
//     NVLS path : multimem.ld_reduce (the switch adds the `world` copies) + multimem.st (the switch
//                 broadcasts the sum into every copy)            -> ~1x bucket bytes per NVLink direction
//     P2P path  : vector loads from every peer, fixed-order fp32 sum, vector stores to every peer
//   block-level end barrier    (all slices of this rank's copy have been written when the kernel ends)
// The sum of a slice is formed once and broadcast, so all ranks hold bit-identical results.
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

constexpr int PEER_MAX_WORLD = 16;
constexpr int PEER_MAX_BLOCKS = 592;  // flag area: one barrier slot set per block
constexpr int PEER_THREADS = 512;
// launch shape (wgs_peer_configure): blocks x 512 threads x U float4 per thread and trip = bytes in flight on NVLink.
// 0 = the measured defaults: the NVLS path is bound by the switch and runs fastest with little in flight (8 x B200,
// profiles/r02o_tune_peer_n8*.json: 48 x 2 -> 59 / 146 / 290 us for 16.8 / 56 / 128 MB, 144 x 4 -> 68 / 170 / 318,
// 592 x 8 -> 102 / 203 / 411; NCCL 100 / 244 / 390); the P2P path keeps every peer's loads in flight with 144 x 4.
static int g_peer_blocks = 0, g_peer_unroll = 0;

constexpr int PEER_MAX_SEGMENTS = 12;

struct PeerParams {
    float* buf[PEER_MAX_WORLD];        // this rank's mapping of every rank's buffer
    uint32_t* flags[PEER_MAX_WORLD];   // ... and of every rank's flag array [PEER_MAX_BLOCKS][PEER_MAX_WORLD]
    float* mc;                         // multicast mapping of the buffer (null: P2P path)
    int rank, world;
    int mode;                          // 0 allreduce, 1 reduce-scatter (sum of my slice -> my copy), 2 allgather (my slice -> all)
    int nseg;
    size_t seg_off4[PEER_MAX_SEGMENTS];   // the exchanged data: `nseg` pieces of the buffer (offset / length in float4),
    size_t seg_end4[PEER_MAX_SEGMENTS];   // laid end to end as one virtual array; seg_end4 = running end in that array
    size_t n4;                         // float4 elements of the virtual array
    uint32_t* epochs;                  // [PEER_MAX_BLOCKS] in THIS rank's flag area: barrier generation of each block
};

// virtual float4 index -> float4 offset inside the buffer
__device__ __forceinline__ size_t seg_addr(const PeerParams& p, size_t i) {
    size_t begin = 0;
#pragma unroll
    for (int s = 0; s < PEER_MAX_SEGMENTS - 1; s++) {
        if (s + 1 < p.nseg && i >= p.seg_end4[s]) begin = p.seg_end4[s];
    }
    int k = 0;
#pragma unroll
    for (int s = 0; s < PEER_MAX_SEGMENTS - 1; s++) k += (s + 1 < p.nseg && i >= p.seg_end4[s]) ? 1 : 0;
    return p.seg_off4[k] + (i - begin);
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Block b of every rank meets block b of every other rank.  The flag values only grow, so no reset is needed.
__device__ __forceinline__ void block_barrier(const PeerParams& p, uint32_t value) {
    __syncthreads();
    if ((int)threadIdx.x < p.world) {
        const int t = threadIdx.x;
        st_release_sys(p.flags[t] + blockIdx.x * PEER_MAX_WORLD + p.rank, value);
        const uint32_t* mine = p.flags[p.rank] + blockIdx.x * PEER_MAX_WORLD + t;
        while ((int32_t)(ld_acquire_sys(mine) - value) < 0) {}
    }
    __syncthreads();
}

__device__ __forceinline__ float4 multimem_ld_reduce(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w) : "memory");
}

template <bool NVLS, int U>   // U independent float4 per thread and trip: keeps U (x world) 16-B requests in flight
__global__ void __launch_bounds__(PEER_THREADS) peer_allreduce_kernel(const __grid_constant__ PeerParams p) {
    // The barrier generation lives on the device (one counter per block, advanced by 2 per launch), so the
    // launch has no per-call host argument and can be replayed from a CUDA graph; all ranks launch equally often.
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = p.epochs[blockIdx.x] + 1u;
    __syncthreads();
    const uint32_t epoch = s_epoch;
    block_barrier(p, epoch);
    const size_t per = (p.n4 + p.world - 1) / p.world;
    const size_t begin = min(p.n4, per * p.rank), end = min(p.n4, begin + per);
    const size_t stride = (size_t)gridDim.x * PEER_THREADS;
    const bool reduce = p.mode != 2, spread = p.mode != 1;
    float4* const mine = reinterpret_cast<float4*>(p.buf[p.rank]);
    for (size_t i0 = begin + (size_t)blockIdx.x * PEER_THREADS + threadIdx.x; i0 < end; i0 += U * stride) {
        float4 s[U];
        size_t a[U];
#pragma unroll
        for (int u = 0; u < U; u++) a[u] = i0 + u * stride < end ? seg_addr(p, i0 + u * stride) : 0;
        if (NVLS) {
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < end) s[u] = reduce ? multimem_ld_reduce(p.mc + 4 * a[u]) : __ldcg(mine + a[u]);
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < end) {
                    if (spread) multimem_st(p.mc + 4 * a[u], s[u]);
                    else __stcg(mine + a[u], s[u]);
                }
        } else {
#pragma unroll
            for (int u = 0; u < U; u++) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (reduce) {
                for (int r = 0; r < p.world; r++) {   // fixed order: identical bits on every run
                    float4 v[U];
#pragma unroll
                    for (int u = 0; u < U; u++)
                        v[u] = i0 + u * stride < end ? __ldcg(reinterpret_cast<const float4*>(p.buf[r]) + a[u])
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int u = 0; u < U; u++) { s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w; }
                }
            } else {
#pragma unroll
                for (int u = 0; u < U; u++)
                    if (i0 + u * stride < end) s[u] = __ldcg(mine + a[u]);
            }
            if (spread) {
                for (int r = 0; r < p.world; r++)
#pragma unroll
                    for (int u = 0; u < U; u++)
                        if (i0 + u * stride < end) __stcg(reinterpret_cast<float4*>(p.buf[r]) + a[u], s[u]);
            } else {
#pragma unroll
                for (int u = 0; u < U; u++)
                    if (i0 + u * stride < end) __stcg(mine + a[u], s[u]);
            }
        }
    }
    block_barrier(p, epoch + 1u);
    if (threadIdx.x == 0) p.epochs[blockIdx.x] = epoch + 1u;
}

}  // namespace wgs

extern "C" {
void wgs_set_error_message(const char* msg);

size_t wgs_peer_flags_bytes(void) {
    return (size_t)wgs::PEER_MAX_BLOCKS * (wgs::PEER_MAX_WORLD + 1) * sizeof(uint32_t);  // flags + per-block generation
}

int wgs_peer_exchange(const uint64_t* buffer_ptrs, const uint64_t* flag_ptrs, uint64_t multicast_ptr, int32_t rank,
                      int32_t world, const uint64_t* seg_offsets, const uint64_t* seg_lengths, int32_t n_segments,
                      int32_t mode, void* cuda_stream) {
    using namespace wgs;
    if (!buffer_ptrs || !flag_ptrs || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world ||
        !seg_offsets || !seg_lengths || n_segments < 1 || n_segments > PEER_MAX_SEGMENTS || mode < 0 || mode > 3) {
        wgs_set_error_message("wgs_peer_exchange: bad world/rank/segment table/mode");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    PeerParams p;
    for (int r = 0; r < PEER_MAX_WORLD; r++) {
        p.buf[r] = r < world ? (float*)buffer_ptrs[r] : nullptr;
        p.flags[r] = r < world ? (uint32_t*)flag_ptrs[r] : nullptr;
        if (r < world && (!p.buf[r] || !p.flags[r] || (buffer_ptrs[r] & 15u))) {
            wgs_set_error_message("wgs_peer_exchange: null or misaligned peer pointer");
            return WGS_ERR_INVALID_ARGUMENT;
        }
    }
    size_t run = 0;
    for (int k = 0; k < PEER_MAX_SEGMENTS; k++) {
        if (k < n_segments) {
            if ((seg_offsets[k] & 3u) || (seg_lengths[k] & 3u)) {
                wgs_set_error_message("wgs_peer_exchange: segment offsets and lengths must be multiples of 4 floats");
                return WGS_ERR_INVALID_ARGUMENT;
            }
            p.seg_off4[k] = seg_offsets[k] / 4;
            run += seg_lengths[k] / 4;
        } else {
            p.seg_off4[k] = 0;
        }
        p.seg_end4[k] = run;
    }
    p.mc = (float*)multicast_ptr; p.rank = rank; p.world = world; p.n4 = run; p.mode = mode; p.nseg = n_segments;
    p.epochs = p.flags[rank] + PEER_MAX_BLOCKS * PEER_MAX_WORLD;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    // mode 3, rendezvous: ONE block runs the two flag barriers and moves nothing.  Put in front of an exchange whose
    // peers may arrive late (e.g. behind their own H2D copies), it does the waiting with one block instead of with the
    // whole exchange grid spinning on SM slots that a concurrent stream's kernels could use.
    const int blocks = g_peer_blocks ? g_peer_blocks : (p.mc ? 48 : 144);
    const int unroll = g_peer_unroll ? g_peer_unroll : (p.mc ? 2 : 4);
    const int grid = mode == 3 ? 1 : blocks;
    if (mode == 3) { p.n4 = 0; p.mode = 0; }
#define WGS_PEER_LAUNCH(U_)                                                                      \
    do {                                                                                         \
        if (p.mc) peer_allreduce_kernel<true, U_><<<grid, PEER_THREADS, 0, s>>>(p);              \
        else peer_allreduce_kernel<false, U_><<<grid, PEER_THREADS, 0, s>>>(p);                  \
    } while (0)
    switch (unroll) {
        case 1: WGS_PEER_LAUNCH(1); break;
        case 2: WGS_PEER_LAUNCH(2); break;
        case 8: WGS_PEER_LAUNCH(8); break;
        default: WGS_PEER_LAUNCH(4); break;
    }
#undef WGS_PEER_LAUNCH
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}

int wgs_peer_configure(int32_t blocks, int32_t unroll) {
    if (blocks < 0 || blocks > wgs::PEER_MAX_BLOCKS || (unroll != 0 && unroll != 1 && unroll != 2 && unroll != 4 && unroll != 8)) {
        wgs_set_error_message("wgs_peer_configure: blocks in [0, 592], unroll in {0, 1, 2, 4, 8} (0 = default)");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    wgs::g_peer_blocks = blocks;
    wgs::g_peer_unroll = unroll;
    return WGS_OK;
}

int wgs_peer_allreduce(const uint64_t* buffer_ptrs, const uint64_t* flag_ptrs, uint64_t multicast_ptr, int32_t rank,
                       int32_t world, uint64_t n_floats, void* cuda_stream) {
    const uint64_t off = 0;
    return wgs_peer_exchange(buffer_ptrs, flag_ptrs, multicast_ptr, rank, world, &off, &n_floats, 1, 0, cuda_stream);
}
}
