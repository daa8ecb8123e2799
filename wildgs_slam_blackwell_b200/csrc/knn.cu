// distCUDA2: mean squared distance to the 3 nearest other points (include/wgs_knn.h).
// Replaces the reference's Morton sort (cub radix sort) + 1024-point boxes + per-point scan over ALL
// boxes (simple_knn.cu:64-222; O(P * P/1024) box tests and whole boxes of candidates) with a uniform
// grid: bounding box (one reduction, no host readback), counting sort of the points by cell, and a
// ring search around each point's cell that stops as soon as the 3rd-best distance is provably final.
// The result is the exact 3-NN, as in the reference (its boxes only prune).
#include <cfloat>

#include "../../include/wgs_knn.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

constexpr int KNN_MAX_G = 128;

struct KnnParams {
    int P, G;
    const float* pts;
    float* out;
    uint32_t* bounds;   // [6] ordered-uint min xyz, max xyz
    uint32_t* cell_of;  // [P]
    uint32_t* start;    // [cells + 1]
    uint32_t* cursor;   // [cells]
    float4* sorted;     // [P] x, y, z, original index (bits)
};

__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void knn_init_kernel(KnnParams p, int cells) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 3) p.bounds[i] = 0xffffffffu;
    else if (i < 6) p.bounds[i] = 0u;
    for (int c = i; c <= cells; c += gridDim.x * blockDim.x) p.start[c] = 0u;
}

__global__ void __launch_bounds__(256) knn_bounds_kernel(KnnParams p) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * 256 + threadIdx.x; i < p.P; i += gridDim.x * 256) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float v = __ldg(p.pts + 3 * (size_t)i + k);
            lo[k] = fminf(lo[k], v);
            hi[k] = fmaxf(hi[k], v);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            atomicMin(p.bounds + k, f2ord(lo[k]));
            atomicMax(p.bounds + 3 + k, f2ord(hi[k]));
        }
    }
}

struct Grid {
    float mn[3], cw[3], inv[3];
};
__device__ __forceinline__ Grid load_grid(const KnnParams& p) {
    Grid g;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        g.mn[k] = ord2f(p.bounds[k]);
        const float ext = ord2f(p.bounds[3 + k]) - g.mn[k];
        g.cw[k] = ext / (float)p.G;
        g.inv[k] = ext > 0.f ? (float)p.G / ext : 0.f;
    }
    return g;
}
__device__ __forceinline__ int cell_coord(float v, float mn, float inv, int G) {
    return min(G - 1, max(0, (int)((v - mn) * inv)));
}

__global__ void __launch_bounds__(256) knn_count_kernel(KnnParams p) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= p.P) return;
    const Grid g = load_grid(p);
    const int cx = cell_coord(__ldg(p.pts + 3 * (size_t)i), g.mn[0], g.inv[0], p.G);
    const int cy = cell_coord(__ldg(p.pts + 3 * (size_t)i + 1), g.mn[1], g.inv[1], p.G);
    const int cz = cell_coord(__ldg(p.pts + 3 * (size_t)i + 2), g.mn[2], g.inv[2], p.G);
    const uint32_t c = (uint32_t)((cz * p.G + cy) * p.G + cx);
    p.cell_of[i] = c;
    atomicAdd(p.start + c + 1, 1u);   // counts shifted by one: the inclusive scan below yields the starts
}

// single-block inclusive scan of start[1..cells] in place (start[0] = 0); copies the starts to cursor
__global__ void __launch_bounds__(1024) knn_scan_kernel(KnnParams p, int cells) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < cells; base += 1024) {
        const int c = base + tid;
        const uint32_t v0 = c < cells ? p.start[c + 1] : 0u;
        uint32_t v = v0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) wsum[warp] = v;
        __syncthreads();
        if (warp == 0) {
            uint32_t s = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += u;
            }
            wsum[lane] = s;
        }
        __syncthreads();
        const uint32_t incl = carry_s + (warp ? wsum[warp - 1] : 0u) + v;
        if (c < cells) {
            p.start[c + 1] = incl;
            p.cursor[c] = incl - v0;
        }
        __syncthreads();
        if (tid == 1023) carry_s = incl;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) knn_scatter_kernel(KnnParams p) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= p.P) return;
    const uint32_t slot = atomicAdd(p.cursor + p.cell_of[i], 1u);
    p.sorted[slot] = make_float4(__ldg(p.pts + 3 * (size_t)i), __ldg(p.pts + 3 * (size_t)i + 1),
                                 __ldg(p.pts + 3 * (size_t)i + 2), __uint_as_float((uint32_t)i));
}

// reference updateKBest<3> (simple_knn.cu:123-137)
__device__ __forceinline__ void update3(float dist, float (&b)[3]) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (b[j] > dist) {
            const float t = b[j];
            b[j] = dist;
            dist = t;
        }
    }
}

__global__ void __launch_bounds__(128) knn_query_kernel(KnnParams p) {
    const int i = blockIdx.x * 128 + threadIdx.x;   // slot in cell order: neighbouring threads search the same cells
    if (i >= p.P) return;
    const Grid g = load_grid(p);
    const int G = p.G;
    const float4 me = p.sorted[i];
    const int c[3] = {cell_coord(me.x, g.mn[0], g.inv[0], G), cell_coord(me.y, g.mn[1], g.inv[1], G),
                      cell_coord(me.z, g.mn[2], g.inv[2], G)};
    const float pos[3] = {me.x, me.y, me.z};
    // slack for the rounding of cell boundaries: only ever delays the stop, never changes the result
    const float slack = 1e-5f * fmaxf(fmaxf(g.cw[0], g.cw[1]), g.cw[2]) * (float)G;
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    for (int r = 0; r < G; r++) {
        const int z0 = max(c[2] - r, 0), z1 = min(c[2] + r, G - 1);
        const int y0 = max(c[1] - r, 0), y1 = min(c[1] + r, G - 1);
        const int x0 = max(c[0] - r, 0), x1 = min(c[0] + r, G - 1);
        for (int z = z0; z <= z1; z++) {
            const bool zf = abs(z - c[2]) == r;
            for (int y = y0; y <= y1; y++) {
                const bool yf = zf || abs(y - c[1]) == r;
                // on a face in z or y: the whole x row belongs to ring r; otherwise only its two ends
                const int step = yf ? 1 : max(x1 - x0, 1);
                for (int x = x0; x <= x1; x += step) {
                    if (!yf && abs(x - c[0]) != r) continue;
                    const uint32_t cell = (uint32_t)((z * G + y) * G + x);
                    const uint32_t s = p.start[cell], e = p.start[cell + 1];
                    for (uint32_t j = s; j < e; j++) {
                        if ((int)j == i) continue;
                        const float4 q = p.sorted[j];
                        const float dx = q.x - me.x, dy = q.y - me.y, dz = q.z - me.z;
                        update3(dx * dx + dy * dy + dz * dz, best);
                    }
                }
            }
        }
        // anything not visited yet lies outside the block of cells [c - r, c + r] (clipped to the grid)
        float dmin = FLT_MAX;
        bool open = false;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (c[k] - r > 0) { dmin = fminf(dmin, pos[k] - (g.mn[k] + (float)(c[k] - r) * g.cw[k])); open = true; }
            if (c[k] + r < G - 1) { dmin = fminf(dmin, (g.mn[k] + (float)(c[k] + r + 1) * g.cw[k]) - pos[k]); open = true; }
        }
        if (!open) break;
        dmin = fmaxf(dmin - slack, 0.f);
        if (best[2] <= dmin * dmin) break;
    }
    p.out[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;   // simple_knn.cu:180
}

static int knn_grid(int P) {
    int G = (int)ceil(cbrt((double)(P > 0 ? P : 1) / 3.0));
    return G < 1 ? 1 : (G > KNN_MAX_G ? KNN_MAX_G : G);
}

}  // namespace wgs

using namespace wgs;
extern "C" {
void wgs_set_error_message(const char* msg);

size_t wgs_knn_workspace_bytes(int32_t P) {
    if (P < 0) return 0;
    const size_t G = (size_t)knn_grid(P), cells = G * G * G, n = (size_t)(P > 0 ? P : 1);
    return 256 + align_up(n * 4) + 2 * align_up((cells + 1) * 4) + align_up(n * 16);
}

int wgs_dist_cuda2(int32_t P, const float* points, float* mean_dists, void* workspace, void* cuda_stream) {
    if (P < 0 || (P > 0 && (!points || !mean_dists || !workspace || ((uintptr_t)workspace & 255u)))) {
        wgs_set_error_message("wgs_dist_cuda2: null/misaligned pointer or negative P");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (P == 0) return WGS_OK;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    KnnParams p;
    p.P = P; p.G = knn_grid(P); p.pts = points; p.out = mean_dists;
    const size_t cells = (size_t)p.G * p.G * p.G;
    char* w = (char*)workspace;
    p.bounds = (uint32_t*)w; w += 256;
    p.cell_of = (uint32_t*)w; w += align_up((size_t)P * 4);
    p.start = (uint32_t*)w; w += align_up((cells + 1) * 4);
    p.cursor = (uint32_t*)w; w += align_up((cells + 1) * 4);
    p.sorted = (float4*)w;
    const int pb = (P + 255) / 256;
    knn_init_kernel<<<(int)min((size_t)1024, (cells + 256) / 256), 256, 0, s>>>(p, (int)cells);
    knn_bounds_kernel<<<min(pb, 1184), 256, 0, s>>>(p);
    knn_count_kernel<<<pb, 256, 0, s>>>(p);
    knn_scan_kernel<<<1, 1024, 0, s>>>(p, (int)cells);
    knn_scatter_kernel<<<pb, 256, 0, s>>>(p);
    knn_query_kernel<<<(P + 127) / 128, 128, 0, s>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}
}
