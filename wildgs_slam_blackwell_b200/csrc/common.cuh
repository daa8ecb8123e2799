// Shared device/host definitions of the B200-native Gaussian rasterizer (sm_100a only).
//
// Data layout in HBM (all workspaces are caller-owned, see include/wgs_rasterizer.h):
//   geom   : per Gaussian  rec[2i], rec[2i+1] (float4 x2 = one 32-B sector):
//              {mean2D.x, mean2D.y, conic.x, conic.y} {conic.z, opacity, depth, ln(255*opacity)}
//            col[i] (float4) {r, g, b, 0};  rect[i] (uint2, packed u16 tile rect);
//            clamped[i] (u8, bit c = SH colour channel c was clamped at 0)
//   tile   : count[tiles] u32, cursor[tiles] u32, ranges[tiles] uint2 [start,end), misc[32] u32
//   binning: bucket[capacity] u64 = depth_bits<<32 | gaussian_id, grouped by tile, sorted
//            ascending inside each tile  (== the reference's stable (tile|depth) radix order)
//   image  : final_T[N] f32, n_contrib[N] u32
//   grad   : acc[3i..3i+2] float4 x3 per Gaussian (dmean2D.xy, dconic.xyw -- both without their
//            constant factors, see preprocess_bwd --, dopacity, dcolor.rgb, ddepth, 2 pad)
//            -- zero between backward passes; pose partials; ticket
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define WGS_TILE 16            // BLOCK_X == BLOCK_Y == 16 (reference config.h:14-16)
#define WGS_TILE_PIX 256
// The per-tile instance histogram / scatter cursors are privatised into up to this many copies
// (copy = blockIdx.x % copies in preprocess_fwd and scatter, which share their block mapping):
// same-line L2 atomics serialise, and 1200 counters are only 38 cache lines.  The copy count
// actually used is hist_copies(tiles): fewer copies when there are many tiles already.
#define WGS_HIST_COPIES 32

// Round-to-nearest primitives that nvcc never re-associates or contracts.  The forward path
// is written in these so that it reproduces, bit for bit, the instruction sequence nvcc 12.9
// emits for the reference sources (default -fmad=true); see DESIGN.md "bit-exactness".
#define FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))
#define DIV(a, b) __fdiv_rn((a), (b))
#define RCP(a) __frcp_rn((a))
#define SQRT(a) __fsqrt_rn((a))

namespace wgs {

// ---- programmatic dependent launch (PDL) -------------------------------------------------------------------------
// The kernels of one forward / backward follow each other on one stream and each needs the previous one's output, so
// between two of them the GPU drains, the next grid is launched and its CTAs ramp up: 2-4 us per boundary, a tenth of
// a 0.43 ms frame.  With PDL the next grid is launched while the previous one is still running: a kernel of the chain
//   * starts with pdl_wait() -- griddepcontrol.wait: returns once the preceding kernel has completed and its writes
//     are visible -- as its FIRST statement, before any early return (a grid whose CTAs all left without waiting would
//     "complete" before its predecessor, and the kernel after it would no longer be ordered behind that one),
//   * then calls pdl_trigger() -- griddepcontrol.launch_dependents: once every CTA of this grid has executed it (i.e.
//     the last wave is resident), the next kernel's CTAs may take the SM slots that free up and sit in their wait.
// Nothing is read or written before the wait, so the ordering of the data is that of a plain stream.  Launched without
// the attribute (WGS_PDL=0, or a stream of a library that does not know PDL) both instructions are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // api.cu: environment WGS_PDL (default on)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                  Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

enum MiscSlot : int {
    MISC_NUM_RENDERED = 0,   // R: total (Gaussian,tile) instances of this forward
    MISC_OVERFLOW = 1,       // set if R > capacity
    MISC_N_LARGE = 2,        // number of tiles handed to the large sort kernel
    MISC_PREFILTER_HIT = 4,  // a point was culled although prefiltered was set
    MISC_TICKET = 5,
    MISC_WORDS = 32
};

struct GeomView {
    float4* rec;       // 2 per Gaussian
    float4* col;       // 1 per Gaussian
    uint2* rect;       // 1 per Gaussian
    uint8_t* clamped;  // 1 per Gaussian
};

struct TileView {
    uint32_t* count;   // [WGS_HIST_COPIES][tiles]
    uint32_t* cursor;  // [WGS_HIST_COPIES][tiles]
    uint2* ranges;
    uint32_t* misc;
    uint32_t* large_list;  // tile ids whose segment is too long for the small sort kernel
    uint32_t* order;       // tile ids, longest segment first (block -> tile map of the compositing kernels)
};

struct ImageView {
    float* final_T;
    uint32_t* n_contrib;
};

struct GradView {
    float4* acc;           // 3 per Gaussian
    float* pose_partials;  // 6 per preprocess-bwd block
    uint32_t* ticket;
};

__host__ __device__ inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }
__host__ __device__ inline int hist_copies(size_t tiles) {
    int k = WGS_HIST_COPIES;
    while (k > 1 && (size_t)k * tiles > 65536) k >>= 1;
    return k;
}

__host__ __device__ inline size_t geom_bytes(size_t P) {
    return align_up(P * 32) + align_up(P * 16) + align_up(P * 8) + align_up(P);
}
__host__ __device__ inline GeomView geom_view(void* base, size_t P) {
    char* p = (char*)base;
    GeomView g;
    g.rec = (float4*)p; p += align_up(P * 32);
    g.col = (float4*)p; p += align_up(P * 16);
    g.rect = (uint2*)p; p += align_up(P * 8);
    g.clamped = (uint8_t*)p;
    return g;
}
// count[] and misc[] are adjacent so that one memset clears both at the start of a forward
__host__ __device__ inline size_t tile_clear_bytes(size_t tiles) {
    return align_up(WGS_HIST_COPIES * tiles * 4) + align_up(MISC_WORDS * 4);
}
__host__ __device__ inline size_t tile_bytes(size_t tiles) {
    return tile_clear_bytes(tiles) + align_up(WGS_HIST_COPIES * tiles * 4) + 2 * align_up(tiles * 4) + align_up(tiles * 8);
}
__host__ __device__ inline TileView tile_view(void* base, size_t tiles) {
    char* p = (char*)base;
    TileView t;
    t.count = (uint32_t*)p; p += align_up(WGS_HIST_COPIES * tiles * 4);
    t.misc = (uint32_t*)p; p += align_up(MISC_WORDS * 4);
    t.cursor = (uint32_t*)p; p += align_up(WGS_HIST_COPIES * tiles * 4);
    t.large_list = (uint32_t*)p; p += align_up(tiles * 4);
    t.order = (uint32_t*)p; p += align_up(tiles * 4);
    t.ranges = (uint2*)p;
    return t;
}
__host__ __device__ inline size_t image_bytes(size_t N) { return 2 * align_up(N * 4); }
__host__ __device__ inline ImageView image_view(void* base, size_t N) {
    char* p = (char*)base;
    ImageView v;
    v.final_T = (float*)p; p += align_up(N * 4);
    v.n_contrib = (uint32_t*)p;
    return v;
}
#define WGS_PRE_BWD_THREADS 256
__host__ __device__ inline size_t grad_bytes(size_t P) {
    size_t blocks = (P + WGS_PRE_BWD_THREADS - 1) / WGS_PRE_BWD_THREADS;
    return align_up(P * 48) + align_up((blocks + 1) * 6 * 4) + 256;
}
__host__ __device__ inline GradView grad_view(void* base, size_t P) {
    char* p = (char*)base;
    size_t blocks = (P + WGS_PRE_BWD_THREADS - 1) / WGS_PRE_BWD_THREADS;
    GradView g;
    g.acc = (float4*)p; p += align_up(P * 48);
    g.pose_partials = (float*)p; p += align_up((blocks + 1) * 6 * 4);
    g.ticket = (uint32_t*)p;
    return g;
}

// SH basis constants (reference auxiliary.h:22-39)
__device__ constexpr float kSH_C0 = 0.28209479177387814f;
__device__ constexpr float kSH_C1 = 0.4886025119029199f;
__device__ constexpr float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                        -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                        0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                        -0.5900435899266435f};

// a*x + b*y + c*z + d exactly as nvcc contracts the reference's transformPoint4x{3,4}
// (auxiliary.h:58-77): mul(y,b) -> fma(x,a,.) -> fma(z,c,.) -> add(d,.)
__device__ __forceinline__ float dot3p(float x, float y, float z, float a, float b, float c, float d) {
    return ADD(d, FMA(z, c, FMA(x, a, MUL(y, b))));
}
// 3-term dot as nvcc contracts GLM's mat3 products: mul(a1,b1) -> fma(a0,b0,.) -> fma(a2,b2,.)
__device__ __forceinline__ float dot3g(float a0, float b0, float a1, float b1, float a2, float b2) {
    return FMA(a2, b2, FMA(a0, b0, MUL(a1, b1)));
}

// reference auxiliary.h:41-44 -- evaluated in double because of the 1.0 / 0.5 literals
__device__ __forceinline__ float ndc2pix(float v, int S) {
    return (float)(__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5));
}

// reference auxiliary.h:46-56 (division by 16 is an exact multiply)
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, uint32_t& x0,
                                         uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    const float r = (float)radius;
    x0 = min((uint32_t)gx, (uint32_t)max(0, (int)MUL(SUB(px, r), 0.0625f)));
    y0 = min((uint32_t)gy, (uint32_t)max(0, (int)MUL(SUB(py, r), 0.0625f)));
    x1 = min((uint32_t)gx, (uint32_t)max(0, (int)MUL(ADD(ADD(ADD(px, r), 16.0f), -1.0f), 0.0625f)));
    y1 = min((uint32_t)gy, (uint32_t)max(0, (int)MUL(ADD(ADD(ADD(py, r), 16.0f), -1.0f), 0.0625f)));
}

// Conservative test "can Gaussian (x,y,conic,thr) reach alpha >= 1/255 anywhere in the pixel
// rectangle [px0,px1]x[py0,py1]?".  Maximises the concave power(d) over the rectangle; the
// slack covers fp32 rounding of the per-pixel evaluation.  NaN inputs answer "yes".
__device__ __forceinline__ bool reaches_rect(float gx, float gy, float A, float B, float C, float thr,
                                             float px0, float py0, float px1, float py1) {
    // d = mean - pixel, so d ranges over [gx-px1, gx-px0] x [gy-py1, gy-py0]
    const float dx_lo = gx - px1, dx_hi = gx - px0, dy_lo = gy - py1, dy_hi = gy - py0;
    const float cx = fminf(fmaxf(0.f, dx_lo), dx_hi);
    const float cy = fminf(fmaxf(0.f, dy_lo), dy_hi);
    // candidate 1: dx = cx, best dy;  candidate 2: dy = cy, best dx
    const float dy1 = fminf(fmaxf(__fdividef(-B * cx, C), dy_lo), dy_hi);
    const float dx2 = fminf(fmaxf(__fdividef(-B * cy, A), dx_lo), dx_hi);
    const float q1 = -0.5f * (A * cx * cx + C * dy1 * dy1) - B * cx * dy1;
    const float q2 = -0.5f * (A * dx2 * dx2 + C * cy * cy) - B * dx2 * cy;
    const float q = fmaxf(q1, q2);
    const float mag = A * (cx * cx + dx2 * dx2) + C * (cy * cy + dy1 * dy1);
    return !(q + 0.05f + 1e-5f * mag < -thr);
}

}  // namespace wgs
