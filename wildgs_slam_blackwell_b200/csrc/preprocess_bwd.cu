// Backward per-Gaussian stage, one fused pass.  Replaces reference computeCov2DCUDA
// (cuda_rasterizer/backward.cu:150-346), preprocessCUDA (backward, :418-539), the SH backward
// (:21-145), the cov3D backward (:350-413) and the torch.sum over the (P,6) dL_dtau buffer in
// diff_gaussian_rasterization/__init__.py:152-154.
//
// Each thread consumes (and re-zeroes) the 48-byte accumulator record render_bwd filled
// (pixel moments of G*dL/dalpha, colour and depth sums),
// recomputes the forward intermediates it needs (cov3D, T = W*J) instead of reading a stored
// copy, writes every per-Gaussian output exactly once (so outputs need no pre-zeroing), and
// the six pose-gradient terms are reduced warp -> block -> grid (per-block partials, last
// block sums them in a fixed order: deterministic, no (P,6) buffer, no atomics on dL_dtau).
#include "common.cuh"
#include "kernels.h"
#include "sh_stage.cuh"

namespace wgs {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

__device__ __forceinline__ float rcpf(float x) { return __fdividef(1.f, x); }

// gradient store (WgsBackwardArgs::accumulate): 0 overwrite, 1 add to what the buffer holds (plain read-modify-write:
// one backward at a time per buffer), 2 add with a reduction atomic (RED.ADD.F32: backwards on several streams may
// accumulate into one buffer concurrently; a warp's consecutive words still travel as whole sectors)
__device__ __forceinline__ void gput(float* __restrict__ ptr, size_t i, float v, int acc) {
    if (acc == 2) atomicAdd(ptr + i, v);
    else ptr[i] = acc ? ptr[i] + v : v;
}

__device__ __forceinline__ void cov3d_natural(const float* s3, float mod, const float* q4, float c[6], float R[9],
                                              float s[3]) {
    // R[3*col + row] as GLM stores it (reference computeCov3D, forward.cu:120-154)
    const float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
    R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
    R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
    R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
    s[0] = mod * s3[0]; s[1] = mod * s3[1]; s[2] = mod * s3[2];
    float M[9];  // M = S*R : M[c][r] = s_r R[c][r]
#pragma unroll
    for (int cc = 0; cc < 3; cc++)
#pragma unroll
        for (int rr = 0; rr < 3; rr++) M[3 * cc + rr] = s[rr] * R[3 * cc + rr];
    // Sigma = M^T M : Sigma[c][r] = dot(M[r][:], M[c][:])
    c[0] = M[0] * M[0] + M[1] * M[1] + M[2] * M[2];
    c[1] = M[0] * M[3] + M[1] * M[4] + M[2] * M[5];
    c[2] = M[0] * M[6] + M[1] * M[7] + M[2] * M[8];
    c[3] = M[3] * M[3] + M[4] * M[4] + M[5] * M[5];
    c[4] = M[3] * M[6] + M[4] * M[7] + M[5] * M[8];
    c[5] = M[6] * M[6] + M[7] * M[7] + M[8] * M[8];
}

// STAGED (1 < M <= 16): every warp moves the SH rows of its 32 Gaussians -- one contiguous span of global memory --
// through shared memory (sh_stage.cuh): coalesced cp.async in (rows of culled Gaussians skipped), each thread reads
// its coefficients and leaves its dL_dsh there in place, coalesced write-back (zeros for culled rows).  Per-thread
// row access to global memory (192-B stride) moved 1.9 GB for 1.2 GB of algorithmic traffic at config 4.
// QUEUED (tile-row strips, where few Gaussians are visible): see below; otherwise every thread takes its own Gaussian.
template <bool STAGED, bool QUEUED>
__global__ void __launch_bounds__(WGS_PRE_BWD_THREADS, 768 / WGS_PRE_BWD_THREADS) preprocess_bwd_kernel(const __grid_constant__ PreBwdParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    extern __shared__ float sh_stage[];
    const int accm = p.accumulate;
    __shared__ float s_part[WGS_PRE_BWD_THREADS / 32][6];   // pose partials of the block's warps, summed over its chunks
    if ((threadIdx.x & 31) < 6) s_part[threadIdx.x >> 5][threadIdx.x & 31] = 0.f;   // (each warp only touches its own row)
    __syncwarp();   // lanes 0-5 wrote the row, lane 0 read-modify-writes it below
    // A forward that overflowed its instance capacity (no-wait mode) rendered the background only: its backward has
    // zero gradients, and the view must not count as an observation in the densification statistics either.
    const bool stats_on = (p.stat_grad_accum || p.stat_denom || p.stat_max_radii) && p.misc[MISC_OVERFLOW] == 0;
    // A block walks chunks of WGS_PRE_BWD_THREADS Gaussians, chunk = blockIdx.x, + gridDim.x, ... (the grid is sized to
    // the SMs, launch_preprocess_bwd), so the block-level epilogue (pose partials, fence, ticket atomic) is paid once per
    // block.  STAGED: a thread works on its own Gaussian of the chunk (the SH rows of a warp must stay one contiguous
    // span); dead threads of the last chunk shadow its last Gaussian and write nothing.  Otherwise the block QUEUES the
    // visible Gaussians (radii > 0 and a non-empty rectangle inside the rendered strip) of as many chunks as it takes
    // to fill WGS_PRE_BWD_THREADS slots and only then runs the expensive part, on full warps: a 3-row strip of a mapping
    // window keeps 1/7 of the Gaussians, and with per-chunk compaction one or two warps of a block worked while the
    // others sat in the next barrier (ncu: 16 % issue utilisation, 24 of 34 warp-cycles in the barrier).  Invisible
    // Gaussians have nothing to add to an accumulating backward; an overwriting one zero-fills their outputs here.
    const int nchunks = (p.P + WGS_PRE_BWD_THREADS - 1) / WGS_PRE_BWD_THREADS;
    __shared__ int s_queue[QUEUED ? 2 * WGS_PRE_BWD_THREADS : 1];
    __shared__ int s_wcount[WGS_PRE_BWD_THREADS / 32];
    int chunk = blockIdx.x, qn = 0;   // qn: queued Gaussians (the same value in every thread)
    // (the STAGED variant is launched with one block per chunk and runs the body once: with the SH rows in flight the
    // looping form cost it 10 % -- config 4: 0.51 vs 0.46 ms; at M = 1 the loop wins -- config 3: 78 vs 87 us)
    for (int pass = 0; !STAGED || pass < 1; pass++) {
    float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // staged SH rows of this warp: packed (P,M,3) rows, or the (P,M-1,3) rest rows in raw-parameter mode
    const int sh_rw = STAGED ? (p.sh_rest ? 3 * (p.M - 1) : 3 * p.M) : 1;
    const int sh_rs = sh_row_stride(sh_rw);
    int wrow0 = 0;  // STAGED: first Gaussian of the warp
    float* const wstage = sh_stage + (threadIdx.x & ~31) * sh_rs;
    unsigned vmask = 0;
    bool live, work;
    int idx, rad = 0;
    uint2 rect = make_uint2(0u, 0u);
    if (STAGED || !QUEUED) {
        if (chunk >= nchunks) break;
        const int gidx = chunk * WGS_PRE_BWD_THREADS + threadIdx.x;
        wrow0 = chunk * WGS_PRE_BWD_THREADS + (threadIdx.x & ~31);
        chunk += gridDim.x;
        live = gidx < p.P;
        work = true;
        idx = live ? gidx : p.P - 1;
        rad = __ldg(p.radii + idx);
        rect = __ldg(p.geom.rect + idx);
    } else {
        const int lane_ = threadIdx.x & 31, warp_ = threadIdx.x >> 5;
        while (qn < WGS_PRE_BWD_THREADS && chunk < nchunks) {
            const int gidx = chunk * WGS_PRE_BWD_THREADS + threadIdx.x;
            chunk += gridDim.x;
            live = gidx < p.P;
            bool vis1 = false;
            if (live) {
                const int r_ = __ldg(p.radii + gidx);
                const uint2 q_ = __ldg(p.geom.rect + gidx);
                vis1 = r_ > 0 && q_.x != q_.y;
            }
            const unsigned bal = __ballot_sync(0xffffffffu, vis1);
            if (lane_ == 0) s_wcount[warp_] = __popc(bal);
            if (live && !vis1 && accm == 0) {
                const size_t i = (size_t)gidx;
                if (p.dL_dmeans3D) { p.dL_dmeans3D[3 * i] = 0.f; p.dL_dmeans3D[3 * i + 1] = 0.f; p.dL_dmeans3D[3 * i + 2] = 0.f; }
                if (p.dL_dmeans2D) { p.dL_dmeans2D[3 * i] = 0.f; p.dL_dmeans2D[3 * i + 1] = 0.f; p.dL_dmeans2D[3 * i + 2] = 0.f; }
                if (p.dL_dopacity) p.dL_dopacity[i] = 0.f;
                if (p.dL_dcolors) { p.dL_dcolors[3 * i] = 0.f; p.dL_dcolors[3 * i + 1] = 0.f; p.dL_dcolors[3 * i + 2] = 0.f; }
                if (p.dL_dcov3D) {
#pragma unroll
                    for (int k = 0; k < 6; k++) p.dL_dcov3D[6 * i + k] = 0.f;
                }
                if (p.dL_dscales) { p.dL_dscales[3 * i] = 0.f; p.dL_dscales[3 * i + 1] = 0.f; p.dL_dscales[3 * i + 2] = 0.f; }
                if (p.dL_drotations) {
#pragma unroll
                    for (int k = 0; k < 4; k++) p.dL_drotations[4 * i + k] = 0.f;
                }
                if (p.sh && p.dL_dsh) {
                    if (p.sh_rest) {
                        float* __restrict__ d0 = p.dL_dsh + i * 3;
                        d0[0] = d0[1] = d0[2] = 0.f;
                        if (p.dL_dsh_rest)
                            for (int k = 0; k < 3 * (p.M - 1); k++) p.dL_dsh_rest[i * (p.M - 1) * 3 + k] = 0.f;
                    } else {
                        for (int k = 0; k < 3 * p.M; k++) p.dL_dsh[i * p.M * 3 + k] = 0.f;
                    }
                }
            }
            __syncthreads();
            int before = 0, total = 0;
#pragma unroll
            for (int w = 0; w < WGS_PRE_BWD_THREADS / 32; w++) {
                const int c = s_wcount[w];
                before += w < warp_ ? c : 0;
                total += c;
            }
            if (vis1) s_queue[qn + before + __popc(bal & ((1u << lane_) - 1u))] = gidx;
            qn += total;
            __syncthreads();
        }
        if (qn == 0) break;
        const int take = min(qn, WGS_PRE_BWD_THREADS);
        work = (int)threadIdx.x < take;
        live = work;
        idx = work ? s_queue[threadIdx.x] : p.P - 1;
        // the queue's tail moves to its head
        const int rest = qn - take;
        const int moved = (int)threadIdx.x < rest ? s_queue[take + threadIdx.x] : 0;
        __syncthreads();
        if ((int)threadIdx.x < rest) s_queue[threadIdx.x] = moved;
        qn = rest;
        if (work) {
            // (radii / rect of the adopted Gaussian: both answered "visible")
            rad = __ldg(p.radii + idx);
            rect = __ldg(p.geom.rect + idx);
        }
    }
    if (work) {
        // Issue every load this Gaussian can need before looking at radii, so that the latencies
        // overlap (the kernel is latency-bound: 80 registers, 3 CTAs/SM).  Accumulator records of
        // culled Gaussians are never written, so they read as zero.
        float4 g0 = __ldcg(p.grad.acc + 3 * idx), g1 = __ldcg(p.grad.acc + 3 * idx + 1), g2 = __ldcg(p.grad.acc + 3 * idx + 2);
        const V3 m = {__ldg(p.means3D + 3 * idx), __ldg(p.means3D + 3 * idx + 1), __ldg(p.means3D + 3 * idx + 2)};
        float s3[3] = {0.f, 0.f, 0.f}, q4[4] = {1.f, 0.f, 0.f, 0.f}, c3[6];
        if (p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = __ldg(p.cov3D_precomp + 6 * idx + k);
        } else {
#pragma unroll
            for (int k = 0; k < 3; k++) s3[k] = __ldg(p.scales + 3 * idx + k);
#pragma unroll
            for (int k = 0; k < 4; k++) q4[k] = __ldg(p.rotations + 4 * idx + k);
        }
        // raw-parameter mode: the forward applied exp / normalize (preprocess_fwd.cu); redo them here
        float qn = 1.f;
        if (p.act && !p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 3; k++) s3[k] = expf(s3[k]);
            qn = fmaxf(sqrtf(q4[0] * q4[0] + q4[1] * q4[1] + q4[2] * q4[2] + q4[3] * q4[3]), 1e-12f);
#pragma unroll
            for (int k = 0; k < 4; k++) q4[k] = q4[k] * rcpf(qn);
        }
        const float4 ra = __ldcg(p.geom.rec + 2 * idx), rb = __ldcg(p.geom.rec + 2 * idx + 1);
        const bool vis = live && rad > 0 && rect.x != rect.y;  // rect is empty outside the rendered strip
        if (STAGED) {
            vmask = __ballot_sync(0xffffffffu, vis);
            sh_stage_load(wstage, (p.sh_rest ? p.sh_rest : p.sh) + (size_t)wrow0 * sh_rw, vmask, sh_rw, sh_rs);
        }
        if (vis) {
            const float4 z4 = make_float4(0, 0, 0, 0);
            p.grad.acc[3 * idx] = z4; p.grad.acc[3 * idx + 1] = z4; p.grad.acc[3 * idx + 2] = z4;
        }
        // render_bwd accumulates the moments of w = G*dL/dalpha over the pixels, d = mean2D - pixel:
        //   g0 = (S w, S w dx, S w dy, S w dx^2), g1.xy = (S w dx dy, S w dy^2).
        // With dL/dG = opacity * dL/dalpha (backward.cu:741-742) the reference's per-pixel terms
        // (backward.cu:744-757) sum to: dL/dmean2D = -(W/2, H/2) * o * (conic.x Sx + conic.y Sy,
        // conic.y Sx + conic.z Sy), dL/dconic = -1/2 * o * (Sxx, Sxy, Syy), dL/dopacity = S w.
        // Records of culled Gaussians are zero, but their geom record is stale: gate on vis.
        const float o = vis ? rb.y : 0.f, kx = vis ? ra.z : 0.f, ky = vis ? ra.w : 0.f, kz = vis ? rb.x : 0.f;
        const float dm2x = -0.5f * p.W * o * (kx * g0.y + ky * g0.z);   // dL/dmean2D (NDC-scaled)
        const float dm2y = -0.5f * p.H * o * (ky * g0.y + kz * g0.z);
        const float dcx = -0.5f * o * g0.w, dcy = -0.5f * o * g1.x, dcz = -0.5f * o * g1.y;  // dL/dconic (x, y, w)
        // dL/dopacity = S w; raw-parameter mode: through the sigmoid, d sigmoid = o (1 - o)
        const float dop = p.act ? g0.x * o * (1.f - o) : g0.x;
        const float dcol[3] = {g1.z, g1.w, g2.x};
        const float ddepth = g2.y;
        V3 dmean = {0.f, 0.f, 0.f};
        float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
        if (vis) {
            const float* __restrict__ V = p.viewmatrix;
            const float* __restrict__ Pm = p.projmatrix;
            float Rq[9], sc[3];
            if (!p.cov3D_precomp) cov3d_natural(s3, p.scale_modifier, q4, c3, Rq, sc);
            // ---- computeCov2DCUDA (backward.cu:150-346) ----
            V3 t = {V[0] * m.x + V[4] * m.y + V[8] * m.z + V[12], V[1] * m.x + V[5] * m.y + V[9] * m.z + V[13],
                    V[2] * m.x + V[6] * m.y + V[10] * m.z + V[14]};
            const float limx = 1.3f * p.tanfovx, limy = 1.3f * p.tanfovy;
            const float rz = rcpf(t.z);  // gradients are a tolerance path: MUFU reciprocal instead of IEEE division
            const float txtz = t.x * rz, tytz = t.y * rz;
            t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
            t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
            const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
            const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
            const float hx = p.focal_x, hy = p.focal_y;
            const float J00 = hx * rz, J02 = -(hx * t.x) * (rz * rz), J11 = hy * rz, J12 = -(hy * t.y) * (rz * rz);
            // T = W*J, T[c][r]; W[c][r] with W = mat3(V0,V4,V8, V1,V5,V9, V2,V6,V10)
            const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
            const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
            // Vrk rows dotted with T rows
            const float A0 = T00 * c3[0] + T01 * c3[1] + T02 * c3[2], A1 = T00 * c3[1] + T01 * c3[3] + T02 * c3[4],
                        A2 = T00 * c3[2] + T01 * c3[4] + T02 * c3[5];
            const float B0 = T10 * c3[0] + T11 * c3[1] + T12 * c3[2], B1 = T10 * c3[1] + T11 * c3[3] + T12 * c3[4],
                        B2 = T10 * c3[2] + T11 * c3[4] + T12 * c3[5];
            const float a = T00 * A0 + T01 * A1 + T02 * A2 + 0.3f;
            const float b = T00 * B0 + T01 * B1 + T02 * B2;
            const float c = T10 * B0 + T11 * B1 + T12 * B2 + 0.3f;
            const float denom = a * c - b * b;
            const float denom2inv = rcpf((denom * denom) + 0.0000001f);
            float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
            if (denom2inv != 0.f) {
                dL_da = denom2inv * (-c * c * dcx + 2.f * b * c * dcy + (denom - a * c) * dcz);
                dL_dc = denom2inv * (-a * a * dcz + 2.f * a * b * dcy + (denom - a * c) * dcx);
                dL_db = denom2inv * 2.f * (b * c * dcx - (denom + 2.f * b * b) * dcy + a * b * dcz);
                dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
                dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
                dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
                dcov[1] = 2.f * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2.f * T10 * T11 * dL_dc;
                dcov[2] = 2.f * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2.f * T10 * T12 * dL_dc;
                dcov[4] = 2.f * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2.f * T11 * T12 * dL_dc;
            }
            // A* = T[0][:]·Vrk[k][:], B* = T[1][:]·Vrk[k][:]  (Vrk symmetric)
            const float dT00 = 2.f * A0 * dL_da + B0 * dL_db, dT01 = 2.f * A1 * dL_da + B1 * dL_db,
                        dT02 = 2.f * A2 * dL_da + B2 * dL_db;
            const float dT10 = 2.f * B0 * dL_dc + A0 * dL_db, dT11 = 2.f * B1 * dL_dc + A1 * dL_db,
                        dT12 = 2.f * B2 * dL_dc + A2 * dL_db;
            // W[c][r]: W[0]=(V0,V4,V8) W[1]=(V1,V5,V9) W[2]=(V2,V6,V10)
            const float dJ00 = V[0] * dT00 + V[4] * dT01 + V[8] * dT02;
            const float dJ02 = V[2] * dT00 + V[6] * dT01 + V[10] * dT02;
            const float dJ11 = V[1] * dT10 + V[5] * dT11 + V[9] * dT12;
            const float dJ12 = V[2] * dT10 + V[6] * dT11 + V[10] * dT12;
            const float tz = rz, tz2 = tz * tz, tz3 = tz2 * tz;
            const float dtx = x_grad_mul * -hx * tz2 * dJ02;
            const float dty = y_grad_mul * -hy * tz2 * dJ12;
            const float dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2.f * hx * t.x) * tz3 * dJ02 + (2.f * hy * t.y) * tz3 * dJ12;
            // pose via camera-space mean (clamped t), (:273-288)
            tau[0] += dtx; tau[1] += dty; tau[2] += dtz;
            tau[3] += dty * (-t.z) + dtz * t.y;
            tau[4] += dtx * t.z + dtz * (-t.x);
            tau[5] += dtx * (-t.y) + dty * t.x;
            dmean = {V[0] * dtx + V[1] * dty + V[2] * dtz, V[4] * dtx + V[5] * dty + V[6] * dtz,
                     V[8] * dtx + V[9] * dty + V[10] * dtz};
            // pose via W (:299-343)
            {
                const V3 dWc1 = {J00 * dT00, J11 * dT10, J02 * dT00 + J12 * dT10};
                const V3 dWc2 = {J00 * dT01, J11 * dT11, J02 * dT01 + J12 * dT11};
                const V3 dWc3 = {J00 * dT02, J11 * dT12, J02 * dT02 + J12 * dT12};
                const V3 c1 = {V[0], V[1], V[2]}, c2 = {V[4], V[5], V[6]}, c3v = {V[8], V[9], V[10]};
                // -skew(c).cols: (0,-cz,cy) (cz,0,-cx) (-cy,cx,0)
                tau[3] += (-c1.z * dWc1.y + c1.y * dWc1.z) + (-c2.z * dWc2.y + c2.y * dWc2.z) + (-c3v.z * dWc3.y + c3v.y * dWc3.z);
                tau[4] += (c1.z * dWc1.x - c1.x * dWc1.z) + (c2.z * dWc2.x - c2.x * dWc2.z) + (c3v.z * dWc3.x - c3v.x * dWc3.z);
                tau[5] += (-c1.y * dWc1.x + c1.x * dWc1.y) + (-c2.y * dWc2.x + c2.x * dWc2.y) + (-c3v.y * dWc3.x + c3v.x * dWc3.y);
            }
            // ---- preprocessCUDA backward (:442-538) ----
            const float mhx = Pm[0] * m.x + Pm[4] * m.y + Pm[8] * m.z + Pm[12];
            const float mhy = Pm[1] * m.x + Pm[5] * m.y + Pm[9] * m.z + Pm[13];
            const float mhw = Pm[3] * m.x + Pm[7] * m.y + Pm[11] * m.z + Pm[15];
            const float m_w = rcpf(mhw + 0.0000001f);
            const float mul1 = mhx * m_w * m_w, mul2 = mhy * m_w * m_w;
            dmean.x += (Pm[0] * m_w - Pm[3] * mul1) * dm2x + (Pm[1] * m_w - Pm[3] * mul2) * dm2y;
            dmean.y += (Pm[4] * m_w - Pm[7] * mul1) * dm2x + (Pm[5] * m_w - Pm[7] * mul2) * dm2y;
            dmean.z += (Pm[8] * m_w - Pm[11] * mul1) * dm2x + (Pm[9] * m_w - Pm[11] * mul2) * dm2y;
            {
                const float alpha = m_w, beta = -mhx * m_w * m_w, gamma = -mhy * m_w * m_w;
                const float ra = __ldg(p.projmatrix_raw), rb = __ldg(p.projmatrix_raw + 5), re = __ldg(p.projmatrix_raw + 11);
                const V3 pC = {V[0] * m.x + V[4] * m.y + V[8] * m.z + V[12], V[1] * m.x + V[5] * m.y + V[9] * m.z + V[13],
                               V[2] * m.x + V[6] * m.y + V[10] * m.z + V[14]};
                const V3 d1 = {alpha * ra, 0.f, beta * re}, d2 = {0.f, alpha * rb, gamma * re};
                const V3 th1 = cross(pC, d1), th2 = cross(pC, d2);
                tau[0] += dm2x * d1.x + dm2y * d2.x; tau[1] += dm2x * d1.y + dm2y * d2.y; tau[2] += dm2x * d1.z + dm2y * d2.z;
                tau[3] += dm2x * th1.x + dm2y * th2.x; tau[4] += dm2x * th1.y + dm2y * th2.y; tau[5] += dm2x * th1.z + dm2y * th2.z;
                // depth = p_view.z (:514-528)
                dmean.x += ddepth * V[2]; dmean.y += ddepth * V[6]; dmean.z += ddepth * V[10];
                tau[2] += ddepth; tau[3] += ddepth * pC.y; tau[4] += -ddepth * pC.x;
            }
            // ---- SH backward (:21-145) ----
            if (p.sh) {
                // packed (P,M,3) layout, or DC / rest split (raw-parameter mode): coefficient k >= 1 lives at shr + 3(k-1)
                const float* __restrict__ sh0 = p.sh_rest ? p.sh + (size_t)idx * 3 : p.sh + (size_t)idx * p.M * 3;
                const float* __restrict__ shr = p.sh_rest ? p.sh_rest + (size_t)idx * (p.M - 1) * 3 : sh0 + 3;
                float* __restrict__ dsh0 = p.sh_rest ? p.dL_dsh + (size_t)idx * 3 : p.dL_dsh + (size_t)idx * p.M * 3;
                float* __restrict__ dshr = p.sh_rest ? p.dL_dsh_rest + (size_t)idx * (p.M - 1) * 3 : dsh0 + 3;
                const uint8_t cb = p.geom.clamped[idx];
                const V3 dRGB = {(cb & 1) ? 0.f : dcol[0], (cb & 2) ? 0.f : dcol[1], (cb & 4) ? 0.f : dcol[2]};
                const V3 dir_orig = {m.x - __ldg(p.campos), m.y - __ldg(p.campos + 1), m.z - __ldg(p.campos + 2)};
                const float il = rsqrtf(dot(dir_orig, dir_orig));
                const float x = dir_orig.x * il, y = dir_orig.y * il, z = dir_orig.z * il;
                // STAGED: coefficient k of this thread's row sits at srow + 3 (k + koff); dL_dsh replaces it in place,
                // so every band reads its coefficients before it puts its gradients
                float* __restrict__ srow = sh_stage + threadIdx.x * sh_rs;  // == wstage + lane * sh_rs
                const int koff = p.sh_rest ? -1 : 0;
                auto SH = [&](int k) {  // only called with k >= 1
                    if (STAGED) return V3{srow[3 * (k + koff)], srow[3 * (k + koff) + 1], srow[3 * (k + koff) + 2]};
                    return V3{__ldg(shr + 3 * (k - 1)), __ldg(shr + 3 * (k - 1) + 1), __ldg(shr + 3 * (k - 1) + 2)};
                };
                auto put = [&](int k, float w) {
                    if (STAGED && (k || !p.sh_rest)) {
                        float* d = srow + 3 * (k + koff);
                        d[0] = w * dRGB.x; d[1] = w * dRGB.y; d[2] = w * dRGB.z;
                        return;
                    }
                    float* d = k ? dshr + 3 * (k - 1) : dsh0;
                    gput(d, 0, w * dRGB.x, accm); gput(d, 1, w * dRGB.y, accm); gput(d, 2, w * dRGB.z, accm);
                };
                V3 dx = {0, 0, 0}, dy = {0, 0, 0}, dz = {0, 0, 0};
                put(0, kSH_C0);
                if (p.D > 0) {
                    dx = -kSH_C1 * SH(3); dy = -kSH_C1 * SH(1); dz = kSH_C1 * SH(2);
                    put(1, -kSH_C1 * y); put(2, kSH_C1 * z); put(3, -kSH_C1 * x);
                    if (p.D > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        dx = dx + (kSH_C2[0] * y) * SH(4) + (kSH_C2[2] * 2.f * -x) * SH(6) + (kSH_C2[3] * z) * SH(7) + (kSH_C2[4] * 2.f * x) * SH(8);
                        dy = dy + (kSH_C2[0] * x) * SH(4) + (kSH_C2[1] * z) * SH(5) + (kSH_C2[2] * 2.f * -y) * SH(6) + (kSH_C2[4] * 2.f * -y) * SH(8);
                        dz = dz + (kSH_C2[1] * y) * SH(5) + (kSH_C2[2] * 2.f * 2.f * z) * SH(6) + (kSH_C2[3] * x) * SH(7);
                        put(4, kSH_C2[0] * xy); put(5, kSH_C2[1] * yz); put(6, kSH_C2[2] * (2.f * zz - xx - yy));
                        put(7, kSH_C2[3] * xz); put(8, kSH_C2[4] * (xx - yy));
                        if (p.D > 2) {
                            dx = dx + (kSH_C3[0] * 3.f * 2.f * xy) * SH(9) + (kSH_C3[1] * yz) * SH(10) + (kSH_C3[2] * -2.f * xy) * SH(11) +
                                 (kSH_C3[3] * -3.f * 2.f * xz) * SH(12) + (kSH_C3[4] * (-3.f * xx + 4.f * zz - yy)) * SH(13) +
                                 (kSH_C3[5] * 2.f * xz) * SH(14) + (kSH_C3[6] * 3.f * (xx - yy)) * SH(15);
                            dy = dy + (kSH_C3[0] * 3.f * (xx - yy)) * SH(9) + (kSH_C3[1] * xz) * SH(10) +
                                 (kSH_C3[2] * (-3.f * yy + 4.f * zz - xx)) * SH(11) + (kSH_C3[3] * -3.f * 2.f * yz) * SH(12) +
                                 (kSH_C3[4] * -2.f * xy) * SH(13) + (kSH_C3[5] * -2.f * yz) * SH(14) + (kSH_C3[6] * -3.f * 2.f * xy) * SH(15);
                            dz = dz + (kSH_C3[1] * xy) * SH(10) + (kSH_C3[2] * 4.f * 2.f * yz) * SH(11) +
                                 (kSH_C3[3] * 3.f * (2.f * zz - xx - yy)) * SH(12) + (kSH_C3[4] * 4.f * 2.f * xz) * SH(13) +
                                 (kSH_C3[5] * (xx - yy)) * SH(14);
                            put(9, kSH_C3[0] * y * (3.f * xx - yy)); put(10, kSH_C3[1] * xy * z);
                            put(11, kSH_C3[2] * y * (4.f * zz - xx - yy)); put(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                            put(13, kSH_C3[4] * x * (4.f * zz - xx - yy)); put(14, kSH_C3[5] * z * (xx - yy));
                            put(15, kSH_C3[6] * x * (xx - 3.f * yy));
                        }
                    }
                }
                for (int k = (p.D + 1) * (p.D + 1); k < p.M; k++) put(k, 0.f);  // inactive bands
                const V3 ddir = {dot(dx, dRGB), dot(dy, dRGB), dot(dz, dRGB)};
                // dnormvdv (auxiliary.h:107-117)
                const V3 v = dir_orig;
                const float sum2 = dot(v, v);
                const float invsum32 = il * il * il;  // 1 / |v|^3
                const V3 dms = {((sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * invsum32,
                                (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * invsum32,
                                (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * invsum32};
                dmean = dmean + dms;
                tau[0] -= dms.x; tau[1] -= dms.y; tau[2] -= dms.z;
            }
            // ---- cov3D backward (:350-413) ----
            if (!p.cov3D_precomp) {
                float M[9];
#pragma unroll
                for (int cc = 0; cc < 3; cc++)
#pragma unroll
                    for (int rr = 0; rr < 3; rr++) M[3 * cc + rr] = sc[rr] * Rq[3 * cc + rr];
                // dL_dSigma (symmetric), GLM [c][r]
                const float S[9] = {dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                                    0.5f * dcov[2], 0.5f * dcov[4], dcov[5]};
                // dL_dM = 2 * M * dL_dSigma : [c][r] = 2 * sum_k M[k][r] * S[c][k]
                float dM[9];
#pragma unroll
                for (int cc = 0; cc < 3; cc++)
#pragma unroll
                    for (int rr = 0; rr < 3; rr++)
                        dM[3 * cc + rr] = 2.f * (M[rr] * S[3 * cc] + M[3 + rr] * S[3 * cc + 1] + M[6 + rr] * S[3 * cc + 2]);
                // Rt = R^T, dMt = dM^T ; Rt[i] (column i of R^T) = row i of R ; dMt[i] = row i of dM
                float dMt[9];  // dMt[3*i + j] = dM[j][i]  (GLM: dMt[i][j])
#pragma unroll
                for (int i = 0; i < 3; i++)
#pragma unroll
                    for (int j = 0; j < 3; j++) dMt[3 * i + j] = dM[3 * j + i];
#pragma unroll
                for (int i = 0; i < 3; i++)
                    dscale[i] = Rq[i] * dMt[3 * i] + Rq[3 + i] * dMt[3 * i + 1] + Rq[6 + i] * dMt[3 * i + 2];
#pragma unroll
                for (int i = 0; i < 3; i++)
#pragma unroll
                    for (int j = 0; j < 3; j++) dMt[3 * i + j] *= sc[i];
                const float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
#define DMT(i, j) dMt[3 * (i) + (j)]
                drot[0] = 2 * z * (DMT(0, 1) - DMT(1, 0)) + 2 * y * (DMT(2, 0) - DMT(0, 2)) + 2 * x * (DMT(1, 2) - DMT(2, 1));
                drot[1] = 2 * y * (DMT(1, 0) + DMT(0, 1)) + 2 * z * (DMT(2, 0) + DMT(0, 2)) + 2 * r * (DMT(1, 2) - DMT(2, 1)) -
                          4 * x * (DMT(2, 2) + DMT(1, 1));
                drot[2] = 2 * x * (DMT(1, 0) + DMT(0, 1)) + 2 * r * (DMT(2, 0) - DMT(0, 2)) + 2 * z * (DMT(1, 2) + DMT(2, 1)) -
                          4 * y * (DMT(2, 2) + DMT(0, 0));
                drot[3] = 2 * r * (DMT(0, 1) - DMT(1, 0)) + 2 * x * (DMT(2, 0) + DMT(0, 2)) + 2 * y * (DMT(1, 2) + DMT(2, 1)) -
                          4 * z * (DMT(1, 1) + DMT(0, 0));
#undef DMT
            }
        } else if (live && p.sh && p.dL_dsh && !accm) {
            if (p.sh_rest) {
                float* __restrict__ d0 = p.dL_dsh + (size_t)idx * 3;
                float* __restrict__ dr = p.dL_dsh_rest + (size_t)idx * (p.M - 1) * 3;
                d0[0] = d0[1] = d0[2] = 0.f;
                if (!STAGED)
                    for (int k = 0; k < 3 * (p.M - 1); k++) dr[k] = 0.f;
            } else if (!STAGED) {
                float* __restrict__ dsh = p.dL_dsh + (size_t)idx * p.M * 3;
                for (int k = 0; k < 3 * p.M; k++) dsh[k] = 0.f;
            }
        }
        if (p.act && vis && !p.cov3D_precomp) {
            // chain through exp (d exp = s) and normalize (d q_hat = (I - q_hat q_hat^T) / |q|)
#pragma unroll
            for (int k = 0; k < 3; k++) dscale[k] *= s3[k];
            const float qd = q4[0] * drot[0] + q4[1] * drot[1] + q4[2] * drot[2] + q4[3] * drot[3];
#pragma unroll
            for (int k = 0; k < 4; k++) drot[k] = (drot[k] - q4[k] * qd) * rcpf(qn);
        }
        if (live && rad > 0 && stats_on) {   // mapper.py:1176-1182 + add_densification_stats (gaussian_model.py:745-749), visibility = radii > 0
            const float gn = sqrtf(dm2x * dm2x + dm2y * dm2y);
            if (accm == 2) {   // views on several streams update the statistics concurrently
                if (p.stat_grad_accum) atomicAdd(p.stat_grad_accum + idx, gn);
                if (p.stat_denom) atomicAdd(p.stat_denom + idx, 1.f);
                // non-negative floats order like their bit patterns
                if (p.stat_max_radii) atomicMax((int*)p.stat_max_radii + idx, __float_as_int((float)rad));
            } else {
                if (p.stat_grad_accum) p.stat_grad_accum[idx] += gn;
                if (p.stat_denom) p.stat_denom[idx] += 1.f;
                if (p.stat_max_radii) p.stat_max_radii[idx] = fmaxf(p.stat_max_radii[idx], (float)rad);
            }
        }
        if (live) {
        if (p.dL_dmeans3D) { gput(p.dL_dmeans3D, 3 * idx, dmean.x, accm); gput(p.dL_dmeans3D, 3 * idx + 1, dmean.y, accm); gput(p.dL_dmeans3D, 3 * idx + 2, dmean.z, accm); }
        if (p.dL_dmeans2D) { gput(p.dL_dmeans2D, 3 * idx, dm2x, accm); gput(p.dL_dmeans2D, 3 * idx + 1, dm2y, accm); gput(p.dL_dmeans2D, 3 * idx + 2, 0.f, accm); }
        if (p.dL_dopacity) gput(p.dL_dopacity, idx, dop, accm);
        if (p.dL_dcolors) { gput(p.dL_dcolors, 3 * idx, dcol[0], accm); gput(p.dL_dcolors, 3 * idx + 1, dcol[1], accm); gput(p.dL_dcolors, 3 * idx + 2, dcol[2], accm); }
        if (p.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) gput(p.dL_dcov3D, 6 * idx + k, dcov[k], accm);
        }
        if (p.dL_dscales) { gput(p.dL_dscales, 3 * idx, dscale[0], accm); gput(p.dL_dscales, 3 * idx + 1, dscale[1], accm); gput(p.dL_dscales, 3 * idx + 2, dscale[2], accm); }
        if (p.dL_drotations) {
#pragma unroll
            for (int k = 0; k < 4; k++) gput(p.dL_drotations, 4 * idx + k, drot[k], accm);
        }
        }
    }
    if (STAGED) {
        float* dst = p.sh_rest ? p.dL_dsh_rest : p.dL_dsh;
        if (dst && wrow0 < p.P)
            sh_stage_store(wstage, dst + (size_t)wrow0 * sh_rw, vmask, (uint32_t)min(32, p.P - wrow0), sh_rw, sh_rs, accm);
        __syncwarp();   // the warp's stage is refilled by its next chunk
    }
    // pose gradient of the chunk: lanes -> the warp's running partial (kept in shared memory, not in six registers
    // that would stay live across the next chunk)
#pragma unroll
    for (int i = 0; i < 6; i++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tau[i] += __shfl_xor_sync(0xffffffffu, tau[i], o);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int i = 0; i < 6; i++) s_part[threadIdx.x >> 5][i] += tau[i];
    }
    }   // fill / process loop
    // ---- pose gradient: warp (over its chunks) -> block -> grid, deterministic for a given grid ----
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (threadIdx.x < 6) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < WGS_PRE_BWD_THREADS / 32; w++) s += s_part[w][threadIdx.x];
        p.grad.pose_partials[6 * blockIdx.x + threadIdx.x] = s;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(p.grad.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) {
        __threadfence();
        // 6 warps, one per component; lanes stride over the blocks, fixed order
        if (warp < 6) {
            double s = 0.0;
            for (unsigned b = lane; b < gridDim.x; b += 32) s += (double)((volatile float*)p.grad.pose_partials)[6 * b + warp];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && p.dL_dtau) p.dL_dtau[warp] = (float)s;
        }
        if (threadIdx.x == 0) *p.grad.ticket = 0;
    }
}

int configure_preprocess_bwd() {
    return (int)cudaFuncSetAttribute(preprocess_bwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sh_stage_bytes(SH_STAGE_MAX_M, WGS_PRE_BWD_THREADS));
}

static_assert(WGS_PRE_BWD_THREADS >= 6 * 32, "the last block sums the six pose components with one warp each");

void launch_preprocess_bwd(const PreBwdParams& p, cudaStream_t s) {
    if (p.P <= 0) return;
    // every resident slot gets one block (768 / WGS_PRE_BWD_THREADS per SM, __launch_bounds__), twice over so that the
    // hardware scheduler can even out blocks whose chunks hold few visible Gaussians
    static int slots = 0;
    if (!slots) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        slots = sms * (768 / WGS_PRE_BWD_THREADS) * 2;
    }
    const int chunks = (p.P + WGS_PRE_BWD_THREADS - 1) / WGS_PRE_BWD_THREADS;
    const int blocks = min(chunks, slots);
    if (p.sh && p.M > 1 && p.M <= SH_STAGE_MAX_M)
        launch_chained(preprocess_bwd_kernel<true, false>, chunks, WGS_PRE_BWD_THREADS, sh_stage_bytes(p.sh_rest ? p.M - 1 : p.M, WGS_PRE_BWD_THREADS), s, p);
    else if (p.strip)   // a strip keeps a fraction of the Gaussians: queue the visible ones (whole views: 32 vs 28 us at config 2)
        launch_chained(preprocess_bwd_kernel<false, true>, blocks, WGS_PRE_BWD_THREADS, 0, s, p);
    else
        launch_chained(preprocess_bwd_kernel<false, false>, blocks, WGS_PRE_BWD_THREADS, 0, s, p);
}

}  // namespace wgs
