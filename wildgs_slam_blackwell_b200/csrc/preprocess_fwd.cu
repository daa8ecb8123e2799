// Forward per-Gaussian stage: projection, near-cull, 3D->2D covariance (EWA), conic, radius,
// tile rectangle, SH->RGB, and the per-tile instance histogram (first radix digit = tile id).
// Replaces reference preprocessCUDA (cuda_rasterizer/forward.cu:157-258) + the tiles_touched
// scan input (rasterizer_impl.cu:280).  Every value that reaches radii / depth bits / tile
// rect / conic is computed with the exact rn-primitive sequence nvcc 12.9 emits for the
// reference (PTX-verified), so those outputs are bit-identical.
#include "common.cuh"
#include "kernels.h"
#include "sh_stage.cuh"

namespace wgs {

// sq = {scale xyz, quaternion rxyz}, loaded by the caller (all loads of a Gaussian are issued up front)
__device__ __forceinline__ void cov3d_from_scale_rot(const float sq[7], float mod, bool act, float c[6]) {
    // reference computeCov3D (forward.cu:120-154); quaternion (r,x,y,z) is NOT normalised
    float s0 = sq[0], s1 = sq[1], s2 = sq[2];
    float r = sq[3], x = sq[4], y = sq[5], z = sq[6];
    if (act) {
        // raw-parameter mode: GaussianModel.get_scaling = exp(_scaling), get_rotation =
        // F.normalize(_rotation) = q / max(|q|, 1e-12)  (gaussian_model.py:54-63,77-83)
        s0 = expf(s0); s1 = expf(s1); s2 = expf(s2);
        const float n = fmaxf(SQRT(ADD(ADD(ADD(MUL(r, r), MUL(x, x)), MUL(y, y)), MUL(z, z))), 1e-12f);
        r = DIV(r, n); x = DIV(x, n); y = DIV(y, n); z = DIV(z, n);
    }
    const float sx = MUL(mod, s0), sy = MUL(mod, s1), sz = MUL(mod, s2);
    // ptxas fuses the second product of each 2-term sum into an FFMA (SASS-verified)
    const float yy = MUL(y, y), zz = MUL(z, z), xz = MUL(x, z), rx = MUL(r, x), rz = MUL(r, z);
    const float yy_zz = ADD(yy, zz), xx_zz = FMA(x, x, zz), xx_yy = FMA(x, x, yy);
    const float xy_m_rz = FMA(x, y, -rz), xy_p_rz = FMA(x, y, rz), ry_p_xz = FMA(r, y, xz), xz_m_ry = FMA(-r, y, xz),
                yz_m_rx = FMA(y, z, -rx), rx_p_yz = FMA(y, z, rx);
    const float R00 = SUB(1.f, ADD(yy_zz, yy_zz)), R01 = ADD(xy_m_rz, xy_m_rz), R02 = ADD(ry_p_xz, ry_p_xz);
    const float R10 = ADD(xy_p_rz, xy_p_rz), R11 = SUB(1.f, ADD(xx_zz, xx_zz)), R12 = ADD(yz_m_rx, yz_m_rx);
    const float R20 = ADD(xz_m_ry, xz_m_ry), R21 = ADD(rx_p_yz, rx_p_yz), R22 = SUB(1.f, ADD(xx_yy, xx_yy));
    // M = S * R (GLM column-major): M[c][r] = s_r * R[c][r]
    const float M00 = MUL(sx, R00), M01 = MUL(sy, R01), M02 = MUL(sz, R02);
    const float M10 = MUL(sx, R10), M11 = MUL(sy, R11), M12 = MUL(sz, R12);
    const float M20 = MUL(sx, R20), M21 = MUL(sy, R21), M22 = MUL(sz, R22);
    // Sigma = M^T M
    c[0] = dot3g(M00, M00, M01, M01, M02, M02);
    c[1] = dot3g(M10, M00, M11, M01, M12, M02);
    c[2] = dot3g(M20, M00, M21, M01, M22, M02);
    c[3] = dot3g(M10, M10, M11, M11, M12, M12);
    c[4] = dot3g(M20, M10, M21, M11, M22, M12);
    c[5] = dot3g(M20, M20, M21, M21, M22, M22);
}

// reference computeColorFromSH (forward.cu:22-73); returns rgb, sets clamp bits
// sh: coefficient 0 (DC); shr: coefficients 1.. (shr = sh + 3 for the packed (P,M,3) layout)
// SMEM: both pointers address the block's staged copy in shared memory (plain loads instead of ld.global.nc)
template <bool SMEM>
__device__ __forceinline__ float3 sh_to_rgb(const float* __restrict__ sh, const float* __restrict__ shr, int deg,
                                            float px, float py, float pz, const float* __restrict__ campos,
                                            uint32_t& clampbits, bool sh_global = false) {
    const float dx = SUB(px, __ldg(campos)), dy = SUB(py, __ldg(campos + 1)), dz = SUB(pz, __ldg(campos + 2));
    const float len = SQRT(FMA(dz, dz, FMA(dx, dx, MUL(dy, dy))));
    const float x = DIV(dx, len), y = DIV(dy, len), z = DIV(dz, len);
    auto LD = [](const float* q) { return SMEM ? *q : __ldg(q); };
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = MUL(SMEM && sh_global ? __ldg(sh + c) : LD(sh + c), kSH_C0);
    if (deg > 0) {
        const float t1 = MUL(y, kSH_C1), t2 = MUL(z, kSH_C1), t3 = MUL(x, kSH_C1);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float v = FMA(-t1, LD(shr + c), res[c]);
            v = FMA(t2, LD(shr + 3 + c), v);
            res[c] = FMA(-t3, LD(shr + 6 + c), v);
        }
        if (deg > 1) {
            const float xx = MUL(x, x), yy = MUL(y, y), zz = MUL(z, z), xy = MUL(x, y), yz = MUL(y, z),
                        xz = MUL(x, z);
            const float zz2 = ADD(zz, zz), xx_yy = SUB(xx, yy);
            const float b4 = MUL(xy, kSH_C2[0]), b5 = MUL(yz, kSH_C2[1]),
                        b6 = MUL(SUB(SUB(zz2, xx), yy), kSH_C2[2]), b7 = MUL(xz, kSH_C2[3]),
                        b8 = MUL(xx_yy, kSH_C2[4]);
#pragma unroll
            for (int c = 0; c < 3; c++) {
                float v = FMA(b4, LD(shr + 9 + c), res[c]);
                v = FMA(b5, LD(shr + 12 + c), v);
                v = FMA(b6, LD(shr + 15 + c), v);
                v = FMA(b7, LD(shr + 18 + c), v);
                res[c] = FMA(b8, LD(shr + 21 + c), v);
            }
            if (deg > 2) {
                const float w = SUB(FMA(zz, 4.f, -xx), yy);
                const float b9 = MUL(MUL(y, kSH_C3[0]), FMA(xx, 3.f, -yy));
                const float b10 = MUL(z, MUL(xy, kSH_C3[1]));
                const float b11 = MUL(MUL(y, kSH_C3[2]), w);
                const float b12 = MUL(MUL(z, kSH_C3[3]), FMA(yy, -3.f, FMA(xx, -3.f, zz2)));
                const float b13 = MUL(MUL(x, kSH_C3[4]), w);
                const float b14 = MUL(MUL(z, kSH_C3[5]), xx_yy);
                const float b15 = MUL(MUL(x, kSH_C3[6]), FMA(yy, -3.f, xx));
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    float v = FMA(b9, LD(shr + 24 + c), res[c]);
                    v = FMA(b10, LD(shr + 27 + c), v);
                    v = FMA(b11, LD(shr + 30 + c), v);
                    v = FMA(b12, LD(shr + 33 + c), v);
                    v = FMA(b13, LD(shr + 36 + c), v);
                    v = FMA(b14, LD(shr + 39 + c), v);
                    res[c] = FMA(b15, LD(shr + 42 + c), v);
                }
            }
        }
    }
    clampbits = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        res[c] = ADD(res[c], 0.5f);
        if (res[c] < 0.f) { clampbits |= 1u << c; res[c] = 0.f; }
    }
    return make_float3(res[0], res[1], res[2]);
}

// STAGED (1 < M <= 16, SH colours): once visibility is known, every warp copies the SH rows of its visible
// Gaussians -- contiguous in global memory -- into shared memory with coalesced cp.async (sh_stage.cuh) and each
// thread reads its row from there; per-thread row reads (192-B stride at M = 16) reached only 2.9 TB/s at config 4.
template <bool STAGED>
__global__ void __launch_bounds__(256) preprocess_fwd_kernel(const __grid_constant__ PreFwdParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    extern __shared__ float sh_stage[];
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    bool visible = false;
    float px = 0.f, py = 0.f, pz = 0.f, depth = 0.f, opac_raw = 0.f;
    float mx = 0.f, my = 0.f, cnx = 0.f, cny = 0.f, cnz = 0.f;
    if (idx < p.P) {
        const float* __restrict__ V = p.viewmatrix;
        const float* __restrict__ Pm = p.projmatrix;
        px = __ldg(p.means3D + 3 * idx); py = __ldg(p.means3D + 3 * idx + 1);
        pz = __ldg(p.means3D + 3 * idx + 2);
        // Issue every load this Gaussian can need before the frustum test, so that the DRAM round trips overlap
        // instead of chaining means -> (depth test) -> scales/rotation -> (visible) -> opacity (the kernel is
        // latency-bound); culled Gaussians pay 32 B of extra traffic.
        float sq[7] = {0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f};
        if (!p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 3; k++) sq[k] = __ldg(p.scales + 3 * idx + k);
#pragma unroll
            for (int k = 0; k < 4; k++) sq[3 + k] = __ldg(p.rotations + 4 * idx + k);
        }
        opac_raw = __ldg(p.opacities + idx);
        depth = dot3p(px, py, pz, V[2], V[6], V[10], V[14]);
        int radius = 0;
        if (depth <= 0.2f) {  // in_frustum (auxiliary.h:139-164)
            if (p.prefiltered) p.tile.misc[MISC_PREFILTER_HIT] = 1;
        } else {
            const float hx = dot3p(px, py, pz, Pm[0], Pm[4], Pm[8], Pm[12]);
            const float hy = dot3p(px, py, pz, Pm[1], Pm[5], Pm[9], Pm[13]);
            const float hw = dot3p(px, py, pz, Pm[3], Pm[7], Pm[11], Pm[15]);
            const float pw = RCP(ADD(hw, 0.0000001f));
            const float projx = MUL(hx, pw), projy = MUL(hy, pw);
            float c[6];
            if (p.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c[k] = __ldg(p.cov3D_precomp + 6 * idx + k);
            } else {
                cov3d_from_scale_rot(sq, p.scale_modifier, p.act != 0, c);
            }
            // computeCov2D (forward.cu:76-115)
            const float tx = dot3p(px, py, pz, V[0], V[4], V[8], V[12]);
            const float ty = dot3p(px, py, pz, V[1], V[5], V[9], V[13]);
            const float tz = depth;
            const float limx = MUL(p.tanfovx, 1.3f), limy = MUL(p.tanfovy, 1.3f);
            const float cx = fminf(limx, fmaxf(-limx, DIV(tx, tz)));
            const float cy = fminf(limy, fmaxf(-limy, DIV(ty, tz)));
            const float tz2 = MUL(tz, tz);
            const float J00 = DIV(p.focal_x, tz), J02 = DIV(MUL(p.focal_x, MUL(cx, -tz)), tz2);
            const float J11 = DIV(p.focal_y, tz), J12 = DIV(MUL(p.focal_y, MUL(cy, -tz)), tz2);
            const float T00 = FMA(V[2], J02, MUL(V[0], J00)), T01 = FMA(V[6], J02, MUL(V[4], J00)),
                        T02 = FMA(J02, V[10], MUL(V[8], J00));
            const float T10 = FMA(V[2], J12, MUL(J11, V[1])), T11 = FMA(V[6], J12, MUL(J11, V[5])),
                        T12 = FMA(J12, V[10], MUL(J11, V[9]));
            const float A0 = dot3g(T00, c[0], T01, c[1], T02, c[2]), A1 = dot3g(T00, c[1], T01, c[3], T02, c[4]),
                        A2 = dot3g(T00, c[2], T01, c[4], T02, c[5]);
            const float B0 = dot3g(T10, c[0], T11, c[1], T12, c[2]), B1 = dot3g(T10, c[1], T11, c[3], T12, c[4]),
                        B2 = dot3g(T10, c[2], T11, c[4], T12, c[5]);
            const float covx = ADD(dot3g(T00, A0, T01, A1, T02, A2), 0.3f);
            const float covy = dot3g(T00, B0, T01, B1, T02, B2);
            const float covz = ADD(dot3g(T10, B0, T11, B1, T12, B2), 0.3f);
            const float det = FMA(covx, covz, -MUL(covy, covy));
            if (det != 0.0f) {
                const float det_inv = RCP(det);
                cnx = MUL(covz, det_inv); cny = MUL(det_inv, -covy); cnz = MUL(covx, det_inv);
                const float mid = MUL(ADD(covx, covz), 0.5f);
                const float s = SQRT(fmaxf(FMA(mid, mid, -det), 0.1f));
                const float rad_f = ceilf(MUL(SQRT(fmaxf(ADD(mid, s), SUB(mid, s))), 3.f));
                mx = ndc2pix(projx, p.W); my = ndc2pix(projy, p.H);
                const int r = (int)rad_f;
                get_rect(mx, my, r, p.gx, p.gy, x0, y0, x1, y1);
                if ((x1 - x0) * (y1 - y0) != 0) {
                    radius = r;  // whole-image meaning (forward.cu:238-239), also in strip mode
                    y0 = max(y0, (uint32_t)p.row0);
                    y1 = min(y1, (uint32_t)p.row1);
                    visible = y1 > y0;  // touches the strip being rendered
                }
            }
        }
        p.radii[idx] = radius;
        p.n_touched[idx] = 0;
    }
    const int sh_rw = STAGED ? (p.sh_rest ? 3 * (p.M - 1) : 3 * p.M) : 1;
    const int sh_rs = sh_row_stride(sh_rw);
    if (STAGED) {
        const int wrow0 = blockIdx.x * 256 + (threadIdx.x & ~31);
        sh_stage_load(sh_stage + (threadIdx.x & ~31) * sh_rs, (p.sh_rest ? p.sh_rest : p.sh) + (size_t)wrow0 * sh_rw,
                      __ballot_sync(0xffffffffu, visible), sh_rw, sh_rs);
    }
    if (idx < p.P) {
        if (visible) {
            float3 rgb;
            uint32_t cb = 0;
            if (p.colors_precomp) {
                rgb = make_float3(__ldg(p.colors_precomp + 3 * idx), __ldg(p.colors_precomp + 3 * idx + 1),
                                  __ldg(p.colors_precomp + 3 * idx + 2));
            } else {
                if (STAGED) {
                    // packed rows hold the DC term too; in the dc | rest split it stays a (near-coalesced) global read
                    const float* srow = sh_stage + threadIdx.x * sh_rs;
                    if (p.sh_rest) rgb = sh_to_rgb<true>(p.sh + (size_t)idx * 3, srow, p.D, px, py, pz, p.campos, cb, true);
                    else rgb = sh_to_rgb<true>(srow, srow + 3, p.D, px, py, pz, p.campos, cb);
                } else {
                    const float* sh0 = p.sh_rest ? p.sh + (size_t)idx * 3 : p.sh + (size_t)idx * p.M * 3;
                    const float* shr = p.sh_rest ? p.sh_rest + (size_t)idx * (p.M - 1) * 3 : sh0 + 3;
                    rgb = sh_to_rgb<false>(sh0, shr, p.D, px, py, pz, p.campos, cb);
                }
            }
            float o = opac_raw;
            if (p.act) o = DIV(1.f, ADD(1.f, expf(-o)));  // torch.sigmoid (gaussian_model.py:59,95-97)
            p.geom.rec[2 * idx] = make_float4(mx, my, cnx, cny);
            p.geom.rec[2 * idx + 1] = make_float4(cnz, o, depth, logf(255.f * o));
            p.geom.col[idx] = make_float4(rgb.x, rgb.y, rgb.z, 0.f);
            p.geom.clamped[idx] = (uint8_t)cb;
            p.geom.rect[idx] = make_uint2(x0 | (y0 << 16), x1 | (y1 << 16));
        } else {
            x0 = y0 = x1 = y1 = 0;
            p.geom.rect[idx] = make_uint2(0, 0);
        }
    }
    // Tile histogram: one RED per (Gaussian, tile).  Wide rectangles are spread over the warp.
    const uint32_t w = x1 - x0, n = w * (y1 - y0);
    const bool wide = n > 32;
    uint32_t* __restrict__ count = p.tile.count + (size_t)(blockIdx.x % p.copies) * (p.gx * p.gy);
    if (visible && !wide) {
        for (uint32_t y = y0; y < y1; y++)
            for (uint32_t x = x0; x < x1; x++) atomicAdd(count + y * p.gx + x, 1u);
    }
    unsigned m = __ballot_sync(0xffffffffu, wide);
    while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const uint32_t sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
        const uint32_t sw = __shfl_sync(0xffffffffu, w, src), sn = __shfl_sync(0xffffffffu, n, src);
        for (uint32_t i = lane; i < sn; i += 32) atomicAdd(count + (sy0 + i / sw) * p.gx + sx0 + i % sw, 1u);
    }
}

// reference checkFrustum (rasterizer_impl.cu:54-66)
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ V,
                                    uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float depth = dot3p(__ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2),
                              V[2], V[6], V[10], V[14]);
    present[idx] = !(depth <= 0.2f);
}

int configure_preprocess_fwd() {
    return (int)cudaFuncSetAttribute(preprocess_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sh_stage_bytes(SH_STAGE_MAX_M, 256));
}

void launch_preprocess_fwd(const PreFwdParams& p, cudaStream_t s) {
    if (p.P <= 0) return;
    if (p.sh && !p.colors_precomp && p.M > 1 && p.M <= SH_STAGE_MAX_M)
        launch_chained(preprocess_fwd_kernel<true>, (p.P + 255) / 256, 256, sh_stage_bytes(p.sh_rest ? p.M - 1 : p.M, 256), s, p);
    else
        launch_chained(preprocess_fwd_kernel<false>, (p.P + 255) / 256, 256, 0, s, p);
}
void launch_mark_visible(int P, const float* means3D, const float* V, uint8_t* present, cudaStream_t s) {
    if (P > 0) mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, V, present);
}

}  // namespace wgs
