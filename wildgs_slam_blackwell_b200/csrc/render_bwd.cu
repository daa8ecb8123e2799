// Backward of the alpha compositing.  Replaces reference renderCUDA (backward)
// + render_cuda_reduce_sum (cuda_rasterizer/backward.cu:541-787).
//
// Same decomposition as the forward (CTA = 16x16 tile, warp = 8x4 pixels, cp.async-staged
// batches, per-warp ellipse/rectangle culling), traversed back to front starting at the
// tile's last contributor.  The kernel is FP32-issue bound, so the design minimises
// instructions per (warp, Gaussian) pair:
//  * the four per-channel "accum_rec" recurrences of the reference (backward.cu:709-722)
//    collapse into ONE scalar recurrence on x = c.dL_dpixel + depth*dL_ddepth (they are linear);
//  * every per-Gaussian gradient of the reference (backward.cu:724-757) is a sum over pixels of
//    one of TWO per-pixel scalars times a factor that does not depend on the Gaussian:
//        w = G * dL/dalpha      -> dL/dopacity = S[w],   and with d = mean2D - pixel
//                                  dL/dmean2D, dL/dconic  from the moments S[w d_x], S[w d_y],
//                                  S[w d_x^2], S[w d_x d_y], S[w d_y^2]   (conic, opacity and the
//                                  constants 0.5*W, 0.5*H, -0.5 are applied once per Gaussian in
//                                  preprocess_bwd)
//        c = alpha * T          -> dL/dcolor = S[c dL_dpixel.rgb], dL/ddepth = S[c dL_ddepth]
//    and d = (mean2D - warp origin) - (lane_x, lane_y), so inside a warp the ten sums are the
//    product  [16 pending Gaussians x 32 lanes] * [32 lanes x 10]  with a matrix that is CONSTANT
//    for the warp (1, u, v, u^2, uv, v^2 of the lane coordinates; the lane's pixel cotangents).
//    The reference's 256-wide shared-memory tree + 9 block barriers per list entry therefore
//    becomes: 2 STS per (lane, Gaussian), and per 16 Gaussians 8 ldmatrix + 16 mma.sync
//    (m16n8k8 TF32, operands split hi/lo so the products carry ~21 mantissa bits, fp32
//    accumulate), a 6-shuffle conversion from lane moments to d-moments and 4 vector REDs.
//    This is the one place of the path where a dense contraction exists; it is tiny (1 mma per
//    pair), the point is to take ~40 instructions per pair off the FP32/LSU issue slots.
// ex2.approx / rcp.approx are used (tolerance on gradients is rtol 1e-3; the forward, which is
// bit-exact, is untouched).
#include "common.cuh"
#include "kernels.h"
#include "render_common.cuh"

namespace wgs {

constexpr int BB = 128;  // list entries staged per batch
constexpr int MG = 16;   // pending Gaussians per tensor-core reduction (the M of m16n8k8)
constexpr int VP = 36;   // floats per pending row (32 lanes + pad: 144-B pitch keeps ldmatrix conflict-free)

struct __align__(16) BwdStage {
    float4 rec0[BB];  // x, y, conic.x, conic.y
    float4 rec1[BB];  // conic.z, opacity, depth, ln(255*opacity)
    float4 col[BB];   // r, g, b, -
    uint32_t id[BB];
};

// per-warp staging of the pending (Gaussian, lane) scalars
struct __align__(16) WarpRed {
    float w[MG][VP];   // G * dL/dalpha
    float c[MG][VP];   // alpha * T
    float4 cst[MG];    // {gaussian id (bits), mean2D.x - warp x0, mean2D.y - warp y0, -}
};

struct __align__(16) BwdSmem {
    BwdStage st[2];
    WarpRed red[8];
    float4 list[8][32];  // per warp: the survivors of the current chunk {gaussian id (bits), mean2D - warp origin, slot (bits)}
    uint32_t hi;
};
// four CTAs per SM (the register budget at 64 registers x 256 threads): 228 KB of shared memory, 1 KB reserved per CTA
static_assert(sizeof(BwdSmem) + 1024 <= 233472 / 4, "BwdSmem no longer fits four CTAs per SM");

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void ldsm4(unsigned addr, unsigned (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void red_add_v2(float* addr, float x, float y) {
    asm volatile("red.relaxed.gpu.global.add.v2.f32 [%0], {%1, %2};\n" ::"l"(addr), "f"(x), "f"(y) : "memory");
}
// tf32 keeps the upper 19 bits; hi + lo reproduces v to ~21 mantissa bits
__device__ __forceinline__ void split_tf32(const unsigned (&v)[4], unsigned (&h)[4], unsigned (&l)[4]) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        h[i] = v[i] & 0xffffe000u;
        l[i] = __float_as_uint(__uint_as_float(v[i]) - __uint_as_float(h[i]));
    }
}

// threads 0..127 copy rec0/rec1 of slot tid, threads 128..255 copy col + id of slot tid-128
__device__ __forceinline__ void bwd_gather(BwdStage& st, int tid, uint32_t id, const GeomView& g) {
    const int slot = tid & (BB - 1);
    if (tid < BB) {
        cp_async16(&st.rec0[slot], g.rec + 2 * (size_t)id);
        cp_async16(&st.rec1[slot], g.rec + 2 * (size_t)id + 1);
    } else {
        cp_async16(&st.col[slot], g.col + id);
        st.id[slot] = id;
    }
}

// Reduce the n (<= MG) pending Gaussians of this warp over its 32 lanes on the tensor pipe and add
// the ten sums of each to its accumulator record:
//   [S w, S w dx, S w dy, S w dx^2 | S w dx dy, S w dy^2, dcolor.r, dcolor.g | dcolor.b, ddepth, -, -]
// bw: B fragments of the lane-moment matrix (columns 1,u,v,u^2,uv,v^2,0,0); bc: B fragments of the
// cotangent matrix (columns dp.r,dp.g,dp.b,dp.depth hi parts, then their four lo parts).
__device__ __forceinline__ void flush_pending(const WarpRed& wr, unsigned lm_addr, int n, int lane,
                                              const unsigned (&bw)[4][2], const unsigned (&bc)[4][2],
                                              float* __restrict__ acc) {
    __syncwarp();
    float mw[4] = {0.f, 0.f, 0.f, 0.f}, mc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < 4; s++) {  // k-step s = pixel row s of the warp's 8x4 rectangle
        unsigned a[4], h[4], l[4];
        ldsm4(lm_addr + 32 * s, a);
        split_tf32(a, h, l);
        mma_tf32(mw, h, bw[s][0], bw[s][1]);
        mma_tf32(mw, l, bw[s][0], bw[s][1]);
        ldsm4(lm_addr + MG * VP * 4 + 32 * s, a);
        split_tf32(a, h, l);
        mma_tf32(mc, h, bc[s][0], bc[s][1]);
        mma_tf32(mc, l, bc[s][0], bc[s][1]);
    }
    // D fragment: lane (q = lane/4, t = lane%4) holds columns 2t, 2t+1 of rows q (regs 0,1) and q+8 (regs 2,3)
    const int q = lane >> 2, t = lane & 3, qb = lane & ~3;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const int row = q + 8 * half;
        const float c0 = mw[2 * half], c1 = mw[2 * half + 1];
        const float S0 = __shfl_sync(0xffffffffu, c0, qb), Su = __shfl_sync(0xffffffffu, c1, qb);
        const float Sv = __shfl_sync(0xffffffffu, c0, qb + 1), Suu = __shfl_sync(0xffffffffu, c1, qb + 1);
        const float Suv = __shfl_sync(0xffffffffu, c0, qb + 2), Svv = __shfl_sync(0xffffffffu, c1, qb + 2);
        // cotangent sums: hi columns (t = 0,1) + lo columns (t = 2,3)
        const float d0 = mc[2 * half] + __shfl_down_sync(0xffffffffu, mc[2 * half], 2);
        const float d1 = mc[2 * half + 1] + __shfl_down_sync(0xffffffffu, mc[2 * half + 1], 2);
        const float4 k = wr.cst[row];
        // d = (e, f) - (u, v): lane moments -> moments of d
        const float e = k.y, f = k.z;
        const float Sx = fmaf(e, S0, -Su), Sy = fmaf(f, S0, -Sv);
        const float Sxx = fmaf(e, Sx - Su, Suu);
        const float Sxy = fmaf(f, Sx, fmaf(-e, Sv, Suv));
        const float Syy = fmaf(f, Sy - Sv, Svv);
        const float o0 = t == 0 ? S0 : (t == 1 ? Sy : Sxy);
        const float o1 = t == 0 ? Sx : (t == 1 ? Sxx : Syy);
        float* rec = acc + 12 * (size_t)__float_as_uint(k.x);
        if (row < n) {
            if (t < 3) red_add_v2(rec + 2 * t, o0, o1);
            if (t < 2) red_add_v2(rec + 6 + 2 * t, d0, d1);
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(256, 4) render_bwd_kernel(const __grid_constant__ RenderBwdParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);
    if (p.tile.misc[MISC_OVERFLOW]) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // longest tiles first (tile_scan's order); a tile-row strip skips the tiles outside it
    const int tile = (int)p.tile.order[blockIdx.x];
    if (tile < p.tile0 || tile >= p.tile0 + p.ntiles) return;
    const int tx = tile % p.gx, ty = tile / p.gx;
    const int wx0 = tx * WGS_TILE + (warp & 1) * 8, wy0 = ty * WGS_TILE + (warp >> 1) * 4;
    const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
    const bool inside = pxi < p.W && pyi < p.H;
    const float pixx = (float)pxi, pixy = (float)pyi;
    const float rx0 = (float)wx0, ry0 = (float)wy0;
    const float rx1 = (float)min(wx0 + 7, p.W - 1), ry1 = (float)min(wy0 + 3, p.H - 1);
    const size_t N = (size_t)p.W * p.H;
    const size_t pix = (size_t)pyi * p.W + pxi;

    const uint2 range = p.tile.ranges[tile];
    const unsigned long long* __restrict__ list = p.bucket + range.x;

    const int last_contributor = inside ? (int)p.image.n_contrib[pix] : 0;
    const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    if (tid == 0) sm.hi = 0;
    __syncthreads();
    if (lane == 0 && warp_last) atomicMax(&sm.hi, (uint32_t)warp_last);
    __syncthreads();
    const int hi = (int)sm.hi;  // entries [hi, total) contributed to no pixel of this tile
    if (hi == 0) return;
    const int rounds = (hi + BB - 1) / BB;

    const float T_final = inside ? p.image.final_T[pix] : 0.f;
    float T = T_final;
    float dpx = 0.f, dpy = 0.f, dpz = 0.f, dpd = 0.f;
    if (inside) {
        dpx = __ldg(p.dL_dcolor + pix); dpy = __ldg(p.dL_dcolor + N + pix); dpz = __ldg(p.dL_dcolor + 2 * N + pix);
        dpd = __ldg(p.dL_ddepth + pix);
    }
    // T_final * (bg . dL_dpixel): the background term of dL_dalpha (backward.cu:733-738)
    const float tfbg = T_final * (__ldg(p.bg) * dpx + __ldg(p.bg + 1) * dpy + __ldg(p.bg + 2) * dpz);
    float last_alpha = 0.f, last_x = 0.f, A = 0.f;
    float* __restrict__ acc = (float*)p.grad.acc;

    WarpRed& wr = sm.red[warp];
    float4* __restrict__ wl = sm.list[warp];
    // ---- constant B fragments of this warp (b0: k = t, b1: k = t + 4 of k-step s; column n = q) ----
    unsigned bw[4][2], bc[4][2];
    {
        const int q = lane >> 2, t = lane & 3;
        // cotangents of all 32 lanes, hi/lo split, as an 8 x 32 table in the (still unused) w rows
        {
            const float v4[4] = {dpx, dpy, dpz, dpd};
#pragma unroll
            for (int n = 0; n < 4; n++) {
                const float h = __uint_as_float(__float_as_uint(v4[n]) & 0xffffe000u);
                wr.w[n][lane] = h;
                wr.w[n + 4][lane] = v4[n] - h;
            }
        }
        __syncwarp();
#pragma unroll
        for (int s = 0; s < 4; s++) {
#pragma unroll
            for (int b = 0; b < 2; b++) {
                const int u = t + 4 * b, v = s;  // lane coordinates of pixel lane j = 8 s + u
                const float m = q == 0 ? 1.f : q == 1 ? (float)u : q == 2 ? (float)v : q == 3 ? (float)(u * u)
                              : q == 4 ? (float)(u * v) : q == 5 ? (float)(v * v) : 0.f;
                bw[s][b] = __float_as_uint(m);
                bc[s][b] = __float_as_uint(wr.w[q][8 * s + u]);
            }
        }
        __syncwarp();
    }
    // ldmatrix row address of this lane: matrix i = lane/8 -> rows (i&1)*8 + lane%8, columns 4*(i>>1)
    const unsigned lm_addr = (unsigned)__cvta_generic_to_shared(&wr.w[((lane >> 3) & 1) * 8 + (lane & 7)][4 * (lane >> 4)]);
    unsigned wcol, wcst;  // shared addresses of this lane's column in the w plane / of cst[0] (32-bit registers)
    asm volatile("mov.u32 %0, %1;" : "=r"(wcol) : "r"((unsigned)__cvta_generic_to_shared(&wr.w[0][lane])));
    asm volatile("mov.u32 %0, %1;" : "=r"(wcst) : "r"((unsigned)__cvta_generic_to_shared(&wr.cst[0])));
    unsigned ca = wcol, cs = wcst;  // where the next pending row goes

    // round r holds list entries k = hi - (r+1)*BB + slot, slot in [0, BB)
    const int slot = tid & (BB - 1);
    uint32_t id_next = 0;
    {
        const int k = hi - BB + slot;
        if (k >= 0) bwd_gather(sm.st[0], tid, (uint32_t)list[k], p.geom);
        cp_async_commit();
        if (k - BB >= 0) id_next = (uint32_t)list[k - BB];
    }
    for (int r = 0; r < rounds; r++) {
        const int buf = r & 1;
        const int base = hi - (r + 1) * BB;
        cp_async_wait<0>();
        __syncthreads();
        if (r + 1 < rounds) {
            const int k = base - BB + slot;
            if (k >= 0) bwd_gather(sm.st[buf ^ 1], tid, id_next, p.geom);
            if (k - BB >= 0) id_next = (uint32_t)list[k - BB];
        }
        cp_async_commit();
        if (base >= warp_last) continue;  // whole batch lies behind this warp's last contributor

        const BwdStage& s = sm.st[buf];
        for (int c0 = BB - 32; c0 >= 0; c0 -= 32) {
            const int mine = c0 + lane;
            const int kmine = base + mine;
            bool hit = false;
            float ex = 0.f, ey = 0.f;
            if (kmine >= 0 && kmine < warp_last) {
                const float4 a = s.rec0[mine];
                const float4 b = s.rec1[mine];
                hit = reaches_rect(a.x, a.y, a.z, a.w, b.x, b.w, rx0, ry0, rx1, ry1);
                ex = a.x - rx0;
                ey = a.y - ry0;
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m == 0) continue;
            // compact the survivors, last list entry first (back to front): the walk below reads slots 0..n-1 at
            // warp-uniform addresses instead of decoding the bit mask once per entry
            if (hit) wl[__popc(m & (0xfffffffeu << lane))] = make_float4(__uint_as_float(s.id[mine]), ex, ey, __int_as_float(mine));
            __syncwarp();
            const int n = __popc(m);
#pragma unroll 2
            for (int q = 0; q < n; q++) {
                const float4 e = wl[q];
                const int j = __float_as_int(e.w);
                const float4 a = s.rec0[j];
                const float4 b = s.rec1[j];
                const float dx = a.x - pixx, dy = a.y - pixy;
                // `power` is the forward's expression bit for bit, so the power > 0 decision agrees with it.  The
                // alpha >= 1/255 decision uses ex2.approx here and expf there (relative difference <= 6e-7): a pair
                // within that band of the threshold (about one in 1e6 blended pairs) can be taken on one side only,
                // which scales T for the entries in front of it at that one pixel by 1/(1 - 1/255) -- 0.4 % of one
                // pixel's share of those Gaussians' gradients, inside the 1e-3 gradient tolerance; deciding with the
                // forward's expf would add 8 instructions to every (warp, entry) pair.
                const float power = FMA(FMA(dx, MUL(dx, a.z), MUL(dy, MUL(dy, b.x))), -0.5f, -MUL(dy, MUL(dx, a.w)));
                const float G = ex2_approx(power * 1.4426950408889634f);
                const float alpha = fminf(b.y * G, 0.99f);
                // backward.cu:685-706: behind this pixel's last contributor, power > 0, alpha < 1/255 -> no contribution
                const bool active = (base + j < last_contributor) & !(power > 0.f) & !(alpha < 1.0f / 255.0f);
                float w = 0.f, c = 0.f;
                if (active) {
                    const float4 col = s.col[j];
                    const float inv = rcp_approx(1.f - alpha);
                    T *= inv;
                    const float x = fmaf(col.x, dpx, fmaf(col.y, dpy, fmaf(col.z, dpz, b.z * dpd)));
                    A = fmaf(last_alpha, last_x - A, A);
                    last_x = x;
                    last_alpha = alpha;
                    const float dLa = fmaf(-tfbg, inv, (x - A) * T);  // dL/dalpha
                    c = alpha * T;
                    w = G * dLa;
                }
                if (__any_sync(0xffffffffu, active)) {
                    // row `pending` of the w / c planes and of cst (every lane stores the same cst record: one
                    // address, no lane predicate to keep)
                    asm volatile("st.shared.f32 [%0], %1;\n\tst.shared.f32 [%0+%3], %2;\n" ::"r"(ca), "f"(w), "f"(c),
                                 "n"(MG * VP * 4) : "memory");
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %3};\n" ::"r"(cs), "f"(e.x), "f"(e.y), "f"(e.z)
                                 : "memory");
                    ca += VP * 4;
                    cs += 16;
                    if (cs == wcst + MG * 16) {
                        flush_pending(wr, lm_addr, MG, lane, bw, bc, acc);
                        ca = wcol;
                        cs = wcst;
                    }
                }
            }
            __syncwarp();  // the list is rewritten by the next chunk
        }
    }
    if (cs != wcst) flush_pending(wr, lm_addr, (int)(cs - wcst) >> 4, lane, bw, bc, acc);
    cp_async_wait<0>();
}

void launch_render_bwd(const RenderBwdParams& p, cudaStream_t s) {
    launch_chained(render_bwd_kernel, p.gx * p.gy, 256, sizeof(BwdSmem), s, p);
}

int configure_render_bwd() {
    return (int)cudaFuncSetAttribute(render_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(BwdSmem));
}

}  // namespace wgs
