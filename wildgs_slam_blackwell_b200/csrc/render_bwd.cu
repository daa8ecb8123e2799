// Backward of the alpha compositing.  Replaces reference renderCUDA (backward)
// + render_cuda_reduce_sum (cuda_rasterizer/backward.cu:541-787).
//
// Same decomposition as the forward (CTA = 16x16 tile, warp = 8x4 pixels, cp.async-staged
// batches, per-warp ellipse/rectangle culling), traversed back to front starting at the
// tile's last contributor.  The kernel is FP32-issue bound (ncu: 77 % issue utilisation), so
// the design minimises instructions per (warp, Gaussian) pair:
//  * the four per-channel "accum_rec" recurrences of the reference (backward.cu:709-722)
//    collapse into ONE scalar recurrence on x = c.dL_dpixel + depth*dL_ddepth (they are linear);
//  * constant factors (0.5*W, 0.5*H, -0.5) are applied once per Gaussian in preprocess_bwd,
//    not once per pixel here;
//  * the reference's 256-wide shared-memory tree + 9 block barriers per list entry becomes a
//    warp-private shared-memory transpose: every lane stores its 10 terms, and once GROUP
//    Gaussians are pending, 10*GROUP lanes each sum one 32-float row (8 LDS.128, conflict-free
//    at a 36-float row pitch) and issue one RED.ADD.F32 into the 48-byte accumulator record.
//    No block barrier per pair at all.
// ex2.approx / rcp.approx are used (tolerance on gradients is rtol 1e-3; the forward, which is
// bit-exact, is untouched).
#include "common.cuh"
#include "kernels.h"
#include "render_common.cuh"

namespace wgs {

constexpr int BB = 128;    // list entries staged per batch
constexpr int GROUP = 3;   // Gaussians per shared-memory reduction
constexpr int ROW = 36;    // floats per reduction row (32 + pad: 144 B pitch, 16-B aligned, conflict-free)

struct __align__(16) BwdStage {
    float4 rec0[BB];  // x, y, conic.x, conic.y
    float4 rec1[BB];  // conic.z, opacity, depth, ln(255*opacity)
    float4 col[BB];   // r, g, b, -
    uint32_t id[BB];
};

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

// Sum the pending rows of this warp and add them to the per-Gaussian accumulators.
__device__ __forceinline__ void flush_group(const float* my_row, uint32_t rid, int n, int lane, int sub,
                                            float* __restrict__ acc) {
    __syncwarp();
    if (lane < 10 * n) {
        const float4* r4 = reinterpret_cast<const float4*>(my_row);
        float4 s0 = r4[0], s1 = r4[1], s2 = r4[2], s3 = r4[3];
        const float4 t0 = r4[4], t1 = r4[5], t2 = r4[6], t3 = r4[7];
        s0.x += t0.x; s0.y += t0.y; s0.z += t0.z; s0.w += t0.w;
        s1.x += t1.x; s1.y += t1.y; s1.z += t1.z; s1.w += t1.w;
        s2.x += t2.x; s2.y += t2.y; s2.z += t2.z; s2.w += t2.w;
        s3.x += t3.x; s3.y += t3.y; s3.z += t3.z; s3.w += t3.w;
        s0.x += s1.x; s0.y += s1.y; s0.z += s1.z; s0.w += s1.w;
        s2.x += s3.x; s2.y += s3.y; s2.z += s3.z; s2.w += s3.w;
        s0.x += s2.x; s0.y += s2.y; s0.z += s2.z; s0.w += s2.w;
        const float sum = (s0.x + s0.y) + (s0.z + s0.w);
        atomicAdd(acc + 12 * (size_t)rid + sub, sum);
    }
    __syncwarp();
}

__global__ void __launch_bounds__(256, 4) render_bwd_kernel(const __grid_constant__ RenderBwdParams p) {
    __shared__ BwdStage st[2];
    __shared__ __align__(16) float red[8][10 * GROUP][ROW];
    __shared__ uint32_t s_hi;
    if (p.tile.misc[MISC_OVERFLOW]) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = p.tile0 + blockIdx.x;
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
    if (tid == 0) s_hi = 0;
    __syncthreads();
    if (lane == 0 && warp_last) atomicMax(&s_hi, (uint32_t)warp_last);
    __syncthreads();
    const int hi = (int)s_hi;  // entries [hi, total) contributed to no pixel of this tile
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
    // shared-window address of this lane's column (written per Gaussian); kept as a 32-bit register
    unsigned wcol;
    asm volatile("mov.u32 %0, %1;" : "=r"(wcol) : "r"((unsigned)__cvta_generic_to_shared(&red[warp][0][lane])));
    const float* const wrow = &red[warp][lane < 10 * GROUP ? lane : 0][0];  // this lane's row (summed at flush)
    const int my_g = lane / 10, my_sub = lane - 10 * my_g;  // which pending Gaussian / term this lane flushes
    uint32_t my_id = 0;                              // id of pending Gaussian my_g
    int pending = 0;

    // round r holds list entries k = hi - (r+1)*BB + slot, slot in [0, BB)
    const int slot = tid & (BB - 1);
    uint32_t id_next = 0;
    {
        const int k = hi - BB + slot;
        if (k >= 0) bwd_gather(st[0], tid, (uint32_t)list[k], p.geom);
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
            if (k >= 0) bwd_gather(st[buf ^ 1], tid, id_next, p.geom);
            if (k - BB >= 0) id_next = (uint32_t)list[k - BB];
        }
        cp_async_commit();
        if (base >= warp_last) continue;  // whole batch lies behind this warp's last contributor

        const BwdStage& s = st[buf];
        for (int c0 = BB - 32; c0 >= 0; c0 -= 32) {
            const int mine = c0 + lane;
            const int kmine = base + mine;
            bool hit = false;
            if (kmine >= 0 && kmine < warp_last) {
                const float4 a = s.rec0[mine];
                const float4 b = s.rec1[mine];
                hit = reaches_rect(a.x, a.y, a.z, a.w, b.x, b.w, rx0, ry0, rx1, ry1);
            }
            unsigned m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int bit = 31 - __clz(m);
                m &= ~(1u << bit);
                const int j = c0 + bit;
                const int k = base + j;
                const float4 a = s.rec0[j];
                const float4 b = s.rec1[j];
                float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f, v9 = 0.f;
                bool active = false;
                if (k < last_contributor) {
                    const float dx = a.x - pixx, dy = a.y - pixy;
                    // same expression as the forward so that skip decisions agree
                    const float power = FMA(FMA(dx, MUL(dx, a.z), MUL(dy, MUL(dy, b.x))), -0.5f, -MUL(dy, MUL(dx, a.w)));
                    if (!(power > 0.f)) {
                        const float G = ex2_approx(power * 1.4426950408889634f);
                        const float alpha = fminf(b.y * G, 0.99f);
                        if (!(alpha < 1.0f / 255.0f)) {
                            active = true;
                            const float4 col = s.col[j];
                            const float inv = rcp_approx(1.f - alpha);
                            T *= inv;
                            const float x = fmaf(col.x, dpx, fmaf(col.y, dpy, fmaf(col.z, dpz, b.z * dpd)));
                            A = fmaf(last_alpha, last_x - A, A);
                            last_x = x;
                            last_alpha = alpha;
                            const float dLa = fmaf(-tfbg, inv, (x - A) * T);  // dL/dalpha
                            const float dch = alpha * T;
                            const float t = b.y * dLa;                       // dL/dG
                            const float aa = G * dx * t, bb = G * dy * t;
                            v0 = fmaf(aa, a.z, bb * a.w);   // -> dL/dmean2D.x after * -(W/2)
                            v1 = fmaf(bb, b.x, aa * a.w);   // -> dL/dmean2D.y after * -(H/2)
                            v2 = aa * dx;                   // -> dL/dconic.x after * -1/2
                            v3 = aa * dy;                   // -> dL/dconic.y after * -1/2
                            v4 = bb * dy;                   // -> dL/dconic.w after * -1/2
                            v5 = G * dLa;                   // dL/dopacity
                            v6 = dch * dpx; v7 = dch * dpy; v8 = dch * dpz;  // dL/dcolor
                            v9 = dch * dpd;                 // dL/ddepth
                        }
                    }
                }
                if (__any_sync(0xffffffffu, active)) {
                    const unsigned c = wcol + pending * (10 * ROW * 4);
                    asm volatile(
                        "st.shared.f32 [%0], %1;\n\tst.shared.f32 [%0+144], %2;\n\tst.shared.f32 [%0+288], %3;\n\t"
                        "st.shared.f32 [%0+432], %4;\n\tst.shared.f32 [%0+576], %5;\n\tst.shared.f32 [%0+720], %6;\n\t"
                        "st.shared.f32 [%0+864], %7;\n\tst.shared.f32 [%0+1008], %8;\n\tst.shared.f32 [%0+1152], %9;\n\t"
                        "st.shared.f32 [%0+1296], %10;\n" ::"r"(c), "f"(v0), "f"(v1), "f"(v2), "f"(v3), "f"(v4), "f"(v5),
                        "f"(v6), "f"(v7), "f"(v8), "f"(v9) : "memory");
                    const uint32_t gid = s.id[j];
                    if (pending == my_g) my_id = gid;
                    if (++pending == GROUP) {
                        flush_group(wrow, my_id, GROUP, lane, my_sub, acc);
                        pending = 0;
                    }
                }
            }
        }
    }
    if (pending) flush_group(wrow, my_id, pending, lane, my_sub, acc);
    cp_async_wait<0>();
}

void launch_render_bwd(const RenderBwdParams& p, cudaStream_t s) {
    render_bwd_kernel<<<p.ntiles, 256, 0, s>>>(p);
}

}  // namespace wgs
