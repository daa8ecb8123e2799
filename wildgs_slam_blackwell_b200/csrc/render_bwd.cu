// Backward of the alpha compositing.  Replaces reference renderCUDA (backward)
// + render_cuda_reduce_sum (cuda_rasterizer/backward.cu:541-787).
//
// Same decomposition as the forward (CTA = 16x16 tile, warp = 8x4 pixels, cp.async-staged
// batches, per-warp ellipse/rectangle culling), traversed back to front starting at the
// tile's last contributor.  The reference reduces every (tile, Gaussian) pair with a 256-wide
// shared-memory tree (9 block barriers) and a contended shared atomic; here the ten per-pixel
// gradient terms are reduced inside the warp with a 12-shuffle reduce-scatter butterfly, after
// which ten lanes each own one total and issue one RED.ADD.F32 into the 48-byte per-Gaussian
// accumulator record -- no block barrier per pair at all.
#include "common.cuh"
#include "kernels.h"
#include "render_common.cuh"

namespace wgs {

// Reduce-scatter N per-lane values across the 32 lanes of a warp: after the call lane L holds
// in v[0] the warp-wide sum of original value index 5*b4 + 3*b3 + 2*b2 + b1 (b_k = bit k of
// L) -- see owner_index() -- using 5+3+2+1+1 = 12 shuffles instead of 10*5 = 50.
template <int N, int OFFSET>
__device__ __forceinline__ void reduce_scatter(float (&v)[10], int lane) {
    if constexpr (N == 1) {
#pragma unroll
        for (int o = OFFSET; o > 0; o >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
    } else {
        constexpr int M = (N + 1) / 2;
        const bool up = (lane & OFFSET) != 0;
#pragma unroll
        for (int i = 0; i < M; i++) {
            const float lo = v[i];
            const float hi = (i + M < N) ? v[i + M] : 0.f;
            const float send = up ? lo : hi;
            const float keep = up ? hi : lo;
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFFSET);
        }
        reduce_scatter<M, OFFSET / 2>(v, lane);
    }
}

// value index owned by a lane after reduce_scatter<10,16>, or -1
__device__ __forceinline__ int owner_index(int lane) {
    const int b4 = (lane >> 4) & 1, b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    const int s2 = 2 * b2 + b1, s3 = 3 * b3 + s2;
    if ((lane & 1) || s2 >= 3 || s3 >= 5) return -1;
    return 5 * b4 + s3;
}

__global__ void __launch_bounds__(256) render_bwd_kernel(const __grid_constant__ RenderBwdParams p) {
    __shared__ FwdStage st[2];
    __shared__ uint32_t s_hi;
    if (p.tile.misc[MISC_OVERFLOW]) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
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
    const int rounds = (hi + BATCH - 1) / BATCH;

    const float T_final = inside ? p.image.final_T[pix] : 0.f;
    float T = T_final;
    float dpx = 0.f, dpy = 0.f, dpz = 0.f, dpd = 0.f;
    if (inside) {
        dpx = __ldg(p.dL_dcolor + pix); dpy = __ldg(p.dL_dcolor + N + pix); dpz = __ldg(p.dL_dcolor + 2 * N + pix);
        dpd = __ldg(p.dL_ddepth + pix);
    }
    const float bg_dot = __ldg(p.bg) * dpx + __ldg(p.bg + 1) * dpy + __ldg(p.bg + 2) * dpz;
    const float ddelx_dx = 0.5f * p.W, ddely_dy = 0.5f * p.H;
    float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f;
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, ard = 0.f;
    const int own = owner_index(lane);
    float* __restrict__ acc = (float*)p.grad.acc;

    // round r holds list entries k = hi - (r+1)*BATCH + slot, slot in [0, BATCH)
    uint32_t id_next = 0;
    {
        const int k = hi - BATCH + tid;
        if (k >= 0) stage_gather(st[0], tid, (uint32_t)list[k], p.geom);
        cp_async_commit();
        if (k - BATCH >= 0) id_next = (uint32_t)list[k - BATCH];
    }
    for (int r = 0; r < rounds; r++) {
        const int buf = r & 1;
        const int base = hi - (r + 1) * BATCH;
        cp_async_wait<0>();
        __syncthreads();
        if (r + 1 < rounds) {
            const int k = base - BATCH + tid;
            if (k >= 0) stage_gather(st[buf ^ 1], tid, id_next, p.geom);
            if (k - BATCH >= 0) id_next = (uint32_t)list[k - BATCH];
        }
        cp_async_commit();
        if (base >= warp_last) continue;  // whole batch lies behind this warp's last contributor

        const FwdStage& s = st[buf];
        for (int c0 = BATCH - 32; c0 >= 0; c0 -= 32) {
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
                float v[10];
#pragma unroll
                for (int i = 0; i < 10; i++) v[i] = 0.f;
                bool active = false;
                if (k < last_contributor) {
                    const float dx = SUB(a.x, pixx), dy = SUB(a.y, pixy);
                    const float power =
                        FMA(FMA(dx, MUL(dx, a.z), MUL(dy, MUL(dy, b.x))), -0.5f, -MUL(dy, MUL(dx, a.w)));
                    if (!(power > 0.f)) {
                        const float G = expf(power);
                        const float alpha = fminf(MUL(b.y, G), 0.99f);
                        if (!(alpha < 1.0f / 255.0f)) {
                            active = true;
                            const float4 col = s.col[j];
                            const float inv = __frcp_rn(1.f - alpha);
                            T = T * inv;
                            const float dch = alpha * T;
                            const float om = 1.f - last_alpha;
                            ar0 = last_alpha * lc0 + om * ar0; lc0 = col.x;
                            ar1 = last_alpha * lc1 + om * ar1; lc1 = col.y;
                            ar2 = last_alpha * lc2 + om * ar2; lc2 = col.z;
                            ard = last_alpha * ld + om * ard;  ld = b.z;
                            float dL_dalpha = (col.x - ar0) * dpx + (col.y - ar1) * dpy + (col.z - ar2) * dpz +
                                              (b.z - ard) * dpd;
                            dL_dalpha *= T;
                            last_alpha = alpha;
                            dL_dalpha += (-T_final * inv) * bg_dot;
                            const float dL_dG = b.y * dL_dalpha;
                            const float gdx = G * dx, gdy = G * dy;
                            const float dG_ddelx = -gdx * a.z - gdy * a.w;
                            const float dG_ddely = -gdy * b.x - gdx * a.w;
                            v[0] = dL_dG * dG_ddelx * ddelx_dx;
                            v[1] = dL_dG * dG_ddely * ddely_dy;
                            v[2] = -0.5f * gdx * dx * dL_dG;
                            v[3] = -0.5f * gdx * dy * dL_dG;
                            v[4] = -0.5f * gdy * dy * dL_dG;
                            v[5] = G * dL_dalpha;
                            v[6] = dch * dpx;
                            v[7] = dch * dpy;
                            v[8] = dch * dpz;
                            v[9] = dch * dpd;
                        }
                    }
                }
                if (__any_sync(0xffffffffu, active)) {
                    reduce_scatter<10, 16>(v, lane);
                    if (own >= 0) atomicAdd(acc + 12 * (size_t)s.id[j] + own, v[0]);
                }
            }
        }
    }
    cp_async_wait<0>();
}

void launch_render_bwd(const RenderBwdParams& p, cudaStream_t s) {
    render_bwd_kernel<<<p.gx * p.gy, 256, 0, s>>>(p);
}

}  // namespace wgs
