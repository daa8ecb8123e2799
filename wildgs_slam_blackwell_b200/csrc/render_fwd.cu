// Forward alpha compositing.  Replaces reference renderCUDA (cuda_rasterizer/forward.cu:263-392).
//
// One CTA per 16x16 tile (the binning granularity is part of the bit-exact contract), 8 warps,
// each warp owning an 8x4 pixel sub-rectangle.  Per batch of 256 sorted instances the CTA
// gathers the 48-byte Gaussian records into shared memory with cp.async (double buffered, ids
// prefetched one batch further ahead); every warp then culls the batch against ITS OWN 8x4
// rectangle (one Gaussian per lane, exact-conservative ellipse/rectangle test, ballot) and
// only walks the survivors.  The per-pixel arithmetic is the reference's rn-sequence, so
// colour, depth, opacity, final_T, n_contrib and n_touched are unchanged by the culling.
#include "common.cuh"
#include "kernels.h"
#include "render_common.cuh"

namespace wgs {

// The survivors of one 32-entry chunk, compacted by the warp that culled them: each hit lane already holds its
// Gaussian's two records in registers and writes them to slot popc(hits below it); the walk then reads slots
// 0..n-1 with warp-uniform addresses (LDS broadcast, immediate offsets once unrolled) instead of decoding a bit mask
// per entry.  b.w (ln(255 opacity), only needed by the cull) is replaced by the entry's index in the stage.
struct __align__(16) WarpList {
    float4 a[32];
    float4 b[32];
};

// Walk the n compacted survivors front to back for this lane's pixel.  The arithmetic is the reference's
// rn-sequence (forward.cu:336-380).  A pixel that is done (the reference's `done` flag: its transmittance test
// failed once, or it lies outside the image) carries T with the sign bit set -- T > 0 while it is live -- so the loop
// has no separate flag and no divergent early-outs: the three conditions become one predicate around the blend.
template <bool TOUCH>
__device__ __forceinline__ void walk_survivors(const FwdStage& s, const WarpList& wl, int n, int lc_base, float pixx,
                                               float pixy, float& T, float& C0, float& C1, float& C2, float& D,
                                               uint32_t& last_contributor, int lane, int* __restrict__ n_touched) {
#pragma unroll 2
    for (int k = 0; k < n; k++) {
        const float4 a = wl.a[k];
        const float4 b = wl.b[k];
        const float dx = SUB(a.x, pixx), dy = SUB(a.y, pixy);
        const float power = FMA(FMA(dx, MUL(dx, a.z), MUL(dy, MUL(dy, b.x))), -0.5f, -MUL(dy, MUL(dx, a.w)));
        const float alpha = fminf(MUL(b.y, expf(power)), 0.99f);
        const float test_T = MUL(T, SUB(1.f, alpha));
        // forward.cu:345-372: skip if power > 0, skip if alpha < 1/255, stop if T (1 - alpha) < 1e-4.  A done pixel
        // has T < 0, so its test_T is <= 0: it "stops" again (idempotent) and never blends.
        const bool valid = !(power > 0.f) & !(alpha < 1.0f / 255.0f);
        const bool stop = test_T < 0.0001f;
        if (valid & stop) T = -fabsf(T);
        bool touched = false;
        if (valid & !stop) {
            const int j = __float_as_int(b.w);
            const float4 col = s.col[j];
            C0 = FMA(T, MUL(alpha, col.x), C0);
            C1 = FMA(T, MUL(alpha, col.y), C1);
            C2 = FMA(T, MUL(alpha, col.z), C2);
            D = FMA(T, MUL(alpha, b.z), D);
            if (TOUCH) touched = test_T > 0.5f;
            T = test_T;
            last_contributor = (uint32_t)(lc_base + j);
        }
        if (TOUCH) {
            const unsigned tm = __ballot_sync(0xffffffffu, touched);
            if (tm && lane == 0) atomicAdd(n_touched + s.id[__float_as_int(b.w)], __popc(tm));
        }
    }
}

__global__ void __launch_bounds__(256) render_fwd_kernel(const __grid_constant__ RenderFwdParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    __shared__ FwdStage st[2];
    __shared__ WarpList lists[8];
    // R exceeded the instance capacity (only reachable in no-wait / graph mode, where no host is there to grow the
    // buffer and retry): every tile composites an EMPTY list, i.e. background colour, zero depth / opacity and
    // n_contrib = 0, so a missed overflow check never hands uninitialised memory to a loss or an optimiser (the
    // backward of such a forward finds no contributor and produces zero gradients); tile_scan recorded the event
    // in the context's sticky overflow word.
    const bool overflow = p.tile.misc[MISC_OVERFLOW] != 0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // longest tiles first (tile_scan's order); a tile-row strip skips the tiles outside it
    const int tile = (int)p.tile.order[blockIdx.x];
    if (tile < p.tile0 || tile >= p.tile0 + p.ntiles) return;
    const int tx = tile % p.gx, ty = tile / p.gx;
    const int wx0 = tx * WGS_TILE + (warp & 1) * 8, wy0 = ty * WGS_TILE + (warp >> 1) * 4;
    const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
    const bool inside = pxi < p.W && pyi < p.H;
    const float pixx = (float)pxi, pixy = (float)pyi;
    // this warp's pixel rectangle (clipped to the image) for culling
    const float rx0 = (float)wx0, ry0 = (float)wy0;
    const float rx1 = (float)min(wx0 + 7, p.W - 1), ry1 = (float)min(wy0 + 3, p.H - 1);
    const bool warp_in_image = wx0 < p.W && wy0 < p.H;

    const uint2 range = p.tile.ranges[tile];
    const int total = overflow ? 0 : (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;
    const unsigned long long* __restrict__ list = p.bucket + range.x;

    float T = inside ? 1.f : -1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
    uint32_t last_contributor = 0;
    WarpList& wl = lists[warp];
    const unsigned lt_mask = (1u << lane) - 1u;

    // prologue: ids of batch 0 -> issue gathers; prefetch ids of batch 1
    uint32_t id_next = 0;
    if (tid < total) {
        const uint32_t id = (uint32_t)list[tid];
        st[0].id[tid] = id;
        cp_async16(&st[0].rec0[tid], p.geom.rec + 2 * (size_t)id);
        cp_async16(&st[0].rec1[tid], p.geom.rec + 2 * (size_t)id + 1);
        cp_async16(&st[0].col[tid], p.geom.col + id);
    }
    cp_async_commit();
    if (BATCH + tid < total) id_next = (uint32_t)list[BATCH + tid];

    for (int r = 0; r < rounds; r++) {
        const int buf = r & 1;
        const int count = min(BATCH, total - r * BATCH);
        cp_async_wait<0>();
        // all threads done with buffer buf^1 (read in round r-1) and batch r visible
        if (__syncthreads_and(T < 0.f)) break;
        if (r + 1 < rounds) {
            const int k = (r + 1) * BATCH + tid;
            if (k < total) {
                FwdStage& nx = st[buf ^ 1];
                nx.id[tid] = id_next;
                cp_async16(&nx.rec0[tid], p.geom.rec + 2 * (size_t)id_next);
                cp_async16(&nx.rec1[tid], p.geom.rec + 2 * (size_t)id_next + 1);
                cp_async16(&nx.col[tid], p.geom.col + id_next);
            }
            if (k + BATCH < total) id_next = (uint32_t)list[k + BATCH];
        }
        cp_async_commit();
        if (!warp_in_image || __all_sync(0xffffffffu, T < 0.f)) continue;

        const FwdStage& s = st[buf];
        for (int c0 = 0; c0 < count; c0 += 32) {
            const int mine = c0 + lane;
            bool hit = false;
            float4 a, b;
            if (mine < count) {
                a = s.rec0[mine];
                b = s.rec1[mine];
                hit = reaches_rect(a.x, a.y, a.z, a.w, b.x, b.w, rx0, ry0, rx1, ry1);
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m == 0) continue;
            if (hit) {
                const int slot = __popc(m & lt_mask);
                b.w = __int_as_float(mine);
                wl.a[slot] = a;
                wl.b[slot] = b;
            }
            __syncwarp();
            // T only decreases: once no pixel of the warp has T > 0.5, n_touched cannot change any more
            // (forward.cu:369-371 needs test_T > 0.5), and the cheaper loop without its vote is used.
            if (__any_sync(0xffffffffu, T > 0.5f))
                walk_survivors<true>(s, wl, __popc(m), r * BATCH + 1, pixx, pixy, T, C0, C1, C2, D, last_contributor,
                                     lane, p.n_touched);
            else
                walk_survivors<false>(s, wl, __popc(m), r * BATCH + 1, pixx, pixy, T, C0, C1, C2, D, last_contributor,
                                      lane, p.n_touched);
            __syncwarp();      // the list is rewritten by the next chunk
            if (__all_sync(0xffffffffu, T < 0.f)) break;
        }
    }
    cp_async_wait<0>();
    T = fabsf(T);
    if (inside) {
        const size_t N = (size_t)p.W * p.H;
        const size_t pix = (size_t)pyi * p.W + pxi;
        p.image.final_T[pix] = T;
        p.image.n_contrib[pix] = last_contributor;
        p.out_color[pix] = FMA(T, __ldg(p.bg), C0);
        p.out_color[N + pix] = FMA(T, __ldg(p.bg + 1), C1);
        p.out_color[2 * N + pix] = FMA(T, __ldg(p.bg + 2), C2);
        p.out_depth[pix] = D;
        p.out_opacity[pix] = SUB(1.f, T);
    }
}

// Debug counting pass (wgs_count_pairs): one thread per pixel re-walks its tile list up to n_contrib with the
// forward's arithmetic and counts the pairs that pass both alpha tests (= the pairs the forward composited: the
// transmittance cut only ever fires behind the last contributor).  counts[0] += n_contrib, counts[1] += blended.
__global__ void __launch_bounds__(256) count_pairs_kernel(const __grid_constant__ RenderFwdParams p,
                                                          unsigned long long* __restrict__ counts) {
    if (p.tile.misc[MISC_OVERFLOW]) return;
    const int tile = blockIdx.x;
    if (tile < p.tile0 || tile >= p.tile0 + p.ntiles) return;
    const int tx = tile % p.gx, ty = tile / p.gx;
    const int pxi = tx * WGS_TILE + (threadIdx.x & 15), pyi = ty * WGS_TILE + (threadIdx.x >> 4);
    unsigned long long n_eval = 0, n_blend = 0;
    if (pxi < p.W && pyi < p.H) {
        const uint2 range = p.tile.ranges[tile];
        const unsigned long long* __restrict__ list = p.bucket + range.x;
        const int last = (int)p.image.n_contrib[(size_t)pyi * p.W + pxi];
        const float pixx = (float)pxi, pixy = (float)pyi;
        n_eval = (unsigned long long)last;
        for (int k = 0; k < last; k++) {
            const uint32_t id = (uint32_t)list[k];
            const float4 a = p.geom.rec[2 * (size_t)id], b = p.geom.rec[2 * (size_t)id + 1];
            const float dx = SUB(a.x, pixx), dy = SUB(a.y, pixy);
            const float power = FMA(FMA(dx, MUL(dx, a.z), MUL(dy, MUL(dy, b.x))), -0.5f, -MUL(dy, MUL(dx, a.w)));
            if (power > 0.f) continue;
            const float alpha = fminf(MUL(b.y, expf(power)), 0.99f);
            if (alpha < 1.0f / 255.0f) continue;
            n_blend++;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        n_eval += __shfl_xor_sync(0xffffffffu, n_eval, o);
        n_blend += __shfl_xor_sync(0xffffffffu, n_blend, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(counts, n_eval);
        atomicAdd(counts + 1, n_blend);
    }
}

void launch_count_pairs(const RenderFwdParams& p, unsigned long long* counts, cudaStream_t s) {
    cudaMemsetAsync(counts, 0, 2 * sizeof(unsigned long long), s);
    count_pairs_kernel<<<p.gx * p.gy, 256, 0, s>>>(p, counts);
}

void launch_render_fwd(const RenderFwdParams& p, cudaStream_t s) {
    launch_chained(render_fwd_kernel, p.gx * p.gy, 256, 0, s, p);
}

}  // namespace wgs
