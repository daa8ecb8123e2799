// Binning: (tile | depth) ordering of the (Gaussian, tile) instances.
//
// Replaces the reference's InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs
// over 64-bit tile<<32|depth keys + identifyTileRanges (rasterizer_impl.cu:70-138, 280-321)
// with an MSD-first formulation sized for B200's shared memory:
//   1. the tile digit is histogrammed by preprocess_fwd (one RED per instance),
//   2. tile_scan_kernel turns the histogram into per-tile [start,end) ranges (this IS
//      identifyTileRanges) and publishes R = num_rendered,
//   3. scatter_kernel drops depth_bits<<32|gaussian_id into its tile's segment,
//   4. each tile segment is sorted in shared memory (up to 24576 entries, 192 KB) by one CTA.
// Because gaussian ids are unique inside a tile, ordering by (depth_bits, id) is a total order
// and equals the order the reference's STABLE radix sort produces (ties keep emission order =
// ascending id), so sorted ids, keys and ranges are bit-identical.
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"

namespace cg = cooperative_groups;

namespace wgs {

constexpr int SMALL_MAX = 4096;    // 32 KB of keys, 512-thread CTA, 4 CTAs/SM
constexpr int LARGE_MAX = 24576;   // 192 KB of keys, 1024-thread CTA, 1 CTA/SM
constexpr int SMALL_THREADS = 512;
constexpr int LARGE_THREADS = 1024;

// R = num_rendered of this forward: to the tile workspace, to the mapped host word the blocking forward waits on,
// and -- when it exceeds the instance capacity -- to the context's STICKY overflow record, which only
// wgs_overflow_status() clears (a no-wait / graph-replayed forward has nobody looking at its own R).
__device__ __forceinline__ void publish_total(const BinParams& p, uint32_t total) {
    const bool over = (long long)total > p.capacity;
    p.tile.misc[MISC_NUM_RENDERED] = total;
    p.tile.misc[MISC_OVERFLOW] = over;
    if (p.host_R) *(volatile unsigned int*)p.host_R = total;
    if (over && p.sticky) {
        atomicAdd(p.sticky, 1u);
        atomicMax(p.sticky + 1, total);
    }
}

// K = number of privatised histogram copies (compile-time so that all K loads are in flight at once)
template <int K>
__global__ void __launch_bounds__(1024) tile_scan_kernel(const __grid_constant__ BinParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    __shared__ uint32_t warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t carry = 0;
    for (int base = 0; base < p.tiles; base += 1024) {
        const int t = base + tid;
        uint32_t c = 0, ck[K];
#pragma unroll
        for (int k = 0; k < K; k++) ck[k] = t < p.tiles ? __ldcg(p.tile.count + (size_t)k * p.tiles + t) : 0u;
#pragma unroll
        for (int k = 0; k < K; k++) c += ck[k];
        uint32_t v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) warp_sums[warp] = v;
        __syncthreads();
        if (warp == 0) {
            uint32_t s = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += u;
            }
            warp_sums[lane] = s;
        }
        __syncthreads();
        const uint32_t excl = carry + (warp ? warp_sums[warp - 1] : 0u) + v - c;
        if (t < p.tiles) {
            p.tile.ranges[t] = make_uint2(excl, excl + c);
            p.tile.order[t] = (uint32_t)t;  // no load-balancing order beyond 8192 tiles (grid >> slots anyway)
            uint32_t run = excl;  // copy k of this tile scatters into [run_k, run_k + count_k)
#pragma unroll
            for (int k = 0; k < K; k++) {
                p.tile.cursor[(size_t)k * p.tiles + t] = run;
                run += ck[k];
            }
            if (c > (uint32_t)SMALL_MAX) p.tile.large_list[atomicAdd(p.tile.misc + MISC_N_LARGE, 1u)] = t;
        }
        carry += warp_sums[31];
        __syncthreads();
    }
    if (tid == 0) publish_total(p, carry);
}

// Same scan on one 8-CTA thread-block cluster (tiles <= 8 * 1024): every CTA sums the K copies of
// its contiguous slice of tiles (one tile per thread), scans the slice, publishes the slice total
// in its shared memory, and after one cluster barrier reads the totals of the lower-ranked CTAs
// through distributed shared memory.  Replaces ~12 us of single-SM latency by ~5 us.
//
// The same kernel emits `order`: the tiles in descending order of their instance count
// (a cluster-wide counting sort over ORDER_CLASSES classes of 16 instances, class histograms
// exchanged through distributed shared memory as well).  The compositing kernels take tile
// order[blockIdx.x], i.e. the hardware block scheduler hands out the longest tiles first
// (longest-processing-time-first), which shortens the tail of the 1.35-wave grids.
constexpr int SCAN_CLUSTER = 8;
constexpr int ORDER_CLASSES = 256;
template <int K>
__global__ void __cluster_dims__(SCAN_CLUSTER, 1, 1) __launch_bounds__(1024)
    tile_scan_cluster_kernel(const __grid_constant__ BinParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t s_total;
    __shared__ uint32_t cls_hist[ORDER_CLASSES];  // tiles of this CTA per class
    __shared__ uint32_t cls_off[ORDER_CLASSES];   // first `order` slot of (class, this CTA)
    __shared__ uint32_t cls_wsum[ORDER_CLASSES / 32];
    cg::cluster_group cluster = cg::this_cluster();
    const int crank = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk = (p.tiles + SCAN_CLUSTER - 1) / SCAN_CLUSTER;  // <= 1024
    const int t = crank * chunk + tid;
    const bool valid = tid < chunk && t < p.tiles;
    uint32_t c = 0, ck[K];
#pragma unroll
    for (int k = 0; k < K; k++) ck[k] = valid ? __ldcg(p.tile.count + (size_t)k * p.tiles + t) : 0u;
#pragma unroll
    for (int k = 0; k < K; k++) c += ck[k];
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
    }
    if (lane == 31) warp_sums[warp] = v;
    if (tid < ORDER_CLASSES) cls_hist[tid] = 0;
    __syncthreads();
    if (warp == 0) {
        uint32_t s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += u;
        }
        warp_sums[lane] = s;
        if (lane == 31) s_total = s;
    }
    // class 0 = longest tiles
    const uint32_t cls = (uint32_t)(ORDER_CLASSES - 1) - min(c >> 4, (uint32_t)(ORDER_CLASSES - 1));
    uint32_t cls_rank = 0;
    if (valid) cls_rank = atomicAdd(&cls_hist[cls], 1u);
    cluster.sync();  // also a CTA barrier: warp_sums / s_total / cls_hist of every CTA are visible
    uint32_t base = 0, total = 0;
#pragma unroll
    for (int r = 0; r < SCAN_CLUSTER; r++) {
        const uint32_t tr = *cluster.map_shared_rank(&s_total, r);
        if (r < crank) base += tr;
        total += tr;
    }
    const uint32_t excl = base + (warp ? warp_sums[warp - 1] : 0u) + v - c;
    if (valid) {
        p.tile.ranges[t] = make_uint2(excl, excl + c);
        uint32_t run = excl;
#pragma unroll
        for (int k = 0; k < K; k++) {
            p.tile.cursor[(size_t)k * p.tiles + t] = run;
            run += ck[k];
        }
        if (c > (uint32_t)SMALL_MAX) p.tile.large_list[atomicAdd(p.tile.misc + MISC_N_LARGE, 1u)] = t;
    }
    // counting sort of the tiles by class: slot = (#tiles of lower classes, all CTAs)
    //   + (#tiles of this class in lower-ranked CTAs) + rank inside (class, CTA)
    if (tid < ORDER_CLASSES) {
        uint32_t tot = 0, below = 0;
#pragma unroll
        for (int r = 0; r < SCAN_CLUSTER; r++) {
            const uint32_t h = cluster.map_shared_rank(cls_hist, r)[tid];
            if (r < crank) below += h;
            tot += h;
        }
        uint32_t s = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += u;
        }
        if (lane == 31) cls_wsum[warp] = s;
        cls_off[tid] = s - tot + below;  // + totals of the lower warps, added below
    }
    __syncthreads();
    if (valid) {
        uint32_t off = cls_off[cls] + cls_rank;
        const int cw = (int)(cls >> 5);
#pragma unroll
        for (int w = 0; w < ORDER_CLASSES / 32; w++)
            if (w < cw) off += cls_wsum[w];
        p.tile.order[off] = (uint32_t)t;
    }
    if (crank == 0 && tid == 0) publish_total(p, total);
    cluster.sync();  // no CTA may exit while its shared memory can still be read remotely
}

// duplicateWithKeys (rasterizer_impl.cu:70-111), tile digit already consumed
__global__ void __launch_bounds__(256) scatter_kernel(const __grid_constant__ BinParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    unsigned long long key = 0;
    if (idx < p.P) {
        const uint2 r = p.geom.rect[idx];
        x0 = r.x & 0xffffu; y0 = r.x >> 16; x1 = r.y & 0xffffu; y1 = r.y >> 16;
        if (r.y != r.x) {
            const float depth = ((const float*)(p.geom.rec + 2 * idx + 1))[2];
            key = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)idx;
        }
    }
    const uint32_t w = x1 - x0, n = w * (y1 - y0);
    const bool wide = n > 32;
    uint32_t* __restrict__ cursor = p.tile.cursor + (size_t)(blockIdx.x % p.copies) * p.tiles;
    if (n && !wide) {
        // four cursor atomics in flight before their dependent stores (the kernel is bound by the
        // round trip of the returning atomics, not by their number)
        // (the tile index advances row-major through the rectangle: no division per instance)
        uint32_t t = y0 * p.gx + x0, x = 0;
        for (uint32_t i = 0; i < n; i += 4) {
            uint32_t slot[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (i + j < n) {
                    slot[j] = atomicAdd(cursor + t, 1u);
                    t++;
                    if (++x == w) { x = 0; t += p.gx - w; }
                }
            }
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (i + j < n && (long long)slot[j] < p.capacity) p.bucket[slot[j]] = key;
        }
    }
    unsigned m = __ballot_sync(0xffffffffu, wide);
    while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const uint32_t sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
        const uint32_t sw = __shfl_sync(0xffffffffu, w, src), sn = __shfl_sync(0xffffffffu, n, src);
        const unsigned long long skey = __shfl_sync(0xffffffffu, key, src);
        for (uint32_t i = lane; i < sn; i += 32) {
            const uint32_t slot = atomicAdd(cursor + (sy0 + i / sw) * p.gx + sx0 + i % sw, 1u);
            if ((long long)slot < p.capacity) p.bucket[slot] = skey;
        }
    }
}

// All-ascending bitonic network (no padding needed: comparators whose upper index is >= n
// compare against +inf and never swap).  `a` may be shared or global memory.
__device__ __forceinline__ void bitonic_sort(unsigned long long* a, int n, int tid, int nthreads) {
    int n_pad = 2;
    while (n_pad < n) n_pad <<= 1;
    const int half = n_pad >> 1;
    for (int k = 2; k <= n_pad; k <<= 1) {
        const int hk = k >> 1;
        for (int t = tid; t < half; t += nthreads) {
            const int i = (t / hk) * k + (t % hk);
            const int l = i ^ (k - 1);
            if (l < n) {
                const unsigned long long u = a[i], v = a[l];
                if (u > v) { a[i] = v; a[l] = u; }
            }
        }
        __syncthreads();
        for (int j = k >> 2; j > 0; j >>= 1) {
            for (int t = tid; t < half; t += nthreads) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int l = i | j;
                if (l < n) {
                    const unsigned long long u = a[i], v = a[l];
                    if (u > v) { a[i] = v; a[l] = u; }
                }
            }
            __syncthreads();
        }
    }
}

// One warp sorts a[0, c), c <= 32 * MAXPL, keys distinct (the Gaussian id sits in the low word): every lane counts,
// for its entries, the keys that are smaller (broadcast reads of a[j]) and then stores each entry at its rank.  For the
// buckets that a cluster of near-equal depths fills (a foreground object in front of a far wall puts two narrow depth
// clusters into a tile's linear bucket map; densification clones start at identical positions): c^2 / 32 steps on a
// warp instead of c^2 / 2 on one thread or a whole-segment comparison network.
template <int MAXPL>
__device__ __forceinline__ void warp_rank_sort(unsigned long long* a, int c, int lane) {
    unsigned long long key[MAXPL];
    int rank[MAXPL];
#pragma unroll
    for (int e = 0; e < MAXPL; e++) {
        const int i = lane + 32 * e;
        key[e] = i < c ? a[i] : ~0ull;
        rank[e] = 0;
    }
    for (int j = 0; j < c; j++) {
        const unsigned long long kj = a[j];
#pragma unroll
        for (int e = 0; e < MAXPL; e++) rank[e] += kj < key[e] ? 1 : 0;
    }
    __syncwarp();
#pragma unroll
    for (int e = 0; e < MAXPL; e++)
        if (lane + 32 * e < c) a[rank[e]] = key[e];
    __syncwarp();
}
// entries per lane sized to the bucket: the inner loop costs 1 + 3 MAXPL instructions per compared key
__device__ __forceinline__ void warp_rank_sort_any(unsigned long long* a, int c, int lane) {
    if (c <= 64) warp_rank_sort<2>(a, c, lane);
    else if (c <= 128) warp_rank_sort<4>(a, c, lane);
    else warp_rank_sort<8>(a, c, lane);
}
constexpr int LIGHT_BUCKET = 16;   // up to here: insertion sort by the bucket's thread
constexpr int HEAVY_BUCKET = 256;  // up to here: warp_rank_sort (<= 8 entries per lane); above: comparison network over the whole segment

// One CTA sorts one tile segment of up to SMALL_MAX entries.  Depth keys inside a tile are spread
// over a range, so instead of a comparison network the segment is bucketed: a monotone linear map
// of the depth bits between the tile's min and max onto NB buckets (shared-memory histogram whose
// atomicAdd return value is the entry's rank inside its bucket), an exclusive scan of the
// histogram, a scatter into shared memory, and an insertion sort inside every bucket that
// received more than one entry (full 64-bit compare, so equal depths end up in ascending-id
// order exactly like the reference's stable radix sort).  Expected bucket occupancy is < 1 for
// the typical tile and <= 4 at SMALL_MAX; buckets above LIGHT_BUCKET entries are sorted by a warp each
// (warp_rank_sort), and a bucket above HEAVY_BUCKET by the CTA's bitonic network over that bucket alone.
constexpr int NB = 1024;
constexpr int EPT = SMALL_MAX / SMALL_THREADS;  // entries per thread

__global__ void __launch_bounds__(SMALL_THREADS, 3) tile_sort_small_kernel(const __grid_constant__ BinParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    __shared__ unsigned long long out[SMALL_MAX];
    __shared__ uint32_t hist[NB];
    __shared__ uint32_t wsum[SMALL_THREADS / 32];
    __shared__ uint32_t s_min, s_max, s_maxcnt, n_heavy, n_medium;
    __shared__ uint16_t heavy[SMALL_MAX / HEAVY_BUCKET];   // buckets above HEAVY_BUCKET entries
    __shared__ uint16_t medium[SMALL_MAX / LIGHT_BUCKET];  // buckets of LIGHT_BUCKET + 1 .. HEAVY_BUCKET entries
    const uint2 r = p.tile.ranges[blockIdx.x];
    const int n = (int)(r.y - r.x);
    if (n < 2 || n > SMALL_MAX || p.tile.misc[MISC_OVERFLOW]) return;
    unsigned long long* seg = p.bucket + r.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    unsigned long long key[EPT];
    uint32_t lo = 0xffffffffu, hi = 0u;
#pragma unroll
    for (int e = 0; e < EPT; e++) {
        const int i = tid + e * SMALL_THREADS;
        if (i < n) {
            key[e] = seg[i];
            const uint32_t d = (uint32_t)(key[e] >> 32);
            lo = min(lo, d);
            hi = max(hi, d);
        }
    }
    for (int i = tid; i < NB; i += SMALL_THREADS) hist[i] = 0;
    if (tid == 0) { s_min = 0xffffffffu; s_max = 0u; s_maxcnt = 0u; n_heavy = 0u; n_medium = 0u; }
    __syncthreads();
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if (lane == 0) { atomicMin(&s_min, lo); atomicMax(&s_max, hi); }
    __syncthreads();
    const uint32_t kmin = s_min;
    const float scale = (float)NB / ((float)(s_max - kmin) + 1.0f);
    uint32_t br[EPT];  // bucket << 12 | rank inside the bucket
#pragma unroll
    for (int e = 0; e < EPT; e++) {
        const int i = tid + e * SMALL_THREADS;
        if (i < n) {
            const uint32_t d = (uint32_t)(key[e] >> 32);
            // monotone non-decreasing in d: u32->f32 conversion, positive scale, truncation, clamp
            const uint32_t bkt = min((uint32_t)((float)(d - kmin) * scale), (uint32_t)(NB - 1));
            br[e] = (bkt << 12) | atomicAdd(&hist[bkt], 1u);
        }
    }
    __syncthreads();
    // exclusive scan of the histogram in place (2 buckets per thread) + largest bucket
    {
        static_assert(NB == 2 * SMALL_THREADS, "scan assumes two buckets per thread");
        const uint32_t c0 = hist[2 * tid], c1 = hist[2 * tid + 1];
        uint32_t v = c0 + c1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) wsum[warp] = v;
        const uint32_t mx = __reduce_max_sync(0xffffffffu, max(c0, c1));
        if (lane == 0) atomicMax(&s_maxcnt, mx);
        __syncthreads();
        if (warp == 0) {
            uint32_t w = lane < SMALL_THREADS / 32 ? wsum[lane] : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += u;
            }
            if (lane < SMALL_THREADS / 32) wsum[lane] = w;
        }
        __syncthreads();
        const uint32_t excl = (warp ? wsum[warp - 1] : 0u) + v - (c0 + c1);
        hist[2 * tid] = excl;
        hist[2 * tid + 1] = excl + c0;
    }
    __syncthreads();
    {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
            const int i = tid + e * SMALL_THREADS;
            if (i < n) out[hist[br[e] >> 12] + (br[e] & 0xfffu)] = key[e];
        }
        __syncthreads();
        for (int bkt = tid; bkt < NB; bkt += SMALL_THREADS) {
            const int s0 = (int)hist[bkt], e0 = bkt + 1 < NB ? (int)hist[bkt + 1] : n;
            if (e0 - s0 > HEAVY_BUCKET) heavy[atomicAdd(&n_heavy, 1u)] = (uint16_t)bkt;
            else if (e0 - s0 > LIGHT_BUCKET) medium[atomicAdd(&n_medium, 1u)] = (uint16_t)bkt;
            if (e0 - s0 > LIGHT_BUCKET) continue;
            for (int i = s0 + 1; i < e0; i++) {  // insertion sort of out[s0, e0)
                const unsigned long long k = out[i];
                int j = i - 1;
                while (j >= s0 && out[j] > k) { out[j + 1] = out[j]; j--; }
                out[j + 1] = k;
            }
        }
        __syncthreads();
        {   // medium buckets: one warp each
            const int nm = (int)n_medium;
            for (int m = warp; m < nm; m += SMALL_THREADS / 32) {
                const int bkt = (int)medium[m];
                const int s0 = (int)hist[bkt], e0 = bkt + 1 < NB ? (int)hist[bkt + 1] : n;
                warp_rank_sort_any(out + s0, e0 - s0, lane);
            }
        }
        __syncthreads();
        // heavy buckets (a pile of clones, a whole tile at one depth): the comparison network, over that bucket only
        const int nh = (int)n_heavy;
        for (int h = 0; h < nh; h++) {
            const int bkt = (int)heavy[h];
            const int s0 = (int)hist[bkt], e0 = bkt + 1 < NB ? (int)hist[bkt + 1] : n;
            bitonic_sort(out + s0, e0 - s0, tid, SMALL_THREADS);
        }
    }
    for (int i = tid; i < n; i += SMALL_THREADS) seg[i] = out[i];
}

// Tile segments longer than SMALL_MAX: persistent CTAs (one per SM, 224 KB of shared memory) take them from the list
// the scan kernel wrote.  Up to LARGE_MAX entries the same bucket sort as above runs with NBL buckets over the keys as
// they lie in global memory (L2-resident: the scatter kernel has just written them) -- min/max pass, histogram pass,
// scan, scatter pass into shared memory through per-bucket cursors, insertion sort inside the buckets (expected
// occupancy <= 3), write-back.  A 2 M-Gaussian scene at 640x480 has ~600 such tiles per view: the bitonic network this
// replaces took 0.5 ms of a 1.3 ms view there (profiles/r02r_c5_rank0_launches.csv).  Buckets above LIGHT_BUCKET /
// HEAVY_BUCKET entries are handled as in the small kernel; a segment above LARGE_MAX is sorted by the network in
// global memory.
constexpr int NBL = 8192;
constexpr size_t LARGE_SMEM = (size_t)LARGE_MAX * 8 + (size_t)NBL * 4;

__global__ void __launch_bounds__(LARGE_THREADS) tile_sort_large_kernel(const __grid_constant__ BinParams p) {
    pdl_wait();     // first statement: see common.cuh
    pdl_trigger();
    extern __shared__ unsigned long long lkeys[];
    uint32_t* const lhist = reinterpret_cast<uint32_t*>(lkeys + LARGE_MAX);
    __shared__ uint32_t wsum[LARGE_THREADS / 32];
    __shared__ uint32_t s_min, s_max, s_maxcnt, n_heavy, n_medium;
    __shared__ uint16_t heavy[LARGE_MAX / HEAVY_BUCKET];
    constexpr int MEDIUM_LIST = 1024;          // (2 KB: what the 227 KB of the SM leave; beyond that the buckets are scanned)
    __shared__ uint16_t medium[MEDIUM_LIST];
    if (p.tile.misc[MISC_OVERFLOW]) return;
    const int n_large = (int)p.tile.misc[MISC_N_LARGE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int q = blockIdx.x; q < n_large; q += gridDim.x) {
        const uint2 r = p.tile.ranges[p.tile.large_list[q]];
        const int n = (int)(r.y - r.x);
        unsigned long long* seg = p.bucket + r.x;
        if (n > LARGE_MAX) {
            __syncthreads();
            bitonic_sort(seg, n, tid, LARGE_THREADS);  // in global memory (L2-resident)
            continue;
        }
        // pass 1: depth range of the segment
        uint32_t lo = 0xffffffffu, hi = 0u;
        for (int i = tid; i < n; i += LARGE_THREADS) {
            const uint32_t d = (uint32_t)(__ldcg(seg + i) >> 32);
            lo = min(lo, d);
            hi = max(hi, d);
        }
        for (int i = tid; i < NBL; i += LARGE_THREADS) lhist[i] = 0;
        if (tid == 0) { s_min = 0xffffffffu; s_max = 0u; s_maxcnt = 0u; n_heavy = 0u; n_medium = 0u; }
        __syncthreads();
        lo = __reduce_min_sync(0xffffffffu, lo);
        hi = __reduce_max_sync(0xffffffffu, hi);
        if (lane == 0) { atomicMin(&s_min, lo); atomicMax(&s_max, hi); }
        __syncthreads();
        const uint32_t kmin = s_min;
        const float scale = (float)NBL / ((float)(s_max - kmin) + 1.0f);
        // pass 2: histogram (monotone non-decreasing bucket map, as in the small kernel)
        for (int i = tid; i < n; i += LARGE_THREADS) {
            const uint32_t d = (uint32_t)(__ldcg(seg + i) >> 32);
            atomicAdd(&lhist[min((uint32_t)((float)(d - kmin) * scale), (uint32_t)(NBL - 1))], 1u);
        }
        __syncthreads();
        // exclusive scan in place (NBL / LARGE_THREADS consecutive buckets per thread) + largest bucket
        {
            constexpr int BPT = NBL / LARGE_THREADS;
            uint32_t c[BPT], v = 0, mx = 0;
#pragma unroll
            for (int k = 0; k < BPT; k++) { c[k] = lhist[BPT * tid + k]; v += c[k]; mx = max(mx, c[k]); }
            const uint32_t mine = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            if (lane == 31) wsum[warp] = v;
            mx = __reduce_max_sync(0xffffffffu, mx);
            if (lane == 0) atomicMax(&s_maxcnt, mx);
            __syncthreads();
            if (warp == 0) {
                uint32_t w = wsum[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t u = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += u;
                }
                wsum[lane] = w;
            }
            __syncthreads();
            uint32_t excl = (warp ? wsum[warp - 1] : 0u) + v - mine;
#pragma unroll
            for (int k = 0; k < BPT; k++) { lhist[BPT * tid + k] = excl; excl += c[k]; }
        }
        __syncthreads();
        {
            // pass 3: scatter through the cursors (lhist[b] ends at the END of bucket b)
            for (int i = tid; i < n; i += LARGE_THREADS) {
                const unsigned long long key = __ldcg(seg + i);
                const uint32_t d = (uint32_t)(key >> 32);
                const uint32_t bkt = min((uint32_t)((float)(d - kmin) * scale), (uint32_t)(NBL - 1));
                lkeys[atomicAdd(&lhist[bkt], 1u)] = key;
            }
            __syncthreads();
            for (int bkt = tid; bkt < NBL; bkt += LARGE_THREADS) {
                const int s0 = bkt ? (int)lhist[bkt - 1] : 0, e0 = (int)lhist[bkt];
                if (e0 - s0 > HEAVY_BUCKET) {
                    heavy[atomicAdd(&n_heavy, 1u)] = (uint16_t)bkt;
                } else if (e0 - s0 > LIGHT_BUCKET) {
                    const uint32_t slot = atomicAdd(&n_medium, 1u);
                    if (slot < (uint32_t)MEDIUM_LIST) medium[slot] = (uint16_t)bkt;
                }
                if (e0 - s0 > LIGHT_BUCKET) continue;
                for (int i = s0 + 1; i < e0; i++) {  // insertion sort of lkeys[s0, e0): full 64-bit compare
                    const unsigned long long k = lkeys[i];
                    int j = i - 1;
                    while (j >= s0 && lkeys[j] > k) { lkeys[j + 1] = lkeys[j]; j--; }
                    lkeys[j + 1] = k;
                }
            }
            __syncthreads();
            if (n_medium <= (uint32_t)MEDIUM_LIST) {   // medium buckets: one warp each
                const int nm = (int)n_medium;
                for (int m = warp; m < nm; m += LARGE_THREADS / 32) {
                    const int bkt = (int)medium[m];
                    const int s0 = bkt ? (int)lhist[bkt - 1] : 0, e0 = (int)lhist[bkt];
                    warp_rank_sort_any(lkeys + s0, e0 - s0, lane);
                }
            } else {                                   // (more of them than the list holds: scan the buckets)
                for (int bkt = warp; bkt < NBL; bkt += LARGE_THREADS / 32) {
                    const int s0 = bkt ? (int)lhist[bkt - 1] : 0, e0 = (int)lhist[bkt];
                    if (e0 - s0 > LIGHT_BUCKET && e0 - s0 <= HEAVY_BUCKET) warp_rank_sort_any(lkeys + s0, e0 - s0, lane);
                }
            }
            __syncthreads();
            const int nh = (int)n_heavy;
            for (int h = 0; h < nh; h++) {   // heavy buckets: the comparison network over that bucket only
                const int bkt = (int)heavy[h];
                const int s0 = bkt ? (int)lhist[bkt - 1] : 0, e0 = (int)lhist[bkt];
                bitonic_sort(lkeys + s0, e0 - s0, tid, LARGE_THREADS);
            }
            __syncthreads();
        }
        for (int i = tid; i < n; i += LARGE_THREADS) seg[i] = lkeys[i];
        __syncthreads();
    }
}

// ---- parity/debug exports in the reference's BinningState / ImageState terms ----
__global__ void export_binning_kernel(BinParams p, unsigned long long* keys, unsigned int* list,
                                      unsigned int* ranges) {
    const int t = blockIdx.x;
    const uint2 r = p.tile.ranges[t];
    if (ranges && threadIdx.x == 0) {
        const bool empty = r.x == r.y;  // the reference leaves empty tiles at the memset value
        ranges[2 * t] = empty ? 0u : r.x;
        ranges[2 * t + 1] = empty ? 0u : r.y;
    }
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x) {
        const unsigned long long e = p.bucket[i];
        if (keys) keys[i] = ((unsigned long long)t << 32) | (e >> 32);
        if (list) list[i] = (unsigned int)e;
    }
}

__global__ void export_geometry_kernel(int P, GeomView g, const int* radii, float* depths, float* means2D,
                                       float* conic_opacity, float* rgb, uint8_t* clamped) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const bool vis = radii[i] > 0;
    const float4 a = vis ? g.rec[2 * i] : make_float4(0, 0, 0, 0);
    const float4 b = vis ? g.rec[2 * i + 1] : make_float4(0, 0, 0, 0);
    const float4 c = vis ? g.col[i] : make_float4(0, 0, 0, 0);
    const uint8_t cb = vis ? g.clamped[i] : 0;
    if (depths) depths[i] = b.z;
    if (means2D) { means2D[2 * i] = a.x; means2D[2 * i + 1] = a.y; }
    if (conic_opacity) {
        conic_opacity[4 * i] = a.z; conic_opacity[4 * i + 1] = a.w;
        conic_opacity[4 * i + 2] = b.x; conic_opacity[4 * i + 3] = b.y;
    }
    if (rgb) { rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z; }
    if (clamped) { clamped[3 * i] = cb & 1; clamped[3 * i + 1] = (cb >> 1) & 1; clamped[3 * i + 2] = (cb >> 2) & 1; }
}

void launch_tile_scan(const BinParams& p, cudaStream_t s) {
    if (p.tiles <= SCAN_CLUSTER * 1024) {
        switch (p.copies) {
            case 32: launch_chained(tile_scan_cluster_kernel<32>, SCAN_CLUSTER, 1024, 0, s, p); break;
            case 16: launch_chained(tile_scan_cluster_kernel<16>, SCAN_CLUSTER, 1024, 0, s, p); break;
            case 8: launch_chained(tile_scan_cluster_kernel<8>, SCAN_CLUSTER, 1024, 0, s, p); break;
            case 4: launch_chained(tile_scan_cluster_kernel<4>, SCAN_CLUSTER, 1024, 0, s, p); break;
            case 2: launch_chained(tile_scan_cluster_kernel<2>, SCAN_CLUSTER, 1024, 0, s, p); break;
            default: launch_chained(tile_scan_cluster_kernel<1>, SCAN_CLUSTER, 1024, 0, s, p); break;
        }
        return;
    }
    switch (p.copies) {
        case 32: launch_chained(tile_scan_kernel<32>, 1, 1024, 0, s, p); break;
        case 16: launch_chained(tile_scan_kernel<16>, 1, 1024, 0, s, p); break;
        case 8: launch_chained(tile_scan_kernel<8>, 1, 1024, 0, s, p); break;
        case 4: launch_chained(tile_scan_kernel<4>, 1, 1024, 0, s, p); break;
        case 2: launch_chained(tile_scan_kernel<2>, 1, 1024, 0, s, p); break;
        default: launch_chained(tile_scan_kernel<1>, 1, 1024, 0, s, p); break;
    }
}

void launch_scatter(const BinParams& p, cudaStream_t s) {
    if (p.P > 0) launch_chained(scatter_kernel, (p.P + 255) / 256, 256, 0, s, p);
}

void launch_tile_sorts(const BinParams& p, int num_sms, cudaStream_t s) {
    launch_chained(tile_sort_small_kernel, p.tiles, SMALL_THREADS, 0, s, p);
    launch_chained(tile_sort_large_kernel, num_sms, LARGE_THREADS, LARGE_SMEM, s, p);
}

void launch_export_binning(const BinParams& p, const ImageView& image, int W, int H, unsigned long long* keys,
                           unsigned int* list, unsigned int* ranges, unsigned int* n_contrib, float* final_T,
                           cudaStream_t s) {
    export_binning_kernel<<<p.tiles, 256, 0, s>>>(p, keys, list, ranges);
    const size_t N = (size_t)W * H;
    if (n_contrib) cudaMemcpyAsync(n_contrib, image.n_contrib, N * 4, cudaMemcpyDeviceToDevice, s);
    if (final_T) cudaMemcpyAsync(final_T, image.final_T, N * 4, cudaMemcpyDeviceToDevice, s);
}

void launch_export_geometry(int P, const GeomView& g, const int* radii, float* depths, float* means2D,
                            float* conic_opacity, float* rgb, uint8_t* clamped, cudaStream_t s) {
    if (P > 0)
        export_geometry_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g, radii, depths, means2D, conic_opacity, rgb,
                                                               clamped);
}

int configure_binning() {
    return (int)cudaFuncSetAttribute(tile_sort_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)LARGE_SMEM);
}

}  // namespace wgs
