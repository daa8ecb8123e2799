// Fused row compaction for prune / densify (include/wgs_compact.h).  HBM-bound data movement:
// plan = 3-kernel exclusive scan of the keep mask (1 B read + 4 B written per row), rows = one launch that
// copies the surviving rows of every tensor (word-granular, consecutive lanes on consecutive words).
#include "../../include/wgs_compact.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"
#include <mutex>

namespace wgs {

constexpr int CP_BLOCK = 1024;

__global__ void __launch_bounds__(CP_BLOCK) compact_count_kernel(int P, const uint8_t* __restrict__ keep,
                                                                 uint32_t* __restrict__ block_sums) {
    const int i = blockIdx.x * CP_BLOCK + threadIdx.x;
    const bool k = i < P && keep[i] != 0;
    const int c = __syncthreads_count(k);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = (uint32_t)c;
}

// single block: exclusive scan of block_sums in place, total -> count
__global__ void __launch_bounds__(1024) compact_scan_kernel(int nblocks, uint32_t* __restrict__ block_sums,
                                                            int32_t* __restrict__ count) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        const int b = base + tid;
        const uint32_t v0 = b < nblocks ? block_sums[b] : 0u;
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
        const uint32_t incl = carry + (warp ? wsum[warp - 1] : 0u) + v;
        if (b < nblocks) block_sums[b] = incl - v0;
        __syncthreads();
        if (tid == 1023) carry = incl;
        __syncthreads();
    }
    if (tid == 0) *count = (int32_t)carry;
}

__global__ void __launch_bounds__(CP_BLOCK) compact_index_kernel(int P, const uint8_t* __restrict__ keep,
                                                                 const uint32_t* __restrict__ block_sums,
                                                                 int32_t* __restrict__ dst_index) {
    __shared__ uint32_t wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int i = blockIdx.x * CP_BLOCK + tid;
    const bool k = i < P && keep[i] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, k);
    if (lane == 0) wsum[warp] = __popc(m);
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
    if (i < P) {
        const uint32_t pos = block_sums[blockIdx.x] + (warp ? wsum[warp - 1] : 0u) + __popc(m & ((1u << lane) - 1u));
        dst_index[i] = k ? (int32_t)pos : -1;
    }
}

struct CompactParams {
    WgsCompactTensor t[WGS_COMPACT_MAX_TENSORS];
    int n, P;
    const int32_t* dst_index;
};

// grid.y = tensor, grid.x = chunk of CR_ROWS rows.  The chunk's destination indices are staged in shared memory and
// the chunk is then copied word by word (consecutive lanes on consecutive words, so reads are fully coalesced and
// writes are coalesced wherever survivors are adjacent) with (row, col) advanced incrementally -- no division
// in the loop, any row size.
constexpr int CR_ROWS = 256;
__global__ void __launch_bounds__(256) compact_rows_kernel(const __grid_constant__ CompactParams p) {
    __shared__ int32_t sidx[CR_ROWS];
    const WgsCompactTensor& t = p.t[blockIdx.y];
    const uint32_t words = t.row_bytes >> 2;
    const int tid = threadIdx.x;
    const size_t row0 = (size_t)blockIdx.x * CR_ROWS;
    const uint32_t nrows = (uint32_t)min((size_t)CR_ROWS, (size_t)p.P - row0);
    if ((uint32_t)tid < nrows) sidx[tid] = __ldg(p.dst_index + row0 + tid);
    __syncthreads();
    const uint32_t* __restrict__ src = (const uint32_t*)t.src + row0 * words;
    uint32_t* __restrict__ dst = (uint32_t*)t.dst;
    const uint32_t nw = nrows * words;
    const uint32_t dr = 256u / words, dc = 256u % words;
    uint32_t row = (uint32_t)tid / words, col = (uint32_t)tid % words;
#pragma unroll 4
    for (uint32_t w = tid; w < nw; w += 256) {
        const int32_t d = sidx[row];
        if (d >= 0) dst[(size_t)d * words + col] = __ldg(src + w);
        row += dr;
        col += dc;
        if (col >= words) { col -= words; row++; }
    }
}

}  // namespace wgs

using namespace wgs;
extern "C" {
void wgs_set_error_message(const char* msg);

size_t wgs_compact_workspace_bytes(int32_t P) {
    if (P < 0) return 0;
    return align_up(((size_t)P / CP_BLOCK + 2) * sizeof(uint32_t));
}

int wgs_compact_plan(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count, void* workspace,
                     void* cuda_stream) {
    if (P < 0 || !count || (P > 0 && (!keep || !dst_index || !workspace || ((uintptr_t)workspace & 255u)))) {
        wgs_set_error_message("wgs_compact_plan: null/misaligned pointer or negative P");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (P == 0) {
        cudaMemsetAsync(count, 0, sizeof(int32_t), s);
        return WGS_OK;
    }
    const int nb = (P + CP_BLOCK - 1) / CP_BLOCK;
    uint32_t* sums = (uint32_t*)workspace;
    compact_count_kernel<<<nb, CP_BLOCK, 0, s>>>(P, keep, sums);
    compact_scan_kernel<<<1, 1024, 0, s>>>(nb, sums, count);
    compact_index_kernel<<<nb, CP_BLOCK, 0, s>>>(P, keep, sums, dst_index);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}

// ---- wgs_compact_plan_static: the same plan with the block sums in library-owned static device memory ----
// For callers that hold no allocator on the device they run on (see include/wgs_compact.h).  One array per GPU
// that loads the library, resident until the process exits; sized for WGS_COMPACT_STATIC_MAX_ROWS rows.
}  // extern "C"
namespace wgs {
constexpr int CP_STATIC_WORDS = WGS_COMPACT_STATIC_MAX_ROWS / CP_BLOCK + 2;
static __device__ __align__(256) uint32_t g_static_block_sums[CP_STATIC_WORDS];

// In host code `g_static_block_sums` is only a handle for the runtime's symbol table, not a device address, so it
// is resolved once per device with cudaGetSymbolAddress (that device current) and the resolved pointer is what the
// kernels receive.  No allocation: the array is part of the module image.
constexpr int CP_MAX_DEVICES = 64;
static uint32_t* g_static_ptr[CP_MAX_DEVICES] = {};
static std::mutex g_static_mutex;

static cudaError_t static_block_sums(uint32_t** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= CP_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_block_sums);
        if (e != cudaSuccess) return e;
        g_static_ptr[dev] = (uint32_t*)p;
    }
    *out = g_static_ptr[dev];
    return cudaSuccess;
}
}  // namespace wgs
extern "C" {

int wgs_compact_plan_static(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count, void* cuda_stream) {
    if (P < 0 || !count || (P > 0 && (!keep || !dst_index))) {
        wgs_set_error_message("wgs_compact_plan_static: null pointer or negative P");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (P > WGS_COMPACT_STATIC_MAX_ROWS) {
        wgs_set_error_message("wgs_compact_plan_static: P exceeds WGS_COMPACT_STATIC_MAX_ROWS (16777216 rows); "
                              "use wgs_compact_plan with a caller-owned workspace");
        return WGS_ERR_CAPACITY;
    }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (P == 0) {
        const cudaError_t e0 = cudaMemsetAsync(count, 0, sizeof(int32_t), s);
        if (e0 != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e0)); return WGS_ERR_CUDA; }
        return WGS_OK;
    }
    uint32_t* sums = nullptr;
    cudaError_t e = static_block_sums(&sums);   // first call on this device resolves the symbol (set-up)
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    // Nothing in the array has to start at zero: compact_count_kernel overwrites words [0, nb) -- every word the
    // scan and the index kernel read -- so whatever the previous call left behind is never observed.
    const int nb = (P + CP_BLOCK - 1) / CP_BLOCK;
    compact_count_kernel<<<nb, CP_BLOCK, 0, s>>>(P, keep, sums);
    compact_scan_kernel<<<1, 1024, 0, s>>>(nb, sums, count);
    compact_index_kernel<<<nb, CP_BLOCK, 0, s>>>(P, keep, sums, dst_index);
    e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}

int wgs_compact_rows(int32_t P, const int32_t* dst_index, const WgsCompactTensor* tensors, int32_t n,
                     void* cuda_stream) {
    if (P < 0 || n < 0 || n > WGS_COMPACT_MAX_TENSORS || (n > 0 && !tensors) || (P > 0 && !dst_index)) {
        wgs_set_error_message("wgs_compact_rows: bad argument");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    CompactParams p;
    p.n = 0; p.P = P; p.dst_index = dst_index;
    for (int i = 0; i < n; i++) {
        if (tensors[i].row_bytes == 0) continue;
        if ((tensors[i].row_bytes & 3u) || !tensors[i].src || !tensors[i].dst) {
            wgs_set_error_message("wgs_compact_rows: row_bytes must be a multiple of 4 and pointers non-null");
            return WGS_ERR_INVALID_ARGUMENT;
        }
        p.t[p.n++] = tensors[i];
    }
    if (P == 0 || p.n == 0) return WGS_OK;
    const unsigned gx = (unsigned)((P + CR_ROWS - 1) / CR_ROWS);
    compact_rows_kernel<<<dim3(gx, p.n), 256, 0, (cudaStream_t)cuda_stream>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}
}
