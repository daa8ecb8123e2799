/*
 * wgs_compact.h — C ABI of the fused row compaction (libwgs_rasterizer.so).  SURVEY.md §8(f) row 3
 * (densification side): GaussianModel.prune_points / _prune_optimizer
 * (thirdparty/gaussian_splatting/scene/gaussian_model.py:526-564) select the surviving rows of 6 parameter
 * tensors, their 12 Adam moment tensors and 3 statistics tensors with one boolean-index expression each
 * (≈4 kernels and a host synchronisation per tensor).  Here: ONE scan of the keep mask (wgs_compact_plan)
 * and ONE gather launch for all tensors (wgs_compact_rows).  Pure data movement: results are bit-identical
 * to `tensor[mask]`.  All pointers are DEVICE pointers unless noted.
 */
#ifndef WGS_COMPACT_H_
#define WGS_COMPACT_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define WGS_COMPACT_MAX_TENSORS 32

typedef struct WgsCompactTensor {
  const void* src;     /* (P, row_bytes) */
  void* dst;           /* (count, row_bytes), caller-allocated (count is known after the plan) */
  uint32_t row_bytes;  /* multiple of 4 */
} WgsCompactTensor;

size_t wgs_compact_workspace_bytes(int32_t P);
/* keep: (P) bytes, non-zero = row survives.  dst_index: (P) int32 out, destination row or -1.
 * count: (1) int32 out (device), number of surviving rows.  workspace: wgs_compact_workspace_bytes(P), 256-B aligned. */
int wgs_compact_plan(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count, void* workspace,
                     void* cuda_stream);
/* wgs_compact_plan without a workspace argument.  No counterpart in the reference (its prune path lets torch
 * allocate freely); it exists for hosts that may not allocate device memory where this runs: a plugin living inside
 * another process's fixed memory budget, a real-time mapping thread that must not take the allocator's lock or risk
 * an implicit device synchronisation from cudaMalloc/cudaFree, or a caller that owns no allocator on that device at
 * all and so cannot provide the workspace wgs_compact_plan asks for.  The block sums live in a static device array
 * inside the library (65 544 B on every GPU that loads it, held until the process exits); the function, and anything
 * it calls, allocates and frees nothing.  Same arguments (minus workspace) and bit-identical dst_index / count.
 *  - P > WGS_COMPACT_STATIC_MAX_ROWS is refused with WGS_ERR_CAPACITY and a message; nothing is launched.
 *  - The first call on each device looks up the array's device address (that device must be current, as for every
 *    call); make that first call before starting any stream capture on that device.
 *  - All calls on one device share the array.  It needs no clearing between calls (each call overwrites what it
 *    reads), but two calls whose kernels can overlap on the GPU -- different streams without an ordering between
 *    them, or two host threads -- race on it and produce wrong indices.  Serialise them, or use wgs_compact_plan. */
#define WGS_COMPACT_STATIC_MAX_ROWS (1 << 24)
int wgs_compact_plan_static(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count, void* cuda_stream);
/* tensors: HOST array of `n` (<= WGS_COMPACT_MAX_TENSORS) descriptors */
int wgs_compact_rows(int32_t P, const int32_t* dst_index, const WgsCompactTensor* tensors, int32_t n,
                     void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_COMPACT_H_ */
