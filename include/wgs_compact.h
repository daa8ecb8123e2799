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
/* tensors: HOST array of `n` (<= WGS_COMPACT_MAX_TENSORS) descriptors */
int wgs_compact_rows(int32_t P, const int32_t* dst_index, const WgsCompactTensor* tensors, int32_t n,
                     void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_COMPACT_H_ */
