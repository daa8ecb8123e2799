/*
 * wgs_uncertainty.h — C ABI of the fused uncertainty loss of the mapper (libwgs_rasterizer.so).
 *
 * SURVEY.md §8(f) row 2, last piece: the patch-resolution half of compute_mapping_loss_components
 * (src/utils/dyn_uncertainty/mapping_utils.py:270-321) and its use in get_loss_mapping_uncertainty
 * (src/utils/slam_utils.py:244-256).  Two launches replace ~50 small torch kernels (resampling, median pooling,
 * elementwise chains and their autograd backward):
 *
 *   ssim_l   = min(opacity * ssim_weight * (1 - lum)(1 - structure)(1 - contrast), 5)              full resolution
 *   d_l1     = min(| depth*dm - ref*dm |, 5),  dm = ref > 0.01 and ref < depth_threshold
 *   s        = bilinear(ssim_l) , o = bilinear(opacity), dl = bicubic(d_l1), dr = bicubic(ref)     -> (Hp, Wp)
 *              (F.interpolate to `size`, align_corners=False, A = -0.75: the formulas of torch's kernels)
 *   f        = median of the k x k neighbourhood of s, reflect padding (MedianPool2d, same=True)
 *   p        = max(uncertainty, 0.1) + 1e-3
 *   u        = f / p^2 + 0.5 log p + depth_mult * (dr > depth_threshold ? 0 : dl) / p^2 ;  u = 0 where o < opacity_th
 *   loss     = weight * mean(u)                      (weight = config["uncertainty_params"]["ssim_mult"])
 *   dL/duncertainty = weight / (Hp Wp) * (-2 f / p^3 + 0.5 / p - 2 depth_mult dl' / p^3) where uncertainty >= 0.1, o >= th
 *
 * All pointers are DEVICE pointers; nothing synchronises; the reduction is one block in a fixed order.
 */
#ifndef WGS_UNCERTAINTY_H_
#define WGS_UNCERTAINTY_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct WgsUncertaintyLossArgs {
  int32_t width, height;        /* full resolution */
  int32_t patch_w, patch_h;     /* resolution of the uncertainty map */
  int32_t median;               /* odd, 1..7 (reference: 5) */
  float ssim_weight;            /* 100 + 900 * bias(ssim_frac, 0.8) */
  float depth_threshold;        /* min(10 * median(ref_depth), 50) */
  float depth_mult;             /* uncer_depth_mult */
  float opacity_th;             /* opacity_th_for_uncer_loss */
  float weight;                 /* ssim_mult */
  const float* uncertainty;     /* (Hp, Wp) network output */
  const float* luminance;       /* (H, W) wgs_ssim_components */
  const float* contrast;        /* (H, W) */
  const float* structure;       /* (H, W) */
  const float* opacity;         /* (H, W) rendered opacity */
  const float* depth;           /* (H, W) rendered depth */
  const float* ref_depth;       /* (H, W) */
  float* loss;                  /* out: 1 float */
  float* dL_duncertainty;       /* out: (Hp, Wp); may be null (freeze_uncertainty_loss) */
  void* workspace;              /* wgs_uncertainty_workspace_bytes(patch_w, patch_h) bytes, 256-B aligned */
} WgsUncertaintyLossArgs;

size_t wgs_uncertainty_workspace_bytes(int32_t patch_w, int32_t patch_h);
/* returns WGS_OK or a negative status of wgs_rasterizer.h; message via wgs_last_error() */
int wgs_uncertainty_loss(const WgsUncertaintyLossArgs* args, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_UNCERTAINTY_H_ */
