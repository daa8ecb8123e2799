/*
 * wgs_loss.h — C ABI of the fused photometric loss + cotangent producer (libwgs_rasterizer.so).
 *
 * the step on the OUTPUT side of the rasterizer.  One launch replaces
 * the chain of ~25 elementwise/reduction torch kernels (forward + autograd backward) of the
 * reference's
 *   get_loss_tracking / get_loss_tracking_rgb   (src/utils/slam_utils.py:47-87)
 *   get_loss_mapping  / get_loss_mapping_rgbd   (src/utils/slam_utils.py:107-143; the SSIM term: wgs_ssim.h)
 *   the L1 colour and depth terms of get_loss_mapping_uncertainty (src/utils/slam_utils.py:146-258,
 *   src/utils/dyn_uncertainty/mapping_utils.py:257-268)
 * and writes dL/dcolor and dL/ddepth in exactly the layout wgs_backward() consumes
 * (WgsBackwardArgs::dL_dout_color / dL_dout_depth), plus dL/d(exposure_a, exposure_b).
 *
 *   image_ab = exp(a) * color + b                                  (exposure; identity if null)
 *   m        = (sum_c gt_color > rgb_boundary_threshold) * pixel_mask
 *   w        = 0.5 / uncertainty^2, set to 0 where < 0.1           (1 if uncertainty null)
 *   l_rgb    = opacity * | image_ab*m - gt_color*m | * w           (opacity = 1 if null)
 *   l_depth  = | depth*dm - gt_depth*dm |,  dm = gt_depth > depth_min  [and gt_depth < depth_max if depth_max > 0]
 *   if depth_weighted:  l_depth *= w  where gt_depth < depth + depth_margin      (get_loss_mapping_uncertainty,
 *                                                                                 src/utils/slam_utils.py:238-242)
 *   loss     = color_weight * mean_{3HW}(l_rgb) + depth_weight * mean_{HW}(l_depth)
 *
 * The gradient w.r.t. the rendered opacity is not produced: the reference rasterizer ignores
 * it (diff_gaussian_rasterization/__init__.py:108,116-138).  All pointers are DEVICE pointers;
 * work is enqueued on the caller's stream; the reduction is deterministic (per-block partials
 * summed by the last block in a fixed order, double accumulation).
 */
#ifndef WGS_LOSS_H_
#define WGS_LOSS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct WgsLossArgs {
  int32_t width, height;
  float rgb_boundary_threshold; /* config["Training"]["rgb_boundary_threshold"] */
  float depth_min;              /* 0.01 in the reference (slam_utils.py:140) */
  float color_weight;           /* alpha (mapping) or 1 (tracking) */
  float depth_weight;           /* 1 - alpha (mapping) or 0 (tracking) */
  float depth_max;              /* > 0: the depth mask also needs gt_depth < depth_max (mapping_utils.py:262-266); 0: off */
  float depth_margin;           /* used when depth_weighted: 1.0 in the reference (slam_utils.py:241) */
  int32_t depth_weighted;       /* != 0: multiply l_depth by w where gt_depth < depth + depth_margin (needs uncertainty) */
  int32_t patch_w, patch_h;     /* both > 0: `uncertainty` is the (patch_h, patch_w) network output and the per-pixel value
                                   is made here (mapping_utils.py:270-277): bilinear(clip(u, 0.1) + 1e-3) to (H, W),
                                   then (r - 0.1) * uncertainty_rate + 0.1.  Both 0: `uncertainty` is (H, W) as given */
  float uncertainty_rate;       /* 1 + compute_bias_factor(train_frac, 0.8); used with patch_w / patch_h */
  const float* color;           /* (3,H,W) rendered colour */
  const float* depth;           /* (1,H,W) rendered depth; may be null iff depth_weight == 0 */
  const float* opacity;         /* (1,H,W) rendered opacity or null */
  const float* gt_color;        /* (3,H,W) */
  const float* gt_depth;        /* (1,H,W); may be null iff depth_weight == 0 */
  const float* pixel_mask;      /* (1,H,W) 0/1 floats (viewpoint.grad_mask) or null */
  const float* uncertainty;     /* (H,W) or null */
  const float* exposure_a;      /* 1 float (viewpoint.exposure_a) or null: no exposure compensation */
  const float* exposure_b;      /* 1 float; required iff exposure_a is given */
  float* loss;                  /* out: 1 float */
  float* dL_dcolor;             /* out: (3,H,W) */
  float* dL_ddepth;             /* out: (1,H,W); may be null iff depth_weight == 0 */
  float* dL_dexposure;          /* out: 2 floats {dL/da, dL/db}; may be null iff exposure_a is null */
  float* weight_mean;           /* out (optional): mean over (H, W) of w -- the factor of the SSIM term (slam_utils.py:226-230) */
  void* workspace;              /* wgs_loss_workspace_bytes(width, height) bytes, 256-B aligned; must be zero before
                                   the first call (the kernel leaves it zero again) */
} WgsLossArgs;

size_t wgs_loss_workspace_bytes(int32_t width, int32_t height);
/* returns WGS_OK or a negative status of wgs_rasterizer.h; message via wgs_last_error() */
int wgs_photometric_loss(const WgsLossArgs* args, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_LOSS_H_ */
