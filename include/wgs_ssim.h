/*
 * wgs_ssim.h — C ABI of the fused SSIM loss term and its cotangent (libwgs_rasterizer.so).
 *
 * SURVEY.md §8(f) row 2, the SSIM half of the mapping loss.  Replaces, with two launches, the chain
 *   ssim(img1, img2, window_size=11, size_average)   thirdparty/gaussian_splatting/utils/loss_utils.py:61-101
 * (five depthwise 11x11 conv2d + ~15 elementwise kernels forward, as many again in autograd's backward) that
 * get_loss_mapping_rgbd (src/utils/slam_utils.py:129-136) and get_loss_mapping_uncertainty (:200, :221-223)
 * evaluate on the rendered image every mapping iteration, and with one launch the detached
 *   compute_ssim_components(gt, rendered, window_size)   src/utils/dyn_uncertainty/mapping_utils.py:99-204
 * (luminance / contrast / structure maps of the uncertainty loss).
 *
 *   w        = normalised 1-D Gaussian, sigma 1.5, `window` taps (odd, <= 11); 2-D window = w w^T, zero padding
 *   mu1, mu2 = w * img1, w * img2;  s11, s22, s12 = w * img1^2, w * img2^2, w * (img1 img2)
 *   ssim_map = (2 mu1 mu2 + C1)(2 (s12 - mu1 mu2) + C2) / ((mu1^2 + mu2^2 + C1)(s11 - mu1^2 + s22 - mu2^2 + C2))
 *   mean[g]  = mean of ssim_map over the `planes_per_mean` consecutive (H,W) planes of group g
 * Backward: dL/dimg1 for given dL/dmean[g] (img2 is the ground truth: no gradient).  The forward leaves three
 * per-pixel maps (d ssim / d mu1, d s11, d s12) in `maps`; the backward convolves them with the same window:
 *   dL/dimg1 = k_g (w * d_mu1 + 2 img1 (w * d_s11) + img2 (w * d_s12)),  k_g = dL/dmean[g] / (planes_per_mean H W)
 * All pointers are DEVICE pointers, work is enqueued on the caller's stream, nothing synchronises; the reduction
 * is deterministic (per-block partials, fixed-order double sum).
 */
#ifndef WGS_SSIM_H_
#define WGS_SSIM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WGS_SSIM_MAX_WINDOW 11

typedef struct WgsSsimArgs {
  int32_t width, height;
  int32_t planes;          /* B*C images of (H,W), contiguous */
  int32_t planes_per_mean; /* C (size_average=False: one mean per batch item) or B*C (size_average=True) */
  int32_t window;          /* odd, 1..11 (reference: 11) */
  const float* img1;       /* (planes,H,W) rendered */
  const float* img2;       /* (planes,H,W) ground truth */
  float* mean;             /* out (forward): planes / planes_per_mean floats */
  float* maps;             /* (3,planes,H,W): written by forward (may be null: no backward wanted), read by backward */
  const float* dL_dmean;   /* backward: planes / planes_per_mean floats */
  float* dL_dimg1;         /* backward out: (planes,H,W) */
  void* workspace;         /* wgs_ssim_workspace_bytes() bytes, 256-B aligned (forward; backward when dL_dexposure is set) */
  /* Exposure compensation fused in (slam_utils.py:111, :185-187): img1 is replaced by exp(a) * img1 + b before the
   * windows are applied (zero padding stays zero).  exposure_a / exposure_b: 1 float each, or both null.  The backward
   * then returns dL/dimg1 of the UNCOMPENSATED image and, if dL_dexposure is given, {dL/da, dL/db} (2 floats). */
  const float* exposure_a;
  const float* exposure_b;
  float* dL_dexposure;
} WgsSsimArgs;

size_t wgs_ssim_workspace_bytes(int32_t width, int32_t height, int32_t planes);   /* covers forward and backward */
/* return WGS_OK or a negative status of wgs_rasterizer.h; message via wgs_last_error() */
int wgs_ssim_forward(const WgsSsimArgs* args, void* cuda_stream);
int wgs_ssim_backward(const WgsSsimArgs* args, void* cuda_stream);

/* compute_ssim_components: img1, img2 (batch, channels, H, W); luminance / contrast / structure (batch, H, W), each the
 * mean over the channels of the clipped components (mapping_utils.py:166-204: sigma^2 >= eps, |s12| <= sqrt(s11 s22),
 * contrast and structure clamped at 0.98).  Forward only (the reference detaches them). */
/* exposure_a / exposure_b (1 float each, or both null) compensate img2 -- the reference passes (gt, rendered) */
int wgs_ssim_components(int32_t width, int32_t height, int32_t batch, int32_t channels, int32_t window,
                        const float* img1, const float* img2, const float* exposure_a, const float* exposure_b,
                        float* luminance, float* contrast, float* structure, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_SSIM_H_ */
