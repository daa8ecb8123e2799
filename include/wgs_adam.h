/*
 * wgs_adam.h — C ABI of the fused multi-tensor Adam step (libwgs_rasterizer.so).  SURVEY.md §8(f) row 3
 * (optimiser side): ONE launch updates every parameter tensor of the Gaussian model (xyz, f_dc, f_rest,
 * opacity, scaling, rotation: thirdparty/gaussian_splatting/scene/gaussian_model.py training_setup, a
 * torch.optim.Adam with one param group per tensor, eps 1e-15) instead of torch's per-op foreach kernels.
 * Arithmetic of torch.optim.Adam (amsgrad off, weight_decay 0, maximize off):
 *   m = m + (1 - beta1) (g - m);  v = beta2 v + (1 - beta2) g g;
 *   p = p - (lr / bias_correction1) * m / (sqrt(v) / sqrt(bias_correction2) + eps)
 * The bias corrections are passed in (the host computes 1 - beta^step in double like torch).
 */
#ifndef WGS_ADAM_H_
#define WGS_ADAM_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define WGS_ADAM_MAX_TENSORS 16

typedef struct WgsAdamTensor {
  float* param;            /* (n) updated in place */
  const float* grad;       /* (n) */
  float* exp_avg;          /* (n) updated in place */
  float* exp_avg_sq;       /* (n) updated in place */
  uint64_t n;
  float step_size;         /* lr / (1 - beta1^step); with step_dev: the learning rate itself (unless lr_dev is given) */
  float bias_correction2_sqrt; /* sqrt(1 - beta2^step); ignored with step_dev */
  /* Graph-capturable mode (torch's `capturable=True`): the step count lives on the device, so a captured launch
   * advances it on every replay.  step_dev: 1 float, the number of steps taken so far -- the kernel forms both bias
   * corrections from step_dev[0] + 1 in double and a second launch increments the counters (descriptors of one call
   * may share a counter: it is advanced once).  lr_dev (optional):
   * 1 float read at run time, for learning-rate schedules under a captured graph.  Both null: host-side mode. */
  float* step_dev;
  const float* lr_dev;
} WgsAdamTensor;

/* tensors: HOST array of `count` (<= WGS_ADAM_MAX_TENSORS) descriptors with DEVICE pointers */
/* betas and eps are doubles so that 1 - beta is formed in double, as torch does (lerp weight / addcmul value) */
int wgs_adam_step(const WgsAdamTensor* tensors, int32_t count, double beta1, double beta2, double eps,
                  void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_ADAM_H_ */
