/*
 * wgs_pose.h — C ABI of the fused pose step (libwgs_rasterizer.so).  SURVEY.md §8(f) row 3 ("pose update
 * from the fused 6-vector"): what the pose-refinement loop does between two renders
 * (src/mapper.py:905-907: pose_optimizer.step(); update_pose(viewpoint)), ~45 tiny torch kernels and a host
 * synchronisation in the reference, as ONE single-warp kernel with no host involvement:
 *   1. torch.optim.Adam step (default betas/eps unless given) on cam_rot_delta (3), cam_trans_delta (3),
 *      exposure_a (1), exposure_b (1), each with its own learning rate (src/mapper.py:861-884); skipped when
 *      `adam` is null;
 *   2. update_pose (src/utils/pose_utils.py:81-100): tau = [trans_delta, rot_delta],
 *      new_w2c = SE3_exp(tau) @ [R T; 0 1] (SO3_exp / V with their angle < 1e-5 branches, :33-63), R and T
 *      overwritten IN PLACE, converged = |tau| < threshold, both deltas zeroed;
 *   3. the camera matrices the next render needs (wgs_camera_matrices), when `viewmatrix` is given.
 * The Adam step counter lives on the device, so the launch is identical every iteration and can be part of a
 * CUDA graph.  All pointers are DEVICE pointers (float32 unless noted).
 */
#ifndef WGS_POSE_H_
#define WGS_POSE_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Adam state of the 8 scalars, order: rot_delta[3], trans_delta[3], exposure_a, exposure_b */
typedef struct WgsPoseAdam {
  float* exp_avg;          /* (8) */
  float* exp_avg_sq;       /* (8) */
  int32_t* step;           /* (1) incremented by the kernel */
  const float* grad_rot;   /* (3) */
  const float* grad_trans; /* (3) */
  const float* grad_exposure_a; /* (1) or null: exposure not optimised */
  const float* grad_exposure_b; /* (1) or null */
  double lr_rot, lr_trans, lr_exposure_a, lr_exposure_b; /* doubles: lr / bias_correction1 is formed in double like torch */
  double beta1, beta2, eps;
} WgsPoseAdam;

typedef struct WgsPoseStepArgs {
  float* cam_rot_delta;    /* (3) theta; zeroed on return */
  float* cam_trans_delta;  /* (3) rho; zeroed on return */
  float* exposure_a;       /* (1) or null */
  float* exposure_b;       /* (1) or null */
  float* R;                /* (3,3) row-major W2C rotation, updated in place */
  float* T;                /* (3)   W2C translation, updated in place */
  float converged_threshold; /* 1e-4 in the reference */
  int32_t* converged;      /* (1) out: 1 if |tau| < threshold */
  const WgsPoseAdam* adam; /* HOST pointer or null (update_pose only) */
  const float* projection_matrix; /* (4,4) = P^T, required iff viewmatrix is given */
  float* viewmatrix;       /* (4,4) out or null */
  float* projmatrix;       /* (4,4) out */
  float* campos;           /* (3) out */
} WgsPoseStepArgs;

int wgs_pose_step(const WgsPoseStepArgs* args, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_POSE_H_ */
