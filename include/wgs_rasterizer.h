/*
 * wgs_rasterizer.h — C ABI of the B200-native differentiable Gaussian rasterizer with
 * camera-pose gradients (libwgs_rasterizer.so, sm_100a only).
 *
 * This is the drop-in boundary for ONE path of yuhang1008/WildGS-SLAM-Blackwell: the
 * native layer under `diff_gaussian_rasterization` (reference:
 * thirdparty/diff-gaussian-rasterization-w-pose, "RAST/" below).  Each entry point names
 * the reference interface it replaces.  No C++/torch types cross the boundary: plain
 * pointers (DEVICE pointers unless the field says host), sizes and scalars.  All work is
 * enqueued on the caller's stream; kernels never allocate.
 *
 * Memory is caller-owned (the Python host keeps it in torch uint8 tensors so autograd
 * lifetime rules match RAST/rasterize_points.cu:27-33,69-82): query sizes with
 * wgs_workspace_sizes() and pass the five workspaces in.  Every workspace pointer must be
 * 256-byte aligned.
 */
#ifndef WGS_RASTERIZER_H_
#define WGS_RASTERIZER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WGS_ABI_VERSION 11

/* status codes (negative = error) */
#define WGS_OK 0
#define WGS_ERR_INVALID_ARGUMENT (-1) /* bad shape/null pointer; reference: AT_ERROR, rasterize_points.cu:58-60 */
#define WGS_ERR_CUDA (-2)             /* a CUDA call failed; see wgs_last_error() */
#define WGS_ERR_CAPACITY (-3)         /* num_rendered > instance capacity; *num_rendered holds the needed size; grow `binning` and call again */
#define WGS_ERR_NO_COLOR_SOURCE (-4)  /* neither sh nor colors_precomp; reference: std::runtime_error, rasterizer_impl.cu:245-248 */
#define WGS_ERR_ARCH (-5)             /* device is not sm_100 */

typedef struct WgsContext WgsContext; /* per-device: pinned readback word, event, SM count */

/* Workspace byte sizes for a problem of P Gaussians, W x H pixels and room for
 * `capacity` (Gaussian,tile) instances.  Replaces required<GeometryState/BinningState/
 * ImageState>() (RAST/cuda_rasterizer/rasterizer_impl.h:64-72, rasterizer_impl.cu:155-194).
 * geom: 57 B/Gaussian; tile: 268 B/tile (up to 32 privatised histogram + cursor copies); binning: 8 B/instance; image: 8 B/pixel;
 * grad: 48 B/Gaussian + pose partials (the reference: 79 B/Gaussian, 24 B/instance + CUB
 * temporaries, 16 B/pixel, and eleven zero-filled gradient tensors per backward). */
typedef struct WgsWorkspaceSizes {
  size_t geom_bytes;
  size_t tile_bytes;
  size_t binning_bytes;
  size_t image_bytes;
  size_t grad_bytes;
} WgsWorkspaceSizes;

/* Arguments of the forward pass.  Replaces RasterizeGaussiansCUDA
 * (RAST/rasterize_points.cu:35-122) + CudaRasterizer::Rasterizer::forward
 * (RAST/cuda_rasterizer/rasterizer.h:32-58, rasterizer_impl.cu:198-344).
 * A null pointer means "absent", as the reference's empty tensors do. */
typedef struct WgsForwardArgs {
  uint32_t struct_size; /* = sizeof(WgsForwardArgs) of the header the caller was built against; a binding made
                           for an older (shorter) layout is rejected with WGS_ERR_INVALID_ARGUMENT instead of
                           being read past its end */
  int32_t P;          /* number of Gaussians (0 is legal: outputs are background-only) */
  int32_t sh_degree;  /* active SH degree D (0..3) */
  int32_t M;          /* SH coefficients stored per Gaussian (sh is (P,M,3)); 0 if sh absent */
  int32_t width, height;
  float tanfovx, tanfovy;
  float scale_modifier;
  int32_t prefiltered; /* reference traps on a culled point when set (auxiliary.h:154-161); we set radii=0 and report via status word */
  int32_t debug;       /* synchronise + check after every stage (CHECK_CUDA, auxiliary.h:166-173) */
  /* inputs (device) */
  const float* background;     /* (3,) */
  const float* means3D;        /* (P,3) */
  const float* sh;             /* (P,M,3) or null */
  const float* colors_precomp; /* (P,3) or null */
  const float* opacities;      /* (P,1) */
  const float* scales;         /* (P,3) or null */
  const float* rotations;      /* (P,4) w,x,y,z; NOT re-normalised (forward.cu:129) or null */
  const float* cov3D_precomp;  /* (P,6) or null */
  const float* viewmatrix;     /* (4,4) = W2C^T, i.e. column-major W2C */
  const float* projmatrix;     /* (4,4) = viewmatrix @ P^T */
  const float* campos;         /* (3,) */
  /* outputs (device, caller-allocated, need NOT be pre-zeroed) */
  float* out_color;   /* (3,H,W) planar */
  float* out_depth;   /* (1,H,W) */
  float* out_opacity; /* (1,H,W) */
  int32_t* radii;     /* (P,) */
  int32_t* n_touched; /* (P,) */
  /* workspaces (device) */
  void* geom;
  void* tile;
  void* binning;
  void* image;
  int64_t capacity;   /* instance capacity the binning workspace was sized for */
  /* Tile-row strip [tile_row_begin, tile_row_end) to render (rows of 16 pixels; both 0 = whole image).
   * New capability for sub-frame sharding of a keyframe window across GPUs (no reference
   * counterpart): only the strip's tiles are binned and composited, only the strip's pixels of the
   * outputs are written, radii keep their whole-image meaning, n_touched counts the strip's pixels.
   * Images, n_touched and (in the backward) all gradients of the strips of a partition of the image sum
   * to the whole-image results. */
  int32_t tile_row_begin, tile_row_end;
  /* Raw-parameter mode (the activations the reference's render() wrapper
   * applies with separate torch kernels before every rasterizer call,
   * thirdparty/gaussian_splatting/scene/gaussian_model.py:54-63,77-97, fused into the per-Gaussian stage).
   * activations != 0: `opacities` holds GaussianModel._opacity (sigmoid applied here), `scales` holds
   * _scaling (exp applied here), `rotations` holds _rotation (normalised here, eps 1e-12 like
   * torch.nn.functional.normalize).  sh_rest != null: `sh` holds _features_dc (P,1,3) and `sh_rest`
   * _features_rest (P,M-1,3) -- the torch.cat of get_features is never materialised; M stays the total. */
  int32_t activations;
  const float* sh_rest;
} WgsForwardArgs;

/* Arguments of the backward pass.  Replaces RasterizeGaussiansBackwardCUDA
 * (RAST/rasterize_points.cu:124-211) + CudaRasterizer::Rasterizer::backward
 * (rasterizer.h:60-91, rasterizer_impl.cu:348-455) + the torch.sum over dL_dtau in
 * RAST/diff_gaussian_rasterization/__init__.py:152-154 (folded into the kernel). */
typedef struct WgsBackwardArgs {
  uint32_t struct_size; /* = sizeof(WgsBackwardArgs), checked like WgsForwardArgs::struct_size */
  int32_t P, sh_degree, M, width, height;
  float tanfovx, tanfovy, scale_modifier;
  int32_t debug;
  int32_t accumulate; /* 0: gradient outputs are overwritten; 1: added to (lets a keyframe window sum the
                         per-view Gaussian gradients in place; dL_dtau is always overwritten); 2: added to with
                         reduction atomics, so that backwards running concurrently on several streams may share
                         the output buffers (the caller zeroes them first) */
  const float* background;
  const float* means3D;
  const float* sh;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* projmatrix_raw; /* (4,4) = P^T */
  const float* campos;
  const int32_t* radii;        /* forward output */
  /* cotangents (device); dL_dopacity-image is ignored like the reference (__init__.py:108) */
  const float* dL_dout_color;  /* (3,H,W) */
  const float* dL_dout_depth;  /* (1,H,W) */
  /* gradient outputs (device, caller-allocated, need NOT be pre-zeroed unless accumulate; any may be null) */
  float* dL_dmeans3D;   /* (P,3) */
  float* dL_dmeans2D;   /* (P,3), z = 0 */
  float* dL_dsh;        /* (P,M,3)   (sh path) */
  float* dL_dcolors;    /* (P,3)     (colors_precomp path) */
  float* dL_dopacity;   /* (P,1) */
  float* dL_dscales;    /* (P,3) */
  float* dL_drotations; /* (P,4) */
  float* dL_dcov3D;     /* (P,6)     (cov3D_precomp path) */
  float* dL_dtau;       /* (6,) rho then theta — already summed over Gaussians */
  /* workspaces written by the matching forward */
  const void* geom;
  const void* tile;
  const void* binning;
  const void* image;
  void* grad;           /* accumulators; must be zero before the FIRST backward that uses it (kernels re-zero it) */
  int64_t capacity;
  int32_t tile_row_begin, tile_row_end; /* must equal the matching forward's */
  /* Raw-parameter mode, must equal the matching forward's.  With activations != 0, dL_dopacity,
   * dL_dscales and dL_drotations are gradients w.r.t. the RAW parameters (chain rule through sigmoid /
   * exp / normalize applied in the kernel).  With sh_rest != null, dL_dsh is (P,1,3) and dL_dsh_rest
   * (P,M-1,3). */
  int32_t activations;
  const float* sh_rest;
  float* dL_dsh_rest;
  /* Densification statistics fused into the per-Gaussian backward (any may be null).  For every Gaussian with
   * radii > 0 -- the visibility_filter of the callers -- this backward does what the mapper runs after it
   * (src/mapper.py:1176-1182, gaussian_model.py:745-749):
   *   stat_grad_accum[i] += || dL_dmeans2D[i, :2] ||      (xyz_gradient_accum, (P,1))
   *   stat_denom[i]      += 1                             (denom, (P,1))
   *   stat_max_radii[i]   = max(stat_max_radii[i], radii[i])   (max_radii2D, (P,) float)
   * Whole-image backward only (tile_row_begin/end covering all rows): the norm is per view. */
  float* stat_grad_accum;
  float* stat_denom;
  float* stat_max_radii;
} WgsBackwardArgs;

/* library / device */
int wgs_abi_version(void);
const char* wgs_last_error(void);
WgsContext* wgs_context_create(int device);  /* null on failure (see wgs_last_error) */
void wgs_context_destroy(WgsContext* ctx);

int wgs_workspace_sizes(int32_t P, int32_t width, int32_t height, int64_t capacity,
                        WgsWorkspaceSizes* out);

/* Forward.  On return *num_rendered (host) holds R = number of (Gaussian,tile) instances
 * (the reference's return value, rasterizer_impl.cu:343).  The only host wait is on the
 * 4-byte R word (written to pinned memory by the tile-scan kernel), and it happens after all
 * kernels of the forward are already enqueued.  The `tile` workspace needs no initialisation.
 * Returns WGS_ERR_CAPACITY (outputs invalid, nothing written out of bounds) when R exceeds
 * args->capacity; the caller grows `binning` to *num_rendered entries and calls again. */
int wgs_forward(WgsContext* ctx, const WgsForwardArgs* args, void* cuda_stream,
                int64_t* num_rendered);

int wgs_backward(WgsContext* ctx, const WgsBackwardArgs* args, void* cuda_stream);

/* No-wait operation.  The reference blocks the host once per forward on a D2H copy of R
 * (rasterizer_impl.cu:283-284); wgs_forward waits on a 4-byte pinned word instead.  When the stream is
 * being captured into a CUDA graph, or after wgs_set_async(ctx, 1), wgs_forward does not wait at all:
 * *num_rendered is -1.  A forward whose instance count exceeds `capacity` then writes the background
 * colour / zero depth and opacity (its backward produces zero gradients) and records the overflow in a
 * STICKY per-context device word: wgs_overflow_status() reports, after the caller has synchronised, how
 * many forwards on this context overflowed since the word was last cleared and the largest instance count
 * any of them needed -- so one check after 100 graph replays, or after a window of several views per
 * replay, sees an overflow of ANY of them (wgs_last_num_rendered only knows the most recent forward).
 * Returns WGS_ERR_CAPACITY when the count is non-zero; `clear` != 0 resets the word. */
int wgs_set_async(WgsContext* ctx, int enable);
int wgs_last_num_rendered(WgsContext* ctx, int64_t* num_rendered, int64_t* capacity);
int wgs_overflow_status(WgsContext* ctx, int64_t* overflowed_forwards, int64_t* max_num_rendered, int clear);

/* Chained launches (programmatic dependent launch; no reference counterpart).  The kernels of a forward and of a
 * backward are launched with cudaLaunchAttributeProgrammaticStreamSerialization: each starts with
 * griddepcontrol.wait (all data of the preceding kernel visible) and lets the next grid take the SM slots its own last
 * wave frees, which removes the drain / launch / ramp-up gap between the 5 + 2 kernels of a frame (config 2: 435 ->
 * 427 us).  The CTAs of the waiting grid hold their registers and shared memory, so when a SECOND stream has work that
 * could use those slots (window.WindowTrainer with two streams per rank) the switch should be off: measured 5.41 ms
 * per window step without, 6.00 ms with.  Process-wide; affects launches (and graph captures) made after the call;
 * returns the previous setting.  Default: on, or the environment's WGS_PDL=0/1. */
int wgs_set_chained_launch(int enable);

/* Debug counting pass for the roofline report (no reference counterpart): replays the
 * compositing of the forward that `args` describes and returns, in two device u64 words,
 *   counts[0] = N_eval  = sum over pixels of n_contrib (list entries a pixel must visit)
 *   counts[1] = N_blend = (pixel, Gaussian) pairs that pass both alpha tests and are composited.
 * Not part of any timed path. */
int wgs_count_pairs(const WgsForwardArgs* args, uint64_t* counts, void* cuda_stream);

/* Replaces markVisible (RAST/rasterize_points.cu:213-232, rasterizer_impl.cu:54-66,141-153):
 * present[i] = view-space z > 0.2.  `present` is (P,) bytes (torch.bool). */
int wgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* cuda_stream);

/* Parity/debug export of the binning state in the REFERENCE's format
 * (BinningState / ImageState, rasterizer_impl.h:42-62): sorted 64-bit keys
 * tile<<32|depth_bits, sorted Gaussian ids, per-tile [start,end) with (0,0) for empty tiles,
 * and the per-pixel n_contrib / final_T, reconstructed from the state of the forward that
 * `args` describes.  Destinations are device pointers; any may be null. */
int wgs_export_binning(const WgsForwardArgs* args, int64_t num_rendered,
                       uint64_t* point_list_keys, uint32_t* point_list, uint32_t* ranges,
                       uint32_t* n_contrib, float* final_T, void* cuda_stream);

/* Parity/debug export of the per-Gaussian forward state in the reference's GeometryState
 * terms (rasterizer_impl.h:21-40): depths (P), means2D (2P), conic_opacity (4P), rgb (3P),
 * clamped (3P bytes).  Entries of culled Gaussians (radii==0) are written as 0. */
int wgs_export_geometry(const WgsForwardArgs* args, float* depths, float* means2D,
                        float* conic_opacity, float* rgb, uint8_t* clamped, void* cuda_stream);

/* Stage timing for the roofline report (no reference counterpart: the reference has no timers,
 * SURVEY.md §5).  When enabled, wgs_forward/wgs_backward bracket every stage with CUDA events
 * on the caller's stream; wgs_stage_times() synchronises on the last event and returns the
 * device time of each stage of the most recent forward (+ backward) in milliseconds:
 *   [0] preprocess_fwd  [1] tile_scan  [2] scatter  [3] tile_sort (small+large)
 *   [4] render_fwd      [5] render_bwd [6] preprocess_bwd   [7] reserved (0)
 * Stages that did not run since the last call report 0. */
#define WGS_NUM_STAGES 8
/* Camera matrices of one view in one launch.  Replaces the reference Camera's
 * per-access properties world_view_transform / full_proj_transform / camera_center
 * (src/utils/camera_utils.py:137-151; graphics_utils.py:33-46): R (3,3) row-major and T (3,) of the W2C
 * pose, projection_matrix (4,4) = P^T as stored by the reference; outputs viewmatrix (4,4) = W2C^T,
 * projmatrix (4,4) = viewmatrix @ projection_matrix, campos (3,) = -R^T T.  All device pointers. */
int wgs_camera_matrices(const float* R, const float* T, const float* projection_matrix, float* viewmatrix,
                        float* projmatrix, float* campos, void* cuda_stream);

/* Multi-GPU: in-place sum of the Gaussian-gradient bucket over the `world` ranks of one
 * NVLink/NVSwitch domain, through peer-mapped memory instead of ncclAllReduce.  buffer_ptrs[r] / flag_ptrs[r]
 * (host arrays of `world` device addresses) are THIS rank's mappings of rank r's bucket and flag array
 * (wgs_peer_flags_bytes() bytes, zeroed once); multicast_ptr is the NVSwitch multicast mapping of the bucket
 * or 0 (then plain peer loads/stores are used).  n_floats must be a multiple of 4.  Every rank enqueues the
 * same sequence of calls (the barrier generation is kept on the device, so the launch can be captured into a
 * CUDA graph); results are bit-identical on all ranks. */
size_t wgs_peer_flags_bytes(void);
int wgs_peer_allreduce(const uint64_t* buffer_ptrs, const uint64_t* flag_ptrs, uint64_t multicast_ptr, int32_t rank,
                       int32_t world, uint64_t n_floats, void* cuda_stream);
/* The general form.  The exchanged data are `n_segments` (<= 12) pieces of the buffer, seg_offsets[k] / seg_lengths[k]
 * in floats (multiples of 4) -- a capacity-headroom arena keeps every leaf's gradient in its own segment and only the
 * used rows travel -- laid end to end as one virtual array of which rank r owns the r-th 1/world slice.
 *   mode 0  allreduce       every copy ends up holding the sum
 *   mode 1  reduce-scatter  this rank's copy of ITS slice receives the sum over all ranks (the other slices are untouched):
 *                           the first half of a sharded optimiser step -- Adam then runs on 1/world of the model per rank
 *   mode 2  allgather       this rank's copy of its slice is written into every rank's copy (on another buffer of the same
 *                           layout, the updated parameters return to all ranks)
 *   mode 3  rendezvous      nothing moves: ONE block runs the two peer barriers.  In front of an exchange whose peers may
 *                           arrive late (behind their own H2D copies, say) it does the waiting with one block, not with
 *                           the whole exchange grid spinning on SM slots a concurrent stream's kernels could use
 * Every call is bracketed by the block-level peer barriers of wgs_peer_allreduce; flag_ptrs may belong to a different
 * allocation than buffer_ptrs (one flag area serves all buffers exchanged in stream order). */
int wgs_peer_exchange(const uint64_t* buffer_ptrs, const uint64_t* flag_ptrs, uint64_t multicast_ptr, int32_t rank,
                      int32_t world, const uint64_t* seg_offsets, const uint64_t* seg_lengths, int32_t n_segments,
                      int32_t mode, void* cuda_stream);

/* Launch shape of the exchange kernel: `blocks` CTAs of 512 threads, `unroll` float4 per thread and trip (bytes in flight
 * on NVLink = blocks x 512 x unroll x 16).  Process-wide; EVERY rank must use the same setting for a given flag area (block
 * b of one rank meets block b of the others).  0 selects the measured default: 48 x 2 on the NVSwitch-multimem path (the
 * switch is the bound: 59 / 146 / 290 us for 16.8 / 56 / 128 MB at 8 GPUs, 775 GB/s bus bandwidth; NCCL 100 / 244 / 390),
 * 144 x 4 on the peer load/store path.  Environment WGS_PEER_BLOCKS / WGS_PEER_UNROLL are read by the Python layer. */
int wgs_peer_configure(int32_t blocks, int32_t unroll);

int wgs_stage_timing(WgsContext* ctx, int enable);
int wgs_stage_times(WgsContext* ctx, float* ms_out /* WGS_NUM_STAGES floats, host */);

#ifdef __cplusplus
}
#endif
#endif /* WGS_RASTERIZER_H_ */
