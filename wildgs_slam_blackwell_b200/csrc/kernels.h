// Kernel parameter blocks and host launchers (internal; the public surface is
// include/wgs_rasterizer.h).
#pragma once
#include "common.cuh"

namespace wgs {

struct PreFwdParams {
    int P, D, M, W, H, gx, gy, prefiltered, copies;
    int row0, row1;  // tile-row strip to bin (whole image: 0, gy)
    int act;         // raw-parameter mode: sigmoid / exp / normalize applied here
    const float* sh_rest;  // non-null: sh = (P,1,3) DC, sh_rest = (P,M-1,3)
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *sh, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *viewmatrix, *projmatrix, *campos;
    int* radii;
    int* n_touched;
    GeomView geom;
    TileView tile;
};

struct BinParams {
    int P, gx, gy, tiles, copies;
    long long capacity;
    GeomView geom;
    TileView tile;
    unsigned long long* bucket;
    unsigned int* host_R;  // mapped pinned word (may be null)
    unsigned int* sticky;  // context-level overflow record {forwards that overflowed, largest R among them} (may be null)
};

struct RenderFwdParams {
    int W, H, gx, gy;
    int tile0, ntiles;  // first tile and number of tiles of the strip (whole image: 0, gx*gy)
    const float* bg;
    GeomView geom;
    TileView tile;
    ImageView image;
    const unsigned long long* bucket;
    float *out_color, *out_depth, *out_opacity;
    int* n_touched;
};

struct RenderBwdParams {
    int W, H, gx, gy;
    int tile0, ntiles;
    const float* bg;
    GeomView geom;
    TileView tile;
    ImageView image;
    const unsigned long long* bucket;
    const float *dL_dcolor, *dL_ddepth;
    GradView grad;
};

struct PreBwdParams {
    int P, D, M, W, H, accumulate;
    int act;
    int strip;                 // the backward of a tile-row strip (few visible Gaussians): the queued kernel variant
    const float* sh_rest;
    float* dL_dsh_rest;
    float *stat_grad_accum, *stat_denom, *stat_max_radii;   // fused densification statistics (may be null)
    const uint32_t* misc;      // the forward's TileView::misc: statistics are skipped when MISC_OVERFLOW is set
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *sh, *colors_precomp, *scales, *rotations, *cov3D_precomp;
    const float *viewmatrix, *projmatrix, *projmatrix_raw, *campos;
    const int* radii;
    GeomView geom;
    GradView grad;
    float *dL_dmeans3D, *dL_dmeans2D, *dL_dsh, *dL_dcolors, *dL_dopacity, *dL_dscales, *dL_drotations,
        *dL_dcov3D, *dL_dtau;
};

void launch_preprocess_fwd(const PreFwdParams& p, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* V, uint8_t* present, cudaStream_t s);
void launch_tile_scan(const BinParams& p, cudaStream_t s);   // ranges + R
void launch_scatter(const BinParams& p, cudaStream_t s);
void launch_tile_sorts(const BinParams& p, int num_sms, cudaStream_t s);
void launch_render_fwd(const RenderFwdParams& p, cudaStream_t s);
void launch_render_bwd(const RenderBwdParams& p, cudaStream_t s);
void launch_preprocess_bwd(const PreBwdParams& p, cudaStream_t s);
void launch_export_binning(const BinParams& p, const ImageView& image, int W, int H, unsigned long long* keys,
                           unsigned int* list, unsigned int* ranges, unsigned int* n_contrib, float* final_T,
                           cudaStream_t s);
void launch_count_pairs(const RenderFwdParams& p, unsigned long long* counts, cudaStream_t s);
void launch_export_geometry(int P, const GeomView& g, const int* radii, float* depths, float* means2D,
                            float* conic_opacity, float* rgb, uint8_t* clamped, cudaStream_t s);
int configure_binning();  // one-time cudaFuncSetAttribute calls; returns cudaError_t
int configure_render_bwd();
int configure_preprocess_bwd();
int configure_preprocess_fwd();

}  // namespace wgs
