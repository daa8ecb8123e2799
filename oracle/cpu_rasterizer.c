/*
 * TEST INFRASTRUCTURE — CPU restatement (plain C + OpenMP) of the reference's differentiable
 * Gaussian rasterizer with pose gradients.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline leg of bench.py may load this; the product never does.
 *
 * Parity: PINNED.  tests/test_cpu_oracle.py checks this file against the golden vectors the
 * UNMODIFIED reference (built for sm_100, oracle/ref_cuda/build_ref.py) produced on a B200
 * (tests/golden/ *.npz): radii / depth bits / means2D / conics / rgb / sorted keys / sorted ids /
 * tile ranges bit-exact; images and gradients within tolerance (host libm expf differs from
 * CUDA's by <= 2 ulp, so a handful of alpha-threshold decisions may flip).
 *
 * Reference files restated (paths under thirdparty/diff-gaussian-rasterization-w-pose/):
 *   cuda_rasterizer/auxiliary.h:22-56,58-97,107-132,139-164   constants, ndc2Pix, getRect, in_frustum, dnormvdv
 *   cuda_rasterizer/forward.cu:22-73     computeColorFromSH        -> sh_to_rgb()
 *   cuda_rasterizer/forward.cu:76-115    computeCov2D              -> preprocess_one()
 *   cuda_rasterizer/forward.cu:120-154   computeCov3D              -> cov3d_fwd()
 *   cuda_rasterizer/forward.cu:157-258   preprocessCUDA            -> preprocess_one()
 *   cuda_rasterizer/rasterizer_impl.cu:70-138  duplicateWithKeys / sort / identifyTileRanges -> bin()
 *   cuda_rasterizer/forward.cu:263-392   renderCUDA (forward)      -> render_fwd_tile()
 *   cuda_rasterizer/backward.cu:563-787  renderCUDA (backward)     -> render_bwd_tile()
 *   cuda_rasterizer/backward.cu:150-346  computeCov2DCUDA          -> preprocess_bwd_one()
 *   cuda_rasterizer/backward.cu:418-539, 21-145, 350-413  preprocess / SH / cov3D backward
 *   diff_gaussian_rasterization/__init__.py:152-154  sum of dL_dtau over Gaussians
 *
 * Arithmetic: float32.  The FORWARD per-Gaussian stage is written with explicit fmaf() in the
 * contraction pattern nvcc 12.9 (-fmad=true) emits for the reference sources, and this file is
 * compiled with -ffp-contract=off, so those values are bit-identical to the GPU reference
 * (IEEE sqrt/div/rcp/ceil are exact on both sides).  The backward is the reference's formulas
 * in float32 with per-Gaussian sums accumulated in double (the more accurate side of a
 * tolerance check).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16
#define EXPORT __attribute__((visibility("default")))

typedef struct OrcScene {
    int32_t P, D, M, W, H;
    float tanfovx, tanfovy, scale_modifier;
    const float *bg, *means3D, *sh, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *viewmatrix, *projmatrix, *projmatrix_raw, *campos;
} OrcScene;

typedef struct OrcState {
    OrcScene sc;
    int gx, gy, tiles;
    int64_t R;
    /* GeometryState (rasterizer_impl.h:21-40) */
    float *depths, *means2D, *conic_opacity, *rgb, *cov3D;
    uint8_t* clamped;
    int32_t* radii;
    uint32_t* rect; /* x0,y0,x1,y1 per Gaussian */
    /* BinningState / ImageState (rasterizer_impl.h:42-62) */
    uint64_t* keys;
    uint32_t* list;
    uint32_t* ranges; /* 2 per tile */
    uint32_t* n_contrib;
    float* final_T;
} OrcState;

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

/* transformPoint4x3 / 4x4 rows (auxiliary.h:58-77) as contracted by nvcc */
static inline float dot3p(float x, float y, float z, float a, float b, float c, float d) {
    return d + fmaf(z, c, fmaf(x, a, y * b));
}
/* 3-term dot of GLM mat3 products as contracted by nvcc */
static inline float dot3g(float a0, float b0, float a1, float b1, float a2, float b2) {
    return fmaf(a2, b2, fmaf(a0, b0, a1 * b1));
}
/* auxiliary.h:41-44 (double because of the 1.0 / 0.5 literals) */
static inline float ndc2pix(float v, int S) { return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5); }

static inline uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:46-56 */
static void get_rect(float px, float py, int radius, int gx, int gy, uint32_t* r) {
    const float rf = (float)radius;
    r[0] = umin((uint32_t)gx, (uint32_t)imax(0, (int)((px - rf) / TILE)));
    r[1] = umin((uint32_t)gy, (uint32_t)imax(0, (int)((py - rf) / TILE)));
    r[2] = umin((uint32_t)gx, (uint32_t)imax(0, (int)((px + rf + TILE - 1) / TILE)));
    r[3] = umin((uint32_t)gy, (uint32_t)imax(0, (int)((py + rf + TILE - 1) / TILE)));
}

/* forward.cu:120-154; quaternion (r,x,y,z) NOT normalised */
static void cov3d_fwd(const float* s3, float mod, const float* q, float* c) {
    const float sx = mod * s3[0], sy = mod * s3[1], sz = mod * s3[2];
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    const float yy = y * y, zz = z * z, xz = x * z, rx = r * x, rz = r * z;
    const float yy_zz = yy + zz, xx_zz = fmaf(x, x, zz), xx_yy = fmaf(x, x, yy);
    const float xy_m_rz = fmaf(x, y, -rz), xy_p_rz = fmaf(x, y, rz), ry_p_xz = fmaf(r, y, xz),
                xz_m_ry = fmaf(-r, y, xz), yz_m_rx = fmaf(y, z, -rx), rx_p_yz = fmaf(y, z, rx);
    const float R00 = 1.f - (yy_zz + yy_zz), R01 = xy_m_rz + xy_m_rz, R02 = ry_p_xz + ry_p_xz;
    const float R10 = xy_p_rz + xy_p_rz, R11 = 1.f - (xx_zz + xx_zz), R12 = yz_m_rx + yz_m_rx;
    const float R20 = xz_m_ry + xz_m_ry, R21 = rx_p_yz + rx_p_yz, R22 = 1.f - (xx_yy + xx_yy);
    const float M00 = sx * R00, M01 = sy * R01, M02 = sz * R02;
    const float M10 = sx * R10, M11 = sy * R11, M12 = sz * R12;
    const float M20 = sx * R20, M21 = sy * R21, M22 = sz * R22;
    c[0] = dot3g(M00, M00, M01, M01, M02, M02);
    c[1] = dot3g(M10, M00, M11, M01, M12, M02);
    c[2] = dot3g(M20, M00, M21, M01, M22, M02);
    c[3] = dot3g(M10, M10, M11, M11, M12, M12);
    c[4] = dot3g(M20, M10, M21, M11, M22, M12);
    c[5] = dot3g(M20, M20, M21, M21, M22, M22);
}

/* forward.cu:22-73 */
static void sh_to_rgb(const float* sh, int deg, float px, float py, float pz, const float* campos, float* rgb,
                      uint8_t* clamped) {
    const float dx = px - campos[0], dy = py - campos[1], dz = pz - campos[2];
    const float len = sqrtf(fmaf(dz, dz, fmaf(dx, dx, dy * dy)));
    const float x = dx / len, y = dy / len, z = dz / len;
    float res[3];
    for (int c = 0; c < 3; c++) res[c] = sh[c] * SH_C0;
    if (deg > 0) {
        const float t1 = y * SH_C1, t2 = z * SH_C1, t3 = x * SH_C1;
        for (int c = 0; c < 3; c++) {
            float v = fmaf(-t1, sh[3 + c], res[c]);
            v = fmaf(t2, sh[6 + c], v);
            res[c] = fmaf(-t3, sh[9 + c], v);
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const float zz2 = zz + zz, xx_yy = xx - yy;
            const float b4 = xy * SH_C2[0], b5 = yz * SH_C2[1], b6 = ((zz2 - xx) - yy) * SH_C2[2], b7 = xz * SH_C2[3],
                        b8 = xx_yy * SH_C2[4];
            for (int c = 0; c < 3; c++) {
                float v = fmaf(b4, sh[12 + c], res[c]);
                v = fmaf(b5, sh[15 + c], v);
                v = fmaf(b6, sh[18 + c], v);
                v = fmaf(b7, sh[21 + c], v);
                res[c] = fmaf(b8, sh[24 + c], v);
            }
            if (deg > 2) {
                const float w = fmaf(zz, 4.f, -xx) - yy;
                const float b9 = (y * SH_C3[0]) * fmaf(xx, 3.f, -yy);
                const float b10 = z * (xy * SH_C3[1]);
                const float b11 = (y * SH_C3[2]) * w;
                const float b12 = (z * SH_C3[3]) * fmaf(yy, -3.f, fmaf(xx, -3.f, zz2));
                const float b13 = (x * SH_C3[4]) * w;
                const float b14 = (z * SH_C3[5]) * xx_yy;
                const float b15 = (x * SH_C3[6]) * fmaf(yy, -3.f, xx);
                for (int c = 0; c < 3; c++) {
                    float v = fmaf(b9, sh[27 + c], res[c]);
                    v = fmaf(b10, sh[30 + c], v);
                    v = fmaf(b11, sh[33 + c], v);
                    v = fmaf(b12, sh[36 + c], v);
                    v = fmaf(b13, sh[39 + c], v);
                    v = fmaf(b14, sh[42 + c], v);
                    res[c] = fmaf(b15, sh[45 + c], v);
                }
            }
        }
    }
    for (int c = 0; c < 3; c++) {
        res[c] += 0.5f;
        clamped[c] = res[c] < 0.f;
        rgb[c] = res[c] < 0.f ? 0.f : res[c];
    }
}

/* forward.cu:157-258 (+ computeCov2D :76-115).  Culled Gaussians keep radius 0 and zeroed state
 * (the reference leaves those entries stale; the golden files zero them too). */
static void preprocess_one(OrcState* st, int idx, float focal_x, float focal_y) {
    const OrcScene* s = &st->sc;
    const float* V = s->viewmatrix;
    const float* Pm = s->projmatrix;
    const float px = s->means3D[3 * idx], py = s->means3D[3 * idx + 1], pz = s->means3D[3 * idx + 2];
    st->radii[idx] = 0;
    uint32_t* rect = st->rect + 4 * idx;
    rect[0] = rect[1] = rect[2] = rect[3] = 0;
    const float depth = dot3p(px, py, pz, V[2], V[6], V[10], V[14]);
    if (depth <= 0.2f) return; /* in_frustum, auxiliary.h:139-164 */
    const float hx = dot3p(px, py, pz, Pm[0], Pm[4], Pm[8], Pm[12]);
    const float hy = dot3p(px, py, pz, Pm[1], Pm[5], Pm[9], Pm[13]);
    const float hw = dot3p(px, py, pz, Pm[3], Pm[7], Pm[11], Pm[15]);
    const float pw = 1.0f / (hw + 0.0000001f);
    const float projx = hx * pw, projy = hy * pw;
    float c[6];
    if (s->cov3D_precomp) {
        for (int k = 0; k < 6; k++) c[k] = s->cov3D_precomp[6 * idx + k];
    } else {
        cov3d_fwd(s->scales + 3 * idx, s->scale_modifier, s->rotations + 4 * idx, c);
    }
    const float tx = dot3p(px, py, pz, V[0], V[4], V[8], V[12]);
    const float ty = dot3p(px, py, pz, V[1], V[5], V[9], V[13]);
    const float tz = depth;
    const float limx = s->tanfovx * 1.3f, limy = s->tanfovy * 1.3f;
    const float cx = fminf(limx, fmaxf(-limx, tx / tz));
    const float cy = fminf(limy, fmaxf(-limy, ty / tz));
    const float tz2 = tz * tz;
    const float J00 = focal_x / tz, J02 = (focal_x * (cx * -tz)) / tz2;
    const float J11 = focal_y / tz, J12 = (focal_y * (cy * -tz)) / tz2;
    const float T00 = fmaf(V[2], J02, V[0] * J00), T01 = fmaf(V[6], J02, V[4] * J00), T02 = fmaf(J02, V[10], V[8] * J00);
    const float T10 = fmaf(V[2], J12, J11 * V[1]), T11 = fmaf(V[6], J12, J11 * V[5]), T12 = fmaf(J12, V[10], J11 * V[9]);
    const float A0 = dot3g(T00, c[0], T01, c[1], T02, c[2]), A1 = dot3g(T00, c[1], T01, c[3], T02, c[4]),
                A2 = dot3g(T00, c[2], T01, c[4], T02, c[5]);
    const float B0 = dot3g(T10, c[0], T11, c[1], T12, c[2]), B1 = dot3g(T10, c[1], T11, c[3], T12, c[4]),
                B2 = dot3g(T10, c[2], T11, c[4], T12, c[5]);
    const float covx = dot3g(T00, A0, T01, A1, T02, A2) + 0.3f;
    const float covy = dot3g(T00, B0, T01, B1, T02, B2);
    const float covz = dot3g(T10, B0, T11, B1, T12, B2) + 0.3f;
    const float det = fmaf(covx, covz, -(covy * covy));
    if (det == 0.0f) return;
    const float det_inv = 1.0f / det;
    const float cnx = covz * det_inv, cny = det_inv * -covy, cnz = covx * det_inv;
    const float mid = (covx + covz) * 0.5f;
    const float sq = sqrtf(fmaxf(fmaf(mid, mid, -det), 0.1f));
    const float rad_f = ceilf(sqrtf(fmaxf(mid + sq, mid - sq)) * 3.f);
    const float mx = ndc2pix(projx, s->W), my = ndc2pix(projy, s->H);
    const int r = (int)rad_f;
    uint32_t rc[4];
    get_rect(mx, my, r, st->gx, st->gy, rc);
    if ((rc[2] - rc[0]) * (rc[3] - rc[1]) == 0) return;
    if (s->colors_precomp) {
        for (int k = 0; k < 3; k++) st->rgb[3 * idx + k] = s->colors_precomp[3 * idx + k];
    } else {
        sh_to_rgb(s->sh + (size_t)idx * s->M * 3, s->D, px, py, pz, s->campos, st->rgb + 3 * idx, st->clamped + 3 * idx);
    }
    if (!s->cov3D_precomp)
        for (int k = 0; k < 6; k++) st->cov3D[6 * idx + k] = c[k];
    st->depths[idx] = depth;
    st->radii[idx] = r;
    st->means2D[2 * idx] = mx;
    st->means2D[2 * idx + 1] = my;
    st->conic_opacity[4 * idx] = cnx;
    st->conic_opacity[4 * idx + 1] = cny;
    st->conic_opacity[4 * idx + 2] = cnz;
    st->conic_opacity[4 * idx + 3] = s->opacities[idx];
    memcpy(rect, rc, sizeof(rc));
}

typedef struct { uint64_t key; uint32_t id; } KV;
static int kv_cmp(const void* a, const void* b) {
    const KV *x = (const KV*)a, *y = (const KV*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->id < y->id ? -1 : (x->id > y->id);
}

/* rasterizer_impl.cu:70-138: keys = tile<<32 | depth bits, STABLE sort, per-tile ranges.  Emission
 * order inside a tile is ascending Gaussian id, so stable order == order by (key, id).  Bucketed
 * by tile first (the tile is the key's high word), then each bucket is sorted independently. */
static void bin(OrcState* st) {
    const int P = st->sc.P, gx = st->gx, tiles = st->tiles;
    uint32_t* count = (uint32_t*)calloc((size_t)tiles + 1, 4);
    for (int i = 0; i < P; i++) {
        const uint32_t* r = st->rect + 4 * i;
        for (uint32_t y = r[1]; y < r[3]; y++)
            for (uint32_t x = r[0]; x < r[2]; x++) count[y * gx + x]++;
    }
    uint32_t* start = (uint32_t*)malloc(((size_t)tiles + 1) * 4);
    uint64_t R = 0;
    for (int t = 0; t < tiles; t++) { start[t] = (uint32_t)R; R += count[t]; }
    start[tiles] = (uint32_t)R;
    st->R = (int64_t)R;
    KV* kv = (KV*)malloc((R ? R : 1) * sizeof(KV));
    uint32_t* cur = (uint32_t*)malloc(((size_t)tiles + 1) * 4);
    memcpy(cur, start, ((size_t)tiles + 1) * 4);
    for (int i = 0; i < P; i++) {
        const uint32_t* r = st->rect + 4 * i;
        uint32_t dbits;
        memcpy(&dbits, st->depths + i, 4);
        for (uint32_t y = r[1]; y < r[3]; y++)
            for (uint32_t x = r[0]; x < r[2]; x++) {
                const uint32_t t = y * gx + x;
                KV e = {((uint64_t)t << 32) | dbits, (uint32_t)i};
                kv[cur[t]++] = e;
            }
    }
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < tiles; t++)
        if (count[t] > 1) qsort(kv + start[t], count[t], sizeof(KV), kv_cmp);
    st->keys = (uint64_t*)malloc((R ? R : 1) * 8);
    st->list = (uint32_t*)malloc((R ? R : 1) * 4);
    for (uint64_t i = 0; i < R; i++) { st->keys[i] = kv[i].key; st->list[i] = kv[i].id; }
    /* identifyTileRanges: empty tiles stay (0,0) from the memset (rasterizer_impl.cu:313) */
    for (int t = 0; t < tiles; t++) {
        st->ranges[2 * t] = count[t] ? start[t] : 0u;
        st->ranges[2 * t + 1] = count[t] ? start[t + 1] : 0u;
    }
    free(kv); free(cur); free(start); free(count);
}

/* power = -0.5f*(A dx^2 + C dy^2) - B dx dy, contracted as nvcc does (forward.cu:345, backward.cu:685) */
static inline float gauss_power(float dx, float dy, float A, float B, float C) {
    return fmaf(fmaf(dx, dx * A, dy * (dy * C)), -0.5f, -(dy * (dx * B)));
}

/* forward.cu:263-392, one tile; pixels sequential, list front to back */
static void render_fwd_tile(OrcState* st, int t, float* out_color, float* out_depth, float* out_opacity,
                            int32_t* n_touched) {
    const OrcScene* s = &st->sc;
    const int W = s->W, H = s->H;
    const size_t N = (size_t)W * H;
    const int tx = t % st->gx, ty = t / st->gx;
    const uint32_t r0 = st->ranges[2 * t], r1 = st->ranges[2 * t + 1];
    for (int ly = 0; ly < TILE; ly++)
        for (int lx = 0; lx < TILE; lx++) {
            const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
            if (pxi >= W || pyi >= H) continue;
            const float pixx = (float)pxi, pixy = (float)pyi;
            float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
            uint32_t contributor = 0, last = 0;
            for (uint32_t k = r0; k < r1; k++) {
                contributor++;
                const uint32_t id = st->list[k];
                const float* co = st->conic_opacity + 4 * id;
                const float dx = st->means2D[2 * id] - pixx, dy = st->means2D[2 * id + 1] - pixy;
                const float power = gauss_power(dx, dy, co[0], co[1], co[2]);
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, co[3] * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1.f - alpha);
                if (test_T < 0.0001f) break;
                const float* f = st->rgb + 3 * id;
                C0 = fmaf(f[0] * alpha, T, C0);
                C1 = fmaf(f[1] * alpha, T, C1);
                C2 = fmaf(f[2] * alpha, T, C2);
                D = fmaf(st->depths[id] * alpha, T, D);
                if (test_T > 0.5f) {
#pragma omp atomic
                    n_touched[id]++;
                }
                T = test_T;
                last = contributor;
            }
            const size_t pix = (size_t)pyi * W + pxi;
            st->final_T[pix] = T;
            st->n_contrib[pix] = last;
            out_color[pix] = fmaf(T, s->bg[0], C0);
            out_color[N + pix] = fmaf(T, s->bg[1], C1);
            out_color[2 * N + pix] = fmaf(T, s->bg[2], C2);
            out_depth[pix] = D;
            out_opacity[pix] = 1.f - T;
        }
}

/* per-Gaussian accumulators of the compositing backward (double: the accurate side) */
typedef struct { double m2x, m2y, cx, cy, cw, op, col[3], dep; } GAcc;

/* backward.cu:563-787, one tile; per pixel back to front from its last contributor */
static void render_bwd_tile(const OrcState* st, int t, const float* dL_dcolor, const float* dL_ddepth, GAcc* acc) {
    const OrcScene* s = &st->sc;
    const int W = s->W, H = s->H;
    const size_t N = (size_t)W * H;
    const int tx = t % st->gx, ty = t / st->gx;
    const uint32_t r0 = st->ranges[2 * t], r1 = st->ranges[2 * t + 1];
    const uint32_t n = r1 - r0;
    if (!n) return;
    GAcc* loc = (GAcc*)calloc(n, sizeof(GAcc));
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    for (int ly = 0; ly < TILE; ly++)
        for (int lx = 0; lx < TILE; lx++) {
            const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
            if (pxi >= W || pyi >= H) continue;
            const size_t pix = (size_t)pyi * W + pxi;
            const float pixx = (float)pxi, pixy = (float)pyi;
            const float T_final = st->final_T[pix];
            float T = T_final;
            const int last_contributor = (int)st->n_contrib[pix];
            const float dpix[3] = {dL_dcolor[pix], dL_dcolor[N + pix], dL_dcolor[2 * N + pix]};
            const float dpd = dL_ddepth[pix];
            float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, accum_rec_depth = 0.f, last_depth = 0.f,
                  last_alpha = 0.f;
            float bg_dot = 0.f;
            for (int i = 0; i < 3; i++) bg_dot += s->bg[i] * dpix[i];
            for (int k = last_contributor - 1; k >= 0; k--) {
                const uint32_t id = st->list[r0 + k];
                const float* co = st->conic_opacity + 4 * id;
                const float dx = st->means2D[2 * id] - pixx, dy = st->means2D[2 * id + 1] - pixy;
                const float power = gauss_power(dx, dy, co[0], co[1], co[2]);
                if (power > 0.0f) continue;
                const float G = expf(power);
                const float alpha = fminf(0.99f, co[3] * G);
                if (alpha < 1.0f / 255.0f) continue;
                T = T / (1.f - alpha);
                const float dch = alpha * T;
                float dL_dalpha = 0.f;
                GAcc* a = loc + k;
                for (int ch = 0; ch < 3; ch++) {
                    const float c = st->rgb[3 * id + ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                    last_color[ch] = c;
                    dL_dalpha += (c - accum_rec[ch]) * dpix[ch];
                    a->col[ch] += (double)(dch * dpix[ch]);
                }
                const float depth = st->depths[id];
                accum_rec_depth = last_alpha * last_depth + (1.f - last_alpha) * accum_rec_depth;
                last_depth = depth;
                dL_dalpha += (depth - accum_rec_depth) * dpd;
                a->dep += (double)(dch * dpd);
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                const float dL_dG = co[3] * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                const float dG_ddely = -gdy * co[2] - gdx * co[1];
                a->m2x += (double)(dL_dG * dG_ddelx * ddelx_dx);
                a->m2y += (double)(dL_dG * dG_ddely * ddely_dy);
                a->cx += (double)(-0.5f * gdx * dx * dL_dG);
                a->cy += (double)(-0.5f * gdx * dy * dL_dG);
                a->cw += (double)(-0.5f * gdy * dy * dL_dG);
                a->op += (double)(G * dL_dalpha);
            }
        }
    for (uint32_t k = 0; k < n; k++) {
        GAcc* g = acc + st->list[r0 + k];
        const GAcc* a = loc + k;
        const double v[10] = {a->m2x, a->m2y, a->cx, a->cy, a->cw, a->op, a->col[0], a->col[1], a->col[2], a->dep};
        double* d[10] = {&g->m2x, &g->m2y, &g->cx, &g->cy, &g->cw, &g->op, &g->col[0], &g->col[1], &g->col[2], &g->dep};
        for (int i = 0; i < 10; i++)
            if (v[i] != 0.0) {
#pragma omp atomic
                *d[i] += v[i];
            }
    }
    free(loc);
}

/* ---- per-Gaussian backward: computeCov2DCUDA (backward.cu:150-346) + preprocessCUDA backward
 * (:418-539) + SH backward (:21-145) + cov3D backward (:350-413).  Evaluated in double from the
 * float32 inputs (accurate side of the tolerance check).  tau[6] += this Gaussian's dL_dtau. ---- */
typedef struct OrcGrads {
    float *dL_dmeans3D, *dL_dmeans2D, *dL_dsh, *dL_dcolors, *dL_dopacity, *dL_dscales, *dL_drotations, *dL_dcov3D,
        *dL_dconic, *dL_ddepths;
    float* dL_dtau; /* (6,) summed over Gaussians */
} OrcGrads;

static void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

static void preprocess_bwd_one(const OrcState* st, int idx, const GAcc* g, const OrcGrads* out, double* tau,
                               double focal_x, double focal_y) {
    const OrcScene* s = &st->sc;
    const float* V = s->viewmatrix;
    const float* Pm = s->projmatrix;
    const int M = s->M, D = s->D;
    double dmean[3] = {0, 0, 0}, dcov[6] = {0, 0, 0, 0, 0, 0}, dscale[3] = {0, 0, 0}, drot[4] = {0, 0, 0, 0};
    const int vis = st->radii[idx] > 0;
    if (vis) {
        const double m[3] = {s->means3D[3 * idx], s->means3D[3 * idx + 1], s->means3D[3 * idx + 2]};
        double c3[6];
        for (int k = 0; k < 6; k++) c3[k] = s->cov3D_precomp ? s->cov3D_precomp[6 * idx + k] : st->cov3D[6 * idx + k];
        const double dcx = g->cx, dcy = g->cy, dcz = g->cw; /* dL/dconic (x, y, w) */
        /* computeCov2DCUDA */
        double t[3] = {V[0] * m[0] + V[4] * m[1] + V[8] * m[2] + V[12], V[1] * m[0] + V[5] * m[1] + V[9] * m[2] + V[13],
                       V[2] * m[0] + V[6] * m[1] + V[10] * m[2] + V[14]};
        const double limx = 1.3 * (double)s->tanfovx, limy = 1.3 * (double)s->tanfovy;
        const double txtz = t[0] / t[2], tytz = t[1] / t[2];
        t[0] = fmin(limx, fmax(-limx, txtz)) * t[2];
        t[1] = fmin(limy, fmax(-limy, tytz)) * t[2];
        const double x_grad_mul = (txtz < -limx || txtz > limx) ? 0 : 1;
        const double y_grad_mul = (tytz < -limy || tytz > limy) ? 0 : 1;
        const double hx = focal_x, hy = focal_y;
        const double J00 = hx / t[2], J02 = -(hx * t[0]) / (t[2] * t[2]), J11 = hy / t[2], J12 = -(hy * t[1]) / (t[2] * t[2]);
        const double T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
        const double T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
        const double A0 = T00 * c3[0] + T01 * c3[1] + T02 * c3[2], A1 = T00 * c3[1] + T01 * c3[3] + T02 * c3[4],
                     A2 = T00 * c3[2] + T01 * c3[4] + T02 * c3[5];
        const double B0 = T10 * c3[0] + T11 * c3[1] + T12 * c3[2], B1 = T10 * c3[1] + T11 * c3[3] + T12 * c3[4],
                     B2 = T10 * c3[2] + T11 * c3[4] + T12 * c3[5];
        const double a = T00 * A0 + T01 * A1 + T02 * A2 + 0.3;
        const double b = T00 * B0 + T01 * B1 + T02 * B2;
        const double c = T10 * B0 + T11 * B1 + T12 * B2 + 0.3;
        const double denom = a * c - b * b;
        const double denom2inv = 1.0 / (denom * denom + 0.0000001);
        double dL_da = 0, dL_db = 0, dL_dc = 0;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
            dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
            dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
        }
        const double dT00 = 2 * A0 * dL_da + B0 * dL_db, dT01 = 2 * A1 * dL_da + B1 * dL_db, dT02 = 2 * A2 * dL_da + B2 * dL_db;
        const double dT10 = 2 * B0 * dL_dc + A0 * dL_db, dT11 = 2 * B1 * dL_dc + A1 * dL_db, dT12 = 2 * B2 * dL_dc + A2 * dL_db;
        const double dJ00 = V[0] * dT00 + V[4] * dT01 + V[8] * dT02;
        const double dJ02 = V[2] * dT00 + V[6] * dT01 + V[10] * dT02;
        const double dJ11 = V[1] * dT10 + V[5] * dT11 + V[9] * dT12;
        const double dJ12 = V[2] * dT10 + V[6] * dT11 + V[10] * dT12;
        const double tz = 1.0 / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const double dtx = x_grad_mul * -hx * tz2 * dJ02;
        const double dty = y_grad_mul * -hy * tz2 * dJ12;
        const double dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * t[0]) * tz3 * dJ02 + (2 * hy * t[1]) * tz3 * dJ12;
        /* pose through the (clamped) camera-space mean: [I | -[t]x] (backward.cu:273-288) */
        tau[0] += dtx; tau[1] += dty; tau[2] += dtz;
        tau[3] += dty * (-t[2]) + dtz * t[1];
        tau[4] += dtx * t[2] + dtz * (-t[0]);
        tau[5] += dtx * (-t[1]) + dty * t[0];
        dmean[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
        dmean[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
        dmean[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
        /* pose through the rotation W (backward.cu:299-343) */
        {
            const double dW[3][3] = {{J00 * dT00, J11 * dT10, J02 * dT00 + J12 * dT10},
                                     {J00 * dT01, J11 * dT11, J02 * dT01 + J12 * dT11},
                                     {J00 * dT02, J11 * dT12, J02 * dT02 + J12 * dT12}};
            for (int k = 0; k < 3; k++) {
                const double cv[3] = {V[4 * k], V[4 * k + 1], V[4 * k + 2]}; /* k-th column of R_cw */
                tau[3] += -cv[2] * dW[k][1] + cv[1] * dW[k][2];
                tau[4] += cv[2] * dW[k][0] - cv[0] * dW[k][2];
                tau[5] += -cv[1] * dW[k][0] + cv[0] * dW[k][1];
            }
        }
        /* preprocessCUDA backward (backward.cu:442-538) */
        const double dm2x = g->m2x, dm2y = g->m2y, ddepth = g->dep;
        const double mhx = Pm[0] * m[0] + Pm[4] * m[1] + Pm[8] * m[2] + Pm[12];
        const double mhy = Pm[1] * m[0] + Pm[5] * m[1] + Pm[9] * m[2] + Pm[13];
        const double mhw = Pm[3] * m[0] + Pm[7] * m[1] + Pm[11] * m[2] + Pm[15];
        const double m_w = 1.0 / (mhw + 0.0000001);
        const double mul1 = mhx * m_w * m_w, mul2 = mhy * m_w * m_w;
        dmean[0] += (Pm[0] * m_w - Pm[3] * mul1) * dm2x + (Pm[1] * m_w - Pm[3] * mul2) * dm2y;
        dmean[1] += (Pm[4] * m_w - Pm[7] * mul1) * dm2x + (Pm[5] * m_w - Pm[7] * mul2) * dm2y;
        dmean[2] += (Pm[8] * m_w - Pm[11] * mul1) * dm2x + (Pm[9] * m_w - Pm[11] * mul2) * dm2y;
        {
            const double alpha = m_w, beta = -mhx * m_w * m_w, gamma = -mhy * m_w * m_w;
            const double ra = s->projmatrix_raw[0], rb = s->projmatrix_raw[5], re = s->projmatrix_raw[11];
            const double pC[3] = {V[0] * m[0] + V[4] * m[1] + V[8] * m[2] + V[12], V[1] * m[0] + V[5] * m[1] + V[9] * m[2] + V[13],
                                  V[2] * m[0] + V[6] * m[1] + V[10] * m[2] + V[14]};
            const double d1[3] = {alpha * ra, 0, beta * re}, d2[3] = {0, alpha * rb, gamma * re};
            double th1[3], th2[3];
            cross3(pC, d1, th1); cross3(pC, d2, th2);
            for (int i = 0; i < 3; i++) {
                tau[i] += dm2x * d1[i] + dm2y * d2[i];
                tau[3 + i] += dm2x * th1[i] + dm2y * th2[i];
            }
            dmean[0] += ddepth * V[2]; dmean[1] += ddepth * V[6]; dmean[2] += ddepth * V[10];
            tau[2] += ddepth; tau[3] += ddepth * pC[1]; tau[4] += -ddepth * pC[0];
        }
        /* SH backward (backward.cu:21-145) */
        if (s->sh && !s->colors_precomp && out->dL_dsh) {
            const float* sh = s->sh + (size_t)idx * M * 3;
            float* dsh = out->dL_dsh + (size_t)idx * M * 3;
            const uint8_t* cb = st->clamped + 3 * idx;
            const double dRGB[3] = {cb[0] ? 0 : g->col[0], cb[1] ? 0 : g->col[1], cb[2] ? 0 : g->col[2]};
            const double v[3] = {m[0] - s->campos[0], m[1] - s->campos[1], m[2] - s->campos[2]};
            const double il = 1.0 / sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
            const double x = v[0] * il, y = v[1] * il, z = v[2] * il;
            double dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
            double w[16];
            for (int k = 0; k < 16; k++) w[k] = 0;
#define SHC(k, c) ((double)sh[3 * (k) + (c)])
#define ACC3(dst, coef, k) for (int c_ = 0; c_ < 3; c_++) dst[c_] += (coef) * SHC(k, c_)
            w[0] = SH_C0;
            if (D > 0) {
                w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x;
                ACC3(dx, -(double)SH_C1, 3); ACC3(dy, -(double)SH_C1, 1); ACC3(dz, (double)SH_C1, 2);
                if (D > 1) {
                    const double xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2 * zz - xx - yy);
                    w[7] = SH_C2[3] * xz; w[8] = SH_C2[4] * (xx - yy);
                    ACC3(dx, SH_C2[0] * y, 4); ACC3(dx, SH_C2[2] * 2 * -x, 6); ACC3(dx, SH_C2[3] * z, 7); ACC3(dx, SH_C2[4] * 2 * x, 8);
                    ACC3(dy, SH_C2[0] * x, 4); ACC3(dy, SH_C2[1] * z, 5); ACC3(dy, SH_C2[2] * 2 * -y, 6); ACC3(dy, SH_C2[4] * 2 * -y, 8);
                    ACC3(dz, SH_C2[1] * y, 5); ACC3(dz, SH_C2[2] * 2 * 2 * z, 6); ACC3(dz, SH_C2[3] * x, 7);
                    if (D > 2) {
                        w[9] = SH_C3[0] * y * (3 * xx - yy); w[10] = SH_C3[1] * xy * z; w[11] = SH_C3[2] * y * (4 * zz - xx - yy);
                        w[12] = SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy); w[13] = SH_C3[4] * x * (4 * zz - xx - yy);
                        w[14] = SH_C3[5] * z * (xx - yy); w[15] = SH_C3[6] * x * (xx - 3 * yy);
                        ACC3(dx, SH_C3[0] * 3 * 2 * xy, 9); ACC3(dx, SH_C3[1] * yz, 10); ACC3(dx, SH_C3[2] * -2 * xy, 11);
                        ACC3(dx, SH_C3[3] * -3 * 2 * xz, 12); ACC3(dx, SH_C3[4] * (-3 * xx + 4 * zz - yy), 13);
                        ACC3(dx, SH_C3[5] * 2 * xz, 14); ACC3(dx, SH_C3[6] * 3 * (xx - yy), 15);
                        ACC3(dy, SH_C3[0] * 3 * (xx - yy), 9); ACC3(dy, SH_C3[1] * xz, 10); ACC3(dy, SH_C3[2] * (-3 * yy + 4 * zz - xx), 11);
                        ACC3(dy, SH_C3[3] * -3 * 2 * yz, 12); ACC3(dy, SH_C3[4] * -2 * xy, 13); ACC3(dy, SH_C3[5] * -2 * yz, 14);
                        ACC3(dy, SH_C3[6] * -3 * 2 * xy, 15);
                        ACC3(dz, SH_C3[1] * xy, 10); ACC3(dz, SH_C3[2] * 4 * 2 * yz, 11); ACC3(dz, SH_C3[3] * 3 * (2 * zz - xx - yy), 12);
                        ACC3(dz, SH_C3[4] * 4 * 2 * xz, 13); ACC3(dz, SH_C3[5] * (xx - yy), 14);
                    }
                }
            }
#undef SHC
#undef ACC3
            for (int k = 0; k < M; k++)
                for (int c_ = 0; c_ < 3; c_++) dsh[3 * k + c_] = (float)(k < (D + 1) * (D + 1) ? w[k] * dRGB[c_] : 0.0);
            const double ddir[3] = {dx[0] * dRGB[0] + dx[1] * dRGB[1] + dx[2] * dRGB[2],
                                    dy[0] * dRGB[0] + dy[1] * dRGB[1] + dy[2] * dRGB[2],
                                    dz[0] * dRGB[0] + dz[1] * dRGB[1] + dz[2] * dRGB[2]};
            /* dnormvdv (auxiliary.h:107-117) */
            const double sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
            const double i32 = 1.0 / sqrt(sum2 * sum2 * sum2);
            const double dms[3] = {((sum2 - v[0] * v[0]) * ddir[0] - v[1] * v[0] * ddir[1] - v[2] * v[0] * ddir[2]) * i32,
                                   (-v[0] * v[1] * ddir[0] + (sum2 - v[1] * v[1]) * ddir[1] - v[2] * v[1] * ddir[2]) * i32,
                                   (-v[0] * v[2] * ddir[0] - v[1] * v[2] * ddir[1] + (sum2 - v[2] * v[2]) * ddir[2]) * i32};
            for (int i = 0; i < 3; i++) { dmean[i] += dms[i]; tau[i] -= dms[i]; } /* backward.cu:131-143 */
        }
        /* cov3D backward (backward.cu:350-413), w.r.t. the un-normalised quaternion */
        if (!s->cov3D_precomp) {
            const float* q = s->rotations + 4 * idx;
            const double r = q[0], x = q[1], y = q[2], z = q[3];
            const double sc[3] = {(double)s->scale_modifier * s->scales[3 * idx], (double)s->scale_modifier * s->scales[3 * idx + 1],
                                  (double)s->scale_modifier * s->scales[3 * idx + 2]};
            /* Rq[3*col+row] as GLM stores the matrix the reference builds (forward.cu:131-137) */
            const double Rq[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)};
            double Mm[9];
            for (int cc = 0; cc < 3; cc++)
                for (int rr = 0; rr < 3; rr++) Mm[3 * cc + rr] = sc[rr] * Rq[3 * cc + rr];
            const double S[9] = {dcov[0], 0.5 * dcov[1], 0.5 * dcov[2], 0.5 * dcov[1], dcov[3], 0.5 * dcov[4],
                                 0.5 * dcov[2], 0.5 * dcov[4], dcov[5]};
            double dM[9], dMt[9];
            for (int cc = 0; cc < 3; cc++)
                for (int rr = 0; rr < 3; rr++)
                    dM[3 * cc + rr] = 2 * (Mm[rr] * S[3 * cc] + Mm[3 + rr] * S[3 * cc + 1] + Mm[6 + rr] * S[3 * cc + 2]);
            for (int i = 0; i < 3; i++)
                for (int j = 0; j < 3; j++) dMt[3 * i + j] = dM[3 * j + i];
            for (int i = 0; i < 3; i++) dscale[i] = Rq[i] * dMt[3 * i] + Rq[3 + i] * dMt[3 * i + 1] + Rq[6 + i] * dMt[3 * i + 2];
            for (int i = 0; i < 3; i++)
                for (int j = 0; j < 3; j++) dMt[3 * i + j] *= sc[i];
#define DMT(i, j) dMt[3 * (i) + (j)]
            drot[0] = 2 * z * (DMT(0, 1) - DMT(1, 0)) + 2 * y * (DMT(2, 0) - DMT(0, 2)) + 2 * x * (DMT(1, 2) - DMT(2, 1));
            drot[1] = 2 * y * (DMT(1, 0) + DMT(0, 1)) + 2 * z * (DMT(2, 0) + DMT(0, 2)) + 2 * r * (DMT(1, 2) - DMT(2, 1)) - 4 * x * (DMT(2, 2) + DMT(1, 1));
            drot[2] = 2 * x * (DMT(1, 0) + DMT(0, 1)) + 2 * r * (DMT(2, 0) - DMT(0, 2)) + 2 * z * (DMT(1, 2) + DMT(2, 1)) - 4 * y * (DMT(2, 2) + DMT(0, 0));
            drot[3] = 2 * r * (DMT(0, 1) - DMT(1, 0)) + 2 * x * (DMT(2, 0) + DMT(0, 2)) + 2 * y * (DMT(1, 2) + DMT(2, 1)) - 4 * z * (DMT(1, 1) + DMT(0, 0));
#undef DMT
        }
    } else if (s->sh && !s->colors_precomp && out->dL_dsh) {
        memset(out->dL_dsh + (size_t)idx * M * 3, 0, (size_t)M * 3 * sizeof(float));
    }
    if (out->dL_dmeans3D) for (int i = 0; i < 3; i++) out->dL_dmeans3D[3 * idx + i] = (float)dmean[i];
    if (out->dL_dmeans2D) { out->dL_dmeans2D[3 * idx] = (float)g->m2x; out->dL_dmeans2D[3 * idx + 1] = (float)g->m2y; out->dL_dmeans2D[3 * idx + 2] = 0.f; }
    if (out->dL_dopacity) out->dL_dopacity[idx] = (float)g->op;
    if (out->dL_dcolors) for (int i = 0; i < 3; i++) out->dL_dcolors[3 * idx + i] = (float)g->col[i];
    if (out->dL_ddepths) out->dL_ddepths[idx] = (float)g->dep;
    if (out->dL_dconic) { out->dL_dconic[4 * idx] = (float)g->cx; out->dL_dconic[4 * idx + 1] = (float)g->cy; out->dL_dconic[4 * idx + 2] = 0.f; out->dL_dconic[4 * idx + 3] = (float)g->cw; }
    if (out->dL_dcov3D) for (int k = 0; k < 6; k++) out->dL_dcov3D[6 * idx + k] = (float)dcov[k];
    if (out->dL_dscales) for (int i = 0; i < 3; i++) out->dL_dscales[3 * idx + i] = (float)dscale[i];
    if (out->dL_drotations) for (int i = 0; i < 4; i++) out->dL_drotations[4 * idx + i] = (float)drot[i];
}

/* ------------------------------------ C entry points ------------------------------------ */
static void set_threads(int threads) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#else
    (void)threads;
#endif
}

EXPORT void orc_free(OrcState* st) {
    if (!st) return;
    free(st->depths); free(st->means2D); free(st->conic_opacity); free(st->rgb); free(st->cov3D); free(st->clamped);
    free(st->radii); free(st->rect); free(st->keys); free(st->list); free(st->ranges); free(st->n_contrib);
    free(st->final_T); free(st);
}

/* Rasterizer::forward (rasterizer_impl.cu:198-344).  Outputs: color (3,H,W), depth, opacity (1,H,W),
 * radii (P), n_touched (P).  Returns the state handle (null on bad arguments). */
EXPORT OrcState* orc_forward(const OrcScene* sc, float* out_color, float* out_depth, float* out_opacity, int32_t* radii,
                             int32_t* n_touched, int threads) {
    if (!sc || sc->P < 0 || sc->W <= 0 || sc->H <= 0) return NULL;
    if (sc->P > 0 && !sc->sh && !sc->colors_precomp) return NULL; /* rasterizer_impl.cu:245-248 */
    set_threads(threads);
    OrcState* st = (OrcState*)calloc(1, sizeof(OrcState));
    st->sc = *sc;
    const int P = sc->P, W = sc->W, H = sc->H;
    const size_t Pn = P > 0 ? (size_t)P : 1, N = (size_t)W * H;
    st->gx = (W + TILE - 1) / TILE; st->gy = (H + TILE - 1) / TILE; st->tiles = st->gx * st->gy;
    st->depths = (float*)calloc(Pn, 4); st->means2D = (float*)calloc(Pn, 8); st->conic_opacity = (float*)calloc(Pn, 16);
    st->rgb = (float*)calloc(Pn, 12); st->cov3D = (float*)calloc(Pn, 24); st->clamped = (uint8_t*)calloc(Pn, 3);
    st->radii = (int32_t*)calloc(Pn, 4); st->rect = (uint32_t*)calloc(Pn, 16);
    st->ranges = (uint32_t*)calloc((size_t)st->tiles, 8); st->n_contrib = (uint32_t*)calloc(N, 4);
    st->final_T = (float*)calloc(N, 4);
    const float focal_y = H / (2.0f * sc->tanfovy), focal_x = W / (2.0f * sc->tanfovx); /* rasterizer_impl.cu:225-226 */
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) preprocess_one(st, i, focal_x, focal_y);
    bin(st);
    if (n_touched) memset(n_touched, 0, Pn * 4);
    int32_t* nt = n_touched ? n_touched : (int32_t*)calloc(Pn, 4);
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < st->tiles; t++) render_fwd_tile(st, t, out_color, out_depth, out_opacity, nt);
    if (!n_touched) free(nt);
    if (radii) memcpy(radii, st->radii, (size_t)(P > 0 ? P : 0) * 4);
    return st;
}

EXPORT int64_t orc_num_rendered(const OrcState* st) { return st ? st->R : -1; }

/* state in the reference's GeometryState / BinningState / ImageState terms; any pointer may be null */
EXPORT void orc_export_state(const OrcState* st, float* depths, float* means2D, float* conic_opacity, float* rgb,
                             uint8_t* clamped, float* cov3D, uint64_t* keys, uint32_t* list, uint32_t* ranges,
                             uint32_t* n_contrib, float* final_T) {
    const size_t P = (size_t)st->sc.P, N = (size_t)st->sc.W * st->sc.H, R = (size_t)st->R;
    if (depths) memcpy(depths, st->depths, P * 4);
    if (means2D) memcpy(means2D, st->means2D, P * 8);
    if (conic_opacity) memcpy(conic_opacity, st->conic_opacity, P * 16);
    if (rgb) memcpy(rgb, st->rgb, P * 12);
    if (clamped) memcpy(clamped, st->clamped, P * 3);
    if (cov3D) memcpy(cov3D, st->cov3D, P * 24);
    if (keys) memcpy(keys, st->keys, R * 8);
    if (list) memcpy(list, st->list, R * 4);
    if (ranges) memcpy(ranges, st->ranges, (size_t)st->tiles * 8);
    if (n_contrib) memcpy(n_contrib, st->n_contrib, N * 4);
    if (final_T) memcpy(final_T, st->final_T, N * 4);
}

/* Rasterizer::backward (rasterizer_impl.cu:348-455) + the dL_dtau sum (__init__.py:152-154) */
EXPORT int orc_backward(const OrcState* st, const float* dL_dcolor, const float* dL_ddepth, const OrcGrads* out,
                        int threads) {
    if (!st || !dL_dcolor || !dL_ddepth || !out) return -1;
    set_threads(threads);
    const OrcScene* sc = &st->sc;
    const int P = sc->P;
    GAcc* acc = (GAcc*)calloc(P > 0 ? (size_t)P : 1, sizeof(GAcc));
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < st->tiles; t++) render_bwd_tile(st, t, dL_dcolor, dL_ddepth, acc);
    const double focal_y = sc->H / (2.0f * sc->tanfovy), focal_x = sc->W / (2.0f * sc->tanfovx);
    double tau[6] = {0, 0, 0, 0, 0, 0};
#pragma omp parallel
    {
        double ltau[6] = {0, 0, 0, 0, 0, 0};
#pragma omp for schedule(static)
        for (int i = 0; i < P; i++) preprocess_bwd_one(st, i, acc + i, out, ltau, focal_x, focal_y);
#pragma omp critical
        for (int k = 0; k < 6; k++) tau[k] += ltau[k];
    }
    if (out->dL_dtau)
        for (int k = 0; k < 6; k++) out->dL_dtau[k] = (float)tau[k];
    free(acc);
    return 0;
}

/* checkFrustum / markVisible (rasterizer_impl.cu:54-66,141-153) */
EXPORT void orc_mark_visible(int P, const float* means3D, const float* V, uint8_t* present) {
    for (int i = 0; i < P; i++)
        present[i] = !(dot3p(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2], V[2], V[6], V[10], V[14]) <= 0.2f);
}
