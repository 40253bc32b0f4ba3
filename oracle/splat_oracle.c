/*
 * oracle/splat_oracle.c  --  TEST INFRASTRUCTURE ONLY.  NOT PART OF THE PRODUCT.
 *
 * Scalar CPU restatement of the differentiable Gaussian-splat rasterizer path that
 * PoplarPoplar/3dgs_study reaches through `diff_gaussian_rasterization`
 * (call site: /root/reference/gaussian_renderer/__init__.py:47-62,98-106).
 *
 * PARITY UNPINNED: the rasterizer's own source is a git submodule that is absent from the
 * reference snapshot (github.com/PoplarPoplar/diff-gaussian-rasterization @ 2cf59156, see
 * /root/reference/.SUBMODULES.json:16-22) and the reference holds no tests or golden vectors
 * for it.  This file therefore restates the published 3D-Gaussian-splatting rasterizer
 * algorithm (EWA projection + 0.3 px^2 low-pass, 16x16 tiles, 64-bit tile|depth keys, stable
 * sort, front-to-back blending with the 1/255 alpha cutoff, 0.99 alpha clamp and 1e-4
 * transmittance exit, reverse-traversal backward) and anchors on what IS in the reference:
 *   - SH basis / constants          /root/reference/utils/sh_utils.py:26-43,57-112
 *   - colour = max(SH + 0.5, 0)     /root/reference/gaussian_renderer/__init__.py:85-90
 *   - quaternion (w,x,y,z) -> R     /root/reference/utils/general_utils.py:86-117
 *   - Sigma = (R S)(R S)^T, packing /root/reference/utils/general_utils.py:72-81,119-128
 *                                   /root/reference/scene/gaussian_model.py:27-32
 *   - row-vector matrix convention  /root/reference/utils/graphics_utils.py:25-32
 *                                   /root/reference/scene/cameras.py:103-121
 *   - means2D.grad consumer         /root/reference/scene/gaussian_model.py:565-581
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared object.  The product (3dgs_study_b200/csrc) never does.
 *
 * Canonical FP32 operation order.  The integer outputs (radii, tiles_touched, keys, ranges)
 * must be bit-exact between this oracle and the CUDA path, so every float expression below
 * is written with its rounding points explicit: `fmaf` where a fused multiply-add is meant,
 * plain operators (compiled with -ffp-contract=off) where separate roundings are meant.
 * DESIGN.md section "Canonical arithmetic" is the normative text; this file and the CUDA
 * kernels are two independent implementations of it.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16
#define ORC_API __attribute__((visibility("default")))

/* Hot loops get an FMA-instruction clone when the host supports it; results are identical
 * because fmaf() is correctly rounded either way. */
#if defined(__x86_64__)
#define HOT __attribute__((target_clones("arch=x86-64-v3", "default")))
#else
#define HOT
#endif

/* Counting experiment: 0 = canonical pixel formula, 1 = fp32 three-rounding order. */
static int g_pixel_fp32 = 0;
ORC_API void orc_set_pixel_fp32(int on) { g_pixel_fp32 = on; }

/* Thread count of the OpenMP loops below, set by the caller (bench.py's CPU legs) rather than inherited from an
 * OMP_NUM_THREADS a launcher exported (torchrun sets it to 1). */
#ifdef _OPENMP
#include <omp.h>
ORC_API void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
#else
ORC_API void orc_set_threads(int n) { (void)n; }
#endif

/* SH constants, /root/reference/utils/sh_utils.py:26-43 (rounded to float as the kernels do) */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

static inline float dot3_fma(float a0, float b0, float a1, float b1, float a2, float b2) {
    /* fma(a2,b2, fma(a1,b1, a0*b0)) */
    return fmaf(a2, b2, fmaf(a1, b1, a0 * b0));
}

/* float -> int conversion, round toward zero, saturating, NaN -> 0 (the CUDA cvt.rzi.s32.f32
 * semantics; a plain C cast is undefined out of range). */
static inline int f2i_rz_sat(float x) {
    if (x != x) return 0;
    if (x >= 2147483648.0f) return 2147483647;
    if (x <= -2147483648.0f) return (-2147483647 - 1);
    return (int)x;
}
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void get_rect(float px, float py, int radius, int gx, int gy, int* r /*minx,miny,maxx,maxy*/) {
    float fr = (float)radius;
    r[0] = clampi(f2i_rz_sat((px - fr) / 16.0f), 0, gx);
    r[1] = clampi(f2i_rz_sat((py - fr) / 16.0f), 0, gy);
    r[2] = clampi(f2i_rz_sat((((px + fr) + 16.0f) - 1.0f) / 16.0f), 0, gx);
    r[3] = clampi(f2i_rz_sat((((py + fr) + 16.0f) - 1.0f) / 16.0f), 0, gy);
}

/* quaternion (w,x,y,z), NOT renormalised here (the caller normalises,
 * /root/reference/scene/gaussian_model.py:111-112) */
static void quat_to_R(const float* q, float R[3][3]) {
    float r = q[0], x = q[1], y = q[2], z = q[3];
    R[0][0] = fmaf(-2.0f, fmaf(y, y, z * z), 1.0f);
    R[0][1] = 2.0f * fmaf(x, y, -(r * z));
    R[0][2] = 2.0f * fmaf(x, z, r * y);
    R[1][0] = 2.0f * fmaf(x, y, r * z);
    R[1][1] = fmaf(-2.0f, fmaf(x, x, z * z), 1.0f);
    R[1][2] = 2.0f * fmaf(y, z, -(r * x));
    R[2][0] = 2.0f * fmaf(x, z, -(r * y));
    R[2][1] = 2.0f * fmaf(y, z, r * x);
    R[2][2] = fmaf(-2.0f, fmaf(x, x, y * y), 1.0f);
}

static void cov3d_from_scale_rot(const float* s, float mod, const float* q, float* c6) {
    float R[3][3], L[3][3];
    quat_to_R(q, R);
    float sx = mod * s[0], sy = mod * s[1], sz = mod * s[2];
    for (int i = 0; i < 3; i++) { L[i][0] = R[i][0] * sx; L[i][1] = R[i][1] * sy; L[i][2] = R[i][2] * sz; }
    c6[0] = dot3_fma(L[0][0], L[0][0], L[0][1], L[0][1], L[0][2], L[0][2]);
    c6[1] = dot3_fma(L[0][0], L[1][0], L[0][1], L[1][1], L[0][2], L[1][2]);
    c6[2] = dot3_fma(L[0][0], L[2][0], L[0][1], L[2][1], L[0][2], L[2][2]);
    c6[3] = dot3_fma(L[1][0], L[1][0], L[1][1], L[1][1], L[1][2], L[1][2]);
    c6[4] = dot3_fma(L[1][0], L[2][0], L[1][1], L[2][1], L[1][2], L[2][2]);
    c6[5] = dot3_fma(L[2][0], L[2][0], L[2][1], L[2][1], L[2][2], L[2][2]);
}

/* SH basis values b[0..(D+1)^2) at unit direction (x,y,z); polynomials follow
 * /root/reference/utils/sh_utils.py:74-100 */
static void sh_basis(int D, float x, float y, float z, float* b) {
    b[0] = SH_C0;
    if (D > 0) {
        b[1] = -(SH_C1 * y); b[2] = SH_C1 * z; b[3] = -(SH_C1 * x);
        if (D > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            b[4] = SH_C2[0] * xy;
            b[5] = SH_C2[1] * yz;
            b[6] = SH_C2[2] * ((2.0f * zz - xx) - yy);
            b[7] = SH_C2[3] * xz;
            b[8] = SH_C2[4] * (xx - yy);
            if (D > 2) {
                b[9]  = (SH_C3[0] * y) * fmaf(3.0f, xx, -yy);
                b[10] = (SH_C3[1] * xy) * z;
                b[11] = (SH_C3[2] * y) * ((4.0f * zz - xx) - yy);
                b[12] = (SH_C3[3] * z) * ((2.0f * zz - 3.0f * xx) - 3.0f * yy);
                b[13] = (SH_C3[4] * x) * ((4.0f * zz - xx) - yy);
                b[14] = (SH_C3[5] * z) * (xx - yy);
                b[15] = (SH_C3[6] * x) * fmaf(-3.0f, yy, xx);
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Preprocess forward (per Gaussian).  Outputs default to 0 for culled Gaussians.              */
/* ------------------------------------------------------------------------------------------ */
ORC_API void orc_preprocess(
    int P, int D, int M, int W, int H,
    const float* means3D, const float* scales, float mod, const float* rotations,
    const float* opacities, const float* shs, const float* colors_precomp,
    const float* cov3D_precomp, const float* view, const float* proj, const float* campos,
    float tanfovx, float tanfovy, int antialiasing,
    int32_t* radii, float* means2D, float* depths, float* cov3D, float* conic_opacity,
    float* rgb, uint8_t* clamped, uint32_t* tiles_touched)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const float focal_x = (float)W / (2.0f * tanfovx);
    const float focal_y = (float)H / (2.0f * tanfovy);
    const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;

#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0; tiles_touched[i] = 0;
        means2D[2 * i] = means2D[2 * i + 1] = 0.0f; depths[i] = 0.0f;
        for (int k = 0; k < 6; k++) cov3D[6 * i + k] = 0.0f;
        for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = 0.0f;
        for (int k = 0; k < 3; k++) { rgb[3 * i + k] = 0.0f; clamped[3 * i + k] = 0; }

        const float x = means3D[3 * i], y = means3D[3 * i + 1], z = means3D[3 * i + 2];
        /* view-space point (row-vector convention: p_view = [p,1] @ view) */
        float tx = fmaf(view[8], z, fmaf(view[4], y, view[0] * x)) + view[12];
        float ty = fmaf(view[9], z, fmaf(view[5], y, view[1] * x)) + view[13];
        float tz = fmaf(view[10], z, fmaf(view[6], y, view[2] * x)) + view[14];
        if (tz <= 0.2f) continue;                                   /* near cull */

        float hx = fmaf(proj[8], z, fmaf(proj[4], y, proj[0] * x)) + proj[12];
        float hy = fmaf(proj[9], z, fmaf(proj[5], y, proj[1] * x)) + proj[13];
        float hw = fmaf(proj[11], z, fmaf(proj[7], y, proj[3] * x)) + proj[15];
        float pw = 1.0f / (hw + 0.0000001f);
        float ndcx = hx * pw, ndcy = hy * pw;

        float c6[6];
        if (cov3D_precomp) { for (int k = 0; k < 6; k++) c6[k] = cov3D_precomp[6 * i + k]; }
        else cov3d_from_scale_rot(scales + 3 * i, mod, rotations + 4 * i, c6);

        /* EWA: cov2D = (J Wv) Sigma (J Wv)^T */
        float txtz = tx / tz, tytz = ty / tz;
        float cx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        float cy = fminf(limy, fmaxf(-limy, tytz)) * tz;
        float tz2 = tz * tz;
        float J00 = focal_x / tz, J02 = -(focal_x * cx) / tz2;
        float J11 = focal_y / tz, J12 = -(focal_y * cy) / tz2;
        /* Wv[r][k] = view[4k + r] */
        float M0[3], M1[3];
        for (int k = 0; k < 3; k++) {
            M0[k] = fmaf(J02, view[4 * k + 2], J00 * view[4 * k + 0]);
            M1[k] = fmaf(J12, view[4 * k + 2], J11 * view[4 * k + 1]);
        }
        float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        float N0[3], N1[3];
        for (int k = 0; k < 3; k++) {
            N0[k] = dot3_fma(M0[0], S[0][k], M0[1], S[1][k], M0[2], S[2][k]);
            N1[k] = dot3_fma(M1[0], S[0][k], M1[1], S[1][k], M1[2], S[2][k]);
        }
        float cov00 = dot3_fma(N0[0], M0[0], N0[1], M0[1], N0[2], M0[2]);
        float cov01 = dot3_fma(N0[0], M1[0], N0[1], M1[1], N0[2], M1[2]);
        float cov11 = dot3_fma(N1[0], M1[0], N1[1], M1[1], N1[2], M1[2]);

        float a = cov00 + 0.3f, b = cov01, c = cov11 + 0.3f;
        float det = fmaf(a, c, -(b * b));
        if (det == 0.0f) continue;
        float hconv = 1.0f;
        if (antialiasing) {
            float det0 = fmaf(cov00, cov11, -(cov01 * cov01));
            hconv = sqrtf(fmaxf(0.000025f, det0 / det));
        }
        float det_inv = 1.0f / det;
        float conA = c * det_inv, conB = -b * det_inv, conC = a * det_inv;

        float mid = 0.5f * (a + c);
        float sq = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        float lam1 = mid + sq, lam2 = mid - sq;
        int radius = f2i_rz_sat(ceilf(3.0f * sqrtf(fmaxf(lam1, lam2))));

        /* ndc -> pixel, evaluated in double and rounded once (exact in double).  g_pixel_fp32 (counting
         * experiment only, orc_set_pixel_fp32): the three-rounding fp32 order of SURVEY.md A.2 step 7. */
        float px = (float)((((double)ndcx + 1.0) * (double)W - 1.0) * 0.5);
        float py = (float)((((double)ndcy + 1.0) * (double)H - 1.0) * 0.5);
        if (g_pixel_fp32) {
            px = ((ndcx + 1.0f) * (float)W - 1.0f) * 0.5f;
            py = ((ndcy + 1.0f) * (float)H - 1.0f) * 0.5f;
        }
        int r[4];
        get_rect(px, py, radius, gx, gy, r);
        int area = (r[2] - r[0]) * (r[3] - r[1]);
        if (area == 0) continue;

        if (colors_precomp) {
            for (int k = 0; k < 3; k++) rgb[3 * i + k] = colors_precomp[3 * i + k];
        } else {
            float dx = x - campos[0], dy = y - campos[1], dz = z - campos[2];
            float len = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
            float bsh[16];
            sh_basis(D, dx / len, dy / len, dz / len, bsh);
            int nb = (D + 1) * (D + 1);
            for (int ch = 0; ch < 3; ch++) {
                const float* sh = shs + (size_t)i * M * 3 + ch;
                float res = bsh[0] * sh[0];
                for (int k = 1; k < nb; k++) res = fmaf(bsh[k], sh[3 * k], res);
                res = res + 0.5f;
                clamped[3 * i + ch] = (res < 0.0f);
                rgb[3 * i + ch] = fmaxf(res, 0.0f);
            }
        }
        depths[i] = tz;
        radii[i] = radius;
        means2D[2 * i] = px; means2D[2 * i + 1] = py;
        for (int k = 0; k < 6; k++) cov3D[6 * i + k] = c6[k];
        conic_opacity[4 * i + 0] = conA; conic_opacity[4 * i + 1] = conB;
        conic_opacity[4 * i + 2] = conC; conic_opacity[4 * i + 3] = opacities[i] * hconv;
        tiles_touched[i] = (uint32_t)area;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Binning: key emission in Gaussian-index order, row-major over the tile rect; stable sort;   */
/* tile ranges.  keys/vals must hold R = sum(tiles_touched) entries.  Returns R.               */
/* ------------------------------------------------------------------------------------------ */
ORC_API int64_t orc_bin(int P, int W, int H, const float* means2D, const float* depths,
                        const int32_t* radii, uint64_t* keys, uint32_t* vals, int64_t cap,
                        int32_t* ranges /* [T][2], zero-filled here */)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    int64_t n = 0;
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        int r[4];
        get_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, r);
        uint32_t dbits; memcpy(&dbits, &depths[i], 4);
        for (int ty = r[1]; ty < r[3]; ty++)
            for (int tx = r[0]; tx < r[2]; tx++) {
                if (n >= cap) return -1;
                keys[n] = ((uint64_t)(uint32_t)(ty * gx + tx) << 32) | dbits;
                vals[n] = (uint32_t)i; n++;
            }
    }
    /* stable LSD radix sort, 8 passes x 8 bits over the full 64-bit key */
    uint64_t* k2 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(n > 0 ? n : 1));
    uint32_t* v2 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    uint64_t *ka = keys, *kb = k2; uint32_t *va = vals, *vb = v2;
    for (int pass = 0; pass < 8; pass++) {
        size_t cnt[257]; memset(cnt, 0, sizeof(cnt));
        int sh = pass * 8;
        for (int64_t j = 0; j < n; j++) cnt[((ka[j] >> sh) & 255) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (int64_t j = 0; j < n; j++) { size_t o = cnt[(ka[j] >> sh) & 255]++; kb[o] = ka[j]; vb[o] = va[j]; }
        uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
    }
    /* 8 passes: result is back in keys/vals */
    free(k2); free(v2);
    memset(ranges, 0, sizeof(int32_t) * 2 * (size_t)gx * gy);
    for (int64_t j = 0; j < n; j++) {
        uint32_t t = (uint32_t)(keys[j] >> 32);
        if (j == 0 || (uint32_t)(keys[j - 1] >> 32) != t) ranges[2 * t] = (int32_t)j;
        if (j == n - 1 || (uint32_t)(keys[j + 1] >> 32) != t) ranges[2 * t + 1] = (int32_t)(j + 1);
    }
    return n;
}

/* Canonical exponential (DESIGN.md "Canonical arithmetic").  The blend loop is discontinuous in alpha = o exp(power)
 * (the 1/255 cutoff) and in the running transmittance (the 1e-4 exit), so two exp implementations that differ in
 * the last bit -- glibc's expf and CUDA's expf do -- flip a decision once in ~10^7 evaluations, i.e. a few pixels
 * per full-size view, each by up to 4e-3.  exp is therefore part of the canonical arithmetic: range reduction by
 * ln 2 in two pieces and a degree-7 polynomial (the Cephes expf scheme), every operation an fp32 fma / add with
 * one rounding, identical here and in render_tiles.cu (canon_exp).  <= 0.89 ulp from the true exp on [-20, 0]. */
static inline float canon_exp(float x) {
    x = fmaxf(x, -80.0f);
    float t = fmaf(x, 1.4426950408889634f, 12582912.0f);      /* 1.5 * 2^23: the low bits of t are round(x log2 e) */
    float n = t - 12582912.0f;
    float r = fmaf(n, -0.693359375f, x);
    r = fmaf(n, 2.12194440e-4f, r);
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    uint32_t tb, pb;
    memcpy(&tb, &t, 4); memcpy(&pb, &p, 4);
    pb += tb << 23;                                            /* times 2^n: n added to the exponent field */
    float out; memcpy(&out, &pb, 4);
    return out;
}
ORC_API float orc_canon_exp(float x) { return canon_exp(x); }

/* ------------------------------------------------------------------------------------------ */
/* Render forward                                                                              */
/* ------------------------------------------------------------------------------------------ */
static inline float gauss_power(float gxp, float gyp, float A, float B, float C, float pxf, float pyf,
                                float* dxo, float* dyo) {
    float dx = gxp - pxf, dy = gyp - pyf;
    float hA = -0.5f * A, hC = -0.5f * C;
    float u = fmaf(hA, dx, -(B * dy));
    *dxo = dx; *dyo = dy;
    return fmaf(dx, u, (hC * dy) * dy);
}

HOT ORC_API void orc_render(int W, int H, const int32_t* ranges, const uint32_t* vals,
                        const float* means2D, const float* conic_opacity, const float* rgb,
                        const float* depths, const float* bg, float* out_color,
                        float* out_invdepth /* nullable */, float* final_T, uint32_t* n_contrib,
                        int tile_stride, int tile_offset /* bounded CPU sample: tiles with tile % stride == offset */)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        if (tile_stride > 1 && tile % tile_stride != tile_offset) continue;
        int ty = tile / gx, tx = tile % gx;
        int s = ranges[2 * tile], e = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
            int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            float pxf = (float)px, pyf = (float)py;
            float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dp = 0.0f;
            uint32_t contributor = 0, last = 0;
            for (int k = s; k < e; k++) {
                contributor++;
                uint32_t id = vals[k];
                const float* co = conic_opacity + 4 * (size_t)id;
                float dx, dy;
                float power = gauss_power(means2D[2 * id], means2D[2 * id + 1], co[0], co[1], co[2], pxf, pyf, &dx, &dy);
                if (power > 0.0f) continue;
                float alpha = fminf(0.99f, co[3] * canon_exp(power));
                if (alpha < 1.0f / 255.0f) continue;
                float test_T = T * (1.0f - alpha);
                if (test_T < 0.0001f) break;
                float w = alpha * T;
                C0 = fmaf(rgb[3 * id + 0], w, C0);
                C1 = fmaf(rgb[3 * id + 1], w, C1);
                C2 = fmaf(rgb[3 * id + 2], w, C2);
                Dp = fmaf(1.0f / depths[id], w, Dp);
                T = test_T; last = contributor;
            }
            size_t pix = (size_t)py * W + px, HW = (size_t)H * W;
            final_T[pix] = T; n_contrib[pix] = last;
            out_color[0 * HW + pix] = fmaf(T, bg[0], C0);
            out_color[1 * HW + pix] = fmaf(T, bg[1], C1);
            out_color[2 * HW + pix] = fmaf(T, bg[2], C2);
            if (out_invdepth) out_invdepth[pix] = Dp;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Render backward: reverse traversal from n_contrib; transmittance recovered by division.     */
/* Accumulates in double (one writer per Gaussian at a time under `omp atomic`).                */
/* Outputs:  dL_dmean2D [P,2] in NDC units (pixel gradient x 0.5*W, 0.5*H) -- the quantity the  */
/* reference consumes as viewspace_points.grad (scene/gaussian_model.py:576-580);               */
/* dL_dconic [P,3] = true partials w.r.t. (A,B,C) of power = -0.5(A dx^2 + C dy^2) - B dx dy.    */
/* ------------------------------------------------------------------------------------------ */
HOT ORC_API void orc_render_backward(int P, int W, int H, const int32_t* ranges, const uint32_t* vals,
                                 const float* means2D, const float* conic_opacity, const float* rgb,
                                 const float* depths, const float* bg, const float* final_T,
                                 const uint32_t* n_contrib, const float* dL_dpix,
                                 const float* dL_dinvdepth /* nullable */,
                                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                                 float* dL_dcolor, float* dL_dinvd /* [P], nullable: d/d(1/depth) */,
                                 int tile_stride, int tile_offset)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    double* acc = (double*)calloc((size_t)P * 10, sizeof(double));
    const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        if (tile_stride > 1 && tile % tile_stride != tile_offset) continue;
        int ty = tile / gx, tx = tile % gx;
        int s = ranges[2 * tile];
        for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
            int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            size_t pix = (size_t)py * W + px;
            float pxf = (float)px, pyf = (float)py;
            const double T_final = final_T[pix];
            double T = T_final;
            double g[3] = {dL_dpix[pix], dL_dpix[HW + pix], dL_dpix[2 * HW + pix]};
            double gD = dL_dinvdepth ? dL_dinvdepth[pix] : 0.0;
            double bg_dot = bg[0] * g[0] + bg[1] * g[1] + bg[2] * g[2];
            double last_alpha = 0.0, last_c[3] = {0, 0, 0}, rec[3] = {0, 0, 0};
            double last_invd = 0.0, recD = 0.0;
            for (int k = s + (int)n_contrib[pix] - 1; k >= s; k--) {
                uint32_t id = vals[k];
                const float* co = conic_opacity + 4 * (size_t)id;
                float dxf, dyf;
                float power = gauss_power(means2D[2 * id], means2D[2 * id + 1], co[0], co[1], co[2], pxf, pyf, &dxf, &dyf);
                if (power > 0.0f) continue;
                float Gf = canon_exp(power);
                float alphaf = fminf(0.99f, co[3] * Gf);
                if (alphaf < 1.0f / 255.0f) continue;
                double G = Gf, alpha = alphaf, dx = dxf, dy = dyf;
                T = T / (1.0 - alpha);
                double dch = alpha * T, dL_dalpha = 0.0;
                double* a = acc + (size_t)id * 10;
                for (int ch = 0; ch < 3; ch++) {
                    double c = rgb[3 * id + ch];
                    rec[ch] = last_alpha * last_c[ch] + (1.0 - last_alpha) * rec[ch];
                    last_c[ch] = c;
                    dL_dalpha += (c - rec[ch]) * g[ch];
                    double v = dch * g[ch];
#pragma omp atomic
                    a[6 + ch] += v;
                }
                {   /* inverse depth channel (no background term) */
                    double invd = 1.0 / (double)depths[id];
                    recD = last_alpha * last_invd + (1.0 - last_alpha) * recD;
                    last_invd = invd;
                    dL_dalpha += (invd - recD) * gD;
                    double v = dch * gD;
#pragma omp atomic
                    a[9] += v;
                }
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (1.0 - alpha)) * bg_dot;
                /* straight-through 0.99 clamp: d alpha = o dG + G do regardless of the clamp */
                double dL_dG = (double)co[3] * dL_dalpha;
                double gdx = G * dx, gdy = G * dy;
                double dG_ddelx = -gdx * co[0] - gdy * co[1];
                double dG_ddely = -gdy * co[2] - gdx * co[1];
                double v0 = dL_dG * dG_ddelx * (0.5 * W), v1 = dL_dG * dG_ddely * (0.5 * H);
                double v2 = -0.5 * gdx * dx * dL_dG, v3 = -gdx * dy * dL_dG, v4 = -0.5 * gdy * dy * dL_dG;
                double v5 = G * dL_dalpha;
#pragma omp atomic
                a[0] += v0;
#pragma omp atomic
                a[1] += v1;
#pragma omp atomic
                a[2] += v2;
#pragma omp atomic
                a[3] += v3;
#pragma omp atomic
                a[4] += v4;
#pragma omp atomic
                a[5] += v5;
            }
        }
    }
    for (int i = 0; i < P; i++) {
        const double* a = acc + (size_t)i * 10;
        dL_dmean2D[2 * i] = (float)a[0]; dL_dmean2D[2 * i + 1] = (float)a[1];
        dL_dconic[3 * i] = (float)a[2]; dL_dconic[3 * i + 1] = (float)a[3]; dL_dconic[3 * i + 2] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        dL_dcolor[3 * i] = (float)a[6]; dL_dcolor[3 * i + 1] = (float)a[7]; dL_dcolor[3 * i + 2] = (float)a[8];
        if (dL_dinvd) dL_dinvd[i] = (float)a[9];
    }
    free(acc);
}

/* ------------------------------------------------------------------------------------------ */
/* Per-Gaussian backward: conic -> cov2D -> (cov3D, mean); mean2D -> mean3D; SH; cov3D ->        */
/* (scale, rotation).  Computed in double from the float inputs.                                */
/* dL_dopacity_io: in = gradient w.r.t. the effective opacity stored in conic_opacity.w,        */
/*                 out = gradient w.r.t. the input opacity (differs only with antialiasing).    */
/* ------------------------------------------------------------------------------------------ */
ORC_API void orc_preprocess_backward(
    int P, int D, int M, int W, int H,
    const float* means3D, const float* scales, float mod, const float* rotations,
    const float* opacities, const float* shs, const float* cov3D_precomp,
    const float* view, const float* proj, const float* campos, float tanfovx, float tanfovy,
    int antialiasing, const int32_t* radii, const float* cov3D, const uint8_t* clamped,
    const float* dL_dmean2D, const float* dL_dconic, float* dL_dopacity_io, const float* dL_dcolor,
    const float* dL_dinvd /* nullable */,
    float* dL_dmeans3D, float* dL_dsh, float* dL_dscales, float* dL_drot, float* dL_dcov3D)
{
    const double focal_x = (double)((float)W / (2.0f * tanfovx));
    const double focal_y = (double)((float)H / (2.0f * tanfovy));
    const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
    const int has_sh = (shs != NULL);

#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        for (int k = 0; k < 3; k++) dL_dmeans3D[3 * i + k] = 0.0f;
        if (has_sh) for (int k = 0; k < 3 * M; k++) dL_dsh[(size_t)i * 3 * M + k] = 0.0f;
        if (dL_dscales) for (int k = 0; k < 3; k++) dL_dscales[3 * i + k] = 0.0f;
        if (dL_drot) for (int k = 0; k < 4; k++) dL_drot[4 * i + k] = 0.0f;
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * i + k] = 0.0f;
        if (radii[i] <= 0) { dL_dopacity_io[i] = 0.0f; continue; }

        const double x = means3D[3 * i], y = means3D[3 * i + 1], z = means3D[3 * i + 2];
        /* float view-space point exactly as in the forward (for the clamp decision) */
        float txf = fmaf(view[8], (float)z, fmaf(view[4], (float)y, view[0] * (float)x)) + view[12];
        float tyf = fmaf(view[9], (float)z, fmaf(view[5], (float)y, view[1] * (float)x)) + view[13];
        float tzf = fmaf(view[10], (float)z, fmaf(view[6], (float)y, view[2] * (float)x)) + view[14];
        float txtz = txf / tzf, tytz = tyf / tzf;
        const double xmul = (txtz < -limx || txtz > limx) ? 0.0 : 1.0;
        const double ymul = (tytz < -limy || tytz > limy) ? 0.0 : 1.0;
        const double tz = tzf;
        const double tx = (double)(fminf(limx, fmaxf(-limx, txtz)) * tzf);
        const double ty = (double)(fminf(limy, fmaxf(-limy, tytz)) * tzf);

        const float* c6 = cov3D + 6 * (size_t)i;
        double S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        double J00 = focal_x / tz, J02 = -(focal_x * tx) / (tz * tz);
        double J11 = focal_y / tz, J12 = -(focal_y * ty) / (tz * tz);
        double Wv[3][3];
        for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) Wv[r][k] = view[4 * k + r];
        double M0[3], M1[3], N0[3], N1[3];
        for (int k = 0; k < 3; k++) { M0[k] = J00 * Wv[0][k] + J02 * Wv[2][k]; M1[k] = J11 * Wv[1][k] + J12 * Wv[2][k]; }
        for (int k = 0; k < 3; k++) {
            N0[k] = M0[0] * S[0][k] + M0[1] * S[1][k] + M0[2] * S[2][k];
            N1[k] = M1[0] * S[0][k] + M1[1] * S[1][k] + M1[2] * S[2][k];
        }
        double cov00 = N0[0] * M0[0] + N0[1] * M0[1] + N0[2] * M0[2];
        double cov01 = N0[0] * M1[0] + N0[1] * M1[1] + N0[2] * M1[2];
        double cov11 = N1[0] * M1[0] + N1[1] * M1[1] + N1[2] * M1[2];
        double a = cov00 + 0.3, b = cov01, c = cov11 + 0.3;
        double det = a * c - b * b;

        double LA = dL_dconic[3 * i], LB = dL_dconic[3 * i + 1], LC = dL_dconic[3 * i + 2];
        double dL_da = 0, dL_db = 0, dL_dc = 0;
        double d2inv = 1.0 / (det * det + 0.0000001);
        dL_da = d2inv * (-c * c * LA + b * c * LB - b * b * LC);   /* det - a c = -b^2 */
        dL_dc = d2inv * (-a * a * LC + a * b * LB - b * b * LA);
        dL_db = d2inv * (2.0 * b * c * LA - (det + 2.0 * b * b) * LB + 2.0 * a * b * LC);

        double dLo = dL_dopacity_io[i];
        if (antialiasing) {
            double det0 = cov00 * cov11 - cov01 * cov01;
            double ratio = det0 / det;
            double h = sqrt(ratio > 0.000025 ? ratio : 0.000025);
            double dL_dh = (double)opacities[i] * dLo;
            dLo = h * dLo;
            if (ratio > 0.000025) {
                double dL_dratio = dL_dh * 0.5 / h;
                double inv2 = 1.0 / (det * det);
                dL_da += dL_dratio * (cov11 * det - det0 * c) * inv2;
                dL_dc += dL_dratio * (cov00 * det - det0 * a) * inv2;
                dL_db += dL_dratio * (-2.0 * cov01 * det + det0 * 2.0 * cov01) * inv2;
            }
        }
        dL_dopacity_io[i] = (float)dLo;

        /* cov2D -> packed cov3D (off-diagonals carry both symmetric entries) */
        double dc6[6];
        dc6[0] = M0[0] * M0[0] * dL_da + M0[0] * M1[0] * dL_db + M1[0] * M1[0] * dL_dc;
        dc6[3] = M0[1] * M0[1] * dL_da + M0[1] * M1[1] * dL_db + M1[1] * M1[1] * dL_dc;
        dc6[5] = M0[2] * M0[2] * dL_da + M0[2] * M1[2] * dL_db + M1[2] * M1[2] * dL_dc;
        dc6[1] = 2 * M0[0] * M0[1] * dL_da + (M0[0] * M1[1] + M0[1] * M1[0]) * dL_db + 2 * M1[0] * M1[1] * dL_dc;
        dc6[2] = 2 * M0[0] * M0[2] * dL_da + (M0[0] * M1[2] + M0[2] * M1[0]) * dL_db + 2 * M1[0] * M1[2] * dL_dc;
        dc6[4] = 2 * M0[1] * M0[2] * dL_da + (M0[1] * M1[2] + M0[2] * M1[1]) * dL_db + 2 * M1[1] * M1[2] * dL_dc;
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * i + k] = (float)dc6[k];

        /* cov2D -> M -> J -> t -> mean */
        double dM0[3], dM1[3];
        for (int k = 0; k < 3; k++) { dM0[k] = 2 * dL_da * N0[k] + dL_db * N1[k]; dM1[k] = dL_db * N0[k] + 2 * dL_dc * N1[k]; }
        double dJ00 = 0, dJ02 = 0, dJ11 = 0, dJ12 = 0;
        for (int k = 0; k < 3; k++) { dJ00 += dM0[k] * Wv[0][k]; dJ02 += dM0[k] * Wv[2][k]; dJ11 += dM1[k] * Wv[1][k]; dJ12 += dM1[k] * Wv[2][k]; }
        double itz = 1.0 / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        double dtx = xmul * (-focal_x * itz2) * dJ02;
        double dty = ymul * (-focal_y * itz2) * dJ12;
        double dtz = -focal_x * itz2 * dJ00 - focal_y * itz2 * dJ11 + (2 * focal_x * tx) * itz3 * dJ02 + (2 * focal_y * ty) * itz3 * dJ12;
        /* inverse depth 1/tz feeds the depth channel */
        if (dL_dinvd) dtz += -(double)dL_dinvd[i] * itz2;
        double dmean[3];
        dmean[0] = Wv[0][0] * dtx + Wv[1][0] * dty + Wv[2][0] * dtz;
        dmean[1] = Wv[0][1] * dtx + Wv[1][1] * dty + Wv[2][1] * dtz;
        dmean[2] = Wv[0][2] * dtx + Wv[1][2] * dty + Wv[2][2] * dtz;

        /* mean2D (NDC gradient) -> mean3D through the full projection */
        {
            double hx = proj[0] * x + proj[4] * y + proj[8] * z + proj[12];
            double hy = proj[1] * x + proj[5] * y + proj[9] * z + proj[13];
            double hw = proj[3] * x + proj[7] * y + proj[11] * z + proj[15];
            double mw = 1.0 / (hw + 0.0000001);
            double mul1 = hx * mw * mw, mul2 = hy * mw * mw;
            double gxn = dL_dmean2D[2 * i], gyn = dL_dmean2D[2 * i + 1];
            dmean[0] += (proj[0] * mw - proj[3] * mul1) * gxn + (proj[1] * mw - proj[3] * mul2) * gyn;
            dmean[1] += (proj[4] * mw - proj[7] * mul1) * gxn + (proj[5] * mw - proj[7] * mul2) * gyn;
            dmean[2] += (proj[8] * mw - proj[11] * mul1) * gxn + (proj[9] * mw - proj[11] * mul2) * gyn;
        }

        /* SH -> dL/dsh and dL/dmean via the normalised view direction */
        if (has_sh) {
            double ddx = x - campos[0], ddy = y - campos[1], ddz = z - campos[2];
            double len = sqrt(ddx * ddx + ddy * ddy + ddz * ddz);
            double ux = ddx / len, uy = ddy / len, uz = ddz / len;
            double gcol[3];
            for (int ch = 0; ch < 3; ch++) gcol[ch] = clamped[3 * i + ch] ? 0.0 : (double)dL_dcolor[3 * i + ch];
            double bb[16], bx[16], by[16], bz[16];
            memset(bb, 0, sizeof(bb)); memset(bx, 0, sizeof(bx)); memset(by, 0, sizeof(by)); memset(bz, 0, sizeof(bz));
            bb[0] = 0.28209479177387814;
            if (D > 0) {
                const double C1 = 0.4886025119029199;
                bb[1] = -C1 * uy; by[1] = -C1;
                bb[2] = C1 * uz;  bz[2] = C1;
                bb[3] = -C1 * ux; bx[3] = -C1;
                if (D > 1) {
                    const double c20 = 1.0925484305920792, c22 = 0.31539156525252005, c24 = 0.5462742152960396;
                    double xx = ux * ux, yy = uy * uy, zz = uz * uz;
                    bb[4] = c20 * ux * uy;  bx[4] = c20 * uy; by[4] = c20 * ux;
                    bb[5] = -c20 * uy * uz; by[5] = -c20 * uz; bz[5] = -c20 * uy;
                    bb[6] = c22 * (2 * zz - xx - yy); bx[6] = c22 * -2 * ux; by[6] = c22 * -2 * uy; bz[6] = c22 * 4 * uz;
                    bb[7] = -c20 * ux * uz; bx[7] = -c20 * uz; bz[7] = -c20 * ux;
                    bb[8] = c24 * (xx - yy); bx[8] = c24 * 2 * ux; by[8] = c24 * -2 * uy;
                    if (D > 2) {
                        const double c30 = -0.5900435899266435, c31 = 2.890611442640554, c32 = -0.4570457994644658,
                                     c33 = 0.3731763325901154, c35 = 1.445305721320277;
                        bb[9] = c30 * uy * (3 * xx - yy);  bx[9] = c30 * uy * 6 * ux; by[9] = c30 * (3 * xx - 3 * yy);
                        bb[10] = c31 * ux * uy * uz;       bx[10] = c31 * uy * uz; by[10] = c31 * ux * uz; bz[10] = c31 * ux * uy;
                        bb[11] = c32 * uy * (4 * zz - xx - yy); bx[11] = c32 * uy * -2 * ux; by[11] = c32 * (4 * zz - xx - 3 * yy); bz[11] = c32 * uy * 8 * uz;
                        bb[12] = c33 * uz * (2 * zz - 3 * xx - 3 * yy); bx[12] = c33 * uz * -6 * ux; by[12] = c33 * uz * -6 * uy; bz[12] = c33 * (6 * zz - 3 * xx - 3 * yy);
                        bb[13] = c32 * ux * (4 * zz - xx - yy); bx[13] = c32 * (4 * zz - 3 * xx - yy); by[13] = c32 * ux * -2 * uy; bz[13] = c32 * ux * 8 * uz;
                        bb[14] = c35 * uz * (xx - yy);     bx[14] = c35 * uz * 2 * ux; by[14] = c35 * uz * -2 * uy; bz[14] = c35 * (xx - yy);
                        bb[15] = c30 * ux * (xx - 3 * yy); bx[15] = c30 * (3 * xx - 3 * yy); by[15] = c30 * ux * -6 * uy;
                    }
                }
            }
            int nb = (D + 1) * (D + 1);
            double dRx = 0, dRy = 0, dRz = 0;   /* dL/d(unit dir) */
            for (int k = 0; k < nb; k++) for (int ch = 0; ch < 3; ch++) {
                double s = shs[(size_t)i * M * 3 + 3 * k + ch];
                dL_dsh[(size_t)i * M * 3 + 3 * k + ch] = (float)(bb[k] * gcol[ch]);
                dRx += bx[k] * s * gcol[ch]; dRy += by[k] * s * gcol[ch]; dRz += bz[k] * s * gcol[ch];
            }
            double dotp = ux * dRx + uy * dRy + uz * dRz;
            dmean[0] += (dRx - ux * dotp) / len;
            dmean[1] += (dRy - uy * dotp) / len;
            dmean[2] += (dRz - uz * dotp) / len;
        }
        for (int k = 0; k < 3; k++) dL_dmeans3D[3 * i + k] = (float)dmean[k];

        /* cov3D -> scale, rotation */
        if (!cov3D_precomp && dL_dscales && dL_drot) {
            const float* qf = rotations + 4 * (size_t)i;
            double r = qf[0], qx = qf[1], qy = qf[2], qz = qf[3];
            double R[3][3] = {
                {1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - r * qz), 2 * (qx * qz + r * qy)},
                {2 * (qx * qy + r * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - r * qx)},
                {2 * (qx * qz - r * qy), 2 * (qy * qz + r * qx), 1 - 2 * (qx * qx + qy * qy)}};
            double s[3] = {(double)mod * scales[3 * i], (double)mod * scales[3 * i + 1], (double)mod * scales[3 * i + 2]};
            double G3[3][3] = {{dc6[0], 0.5 * dc6[1], 0.5 * dc6[2]}, {0.5 * dc6[1], dc6[3], 0.5 * dc6[4]}, {0.5 * dc6[2], 0.5 * dc6[4], dc6[5]}};
            double L[3][3], dLm[3][3];
            for (int a2 = 0; a2 < 3; a2++) for (int j = 0; j < 3; j++) L[a2][j] = R[a2][j] * s[j];
            for (int a2 = 0; a2 < 3; a2++) for (int j = 0; j < 3; j++) {
                double v = 0; for (int k = 0; k < 3; k++) v += G3[a2][k] * L[k][j];
                dLm[a2][j] = 2.0 * v;
            }
            double g[3][3];
            for (int j = 0; j < 3; j++) {
                double ds = 0;
                for (int a2 = 0; a2 < 3; a2++) { ds += R[a2][j] * dLm[a2][j]; g[a2][j] = dLm[a2][j] * s[j]; }
                dL_dscales[3 * i + j] = (float)((double)mod * ds);
            }
            double dr = 2 * (-qz * g[0][1] + qy * g[0][2] + qz * g[1][0] - qx * g[1][2] - qy * g[2][0] + qx * g[2][1]);
            double dx = 2 * (qy * g[0][1] + qz * g[0][2] + qy * g[1][0] - 2 * qx * g[1][1] - r * g[1][2] + qz * g[2][0] + r * g[2][1] - 2 * qx * g[2][2]);
            double dy = 2 * (-2 * qy * g[0][0] + qx * g[0][1] + r * g[0][2] + qx * g[1][0] + qz * g[1][2] - r * g[2][0] + qz * g[2][1] - 2 * qy * g[2][2]);
            double dz = 2 * (-2 * qz * g[0][0] - r * g[0][1] + qx * g[0][2] + r * g[1][0] - 2 * qz * g[1][1] + qy * g[1][2] + qx * g[2][0] + qy * g[2][1]);
            dL_drot[4 * i] = (float)dr; dL_drot[4 * i + 1] = (float)dx; dL_drot[4 * i + 2] = (float)dy; dL_drot[4 * i + 3] = (float)dz;
        }
    }
}

ORC_API int orc_version(void) { return 1; }
