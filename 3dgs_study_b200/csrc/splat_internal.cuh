// splat_internal.cuh -- shared declarations of libb200splat (sm_100a).  Not part of the C ABI.
//
// State layout in HBM (all chunks 256-byte aligned inside the caller-allocated opaque buffers):
//
//   geometry buffer (per Gaussian, P entries)
//     depths   f32[P]        view-space z; culled Gaussians hold the bit pattern 0xFFFFFFFF.  Read as u32 it is the
//                            depth-sort key (positive floats order like their bit patterns)
//     rec      float4[3P]    packed 48-byte splat record, gathered by the render kernels:
//                              rec[3i+0] = (px, py, ex, ey)   pixel centre, half extents of {alpha >= 1/255}
//                              rec[3i+1] = (hA, B, hC, thr)   hA = -A/2, hC = -C/2 (exact), power threshold
//                              rec[3i+2] = (o, r, g, b)       effective opacity, colour
//     rect     uint2[P]      tile rectangle (minx | miny << 16, w | h << 16); w*h = tiles touched (0 = culled)
//     dkey_a/b u32[P] x2, dval_a/b u32[P] x2   ping-pong arrays of the depth sort; after its four passes
//                            dval_b = `order`: the ids of the V visible Gaussians by (depth bits, id)
//     emit_part u32[P/256+1] pairs emitted before each block of 256 depth positions
//     scan_part, dmatrix, dtotals   block partials of the export scan; digit-count matrix / totals of the depth passes
//   binning buffer (per (tile, Gaussian) pair, `cap` entries -- the pair CAPACITY, >= the pair count R unless the
//   overflow flag is raised; tile keys are u16 when T <= 65536, else u32)
//     key_a[cap], val_a[cap]   pairs as emitted (depth order); val_a is overwritten by the final, sorted ids
//     key_b[cap], val_b[cap]   pairs after the first (tile column) pass
//     tmatrix, ttotals         digit-count matrix / totals of the two tile passes
//   image buffer
//     ranges   int2[T]       [start, end) of each tile in the sorted list
//     final_T  f32[H*W], n_contrib u32[H*W], tile_order u32[T]
//     counters u32[256]      control block, every hot word on its own 128-byte line (CTR_* below): pair count R
//                            (u64), visible count V, min / max depth key, capacity + overflow status
//
// Binning order.  The list of every tile must end up sorted by (depth bits, Gaussian id), i.e. exactly
// the order a stable sort of the 64-bit (tile << 32 | depth) keys emitted in Gaussian order gives.  We
// get the same order with far less traffic: Gaussians are sorted ONCE by depth (stable, V elements),
// pairs are emitted in that order with only the tile id as key, and a stable two-pass sort on the tile
// id (by tile column, then by tile row) finishes the job.  Every sort pass and every scan is a kernel of
// this library (binning.cu); nothing on this path calls a library sort or scan, and nothing waits on the
// host for R (b200splat.h "pair capacity").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/b200splat.h"

namespace b200splat {

constexpr int TILE = 16;
constexpr float NEAR_Z = 0.2f;
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float ALPHA_MAX = 0.99f;
constexpr float T_EPS = 0.0001f;
// power-threshold safety margin (natural-log units): a pair is skipped before the exp only when
// o*exp(power) < (1/255)*exp(-THR_MARGIN), far outside any fp32 rounding of the exact test.
constexpr float THR_MARGIN = 1.0e-3f;

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// sets the thread-local string b200splat_last_error() returns; returns `code` (api.cu)
int report_error(int code, const char* msg);

struct GeomView {
    float* depths; float4* rec; uint2* rect;
    uint32_t* dkey_a; uint32_t* dval_a; uint32_t* dkey_b; uint32_t* dval_b;
    uint32_t* order;            // = dval_b: result of the depth sort
    uint32_t* emit_part;        // pairs per emit block (256 depth positions) -> after the spine scan, their offsets
    uint32_t* scan_part;        // block partials of the export-only scan
    uint32_t* dmatrix; uint32_t* dtotals;
    int dblocks;                // row stride of dmatrix = radix blocks covering P keys
    int emit_blocks;
    size_t total;
};
struct BinView {
    void* key_a; uint32_t* val_a; void* key_b; uint32_t* val_b;
    uint32_t* vals_sorted;      // = val_a
    uint32_t* tmatrix; uint32_t* ttotals;
    int tblocks;                // row stride of tmatrix = radix blocks covering cap pairs
    int cap;
    size_t total; int key_bytes;
};
struct ImgView {
    int2* ranges; float* final_T; uint32_t* n_contrib; unsigned int* counters;
    uint32_t* tile_order;       // tiles by decreasing list length (NULL: index order)
    size_t clear_bytes;         // ranges + counters are adjacent at the head of the buffer: the bytes a forward clears
    size_t total;
};

GeomView carve_geom(void* base, int P);
BinView carve_binning(void* base, int cap, int num_tiles);
ImgView carve_image(void* base, int W, int H);
inline int tiles_of(int W, int H) { return ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE); }
// slots of ImgView::counters (u32 index); each group sits on its own 128-byte line
constexpr int CTR_WORDS = 256;
constexpr int CTR_PAIRS = 32;      // u64: pair count R = sum of tiles touched
constexpr int CTR_VISIBLE = 64;    // V = Gaussians with tiles touched > 0
constexpr int CTR_KMIN_NEG = 96;   // max over visible of ~depth key  (min key = ~this; 0 = none)
constexpr int CTR_KMAX = 128;      // max over visible of depth key
constexpr int CTR_STATUS = 160;    // written by the emit stage: [0..1] R (u64), [2] pair capacity of the binning buffer,
                                   // [3] 1 if R > capacity: the view's outputs are unspecified (never out of bounds)
int max_tile_bins();               // tile columns / rows the tile sort supports

// ---- stage launchers (each enqueues on `s`, returns the launch status) ---------------------------
// preprocess also accumulates R, V and the min / max depth key into iv.counters (cleared by the caller)
cudaError_t launch_preprocess_fwd(const b200splat_view& v, const b200splat_gaussians& g, GeomView gv,
                                  int32_t* radii, unsigned int* counters, cudaStream_t s);
// stable sort of the visible Gaussians by (depth bits, id) + inclusive scan of tiles touched in that order
cudaError_t launch_depth_order(GeomView gv, int P, const unsigned int* counters, cudaStream_t s);
cudaError_t launch_emit(GeomView gv, int P, int W, int H, BinView bv, unsigned int* counters, cudaStream_t s);
// stable sort of the pairs by tile id (two passes) + tile ranges + tile order; iv.ranges cleared by the caller
cudaError_t launch_tile_sort(BinView bv, int W, int H, ImgView iv, cudaStream_t s);
cudaError_t launch_tile_order(ImgView iv, int W, int H, cudaStream_t s);
cudaError_t launch_render_fwd(const b200splat_view& v, GeomView gv, BinView bv, ImgView iv, float* out_color,
                              float* out_invdepth, cudaStream_t s);
// packed per-Gaussian gradient accumulator written by the render backward, 12 floats / Gaussian:
//   (d_ndc_x, d_ndc_y are produced later)   acc[12i+0..] = (dpx, dpy, dA, dB, dC, do, dr, dg, db, dinvd, -, -)
// touched[P] (bytes, cleared by the caller): set to 1 for every Gaussian whose accumulator row received a sum
cudaError_t launch_render_bwd(const b200splat_view& v, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                              const float* dL_dpix, const float* dL_dinvdepth, float* acc, uint8_t* touched, cudaStream_t s);
// indices of the touched Gaussians, in no particular order; *count (cleared by the caller) receives their number
cudaError_t launch_compact_active(int P, uint8_t* touched, bool clear_flags, uint32_t* active_idx, unsigned* count, cudaStream_t s);
// Per-Gaussian backward over the touched Gaussians only: writes the rows of `out` they own and nothing else --
// the caller has zero-filled `out` (every gradient of an untouched Gaussian is exactly zero).
cudaError_t launch_preprocess_bwd(const b200splat_view& v, const b200splat_gaussians& g, const int32_t* radii,
                                  GeomView gv, float* acc, bool rezero_acc, const b200splat_grads& out, const uint32_t* active_idx,
                                  const unsigned* active_count, float* packed_rows, unsigned* packed_count, float* norm_accum,
                                  bool write_header, cudaStream_t s);
cudaError_t launch_densify_stats_visible(int P, const int32_t* radii, float* denom, float* max_radii2D, cudaStream_t s);
// packed_rows != NULL: instead of the dense outputs (only out.dL_dmeans2D is still written), one 64-byte row per
// touched Gaussian and their number; launch_accumulate_packed adds the rows of any number of views (local or in
// a peer GPU's memory) into dense gradients.
// Zeroes the rows of dL_dsh / dL_dgeom that launch_accumulate_packed wrote (it sets marks[i] for each), and the marks:
// the dense gradients of a step are cleared for what they cost, not for P.
cudaError_t launch_clear_accumulated(int P, int M, float* dL_dsh, float* dL_dgeom, uint8_t* marks, cudaStream_t s);
cudaError_t launch_accumulate_packed(int P, int M, int D, const float* means3D, int n_src, const float* const* rows,
                                     const unsigned* const* counts, const float* campos, float* dL_dsh, float* dL_dgeom,
                                     uint8_t* marks, int store, cudaStream_t s);
cudaError_t launch_render_stats(int W, int H, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                                unsigned long long* out16, cudaStream_t s);
cudaError_t launch_masked_colour_grad(int P, const int32_t* radii, GeomView gv, const float* acc, float* dcol, cudaStream_t s);
cudaError_t launch_sh_grad_from_views(int P, int M, int D, const float* means3D, int n_views, const float* const* dcol,
                                      const float* campos, float* dL_dsh, cudaStream_t s);
cudaError_t launch_densify_stats(int P, const float* dmeans2D, const int32_t* radii, float* accum, float* denom,
                                 float* max_radii, cudaStream_t s);
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);
// Background clear: zero up to 8 device ranges with ONE block per SM, so that the clear runs BESIDE a compute-bound
// kernel (the render backward) instead of taking every thread slot of the GPU for its 0.1 ms the way a full-grid
// memset does (measured in profiles/r02_timeline_n8_rank0_graph_v1.txt: the two ran back to back).
struct ZeroRanges { int n; void* ptr[8]; size_t bytes[8]; };
cudaError_t launch_zero_fill(const ZeroRanges& zr, cudaStream_t s);
cudaError_t launch_export_geom(int P, const int32_t* radii, GeomView gv, const b200splat_state_out& o, cudaStream_t s);
cudaError_t launch_export_binning(int P, int W, int H, GeomView gv, BinView bv, ImgView iv, const b200splat_state_out& o, cudaStream_t s);

// ---- canonical fp32 arithmetic (DESIGN.md "Canonical arithmetic") --------------------------------
// Every rounding point is an explicit intrinsic, so the compiler can neither contract nor
// reassociate; oracle/splat_oracle.c implements the same sequence with fmaf and -ffp-contract=off.
#ifdef __CUDACC__
__device__ __forceinline__ float cmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float cadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float csub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float cfma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float cdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float csqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float cdot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return cfma(a2, b2, cfma(a1, b1, cmul(a0, b0)));
}

struct ViewConsts {   // staged in shared memory once per block
    float view[16];
    float proj[16];
    float campos[3];
};

__device__ __forceinline__ void load_view_consts(ViewConsts& vc, const float* view, const float* proj,
                                                 const float* campos) {
    int t = threadIdx.x;
    if (t < 16) vc.view[t] = view[t];
    else if (t < 32) vc.proj[t - 16] = proj[t - 16];
    else if (t < 35 && campos) vc.campos[t - 32] = campos[t - 32];
    __syncthreads();
}

// tile rectangle of a splat: truncation toward zero (saturating), then clamp to the grid
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, int& minx, int& miny,
                                         int& maxx, int& maxy) {
    float fr = (float)radius;
    minx = min(gx, max(0, (int)cmul(csub(px, fr), 0.0625f)));
    miny = min(gy, max(0, (int)cmul(csub(py, fr), 0.0625f)));
    maxx = min(gx, max(0, (int)cmul(csub(cadd(cadd(px, fr), 16.0f), 1.0f), 0.0625f)));
    maxy = min(gy, max(0, (int)cmul(csub(cadd(cadd(py, fr), 16.0f), 1.0f), 0.0625f)));
}

// 3D covariance (xx,xy,xz,yy,yz,zz) from scale * modifier and a unit quaternion (w,x,y,z)
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* c6) {
    float r = q.x, x = q.y, y = q.z, z = q.w;
    float R00 = cfma(-2.0f, cfma(y, y, cmul(z, z)), 1.0f);
    float R01 = cmul(2.0f, cfma(x, y, -cmul(r, z)));
    float R02 = cmul(2.0f, cfma(x, z, cmul(r, y)));
    float R10 = cmul(2.0f, cfma(x, y, cmul(r, z)));
    float R11 = cfma(-2.0f, cfma(x, x, cmul(z, z)), 1.0f);
    float R12 = cmul(2.0f, cfma(y, z, -cmul(r, x)));
    float R20 = cmul(2.0f, cfma(x, z, -cmul(r, y)));
    float R21 = cmul(2.0f, cfma(y, z, cmul(r, x)));
    float R22 = cfma(-2.0f, cfma(x, x, cmul(y, y)), 1.0f);
    float sx = cmul(mod, s0), sy = cmul(mod, s1), sz = cmul(mod, s2);
    float L00 = cmul(R00, sx), L01 = cmul(R01, sy), L02 = cmul(R02, sz);
    float L10 = cmul(R10, sx), L11 = cmul(R11, sy), L12 = cmul(R12, sz);
    float L20 = cmul(R20, sx), L21 = cmul(R21, sy), L22 = cmul(R22, sz);
    c6[0] = cdot3(L00, L00, L01, L01, L02, L02);
    c6[1] = cdot3(L00, L10, L01, L11, L02, L12);
    c6[2] = cdot3(L00, L20, L01, L21, L02, L22);
    c6[3] = cdot3(L10, L10, L11, L11, L12, L12);
    c6[4] = cdot3(L10, L20, L11, L21, L12, L22);
    c6[5] = cdot3(L20, L20, L21, L21, L22, L22);
}

// SH basis at a unit direction; polynomial forms of /root/reference/utils/sh_utils.py:74-100
__device__ __forceinline__ void sh_basis(int D, float x, float y, float z, float* b) {
    b[0] = 0.28209479177387814f;
    if (D > 0) {
        const float C1 = 0.4886025119029199f;
        b[1] = -cmul(C1, y); b[2] = cmul(C1, z); b[3] = -cmul(C1, x);
        if (D > 1) {
            float xx = cmul(x, x), yy = cmul(y, y), zz = cmul(z, z);
            float xy = cmul(x, y), yz = cmul(y, z), xz = cmul(x, z);
            b[4] = cmul(1.0925484305920792f, xy);
            b[5] = cmul(-1.0925484305920792f, yz);
            b[6] = cmul(0.31539156525252005f, csub(csub(cmul(2.0f, zz), xx), yy));
            b[7] = cmul(-1.0925484305920792f, xz);
            b[8] = cmul(0.5462742152960396f, csub(xx, yy));
            if (D > 2) {
                b[9]  = cmul(cmul(-0.5900435899266435f, y), cfma(3.0f, xx, -yy));
                b[10] = cmul(cmul(2.890611442640554f, xy), z);
                b[11] = cmul(cmul(-0.4570457994644658f, y), csub(csub(cmul(4.0f, zz), xx), yy));
                b[12] = cmul(cmul(0.3731763325901154f, z), csub(csub(cmul(2.0f, zz), cmul(3.0f, xx)), cmul(3.0f, yy)));
                b[13] = cmul(cmul(-0.4570457994644658f, x), csub(csub(cmul(4.0f, zz), xx), yy));
                b[14] = cmul(cmul(1.445305721320277f, z), csub(xx, yy));
                b[15] = cmul(cmul(-0.5900435899266435f, x), cfma(-3.0f, yy, xx));
            }
        }
    }
}
#endif  // __CUDACC__

}  // namespace b200splat
