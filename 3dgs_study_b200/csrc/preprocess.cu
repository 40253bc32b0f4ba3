// preprocess.cu -- per-Gaussian forward (cull, 3D covariance, EWA projection, conic, radius, tile
// rectangle, SH -> RGB, packed splat record) and per-Gaussian backward (conic -> cov2D -> cov3D /
// mean, mean2D -> mean3D, SH, cov3D -> scale / rotation), sm_100a.
//
// Replaces the per-Gaussian stages of _C.rasterize_gaussians / _C.rasterize_gaussians_backward
// named by BASELINE.json's north_star (call site /root/reference/gaussian_renderer/__init__.py:98-106).
// Both kernels are HBM-bound: one thread per Gaussian, cull before the 12*M-byte SH row is touched,
// 16-byte vector loads/stores for the SH rows and the packed record.
#include "splat_internal.cuh"

namespace b200splat {

namespace {

#ifndef PRE_T
#define PRE_T 128
#endif
constexpr int PRE_THREADS = PRE_T;
#ifndef PRE_MINB_FWD
#define PRE_MINB_FWD 8
#endif
#ifndef PRE_MINB_BWD
#define PRE_MINB_BWD 3     // the fp64 geometry chain wants registers; the kernel visits few Gaussians and is latency-bound
#endif

// row stride (floats) of an SH row staged in shared memory: an odd number of 16-byte words
__device__ __forceinline__ int sh_row_stride(int M3) { return ((M3 >> 2) & 1) ? M3 : M3 + 4; }

struct PreParams {
    int P, D, M, W, H, gx, gy;
    float tanfovx, tanfovy, focal_x, focal_y, mod;
    int antialiasing;
    const float* means3D; const float* scales; const float* rotations; const float* opacities;
    const float* shs; const float* colors_precomp; const float* cov3D_precomp;
    const float* view; const float* proj; const float* campos;
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// colour of one Gaussian from its SH row (canonical chain: res = b0*sh0; res = fma(bk, shk, res))
template <bool GLOBAL = true>
__device__ __forceinline__ void eval_color_sh(const float* __restrict__ sh_row, int M, int D, float dx, float dy,
                                              float dz, float* rgb, unsigned& clamped) {
    float len = csqrt(cfma(dz, dz, cfma(dy, dy, cmul(dx, dx))));
    float b[16];
    sh_basis(D, cdiv(dx, len), cdiv(dy, len), cdiv(dz, len), b);
    const int nb = (D + 1) * (D + 1);
    float r0, r1, r2;
    if ((M & 3) == 0) {
        // 3*M floats, 16-byte aligned row: vector loads
        const float4* v = reinterpret_cast<const float4*>(sh_row);
        float f[48];
        const int nv = (3 * nb + 3) >> 2;
#pragma unroll
        for (int k = 0; k < 12; k++) {
            if (k < nv) {
                float4 t = GLOBAL ? __ldg(v + k) : v[k];
                f[4 * k] = t.x; f[4 * k + 1] = t.y; f[4 * k + 2] = t.z; f[4 * k + 3] = t.w;
            }
        }
        r0 = cmul(b[0], f[0]); r1 = cmul(b[0], f[1]); r2 = cmul(b[0], f[2]);
#pragma unroll
        for (int k = 1; k < 16; k++) {
            if (k < nb) {
                r0 = cfma(b[k], f[3 * k], r0); r1 = cfma(b[k], f[3 * k + 1], r1); r2 = cfma(b[k], f[3 * k + 2], r2);
            }
        }
    } else {
        r0 = cmul(b[0], sh_row[0]); r1 = cmul(b[0], sh_row[1]); r2 = cmul(b[0], sh_row[2]);
#pragma unroll
        for (int k = 1; k < 16; k++) {       // static indices keep the basis in registers
            if (k < nb) {
                r0 = cfma(b[k], sh_row[3 * k], r0);
                r1 = cfma(b[k], sh_row[3 * k + 1], r1);
                r2 = cfma(b[k], sh_row[3 * k + 2], r2);
            }
        }
    }
    r0 = cadd(r0, 0.5f); r1 = cadd(r1, 0.5f); r2 = cadd(r2, 0.5f);
    clamped = (r0 < 0.0f ? 1u : 0u) | (r1 < 0.0f ? 2u : 0u) | (r2 < 0.0f ? 4u : 0u);
    // a clamped channel is stored as -0.0f: blending with it is bit-identical to +0.0f, and the sign bit
    // carries the clamp flag to the backward pass (masked_colour_grad_kernel) for free
    const float nz = __int_as_float(0x80000000);
    rgb[0] = r0 < 0.0f ? nz : r0; rgb[1] = r1 < 0.0f ? nz : r1; rgb[2] = r2 < 0.0f ? nz : r2;
}

// Geometry shared by forward and backward: everything up to the 2D covariance.
struct Proj2D {
    float tz, txc, tyc;        // view depth, clamped view x / y
    float txtz, tytz;
    float M0[3], M1[3];        // rows of J * Wv
    float N0[3], N1[3];        // rows of (J Wv) Sigma
    float cov00, cov01, cov11;
};

__device__ __forceinline__ void project_cov(const ViewConsts& vc, float tx, float ty, float tz, const float* c6,
                                            float focal_x, float focal_y, float limx, float limy, Proj2D& o) {
    o.tz = tz;
    o.txtz = cdiv(tx, tz); o.tytz = cdiv(ty, tz);
    o.txc = cmul(fminf(limx, fmaxf(-limx, o.txtz)), tz);
    o.tyc = cmul(fminf(limy, fmaxf(-limy, o.tytz)), tz);
    float tz2 = cmul(tz, tz);
    float J00 = cdiv(focal_x, tz), J02 = cdiv(-cmul(focal_x, o.txc), tz2);
    float J11 = cdiv(focal_y, tz), J12 = cdiv(-cmul(focal_y, o.tyc), tz2);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        o.M0[k] = cfma(J02, vc.view[4 * k + 2], cmul(J00, vc.view[4 * k + 0]));
        o.M1[k] = cfma(J12, vc.view[4 * k + 2], cmul(J11, vc.view[4 * k + 1]));
    }
    const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        o.N0[k] = cdot3(o.M0[0], S[0][k], o.M0[1], S[1][k], o.M0[2], S[2][k]);
        o.N1[k] = cdot3(o.M1[0], S[0][k], o.M1[1], S[1][k], o.M1[2], S[2][k]);
    }
    o.cov00 = cdot3(o.N0[0], o.M0[0], o.N0[1], o.M0[1], o.N0[2], o.M0[2]);
    o.cov01 = cdot3(o.N0[0], o.M1[0], o.N0[1], o.M1[1], o.N0[2], o.M1[2]);
    o.cov11 = cdot3(o.N1[0], o.M1[0], o.N1[1], o.M1[1], o.N1[2], o.M1[2]);
}

__global__ void __launch_bounds__(PRE_THREADS, PRE_MINB_FWD) preprocess_fwd_kernel(PreParams p, float* __restrict__ depths,
                                                                     float4* __restrict__ rec,
                                                                     uint2* __restrict__ rect,
                                                                     int32_t* __restrict__ radii,
                                                                     unsigned int* __restrict__ counters,
                                                                     int coop, int qs) {
    extern __shared__ float4 s_sh4[];
    __shared__ ViewConsts vc;
    __shared__ uint32_t s_pairs[PRE_THREADS / 32], s_vis[PRE_THREADS / 32], s_kmax[PRE_THREADS / 32], s_kminn[PRE_THREADS / 32];
    load_view_consts(vc, p.view, p.proj, p.campos);
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    const bool in_range = i < p.P;

    float x = 0.f, y = 0.f, z = 0.f, tz = 0.f;
    if (in_range) {
        x = __ldg(p.means3D + 3 * i); y = __ldg(p.means3D + 3 * i + 1); z = __ldg(p.means3D + 3 * i + 2);
        tz = cadd(cfma(vc.view[10], z, cfma(vc.view[6], y, cmul(vc.view[2], x))), vc.view[14]);
    }
    int radius = 0;
    uint32_t area = 0;
    float o_eff = 0.f;
    uint2 rc = make_uint2(0u, 0u);
    float ndcx = 0.f, ndcy = 0.f;
    const bool front = in_range && tz > NEAR_Z;
    if (front) {
        const float hx = cadd(cfma(vc.proj[8], z, cfma(vc.proj[4], y, cmul(vc.proj[0], x))), vc.proj[12]);
        const float hy = cadd(cfma(vc.proj[9], z, cfma(vc.proj[5], y, cmul(vc.proj[1], x))), vc.proj[13]);
        const float hw = cadd(cfma(vc.proj[11], z, cfma(vc.proj[7], y, cmul(vc.proj[3], x))), vc.proj[15]);
        const float pw = __frcp_rn(cadd(hw, 0.0000001f));
        ndcx = cmul(hx, pw); ndcy = cmul(hy, pw);
    }
    // ---- SH rows, first wave: the rows of Gaussians whose centre is on or near the screen start their way to
    // shared memory now (asynchronous copies) and land while the covariance work below runs.  The exact
    // visibility is only known after that work; the few visible Gaussians this guess misses (large splats
    // centred off screen) are fetched in a second wave, the few it over-fetches (2-3 %) are ignored. ----
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int M3 = 3 * p.M, nq = M3 >> 2, RS4 = sh_row_stride(M3) >> 2;
    float4* wrows = s_sh4 + (size_t)warp * 32 * RS4;
    const int i0 = blockIdx.x * PRE_THREADS + warp * 32;
    const float4* sh_src = reinterpret_cast<const float4*>(p.shs) + (size_t)i0 * nq;
    const int cq = lane & ((1 << qs) - 1), rstep = 32 >> qs;   // 2^qs >= nq lanes per row: no division
    unsigned emask = 0;
    if (coop) {
        emask = __ballot_sync(0xffffffffu, front && fabsf(ndcx) < 1.15f && fabsf(ndcy) < 1.15f);
        if (cq < nq)
            for (int row = lane >> qs; row < 32; row += rstep)
                if ((emask >> row) & 1u) cp_async16(wrows + row * RS4 + cq, sh_src + row * nq + cq);
        cp_async_commit();
    }
    if (front) {
        const float tx = cadd(cfma(vc.view[8], z, cfma(vc.view[4], y, cmul(vc.view[0], x))), vc.view[12]);
        const float ty = cadd(cfma(vc.view[9], z, cfma(vc.view[5], y, cmul(vc.view[1], x))), vc.view[13]);

        float c6[6];
        if (p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) c6[k] = __ldg(p.cov3D_precomp + 6 * (size_t)i + k);
        } else {
            const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + i);
            cov3d_from_scale_rot(__ldg(p.scales + 3 * i), __ldg(p.scales + 3 * i + 1), __ldg(p.scales + 3 * i + 2),
                                 p.mod, q, c6);
        }
        Proj2D pj;
        project_cov(vc, tx, ty, tz, c6, p.focal_x, p.focal_y, cmul(1.3f, p.tanfovx), cmul(1.3f, p.tanfovy), pj);
        const float a = cadd(pj.cov00, 0.3f), b = pj.cov01, c = cadd(pj.cov11, 0.3f);
        const float det = cfma(a, c, -cmul(b, b));
        if (det != 0.0f) {
            float hconv = 1.0f;
            if (p.antialiasing) {
                float det0 = cfma(pj.cov00, pj.cov11, -cmul(pj.cov01, pj.cov01));
                hconv = csqrt(fmaxf(0.000025f, cdiv(det0, det)));
            }
            const float det_inv = __frcp_rn(det);
            const float conA = cmul(c, det_inv), conB = cmul(-b, det_inv), conC = cmul(a, det_inv);
            const float mid = cmul(0.5f, cadd(a, c));
            const float sq = csqrt(fmaxf(0.1f, cfma(mid, mid, -det)));
            const float lam = fmaxf(cadd(mid, sq), csub(mid, sq));
            const int rad = (int)ceilf(cmul(3.0f, csqrt(lam)));
            // ndc -> pixel in double (exact), rounded once
            const float px = (float)__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn((double)ndcx, 1.0), (double)p.W), -1.0), 0.5);
            const float py = (float)__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn((double)ndcy, 1.0), (double)p.H), -1.0), 0.5);
            int minx, miny, maxx, maxy;
            get_rect(px, py, rad, p.gx, p.gy, minx, miny, maxx, maxy);
            const int ar = (maxx - minx) * (maxy - miny);
            if (ar != 0) {
                const float o = cmul(__ldg(p.opacities + i), hconv);
                // conservative early-out data for the render kernels (not part of the canonical result):
                //   thr: power below which o*exp(power) is certainly < 1/255
                //   ex, ey: half extents of the axis-aligned box of {power >= thr}
                float thr, ex, ey;
                if (o > 0.0f) {
                    thr = fmaxf(-80.0f, -logf(255.0f * o) - THR_MARGIN);     // >= -80: the argument range of canon_exp
                    const float tau2 = -2.0f * thr;
                    if (tau2 > 0.0f) {
                        // extents of {A dx^2 + 2 B dx dy + C dy^2 <= tau2} for the STORED conic (the one the render
                        // kernels evaluate), not for the covariance it was rounded from: max |dx| = sqrt(tau2 C / D),
                        // max |dy| = sqrt(tau2 A / D), D = A C - B^2 in double (fp32 products are exact in double, so
                        // a nearly degenerate conic loses nothing to cancellation).  D <= 0: unbounded, box = everything.
                        const double dd = (double)conA * (double)conC - (double)conB * (double)conB;
                        if (dd > 0.0) {
                            ex = (float)sqrt((double)tau2 * (double)conC / dd) * 1.001f + 0.01f;
                            ey = (float)sqrt((double)tau2 * (double)conA / dd) * 1.001f + 0.01f;
                        } else { ex = 3.0e38f; ey = 3.0e38f; }
                    } else { ex = -1.0f; ey = -1.0f; }
                } else { thr = __int_as_float(0x7f800000); ex = -1.0f; ey = -1.0f; }
                rc = make_uint2((uint32_t)minx | ((uint32_t)miny << 16), (uint32_t)(maxx - minx) | ((uint32_t)(maxy - miny) << 16));
                rec[3 * (size_t)i + 0] = make_float4(px, py, ex, ey);
                rec[3 * (size_t)i + 1] = make_float4(cmul(-0.5f, conA), conB, cmul(-0.5f, conC), thr);
                o_eff = o;
                radius = rad;
                area = (uint32_t)ar;
            }
        }
    }
    // ---- colour (after the cull, so the 12*M-byte SH row of a culled Gaussian is never touched) ----
    if (coop) {   // second wave (rows the guess missed), then wait for everything and publish to the warp
        const unsigned vmask = __ballot_sync(0xffffffffu, area != 0);
        const unsigned late = vmask & ~emask;
        if (late && cq < nq)
            for (int row = lane >> qs; row < 32; row += rstep)
                if ((late >> row) & 1u) cp_async16(wrows + row * RS4 + cq, sh_src + row * nq + cq);
        cp_async_commit();
        cp_async_wait_all();
        __syncwarp();
        if (area) {
            float rgb[3]; unsigned cl;
            eval_color_sh<false>(reinterpret_cast<const float*>(wrows + lane * RS4), p.M, p.D, csub(x, vc.campos[0]),
                                 csub(y, vc.campos[1]), csub(z, vc.campos[2]), rgb, cl);
            rec[3 * (size_t)i + 2] = make_float4(o_eff, rgb[0], rgb[1], rgb[2]);
        }
    } else if (area) {
        float rgb[3];
        if (p.colors_precomp) {
            rgb[0] = __ldg(p.colors_precomp + 3 * i); rgb[1] = __ldg(p.colors_precomp + 3 * i + 1);
            rgb[2] = __ldg(p.colors_precomp + 3 * i + 2);
        } else {
            unsigned cl;
            eval_color_sh<true>(p.shs + (size_t)i * p.M * 3, p.M, p.D, csub(x, vc.campos[0]), csub(y, vc.campos[1]),
                                csub(z, vc.campos[2]), rgb, cl);
        }
        rec[3 * (size_t)i + 2] = make_float4(o_eff, rgb[0], rgb[1], rgb[2]);
    }
    if (in_range) {
        radii[i] = radius;
        rect[i] = rc;
        depths[i] = area ? tz : __uint_as_float(0xffffffffu);   // as u32: the depth-sort key; 0xffffffff = culled
    }
    // Control block of the view (one atomic per quantity and block, each on its own 128-byte line): pair count R,
    // visible count V and the range of the depth keys, from which the depth sort derives its digit width.
    const uint32_t key = __float_as_uint(tz);
    uint32_t sum = __reduce_add_sync(0xffffffffu, area);
    uint32_t nvis = __popc(__ballot_sync(0xffffffffu, area != 0));
    uint32_t kmax = __reduce_max_sync(0xffffffffu, area ? key : 0u);
    uint32_t kminn = __reduce_max_sync(0xffffffffu, area ? ~key : 0u);
    if ((threadIdx.x & 31) == 0) {
        s_pairs[threadIdx.x >> 5] = sum; s_vis[threadIdx.x >> 5] = nvis;
        s_kmax[threadIdx.x >> 5] = kmax; s_kminn[threadIdx.x >> 5] = kminn;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        const bool w = threadIdx.x < PRE_THREADS / 32;
        sum = __reduce_add_sync(0xffffffffu, w ? s_pairs[threadIdx.x] : 0u);
        nvis = __reduce_add_sync(0xffffffffu, w ? s_vis[threadIdx.x] : 0u);
        kmax = __reduce_max_sync(0xffffffffu, w ? s_kmax[threadIdx.x] : 0u);
        kminn = __reduce_max_sync(0xffffffffu, w ? s_kminn[threadIdx.x] : 0u);
        if (nvis) {
            if (threadIdx.x == 0) atomicAdd(reinterpret_cast<unsigned long long*>(counters + CTR_PAIRS), (unsigned long long)sum);
            else if (threadIdx.x == 1) atomicAdd(counters + CTR_VISIBLE, nvis);
            else if (threadIdx.x == 2) atomicMax(counters + CTR_KMAX, kmax);
            else if (threadIdx.x == 3) atomicMax(counters + CTR_KMIN_NEG, kminn);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Per-Gaussian backward.  fp32 with plain operators (gradient tolerance is 1e-3 relative).
// ------------------------------------------------------------------------------------------------
struct BwdParams {
    PreParams p;
    const int32_t* radii;
    const float* acc;    // [P,12] packed render-backward sums
    int rezero;          // clean-scratch protocol: every accumulator row read here is written back as zeros
    int write_header;    // packed mode: packed_count heads a 64-byte header; words 4..6 of it receive the camera centre
    float* norm_accum;   // packed mode, nullable: xyz_gradient_accum[i] += ||dL/dmean2D[i]|| for the Gaussians visited here
    b200splat_grads out;
    const uint32_t* active_idx;     // Gaussians the render backward touched (compact_active_kernel)
    const unsigned* active_count;
    float4* packed_rows;            // packed mode (view-parallel exchange): one 64-byte row per touched Gaussian ...
    unsigned* packed_count;         // ... and their number, instead of the dense outputs
};

__device__ __forceinline__ void store3(float* base, size_t i, float a, float b, float c) {
    base[3 * i] = a; base[3 * i + 1] = b; base[3 * i + 2] = c;
}

// SH rows are staged per warp through shared memory (COOP): the 32 rows of a warp's Gaussians are
// contiguous in HBM, so the warp loads them with fully coalesced 16-byte accesses (12 independent loads
// in flight per lane at SH3), each lane then works on ITS row in shared memory (row stride padded to an
// odd number of 16-byte words: conflict-free), overwrites it in place with the dL/dsh row, and the warp
// streams the 32 result rows back with coalesced 16-byte stores (zeros for culled Gaussians).
// write_dsh = 0 (view-parallel exchange, see sh_grad_from_views_kernel): the dL/dsh rows are NOT written; the
// colour gradient, already multiplied by the SH clamp mask, goes to out.dL_dcolors instead.
__global__ void __launch_bounds__(PRE_THREADS, PRE_MINB_BWD) preprocess_bwd_kernel(BwdParams bp, int coop, int qs, int write_dsh) {
    extern __shared__ float4 s_sh4[];
    __shared__ ViewConsts vc;
    const PreParams& p = bp.p;
    if (blockIdx.x > 0 && blockIdx.x * PRE_THREADS >= (int)__ldg(bp.active_count)) return;   // grid sized for the worst case
    load_view_consts(vc, p.view, p.proj, p.campos);
    // One thread per TOUCHED Gaussian.  A Gaussian that is on screen but never blended (hidden behind opaque splats,
    // or below 1/255 everywhere) has an all-zero accumulator row and every gradient is linear in that row: the
    // caller zero-fills the outputs and this kernel visits only the rows the render backward wrote to (bench
    // scene: 3 M Gaussians, 1.97 M on screen, 0.22 M blended), dense in the warp instead of 2 live lanes of 32.
    const int t = blockIdx.x * PRE_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool in_range = t < (int)__ldg(bp.active_count);
    const int i = in_range ? (int)__ldg(bp.active_idx + t) : 0;
    const b200splat_grads& o = bp.out;
    const bool visible = in_range;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0;
    if (visible) {
        a0 = __ldg(reinterpret_cast<const float4*>(bp.acc) + 3 * (size_t)i);
        a1 = __ldg(reinterpret_cast<const float4*>(bp.acc) + 3 * (size_t)i + 1);
        a2 = __ldg(reinterpret_cast<const float4*>(bp.acc) + 3 * (size_t)i + 2);
    }
    const int M3 = 3 * p.M;
    // ---- COOP: coalesced load of the visible rows of this warp ----
    const int nq = M3 >> 2, RS4 = sh_row_stride(M3) >> 2;
    float4* wrows = s_sh4 + (size_t)warp * 32 * RS4;
    const unsigned vmask = __ballot_sync(0xffffffffu, visible);
    if (coop) {
        // the 32 rows of a warp are gathered (row r belongs to Gaussian i of lane r); each is one contiguous
        // 12 M-byte run, fetched by 2^qs lanes with 16-byte asynchronous copies that land while the geometry runs
        const float4* src = reinterpret_cast<const float4*>(p.shs);
        const int q = lane & ((1 << qs) - 1), rstep = 32 >> qs;
        for (int row = lane >> qs; row < 32; row += rstep) {
            const int ri = __shfl_sync(0xffffffffu, i, row);
            if (q < nq && ((vmask >> row) & 1u)) cp_async16(wrows + row * RS4 + q, src + (size_t)ri * nq + q);
        }
        cp_async_commit();
    }
    float* my_row = reinterpret_cast<float*>(wrows + lane * RS4);

    float dmean[3] = {0.f, 0.f, 0.f};
    float dsc[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f}, dc6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dop = 0.f, dcol[3] = {0.f, 0.f, 0.f}, dndc[2] = {0.f, 0.f};

    float x = 0.f, y = 0.f, z = 0.f;
    if (visible) {
        const float LA = a0.z, LB = a0.w, LC = a1.x;
        float dLo = a1.y;
        dcol[0] = a1.z; dcol[1] = a1.w; dcol[2] = a2.x;
        dndc[0] = a0.x * (0.5f * p.W); dndc[1] = a0.y * (0.5f * p.H);

        x = __ldg(p.means3D + 3 * i); y = __ldg(p.means3D + 3 * i + 1); z = __ldg(p.means3D + 3 * i + 2);
        const float tx = cadd(cfma(vc.view[8], z, cfma(vc.view[4], y, cmul(vc.view[0], x))), vc.view[12]);
        const float ty = cadd(cfma(vc.view[9], z, cfma(vc.view[5], y, cmul(vc.view[1], x))), vc.view[13]);
        const float tz = cadd(cfma(vc.view[10], z, cfma(vc.view[6], y, cmul(vc.view[2], x))), vc.view[14]);

        float c6[6];
        float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
        float s3[3] = {0.f, 0.f, 0.f};
        if (p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) c6[k] = __ldg(p.cov3D_precomp + 6 * (size_t)i + k);
        } else {
            q = __ldg(reinterpret_cast<const float4*>(p.rotations) + i);
            s3[0] = __ldg(p.scales + 3 * i); s3[1] = __ldg(p.scales + 3 * i + 1); s3[2] = __ldg(p.scales + 3 * i + 2);
            cov3d_from_scale_rot(s3[0], s3[1], s3[2], p.mod, q, c6);
        }
        // The chain conic -> cov2D -> (cov3D, mean) -> (scale, rotation) runs in DOUBLE from the fp32 inputs: for an
        // elongated splat a c ~ b^2, and the determinant, its square and the combinations of (LA, LB, LC) below lose
        // 3-4 digits to cancellation in fp32 -- enough to break the 1e-3 gradient gate on the largest splats of a
        // full-size view (measured: 1.06e-3 on dL/drotation).  Only the ~7 % of Gaussians a view touches come here,
        // so the fp64 work is a few microseconds.  The clamp decisions use the forward's fp32 values.
        const float limx = cmul(1.3f, p.tanfovx), limy = cmul(1.3f, p.tanfovy);
        const float txtz = cdiv(tx, tz), tytz = cdiv(ty, tz);
        const double xmul = (txtz < -limx || txtz > limx) ? 0.0 : 1.0;
        const double ymul = (tytz < -limy || tytz > limy) ? 0.0 : 1.0;
        const double dtzv = tz;
        const double txc = (double)cmul(fminf(limx, fmaxf(-limx, txtz)), tz), tyc = (double)cmul(fminf(limy, fmaxf(-limy, tytz)), tz);
        const double fx = p.focal_x, fy = p.focal_y;
        const double J00 = fx / dtzv, J02 = -(fx * txc) / (dtzv * dtzv), J11 = fy / dtzv, J12 = -(fy * tyc) / (dtzv * dtzv);
        double M0[3], M1[3], N0[3], N1[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            M0[k] = J00 * (double)vc.view[4 * k + 0] + J02 * (double)vc.view[4 * k + 2];
            M1[k] = J11 * (double)vc.view[4 * k + 1] + J12 * (double)vc.view[4 * k + 2];
        }
        {
            const double S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
#pragma unroll
            for (int k = 0; k < 3; k++) {
                N0[k] = M0[0] * S[0][k] + M0[1] * S[1][k] + M0[2] * S[2][k];
                N1[k] = M1[0] * S[0][k] + M1[1] * S[1][k] + M1[2] * S[2][k];
            }
        }
        const double cov00 = N0[0] * M0[0] + N0[1] * M0[1] + N0[2] * M0[2];
        const double cov01 = N0[0] * M1[0] + N0[1] * M1[1] + N0[2] * M1[2];
        const double cov11 = N1[0] * M1[0] + N1[1] * M1[1] + N1[2] * M1[2];
        const double a = cov00 + 0.3, b = cov01, c = cov11 + 0.3;
        const double det = a * c - b * b;
        const double d2inv = 1.0 / (det * det + 0.0000001);
        const double dLA = LA, dLB = LB, dLC = LC;
        // det - a c = -b^2
        double dL_da = d2inv * (-c * c * dLA + b * c * dLB - b * b * dLC);
        double dL_dc = d2inv * (-a * a * dLC + a * b * dLB - b * b * dLA);
        double dL_db = d2inv * (2.0 * b * c * dLA - (det + 2.0 * b * b) * dLB + 2.0 * a * b * dLC);
        if (p.antialiasing) {
            const double det0 = cov00 * cov11 - cov01 * cov01;
            const double ratio = det0 / det;
            const double h = sqrt(ratio > 0.000025 ? ratio : 0.000025);
            const double dL_dh = (double)__ldg(p.opacities + i) * (double)dLo;
            dLo = (float)(h * (double)dLo);
            if (ratio > 0.000025) {
                const double dL_dratio = dL_dh * 0.5 / h;
                const double inv2 = 1.0 / (det * det);
                dL_da += dL_dratio * (cov11 * det - det0 * c) * inv2;
                dL_dc += dL_dratio * (cov00 * det - det0 * a) * inv2;
                dL_db += dL_dratio * (-2.0 * cov01 * det + det0 * 2.0 * cov01) * inv2;
            }
        }
        dop = dLo;

        double dc6d[6];
        dc6d[0] = M0[0] * M0[0] * dL_da + M0[0] * M1[0] * dL_db + M1[0] * M1[0] * dL_dc;
        dc6d[3] = M0[1] * M0[1] * dL_da + M0[1] * M1[1] * dL_db + M1[1] * M1[1] * dL_dc;
        dc6d[5] = M0[2] * M0[2] * dL_da + M0[2] * M1[2] * dL_db + M1[2] * M1[2] * dL_dc;
        dc6d[1] = 2.0 * M0[0] * M0[1] * dL_da + (M0[0] * M1[1] + M0[1] * M1[0]) * dL_db + 2.0 * M1[0] * M1[1] * dL_dc;
        dc6d[2] = 2.0 * M0[0] * M0[2] * dL_da + (M0[0] * M1[2] + M0[2] * M1[0]) * dL_db + 2.0 * M1[0] * M1[2] * dL_dc;
        dc6d[4] = 2.0 * M0[1] * M0[2] * dL_da + (M0[1] * M1[2] + M0[2] * M1[1]) * dL_db + 2.0 * M1[1] * M1[2] * dL_dc;
#pragma unroll
        for (int k = 0; k < 6; k++) dc6[k] = (float)dc6d[k];

        // cov2D -> (J Wv) -> J -> view-space point -> mean
        double dJ00 = 0.0, dJ02 = 0.0, dJ11 = 0.0, dJ12 = 0.0;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const double dM0 = 2.0 * dL_da * N0[k] + dL_db * N1[k];
            const double dM1 = dL_db * N0[k] + 2.0 * dL_dc * N1[k];
            dJ00 += dM0 * (double)vc.view[4 * k + 0]; dJ02 += dM0 * (double)vc.view[4 * k + 2];
            dJ11 += dM1 * (double)vc.view[4 * k + 1]; dJ12 += dM1 * (double)vc.view[4 * k + 2];
        }
        const double itz = 1.0 / dtzv, itz2 = itz * itz, itz3 = itz2 * itz;
        const double dtx = xmul * (-fx * itz2) * dJ02;
        const double dty = ymul * (-fy * itz2) * dJ12;
        double dtz = -fx * itz2 * dJ00 - fy * itz2 * dJ11 + (2.0 * fx * txc) * itz3 * dJ02 + (2.0 * fy * tyc) * itz3 * dJ12;
        dtz += -(double)a2.y * itz2;   // inverse-depth channel (zero unless a depth gradient was supplied)
        double dmd[3];
        dmd[0] = (double)vc.view[0] * dtx + (double)vc.view[1] * dty + (double)vc.view[2] * dtz;
        dmd[1] = (double)vc.view[4] * dtx + (double)vc.view[5] * dty + (double)vc.view[6] * dtz;
        dmd[2] = (double)vc.view[8] * dtx + (double)vc.view[9] * dty + (double)vc.view[10] * dtz;

        // mean2D (NDC) -> mean3D through the full projection
        {
            const double dx = x, dy = y, dz = z;
            const double hx = vc.proj[0] * dx + vc.proj[4] * dy + vc.proj[8] * dz + vc.proj[12];
            const double hy = vc.proj[1] * dx + vc.proj[5] * dy + vc.proj[9] * dz + vc.proj[13];
            const double hw = vc.proj[3] * dx + vc.proj[7] * dy + vc.proj[11] * dz + vc.proj[15];
            const double mw = 1.0 / (hw + 0.0000001);
            const double mul1 = hx * mw * mw, mul2 = hy * mw * mw;
            const double gxn = dndc[0], gyn = dndc[1];
            dmd[0] += (vc.proj[0] * mw - vc.proj[3] * mul1) * gxn + (vc.proj[1] * mw - vc.proj[3] * mul2) * gyn;
            dmd[1] += (vc.proj[4] * mw - vc.proj[7] * mul1) * gxn + (vc.proj[5] * mw - vc.proj[7] * mul2) * gyn;
            dmd[2] += (vc.proj[8] * mw - vc.proj[11] * mul1) * gxn + (vc.proj[9] * mw - vc.proj[11] * mul2) * gyn;
        }
        dmean[0] = (float)dmd[0]; dmean[1] = (float)dmd[1]; dmean[2] = (float)dmd[2];

        // cov3D -> scale, rotation
        if (!p.cov3D_precomp) {
            const double r = q.x, qx = q.y, qy = q.z, qz = q.w;
            const double R[3][3] = {
                {1.0 - 2.0 * (qy * qy + qz * qz), 2.0 * (qx * qy - r * qz), 2.0 * (qx * qz + r * qy)},
                {2.0 * (qx * qy + r * qz), 1.0 - 2.0 * (qx * qx + qz * qz), 2.0 * (qy * qz - r * qx)},
                {2.0 * (qx * qz - r * qy), 2.0 * (qy * qz + r * qx), 1.0 - 2.0 * (qx * qx + qy * qy)}};
            const double sm = p.mod;
            const double sd[3] = {sm * s3[0], sm * s3[1], sm * s3[2]};
            const double G3[3][3] = {{dc6d[0], 0.5 * dc6d[1], 0.5 * dc6d[2]},
                                     {0.5 * dc6d[1], dc6d[3], 0.5 * dc6d[4]},
                                     {0.5 * dc6d[2], 0.5 * dc6d[4], dc6d[5]}};
            double g[3][3];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                double ds = 0.0;
#pragma unroll
                for (int a2i = 0; a2i < 3; a2i++) {
                    double vsum = 0.0;
#pragma unroll
                    for (int k = 0; k < 3; k++) vsum += G3[a2i][k] * (R[k][j] * sd[j]);
                    const double dLm = 2.0 * vsum;
                    ds += R[a2i][j] * dLm;
                    g[a2i][j] = dLm * sd[j];
                }
                dsc[j] = (float)(sm * ds);
            }
            drot[0] = (float)(2.0 * (-qz * g[0][1] + qy * g[0][2] + qz * g[1][0] - qx * g[1][2] - qy * g[2][0] + qx * g[2][1]));
            drot[1] = (float)(2.0 * (qy * g[0][1] + qz * g[0][2] + qy * g[1][0] - 2.0 * qx * g[1][1] - r * g[1][2] + qz * g[2][0] +
                                     r * g[2][1] - 2.0 * qx * g[2][2]));
            drot[2] = (float)(2.0 * (-2.0 * qy * g[0][0] + qx * g[0][1] + r * g[0][2] + qx * g[1][0] + qz * g[1][2] - r * g[2][0] +
                                     qz * g[2][1] - 2.0 * qy * g[2][2]));
            drot[3] = (float)(2.0 * (-2.0 * qz * g[0][0] - r * g[0][1] + qx * g[0][2] + r * g[1][0] - 2.0 * qz * g[1][1] +
                                     qy * g[1][2] + qx * g[2][0] + qy * g[2][1]));
        }
    }
    if (coop) {   // the rows have landed; make every lane's copies visible to the warp
        cp_async_wait_all();
        __syncwarp();
    }
    if (visible) {
        // SH: dL/dsh rows and dL/dmean through the normalised view direction
        if (p.shs) {
            const float ddx = x - vc.campos[0], ddy = y - vc.campos[1], ddz = z - vc.campos[2];
            const float len = csqrt(cfma(ddz, ddz, cfma(ddy, ddy, cmul(ddx, ddx))));
            const float ux = cdiv(ddx, len), uy = cdiv(ddy, len), uz = cdiv(ddz, len);
            // recompute the clamp mask with the forward's exact arithmetic
            float rgbf[3]; unsigned cl;
            const float* sh_row = p.shs + (size_t)i * M3;
            if (coop) eval_color_sh<false>(my_row, p.M, p.D, ddx, ddy, ddz, rgbf, cl);
            else eval_color_sh<true>(sh_row, p.M, p.D, ddx, ddy, ddz, rgbf, cl);
            const float g0 = (cl & 1u) ? 0.f : dcol[0], g1 = (cl & 2u) ? 0.f : dcol[1], g2 = (cl & 4u) ? 0.f : dcol[2];
            float bb[16], bx[16], by[16], bz[16];
#pragma unroll
            for (int k = 0; k < 16; k++) { bb[k] = 0.f; bx[k] = 0.f; by[k] = 0.f; bz[k] = 0.f; }
            bb[0] = 0.28209479177387814f;
            if (p.D > 0) {
                const float C1 = 0.4886025119029199f;
                bb[1] = -C1 * uy; by[1] = -C1;
                bb[2] = C1 * uz;  bz[2] = C1;
                bb[3] = -C1 * ux; bx[3] = -C1;
                if (p.D > 1) {
                    const float c20 = 1.0925484305920792f, c22 = 0.31539156525252005f, c24 = 0.5462742152960396f;
                    const float xx = ux * ux, yy = uy * uy, zz = uz * uz;
                    bb[4] = c20 * ux * uy;  bx[4] = c20 * uy; by[4] = c20 * ux;
                    bb[5] = -c20 * uy * uz; by[5] = -c20 * uz; bz[5] = -c20 * uy;
                    bb[6] = c22 * (2.f * zz - xx - yy); bx[6] = -2.f * c22 * ux; by[6] = -2.f * c22 * uy; bz[6] = 4.f * c22 * uz;
                    bb[7] = -c20 * ux * uz; bx[7] = -c20 * uz; bz[7] = -c20 * ux;
                    bb[8] = c24 * (xx - yy); bx[8] = 2.f * c24 * ux; by[8] = -2.f * c24 * uy;
                    if (p.D > 2) {
                        const float c30 = -0.5900435899266435f, c31 = 2.890611442640554f, c32 = -0.4570457994644658f,
                                    c33 = 0.3731763325901154f, c35 = 1.445305721320277f;
                        bb[9] = c30 * uy * (3.f * xx - yy); bx[9] = c30 * uy * 6.f * ux; by[9] = c30 * (3.f * xx - 3.f * yy);
                        bb[10] = c31 * ux * uy * uz; bx[10] = c31 * uy * uz; by[10] = c31 * ux * uz; bz[10] = c31 * ux * uy;
                        bb[11] = c32 * uy * (4.f * zz - xx - yy); bx[11] = c32 * uy * -2.f * ux;
                        by[11] = c32 * (4.f * zz - xx - 3.f * yy); bz[11] = c32 * uy * 8.f * uz;
                        bb[12] = c33 * uz * (2.f * zz - 3.f * xx - 3.f * yy); bx[12] = c33 * uz * -6.f * ux;
                        by[12] = c33 * uz * -6.f * uy; bz[12] = c33 * (6.f * zz - 3.f * xx - 3.f * yy);
                        bb[13] = c32 * ux * (4.f * zz - xx - yy); bx[13] = c32 * (4.f * zz - 3.f * xx - yy);
                        by[13] = c32 * ux * -2.f * uy; bz[13] = c32 * ux * 8.f * uz;
                        bb[14] = c35 * uz * (xx - yy); bx[14] = c35 * uz * 2.f * ux; by[14] = c35 * uz * -2.f * uy;
                        bz[14] = c35 * (xx - yy);
                        bb[15] = c30 * ux * (xx - 3.f * yy); bx[15] = c30 * (3.f * xx - 3.f * yy); by[15] = c30 * ux * -6.f * uy;
                    }
                }
            }
            const int nb = (p.D + 1) * (p.D + 1);
            float dRx = 0.f, dRy = 0.f, dRz = 0.f;
            float* dsh_row = o.dL_dsh + (size_t)i * M3;
            if (coop) {
                float4* w = reinterpret_cast<float4*>(my_row);      // in place: row in, dL/dsh row out
                const int nvec = M3 >> 2;
#pragma unroll
                for (int kv = 0; kv < 12; kv++) {
                    if (kv < nvec) {
                        const float4 t = w[kv];
                        const float f[4] = {t.x, t.y, t.z, t.w};
                        float out[4];
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const int idx = 4 * kv + e, k = idx / 3, ch = idx - 3 * k;
                            const float gch = ch == 0 ? g0 : (ch == 1 ? g1 : g2);
                            const bool act = k < nb;
                            out[e] = act ? bb[k] * gch : 0.f;
                            if (act) { const float sg = f[e] * gch; dRx += bx[k] * sg; dRy += by[k] * sg; dRz += bz[k] * sg; }
                        }
                        if (write_dsh) w[kv] = make_float4(out[0], out[1], out[2], out[3]);
                    }
                }
            } else {
                // static indices only: a dynamic index would push the basis arrays into local memory
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    if (k < p.M) {
                        const bool act = k < nb;
                        const float bk = act ? bb[k] : 0.f;
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            const float gch = ch == 0 ? g0 : (ch == 1 ? g1 : g2);
                            if (write_dsh) dsh_row[3 * k + ch] = bk * gch;
                            if (act) {
                                const float sg = __ldg(sh_row + 3 * k + ch) * gch;
                                dRx += bx[k] * sg; dRy += by[k] * sg; dRz += bz[k] * sg;
                            }
                        }
                    }
                }
                if (write_dsh) for (int k = 16 * 3; k < M3; k++) dsh_row[k] = 0.f;
            }
            if (!write_dsh) { dcol[0] = g0; dcol[1] = g1; dcol[2] = g2; }
            const float dotp = ux * dRx + uy * dRy + uz * dRz;
            const float il = 1.0f / len;
            dmean[0] += (dRx - ux * dotp) * il;
            dmean[1] += (dRy - uy * dotp) * il;
            dmean[2] += (dRz - uz * dotp) * il;
        }
    }

    if (coop && write_dsh) {   // the warp streams its 32 dL/dsh rows out, 16 bytes per lane, each row one contiguous run
        __syncwarp();
        float4* dst = reinterpret_cast<float4*>(o.dL_dsh);
        const int q = lane & ((1 << qs) - 1), rstep = 32 >> qs;
        for (int row = lane >> qs; row < 32; row += rstep) {
            const int ri = __shfl_sync(0xffffffffu, i, row);
            if (q < nq && ((vmask >> row) & 1u)) dst[(size_t)ri * nq + q] = wrows[row * RS4 + q];
        }
    }
    if (bp.rezero && in_range) {
        // clean-scratch protocol: the rows this kernel visits are exactly the rows the render backward wrote to.  Down
        // here, not next to the loads: a store behind a read-only load of the same address makes the thread wait for
        // the load, and the kernel lives on having all its loads in flight together (measured: 62 -> 131 us up there)
        float4* row = reinterpret_cast<float4*>(const_cast<float*>(bp.acc)) + 3 * (size_t)i;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        row[0] = z; row[1] = z; row[2] = z;
    }
    if (bp.packed_rows) {
        // row t of the packed buffer (b200splat.h "packed gradient rows"): everything the other ranks need from
        // this view for Gaussian i, 64 bytes, written by one lane as four 16-byte stores
        if (t == 0) {
            *bp.packed_count = __ldg(bp.active_count);
            if (bp.write_header) {      // how the other ranks learn this view's camera (accumulate_packed, campos = NULL)
                float* hdr = reinterpret_cast<float*>(bp.packed_count);
                hdr[4] = __ldg(p.campos); hdr[5] = __ldg(p.campos + 1); hdr[6] = __ldg(p.campos + 2);
            }
        }
        if (in_range) {
            float4* row = bp.packed_rows + 4 * (size_t)t;
            row[0] = make_float4(__uint_as_float((unsigned)i), dcol[0], dcol[1], dcol[2]);
            row[1] = make_float4(drot[0], drot[1], drot[2], drot[3]);
            row[2] = make_float4(dmean[0], dmean[1], dmean[2], dsc[0]);
            row[3] = make_float4(dsc[1], dsc[2], dop, 0.f);
            if (o.dL_dmeans2D) store3(o.dL_dmeans2D, i, dndc[0], dndc[1], 0.f);
            // the gradient-norm part of the densification statistics (densify_stats_kernel's first line): a Gaussian this
            // kernel does not visit has a zero gradient and would add nothing
            if (bp.norm_accum) bp.norm_accum[i] += sqrtf(dndc[0] * dndc[0] + dndc[1] * dndc[1]);
        }
        return;
    }
    if (!in_range) return;
    store3(o.dL_dmeans3D, i, dmean[0], dmean[1], dmean[2]);
    if (o.dL_dmeans2D) store3(o.dL_dmeans2D, i, dndc[0], dndc[1], 0.f);
    if (o.dL_dcolors) store3(o.dL_dcolors, i, dcol[0], dcol[1], dcol[2]);
    if (o.dL_dopacity) o.dL_dopacity[i] = dop;
    if (o.dL_dscales) store3(o.dL_dscales, i, dsc[0], dsc[1], dsc[2]);
    if (o.dL_drotations) reinterpret_cast<float4*>(o.dL_drotations)[i] = make_float4(drot[0], drot[1], drot[2], drot[3]);
    if (o.dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) o.dL_dcov3D[6 * (size_t)i + k] = dc6[k];
    }
}

// Indices of the Gaussians the render backward touched, in block-arrival order (the consumers write by index, so
// the order does not matter): one atomic per 1024-thread block.
__global__ void __launch_bounds__(1024) compact_active_kernel(int P, uint8_t* __restrict__ touched, int clear_flags,
                                                              uint32_t* __restrict__ active_idx, unsigned* __restrict__ count) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_base;
    const int i = blockIdx.x * 1024 + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool on = i < P && touched[i] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, on);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    if (warp == 0) {
        const unsigned c = s_warp[lane];
        unsigned incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned n = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += n;
        }
        s_warp[lane] = incl - c;
        if (lane == 31) s_base = incl ? atomicAdd(count, incl) : 0u;
    }
    __syncthreads();
    if (on) {
        active_idx[s_base + s_warp[warp] + __popc(m & ((1u << lane) - 1u))] = (uint32_t)i;
        if (clear_flags) touched[i] = 0;      // clean-scratch protocol: the flags are all zero again after this kernel
    }
}

// Colour gradient of every Gaussian times the SH clamp mask, straight from the render backward's sums: the
// quantity the view-parallel exchange ships.  Runs BEFORE the per-Gaussian backward so that the exchange
// overlaps it.
__global__ void __launch_bounds__(256) masked_colour_grad_kernel(int P, const int32_t* __restrict__ radii,
                                                                 const float4* __restrict__ rec, const float4* __restrict__ acc,
                                                                 float* __restrict__ dcol) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float d0 = 0.f, d1 = 0.f, d2 = 0.f;
    if (radii[i] > 0) {
        const float4 c = __ldg(rec + 3 * (size_t)i + 2);
        const float4 a1 = __ldg(acc + 3 * (size_t)i + 1), a2 = __ldg(acc + 3 * (size_t)i + 2);
        d0 = signbit(c.y) ? 0.f : a1.z;
        d1 = signbit(c.z) ? 0.f : a1.w;
        d2 = signbit(c.w) ? 0.f : a2.x;
    }
    dcol[3 * (size_t)i] = d0; dcol[3 * (size_t)i + 1] = d1; dcol[3 * (size_t)i + 2] = d2;
}

// ------------------------------------------------------------------------------------------------
// View-parallel exchange (DESIGN.md "Multi-GPU"): dL/dsh of one view is the outer product
//   dL/dsh[i][k][c] = Y_k(dir(mean_i - campos_v)) * dcol_v[i][c]        (dcol already clamp-masked)
// and every rank holds the means and can be told every view's camera centre, so the ranks exchange the
// 12-byte colour gradients (all-gather, or peer loads) instead of all-reducing the 12*M-byte rows, and
// each rank rebuilds sum_v dL/dsh_v locally.  dcol[v] may point into a peer GPU's memory (NVLink loads).
// ------------------------------------------------------------------------------------------------
struct ViewGrads {
    int n;
    const float* dcol[B200SPLAT_MAX_VIEWS];   // [P,3] each
};

__global__ void __launch_bounds__(PRE_THREADS) sh_grad_from_views_kernel(int P, int M, int D, const float* __restrict__ means3D,
                                                                         ViewGrads vg, const float* __restrict__ campos,
                                                                         float* __restrict__ dL_dsh, int coop, int qs, int dc_off4, int vec_ok) {
    extern __shared__ float4 s_sh4[];
    __shared__ float s_cam[B200SPLAT_MAX_VIEWS * 3];
    if (threadIdx.x < vg.n * 3) s_cam[threadIdx.x] = campos[threadIdx.x];
    __syncthreads();
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool in_range = i < P;
    const int M3 = 3 * M, nb = (D + 1) * (D + 1);
    const int i0 = blockIdx.x * PRE_THREADS + warp * 32;
    // Colour gradients of the warp's 32 Gaussians, all views: 384 contiguous bytes per view, fetched as 24 16-byte
    // loads per view, all views in flight together (what matters when dcol[v] is a PEER pointer: one NVLink round
    // trip per warp instead of 3 x n dependent 4-byte loads per lane), then handed to the lanes through shared memory.
    float* s_dc = reinterpret_cast<float*>(s_sh4 + dc_off4) + (size_t)warp * vg.n * 96;
    const bool full = vec_ok && i0 + 32 <= P;
    if (full) {
        if (lane < 24) {
#pragma unroll 4
            for (int v = 0; v < vg.n; v++)
                reinterpret_cast<float4*>(s_dc + v * 96)[lane] = reinterpret_cast<const float4*>(vg.dcol[v] + 3 * (size_t)i0)[lane];
        }
    } else if (in_range) {
        for (int v = 0; v < vg.n; v++)
            for (int c = 0; c < 3; c++) s_dc[v * 96 + 3 * lane + c] = vg.dcol[v][3 * (size_t)i + c];
    }
    __syncwarp();
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; k++) acc[k] = 0.f;
    if (in_range) {
        const float x = __ldg(means3D + 3 * i), y = __ldg(means3D + 3 * i + 1), z = __ldg(means3D + 3 * i + 2);
        for (int v = 0; v < vg.n; v++) {
            const float g0 = s_dc[v * 96 + 3 * lane], g1 = s_dc[v * 96 + 3 * lane + 1], g2 = s_dc[v * 96 + 3 * lane + 2];
            if (g0 == 0.f && g1 == 0.f && g2 == 0.f) continue;     // culled or clamped in that view
            const float ddx = csub(x, s_cam[3 * v]), ddy = csub(y, s_cam[3 * v + 1]), ddz = csub(z, s_cam[3 * v + 2]);
            const float len = csqrt(cfma(ddz, ddz, cfma(ddy, ddy, cmul(ddx, ddx))));
            float b[16];
            sh_basis(D, cdiv(ddx, len), cdiv(ddy, len), cdiv(ddz, len), b);
#pragma unroll
            for (int k = 0; k < 16; k++) {
                if (k < nb) {
                    acc[3 * k] = fmaf(b[k], g0, acc[3 * k]);
                    acc[3 * k + 1] = fmaf(b[k], g1, acc[3 * k + 1]);
                    acc[3 * k + 2] = fmaf(b[k], g2, acc[3 * k + 2]);
                }
            }
        }
    }
    if (coop) {
        const int nq = M3 >> 2, RS4 = sh_row_stride(M3) >> 2;
        float4* wrows = s_sh4 + (size_t)warp * 32 * RS4;
        float4* w = wrows + lane * RS4;
#pragma unroll
        for (int kv = 0; kv < 12; kv++)
            if (kv < nq) w[kv] = make_float4(acc[4 * kv], acc[4 * kv + 1], acc[4 * kv + 2], acc[4 * kv + 3]);
        __syncwarp();
        float4* dst = reinterpret_cast<float4*>(dL_dsh) + (size_t)i0 * nq;
        const int rows = min(32, P - i0);
        const int q = lane & ((1 << qs) - 1), rstep = 32 >> qs;
        if (q < nq)
            for (int row = lane >> qs; row < rows; row += rstep) dst[row * nq + q] = wrows[row * RS4 + q];
    } else if (in_range) {
        float* row = dL_dsh + (size_t)i * M3;
#pragma unroll
        for (int k = 0; k < 48; k++)
            if (k < M3) row[k] = acc[k];
        for (int k = 48; k < M3; k++) row[k] = 0.f;
    }
}

// ------------------------------------------------------------------------------------------------
// Sparse view-parallel exchange.  A view touches a small fraction of the Gaussians (bench scene: 7 %), so each
// rank publishes ONE 64-byte row per Gaussian its view touched (preprocess_bwd_kernel, packed mode) in memory
// its peers can read, and every rank adds all rows of all views into its dense gradients: the row carries the
// clamp-masked colour gradient, from which dL/dsh = Y(dir(mean_i - campos_v)) (x) dcol is rebuilt on the spot.
// One warp per row; the 16 words of the row are one 64-byte load (over NVLink when the row is a peer's), the
// 48 + 11 sums go out as red.global.add.f32, consecutive lanes to consecutive addresses.
//   row: [0] index  [1..3] dcol  [4..7] d rotation  [8..10] d mean  [11..13] d scale  [14] d opacity  [15] -
// ------------------------------------------------------------------------------------------------
struct PackedSources {
    int n;
    const float* rows[B200SPLAT_MAX_VIEWS];
    const unsigned* count[B200SPLAT_MAX_VIEWS];
};

// The row counts live next to the rows, possibly in a peer's memory: one warp fetches them once (n_src remote
// loads) instead of every warp of the accumulate kernel hammering the same few remote words.
// With fetch_cam, the camera centre of every source view comes the same way: words 4..6 of the 64-byte header the
// row count sits in (the caller put it there), so no collective is needed to tell the ranks each other's cameras.
__global__ void gather_counts_kernel(PackedSources src, unsigned* __restrict__ local_counts, int fetch_cam) {
    const int v = threadIdx.x >> 2, k = threadIdx.x & 3;         // 4 threads per source: count, cx, cy, cz
    if (v >= src.n) return;
    const volatile unsigned* hdr = reinterpret_cast<const volatile unsigned*>(src.count[v]);
    if (k == 0) local_counts[v] = hdr[0];
    else if (fetch_cam) local_counts[32 + 3 * v + (k - 1)] = hdr[3 + k];
}

__device__ unsigned g_packed_counts[256]; // two slots of: 32 row counts + 16 x 3 camera-centre words of the sources (gather_counts_kernel)

constexpr int ACC_WARPS = 4;              // warps per block of accumulate_packed_kernel
constexpr int ACC_RS4 = 13;               // 16-byte words per staged dL/dsh row (odd: conflict-free)

// STORE: the targets are known to hold zero and nobody else writes them while this launch runs (one source, whose
// rows have distinct indices): plain 16-byte stores instead of reductions.
template <bool STORE>
__global__ void __launch_bounds__(ACC_WARPS * 32) accumulate_packed_kernel(int P, int M, int D, const float* __restrict__ means3D,
                                                                           PackedSources src, const unsigned* __restrict__ local_counts,
                                                                           const float* __restrict__ campos,
                                                                           float* __restrict__ dL_dsh, float* __restrict__ dL_dgeom,
                                                                           uint8_t* __restrict__ marks) {
    // A warp takes 32 consecutive rows at a time.  Phase A, one row per lane: the 64-byte row (2 KB contiguous per
    // warp, all loads in flight together -- one NVLink round trip per 32 rows when the rows are a peer's), the mean,
    // the basis and the 3 M products of ITS row, staged in shared memory.  Phase B, one row per iteration, ONE
    // instruction: lanes 0..11 add the dL/dsh row, lanes 12..14 the 48-byte geometry row, as 16-byte reductions
    // (red.global.add.v4.f32).  The L2 retires a bounded number of atomic sectors per second (measured ~60 G/s), so
    // the targets are laid out to touch as few sectors as possible: 6 + 2 per row.
    __shared__ float4 s_row4[ACC_WARPS][32 * ACC_RS4];
    __shared__ float4 s_g4[ACC_WARPS][32 * 3];
    __shared__ unsigned s_idx[ACC_WARPS][32];            // bit 31: the row has a colour gradient
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    const int M3 = 3 * M, nb = (D + 1) * (D + 1), nq = M3 >> 2;
    const bool vec = (M3 & 3) == 0 && nq <= 12 && (reinterpret_cast<size_t>(dL_dsh) & 15) == 0;
    unsigned my_cnt = 0;                  // every source's row count (a local copy: gather_counts_kernel)
    if (lane < src.n) my_cnt = __ldg(local_counts + lane);
    // Batch-major, source-minor: the touched Gaussians of every view come in (roughly) increasing index order, so
    // batch k of all sources lands in the same neighbourhood of the dense arrays -- a line filled for one view's
    // row is still in L2 when the other views' rows for the same Gaussian arrive.
    unsigned max_cnt = min(my_cnt, (unsigned)P);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) max_cnt = max(max_cnt, __shfl_xor_sync(0xffffffffu, max_cnt, d));
    // The (batch, source) items of this warp in order; the 64-byte rows of the NEXT item are fetched before the
    // reductions of the current one are issued, so that their latency -- an NVLink round trip when the rows are a
    // peer's -- is hidden behind them (the kernel is latency-bound: 16-24 warps per SM).
    const unsigned stride = (unsigned)n_warps * 32u;
    auto advance = [&](unsigned& e0, int& v) -> bool {       // next item with rows in it; false: none left
        for (;;) {
            if (++v >= src.n) { v = 0; e0 += stride; }
            if (e0 >= max_cnt) return false;
            if (e0 < min(__shfl_sync(0xffffffffu, my_cnt, v), (unsigned)P)) return true;
        }
    };
    // The 32 rows of an item are one contiguous 2 KB run: fetched as four fully coalesced 512-byte accesses (chunk t of
    // the run = 16-byte words 32 t .. 32 t + 31), not as one 64-byte row per lane -- over NVLink every 32-byte sector
    // would otherwise be requested twice.  The rows are transposed through shared memory when the item is consumed.
    auto fetch = [&](unsigned e0, int v, float4& c0, float4& c1, float4& c2, float4& c3) {
        const unsigned cnt = min(__shfl_sync(0xffffffffu, my_cnt, v), (unsigned)P);
        const unsigned n16 = 4u * min(32u, cnt - e0);                 // 16-byte words in the run
        const float4* run = reinterpret_cast<const float4*>(src.rows[v]) + 4 * (size_t)e0;
        if ((unsigned)lane < n16) c0 = run[lane];
        if ((unsigned)lane + 32u < n16) c1 = run[lane + 32];
        if ((unsigned)lane + 64u < n16) c2 = run[lane + 64];
        if ((unsigned)lane + 96u < n16) c3 = run[lane + 96];
    };
    unsigned e0 = (unsigned)warp * 32u - stride;             // advance() steps onto the first item
    int v = src.n - 1;
    bool have = advance(e0, v);
    float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0, q2 = q0, q3 = q0;
    if (have) fetch(e0, v, q0, q1, q2, q3);
    while (have) {
        {
            const unsigned cnt = min(__shfl_sync(0xffffffffu, my_cnt, v), (unsigned)P);
            const float cx = __ldg(campos + 3 * v), cy = __ldg(campos + 3 * v + 1), cz = __ldg(campos + 3 * v + 2);
            const unsigned e = e0 + lane;
            const int nrow = (int)min(32u, cnt - e0);
            __syncwarp();
            {   // transpose: chunks -> the staging area (free until the products are written) -> this lane's row
                float4* raw = s_row4[wib];
                raw[lane] = q0; raw[lane + 32] = q1; raw[lane + 64] = q2; raw[lane + 96] = q3;
                __syncwarp();
                q0 = raw[4 * lane]; q1 = raw[4 * lane + 1]; q2 = raw[4 * lane + 2]; q3 = raw[4 * lane + 3];
                __syncwarp();
            }
            if (e < cnt) {
                unsigned idx = __float_as_uint(q0.x);
                const bool ok = idx < (unsigned)P;       // always true for a well-formed row
                if (!ok) idx = 0;
                const bool has_col = !(q0.y == 0.f && q0.z == 0.f && q0.w == 0.f);   // false: all channels clamped
                s_idx[wib][lane] = ok ? (idx | (has_col ? 0x80000000u : 0u)) : 0xffffffffu;
                if (marks && ok) marks[idx] = 1;          // what launch_clear_accumulated goes by
                s_g4[wib][lane * 3] = q1; s_g4[wib][lane * 3 + 1] = q2; s_g4[wib][lane * 3 + 2] = q3;
                if (has_col) {
                    const float x = __ldg(means3D + 3 * (size_t)idx), y = __ldg(means3D + 3 * (size_t)idx + 1),
                                z = __ldg(means3D + 3 * (size_t)idx + 2);
                    const float ddx = csub(x, cx), ddy = csub(y, cy), ddz = csub(z, cz);
                    const float len = csqrt(cfma(ddz, ddz, cfma(ddy, ddy, cmul(ddx, ddx))));
                    float b[16];
                    sh_basis(D, cdiv(ddx, len), cdiv(ddy, len), cdiv(ddz, len), b);
                    float pr[48];
#pragma unroll
                    for (int k = 0; k < 16; k++) {
                        const float bk = k < nb ? b[k] : 0.f;
                        pr[3 * k] = bk * q0.y; pr[3 * k + 1] = bk * q0.z; pr[3 * k + 2] = bk * q0.w;
                    }
                    float4* dstrow = s_row4[wib] + lane * ACC_RS4;
#pragma unroll
                    for (int q = 0; q < 12; q++)
                        if (4 * q < M3) dstrow[q] = make_float4(pr[4 * q], pr[4 * q + 1], pr[4 * q + 2], pr[4 * q + 3]);
                }
            }
            // next item's rows on their way while this item's reductions are issued
            unsigned ne0 = e0;
            int nv = v;
            const bool nhave = advance(ne0, nv);
            float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0, n3 = n0;
            if (nhave) fetch(ne0, nv, n0, n1, n2, n3);
            __syncwarp();
            for (int r = 0; r < nrow; r++) {
                const unsigned tag = s_idx[wib][r];
                if (tag == 0xffffffffu) continue;
                const unsigned idx = tag & 0x7fffffffu;
                const bool has_col = (tag >> 31) != 0u;
                if (vec) {
                    const bool sh_lane = lane < nq;
                    if ((sh_lane && has_col) || (lane >= 12 && lane < 15)) {
                        float4* dst = sh_lane ? reinterpret_cast<float4*>(dL_dsh + (size_t)idx * M3) + lane
                                              : reinterpret_cast<float4*>(dL_dgeom) + 3 * (size_t)idx + (lane - 12);
                        const float4 val = sh_lane ? s_row4[wib][r * ACC_RS4 + lane] : s_g4[wib][r * 3 + (lane - 12)];
                        if (STORE) *dst = val; else atomicAdd(dst, val);
                    }
                } else {
                    if (lane >= 12 && lane < 15)
                        atomicAdd(reinterpret_cast<float4*>(dL_dgeom) + 3 * (size_t)idx + (lane - 12), s_g4[wib][r * 3 + (lane - 12)]);
                    if (!has_col) continue;
                    float* drow = dL_dsh + (size_t)idx * M3;
                    const float* srow = reinterpret_cast<const float*>(s_row4[wib] + r * ACC_RS4);
                    for (int el = lane; el < M3 && el < 48; el += 32) atomicAdd(drow + el, srow[el]);
                }
            }
            q0 = n0; q1 = n1; q2 = n2; q3 = n3;
            e0 = ne0; v = nv; have = nhave;
        }
    }
}

// Sparse clear of the dense gradients accumulate_packed_kernel filled: it set marks[i] for every Gaussian it added a row
// into.  One lane per Gaussian reads its mark (3 MB for 3 M Gaussians); the warp then zeroes each marked Gaussian's dL/dsh
// row (lanes 0..nq-1, 16 bytes each), its 48-byte geometry row (lanes 12..14) and the mark -- the mirror image of the
// accumulate kernel's phase B.
__global__ void __launch_bounds__(256) clear_accumulated_kernel(int P, int M3, float* __restrict__ dL_dsh, float* __restrict__ dL_dgeom,
                                                                uint8_t* __restrict__ marks) {
    const int i = blockIdx.x * 256 + threadIdx.x, lane = threadIdx.x & 31;
    const bool dirty = i < P && marks[i] != 0;
    unsigned m = __ballot_sync(0xffffffffu, dirty);
    if (dirty) marks[i] = 0;
    const int i0 = i - lane, nq = M3 >> 2;
    const bool vec = (M3 & 3) == 0 && nq <= 12 && (reinterpret_cast<size_t>(dL_dsh) & 15) == 0;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    while (m) {
        const int r = __ffs(m) - 1;
        m &= m - 1;
        const size_t idx = (size_t)(i0 + r);
        if (lane >= 12 && lane < 15) reinterpret_cast<float4*>(dL_dgeom)[3 * idx + (lane - 12)] = z;
        if (vec) {
            if (lane < nq) reinterpret_cast<float4*>(dL_dsh + idx * M3)[lane] = z;
        } else {
            for (int el = lane; el < M3; el += 32) dL_dsh[idx * M3 + el] = 0.0f;
        }
    }
}

// Densification statistics of one view in ONE pass (the reference does this with five torch ops,
// /root/reference/scene/gaussian_model.py:576-581 and /root/reference/train.py:127):
//   accum[i] += ||dL_dmeans2D[i, :2]||  and  denom[i] += 1   for visible Gaussians,  max_radii[i] = max(., radii[i])
__global__ void densify_stats_kernel(int P, const float* __restrict__ dmeans2D, const int32_t* __restrict__ radii,
                                     float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r > 0) {
        const float gx = dmeans2D[3 * (size_t)i], gy = dmeans2D[3 * (size_t)i + 1];
        accum[i] += sqrtf(gx * gx + gy * gy);
        denom[i] += 1.0f;
        max_radii[i] = fmaxf(max_radii[i], (float)r);
    }
}

// The part of densify_stats_kernel that needs the radii only (known after the forward): denom and max_radii2D.
__global__ void densify_stats_visible_kernel(int P, const int32_t* __restrict__ radii, float* __restrict__ denom,
                                             float* __restrict__ max_radii) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r > 0) {
        denom[i] += 1.0f;
        max_radii[i] = fmaxf(max_radii[i], (float)r);
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float x = means3D[3 * i], y = means3D[3 * i + 1], z = means3D[3 * i + 2];
    const float tz = cadd(cfma(view[10], z, cfma(view[6], y, cmul(view[2], x))), view[14]);
    present[i] = tz > NEAR_Z ? 1 : 0;
}

__global__ void export_geom_kernel(int P, const int32_t* __restrict__ radii, GeomView gv, b200splat_state_out o) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const bool vis = radii[i] > 0;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
    float d = 0.f;
    if (vis) { a = gv.rec[3 * (size_t)i]; b = gv.rec[3 * (size_t)i + 1]; c = gv.rec[3 * (size_t)i + 2]; d = gv.depths[i]; }
    if (o.means2D) { o.means2D[2 * i] = a.x; o.means2D[2 * i + 1] = a.y; }
    if (o.depths) o.depths[i] = d;
    if (o.conic_opacity) {
        o.conic_opacity[4 * i] = -2.0f * b.x; o.conic_opacity[4 * i + 1] = b.y;
        o.conic_opacity[4 * i + 2] = -2.0f * b.z; o.conic_opacity[4 * i + 3] = c.x;
    }
    if (o.rgb) { o.rgb[3 * i] = c.y; o.rgb[3 * i + 1] = c.z; o.rgb[3 * i + 2] = c.w; }
}

// Function attributes are set once per kernel and device (not on every launch: attribute calls have no place inside a
// stream capture, and a captured step replays these launches thousands of times).
template <typename K>
cudaError_t configure_once(K kernel, bool (&done)[64], bool max_carveout) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (done[dev & 63]) return cudaSuccess;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024)) != cudaSuccess) return e;
    if (max_carveout) cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    done[dev & 63] = true;
    return cudaSuccess;
}

int coop_shift(int nq) {   // smallest qs with 2^qs >= nq (nq <= 12)
    int qs = 0;
    while ((1 << qs) < nq) qs++;
    return qs;
}

PreParams make_params(const b200splat_view& v, const b200splat_gaussians& g) {
    PreParams p;
    p.P = g.P; p.D = v.sh_degree; p.M = g.M; p.W = v.image_width; p.H = v.image_height;
    p.gx = (p.W + TILE - 1) / TILE; p.gy = (p.H + TILE - 1) / TILE;
    p.tanfovx = v.tanfovx; p.tanfovy = v.tanfovy;
    p.focal_x = (float)p.W / (2.0f * v.tanfovx);
    p.focal_y = (float)p.H / (2.0f * v.tanfovy);
    p.mod = v.scale_modifier; p.antialiasing = v.antialiasing;
    p.means3D = g.means3D; p.scales = g.scales; p.rotations = g.rotations; p.opacities = g.opacities;
    p.shs = g.shs; p.colors_precomp = g.colors_precomp; p.cov3D_precomp = g.cov3D_precomp;
    p.view = v.viewmatrix; p.proj = v.projmatrix; p.campos = v.campos;
    return p;
}

}  // namespace

cudaError_t launch_preprocess_fwd(const b200splat_view& v, const b200splat_gaussians& g, GeomView gv,
                                  int32_t* radii, unsigned int* counters, cudaStream_t s) {
    if (g.P == 0) return cudaSuccess;
    PreParams p = make_params(v, g);
    const int M3 = 3 * g.M;
    const int coop = (g.shs && g.M > 0 && (M3 & 3) == 0 && g.M <= 16) ? 1 : 0;
    const size_t smem = coop ? (size_t)(PRE_THREADS / 32) * 32 * (((M3 >> 2) & 1) ? M3 : M3 + 4) * sizeof(float) : 0;
    static bool configured[64] = {false};
    if (cudaError_t e = configure_once(preprocess_fwd_kernel, configured, true); e != cudaSuccess) return e;
    preprocess_fwd_kernel<<<(g.P + PRE_THREADS - 1) / PRE_THREADS, PRE_THREADS, smem, s>>>(p, gv.depths, gv.rec, gv.rect,
                                                                                         radii, counters, coop, coop_shift(M3 >> 2));
    return cudaGetLastError();
}

cudaError_t launch_densify_stats_visible(int P, const int32_t* radii, float* denom, float* max_radii2D, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    densify_stats_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, denom, max_radii2D);
    return cudaGetLastError();
}

cudaError_t launch_compact_active(int P, uint8_t* touched, bool clear_flags, uint32_t* active_idx, unsigned* count, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    compact_active_kernel<<<(P + 1023) / 1024, 1024, 0, s>>>(P, touched, clear_flags ? 1 : 0, active_idx, count);
    return cudaGetLastError();
}

cudaError_t launch_preprocess_bwd(const b200splat_view& v, const b200splat_gaussians& g, const int32_t* radii,
                                  GeomView gv, float* acc, bool rezero_acc, const b200splat_grads& out, const uint32_t* active_idx,
                                  const unsigned* active_count, float* packed_rows, unsigned* packed_count, float* norm_accum,
                                  bool write_header, cudaStream_t s) {
    (void)gv;
    if (g.P == 0) return cudaSuccess;
    BwdParams bp;
    bp.p = make_params(v, g);
    bp.radii = radii; bp.acc = acc; bp.rezero = rezero_acc ? 1 : 0; bp.out = out; bp.active_idx = active_idx; bp.active_count = active_count;
    bp.packed_rows = reinterpret_cast<float4*>(packed_rows); bp.packed_count = packed_count; bp.norm_accum = norm_accum;
    bp.write_header = write_header ? 1 : 0;
    const int M3 = 3 * g.M;
    const int coop = (g.shs && g.M > 0 && (M3 & 3) == 0 && g.M <= 16) ? 1 : 0;
    const size_t smem = coop ? (size_t)(PRE_THREADS / 32) * 32 * (((M3 >> 2) & 1) ? M3 : M3 + 4) * sizeof(float) : 0;
    static bool configured[64] = {false};
    if (cudaError_t e = configure_once(preprocess_bwd_kernel, configured, true); e != cudaSuccess) return e;
    preprocess_bwd_kernel<<<(g.P + PRE_THREADS - 1) / PRE_THREADS, PRE_THREADS, smem, s>>>(bp, coop, coop_shift(M3 >> 2), out.dL_dsh ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_accumulate_packed(int P, int M, int D, const float* means3D, int n_src, const float* const* rows,
                                     const unsigned* const* counts, const float* campos, float* dL_dsh, float* dL_dgeom,
                                     uint8_t* marks, int store, cudaStream_t s) {
    if (P == 0 || n_src == 0) return cudaSuccess;
    PackedSources src;
    src.n = n_src;
    for (int v = 0; v < n_src; v++) { src.rows[v] = rows[v]; src.count[v] = counts[v]; }
    int dev = 0, sms = 148;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // two slots of 32 words (store / add launches may be queued back to back) in the module's own device memory:
    // nothing is allocated, nothing to free
    static unsigned* symbol[64] = {nullptr};                 // looked up once per device (not inside a stream capture)
    if (!symbol[dev & 63] && (e = cudaGetSymbolAddress((void**)&symbol[dev & 63], g_packed_counts)) != cudaSuccess) return e;
    unsigned* lc = symbol[dev & 63] + (store ? 128 : 0);
    gather_counts_kernel<<<1, 64, 0, s>>>(src, lc, campos == nullptr);
    if (!campos) campos = reinterpret_cast<const float*>(lc + 32);
    // Add flavour: 4 blocks per SM, each warp walking its items in position order.  The targets of a step (0.2 GB at
    // 8 x 7 % touched) exceed nothing by much, but with every position in flight at once (16 blocks per SM) the L2
    // holds no line long enough for the second view's row of the same Gaussian: measured on 7 local sources
    // 170 us / 634 MB of DRAM traffic at 16 blocks per SM, 158 us / 585 MB at 4, 153 us at 3 (fewer: too few warps).
    const int bps = n_src <= 2 ? 16 : 4;        // (one or two sources: nothing to keep in L2, more warps hide more latency)
    const bool vec_ok = ((3 * M) & 3) == 0;     // the store flavour relies on the 16-byte path
    if (store && vec_ok && n_src == 1)
        accumulate_packed_kernel<true><<<sms * bps, ACC_WARPS * 32, 0, s>>>(P, M, D, means3D, src, lc, campos, dL_dsh, dL_dgeom, marks);
    else
        accumulate_packed_kernel<false><<<sms * bps, ACC_WARPS * 32, 0, s>>>(P, M, D, means3D, src, lc, campos, dL_dsh, dL_dgeom, marks);
    return cudaGetLastError();
}

cudaError_t launch_masked_colour_grad(int P, const int32_t* radii, GeomView gv, const float* acc, float* dcol, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    masked_colour_grad_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, gv.rec, reinterpret_cast<const float4*>(acc), dcol);
    return cudaGetLastError();
}

cudaError_t launch_sh_grad_from_views(int P, int M, int D, const float* means3D, int n_views, const float* const* dcol,
                                      const float* campos, float* dL_dsh, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    ViewGrads vg;
    vg.n = n_views;
    int vec_ok = 1;   // 16-byte loads need every view's array 16-byte aligned
    for (int v = 0; v < B200SPLAT_MAX_VIEWS; v++) {
        vg.dcol[v] = v < n_views ? dcol[v] : nullptr;
        if (v < n_views && (reinterpret_cast<uintptr_t>(dcol[v]) & 15u)) vec_ok = 0;
    }
    const int M3 = 3 * M;
    const int coop = (M > 0 && (M3 & 3) == 0 && M <= 16) ? 1 : 0;
    const size_t rows_bytes = coop ? (size_t)(PRE_THREADS / 32) * 32 * (((M3 >> 2) & 1) ? M3 : M3 + 4) * sizeof(float) : 0;
    const size_t smem = rows_bytes + (size_t)(PRE_THREADS / 32) * n_views * 96 * sizeof(float);
    static bool configured[64] = {false};
    if (cudaError_t e = configure_once(sh_grad_from_views_kernel, configured, false); e != cudaSuccess) return e;
    sh_grad_from_views_kernel<<<(P + PRE_THREADS - 1) / PRE_THREADS, PRE_THREADS, smem, s>>>(P, M, D, means3D, vg, campos, dL_dsh,
                                                                                          coop, coop_shift(M3 >> 2),
                                                                                          (int)(rows_bytes / 16), vec_ok);
    return cudaGetLastError();
}

cudaError_t launch_densify_stats(int P, const float* dmeans2D, const int32_t* radii, float* accum, float* denom,
                                 float* max_radii, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, dmeans2D, radii, accum, denom, max_radii);
    return cudaGetLastError();
}

namespace {
__global__ void __launch_bounds__(256) zero_fill_kernel(ZeroRanges zr) {
    const size_t tid = (size_t)blockIdx.x * 256 + threadIdx.x, nthr = (size_t)gridDim.x * 256;
    for (int r = 0; r < zr.n; r++) {
        char* base = static_cast<char*>(zr.ptr[r]);
        const size_t bytes = zr.bytes[r];
        // head up to 16-byte alignment, body in 16-byte stores (four in flight per thread), tail
        const size_t head = min(bytes, (size_t)((16 - (reinterpret_cast<uintptr_t>(base) & 15)) & 15));
        for (size_t i = tid; i < head; i += nthr) base[i] = 0;
        uint4* body = reinterpret_cast<uint4*>(base + head);
        const size_t n16 = (bytes - head) >> 4;
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        size_t i = tid;
        // streaming stores (evict-first): 850 MB of zeros must not push the render backward's working set out of L2
        for (; i + 3 * nthr < n16; i += 4 * nthr) { __stcs(body + i, z); __stcs(body + i + nthr, z); __stcs(body + i + 2 * nthr, z); __stcs(body + i + 3 * nthr, z); }
        for (; i < n16; i += nthr) __stcs(body + i, z);
        for (size_t j = head + (n16 << 4) + tid; j < bytes; j += nthr) base[j] = 0;
    }
}
}  // namespace

cudaError_t launch_zero_fill(const ZeroRanges& zr, cudaStream_t s) {
    if (zr.n <= 0) return cudaSuccess;
    static int sms[64] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (!sms[dev & 63]) cudaDeviceGetAttribute(&sms[dev & 63], cudaDevAttrMultiProcessorCount, dev);
    zero_fill_kernel<<<2 * (sms[dev & 63] > 0 ? sms[dev & 63] : 148), 256, 0, s>>>(zr);   // two blocks per SM
    return cudaGetLastError();
}

cudaError_t launch_clear_accumulated(int P, int M, float* dL_dsh, float* dL_dgeom, uint8_t* marks, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    clear_accumulated_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, 3 * M, dL_dsh, dL_dgeom, marks);
    return cudaGetLastError();
}

cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
    return cudaGetLastError();
}

cudaError_t launch_export_geom(int P, const int32_t* radii, GeomView gv, const b200splat_state_out& o, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    export_geom_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, gv, o);
    return cudaGetLastError();
}

}  // namespace b200splat
