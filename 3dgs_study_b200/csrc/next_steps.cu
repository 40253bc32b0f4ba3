// next_steps.cu -- the steps either side of the rasterizer (include/b200splat_next.h): fused L1 + SSIM loss
// forward / backward (and the training loss built on it), activations of the raw parameters, one-launch Adam over
// all parameter groups, exact 3-nearest-neighbour mean squared distance.  sm_100a; all of it HBM / shared-memory
// bound elementwise and stencil work.
//
// L1 + SSIM.  A block owns a 32x32 output tile of one plane: the two input tiles with a 5-pixel halo are staged
// in shared memory once (zero outside the image = the zero padding of F.conv2d), the five products are filtered
// along x into shared memory, then along y in registers (each thread produces a 4-row column strip from 14
// staged values per quantity), and the SSIM map, its three partial derivatives and the block's sums leave the
// SM once.  Algorithmic HBM bytes per pixel and plane: forward 8 in + 12 out, backward 20 in + 4 out.
// The sums are reduced deterministically: per-block partials, then one block (per plane) in double.
// Measured and profiled: DESIGN.md section 7.
#include <cfloat>
#include <cmath>
#include <cub/cub.cuh>
#include <mutex>
#include <string>

#include "../../include/b200splat_next.h"
#include "splat_internal.cuh"

using namespace b200splat;

namespace {

constexpr int LT = 32;              // output tile edge
constexpr int LR = 5;               // window radius (11 taps)
constexpr int LIN = LT + 2 * LR;    // staged tile edge (42)
constexpr int LPITCH = LIN + 2;     // shared-memory row pitch
constexpr int LTHREADS = 256;
constexpr int LROWS = LT / (LTHREADS / 32);   // rows per thread in the y pass (4)

struct Taps { float w[2 * LR + 1]; };

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- packed fp32 pairs (one FFMA2 / FMUL2 per two lanes' worth of work; each half is the IEEE result) ----------
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 mk2(float lo, float hi) {
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void un2(f2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    f2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
    return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    f2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f2 ld2(const float2* p) { float2 t = *p; return mk2(t.x, t.y); }

struct Taps2 { f2 w[2 * LR + 1]; };   // (w, w) pairs
__device__ __forceinline__ Taps2 pair_taps(const Taps& tw) {
    Taps2 t;
#pragma unroll
    for (int k = 0; k < 2 * LR + 1; ++k) t.w[k] = mk2(tw.w[k], tw.w[k]);
    return t;
}

constexpr int XTASKS = LIN * (LT / 2);   // x pass: one task = one staged row x two adjacent output columns

// Forward.  Shared memory: the (img, gt) pair of every staged pixel interleaved (float2), and the x-filtered
// quantities as (mu1', mu2') pairs, (E11', E22') pairs and E12' scalars.  x pass: a task owns two adjacent output
// columns of one row, loads its 12 staged pairs with six 16-byte loads (conflict-free: consecutive tasks are 16 bytes
// apart), forms the products once per staged pixel and runs the 11 taps on packed pairs.  y pass: a thread owns one
// column x four rows and walks 14 staged rows.  Same tap order as the oracle (x then y, ascending), and each half of
// a packed FMA is the IEEE fma, so results are those of the scalar formulation.
__global__ void __launch_bounds__(LTHREADS)
l1_ssim_fwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Taps tw,
                   float* __restrict__ maps, size_t map_stride, float2* __restrict__ partials) {
    __shared__ __align__(16) float2 sab[LIN][LPITCH];
    __shared__ __align__(16) float2 sh01[LIN][LT];
    __shared__ __align__(16) float2 sh23[LIN][LT];
    __shared__ __align__(16) float sh4[LIN][LT];
    __shared__ float red[2][LTHREADS / 32];

    const int plane = blockIdx.z;
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t pbase = (size_t)plane * H * W;
    const float* A = img + pbase;
    const float* B = gt + pbase;

    // staging: a warp owns staged rows w, w+8, ...; a lane owns columns lane and lane+32 of a row (no divisions, one
    // row pointer per row); all of a thread's global loads are issued before its first shared-memory store
    {
        constexpr int NR = (LIN + LTHREADS / 32 - 1) / (LTHREADS / 32);      // 6 rows per warp
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        const int xa = x0 + lane - LR, xb = xa + 32;
        const bool xa_in = (xa >= 0) && (xa < W);
        const bool xb_in = (lane < LIN - 32) && (xb >= 0) && (xb < W);
        float va[NR][2], vb[NR][2];
#pragma unroll
        for (int u = 0; u < NR; ++u) {
            const int r = wrp + (LTHREADS / 32) * u;
            const int y = y0 + r - LR;
            const bool yin = (r < LIN) && (y >= 0) && (y < H);
            const size_t ro = (size_t)(yin ? y : 0) * W;
            va[u][0] = (yin && xa_in) ? __ldg(A + ro + xa) : 0.0f;
            vb[u][0] = (yin && xa_in) ? __ldg(B + ro + xa) : 0.0f;
            va[u][1] = (yin && xb_in) ? __ldg(A + ro + xb) : 0.0f;
            vb[u][1] = (yin && xb_in) ? __ldg(B + ro + xb) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < NR; ++u) {
            const int r = wrp + (LTHREADS / 32) * u;
            if (r < LIN) {
                sab[r][lane] = make_float2(va[u][0], vb[u][0]);
                if (lane < LIN - 32) sab[r][lane + 32] = make_float2(va[u][1], vb[u][1]);
            }
        }
    }
    __syncthreads();

    const Taps2 t2 = pair_taps(tw);
    for (int t = threadIdx.x; t < XTASKS; t += LTHREADS) {
        const int r = t >> 4, c = (t & 15) * 2;
        f2 ab[12], qq[12];
        float pq[12];
        const float4* src = reinterpret_cast<const float4*>(&sab[r][c]);
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            float4 v = src[j];
            ab[2 * j] = mk2(v.x, v.y);
            ab[2 * j + 1] = mk2(v.z, v.w);
            pq[2 * j] = __fmul_rn(v.x, v.y);
            pq[2 * j + 1] = __fmul_rn(v.z, v.w);
        }
#pragma unroll
        for (int j = 0; j < 12; ++j) qq[j] = mul2(ab[j], ab[j]);
        f2 m0 = mk2(0.f, 0.f), m1 = m0, e0 = m0, e1 = m0;
        float x0s = 0.f, x1s = 0.f;
#pragma unroll
        for (int k = 0; k < 2 * LR + 1; ++k) {
            m0 = fma2(t2.w[k], ab[k], m0);
            m1 = fma2(t2.w[k], ab[k + 1], m1);
            e0 = fma2(t2.w[k], qq[k], e0);
            e1 = fma2(t2.w[k], qq[k + 1], e1);
            x0s = fmaf(tw.w[k], pq[k], x0s);
            x1s = fmaf(tw.w[k], pq[k + 1], x1s);
        }
        float a0, a1, a2, a3;
        un2(m0, a0, a1); un2(m1, a2, a3);
        *reinterpret_cast<float4*>(&sh01[r][c]) = make_float4(a0, a1, a2, a3);
        un2(e0, a0, a1); un2(e1, a2, a3);
        *reinterpret_cast<float4*>(&sh23[r][c]) = make_float4(a0, a1, a2, a3);
        *reinterpret_cast<float2*>(&sh4[r][c]) = make_float2(x0s, x1s);
    }
    __syncthreads();

    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int r0 = ty * LROWS;
    f2 mu[LROWS], ee[LROWS];
    float e12[LROWS];
#pragma unroll
    for (int j = 0; j < LROWS; ++j) { mu[j] = mk2(0.f, 0.f); ee[j] = mu[j]; e12[j] = 0.f; }
    // 14 staged rows feed 4 output rows: row i contributes to output j with tap k = i - j
#pragma unroll
    for (int i = 0; i < LROWS + 2 * LR; ++i) {
        f2 v01 = ld2(&sh01[r0 + i][tx]);
        f2 v23 = ld2(&sh23[r0 + i][tx]);
        float v4 = sh4[r0 + i][tx];
#pragma unroll
        for (int j = 0; j < LROWS; ++j) {
            const int k = i - j;
            if (k >= 0 && k < 2 * LR + 1) {
                mu[j] = fma2(t2.w[k], v01, mu[j]);
                ee[j] = fma2(t2.w[k], v23, ee[j]);
                e12[j] = fmaf(tw.w[k], v4, e12[j]);
            }
        }
    }

    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    float l1 = 0.0f, ss = 0.0f;
    const int x = x0 + tx;
#pragma unroll
    for (int j = 0; j < LROWS; ++j) {
        int y = y0 + r0 + j;
        if (x < W && y < H) {
            float m1, m2, e11, e22;
            un2(mu[j], m1, m2);
            un2(ee[j], e11, e22);
            float m11 = m1 * m1, m22 = m2 * m2, m12 = m1 * m2;
            float s1 = e11 - m11, s2 = e22 - m22, s12 = e12[j] - m12;
            float An = 2.0f * m12 + C1, Bn = 2.0f * s12 + C2;
            float Cd = m11 + m22 + C1, Dd = s1 + s2 + C2;
            float inv = 1.0f / (Cd * Dd);
            float m = An * Bn * inv;
            ss += m;
            float2 px = sab[r0 + j + LR][tx + LR];
            l1 += fabsf(px.x - px.y);
            if (maps) {
                size_t o = pbase + (size_t)y * W + x;
                maps[o] = 2.0f * inv * (m2 * (Bn - An) - m1 * m * (Dd - Cd));
                maps[map_stride + o] = -(m * (inv * Cd));     // -m / Dd with 1/Dd = inv * Cd: one division per pixel
                maps[2 * map_stride + o] = 2.0f * An * inv;
            }
        }
    }
    l1 = warp_sum(l1);
    ss = warp_sum(ss);
    if (tx == 0) { red[0][ty] = l1; red[1][ty] = ss; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int i = 0; i < LTHREADS / 32; ++i) { a += red[0][i]; b += red[1][i]; }
        size_t blk = ((size_t)plane * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partials[blk] = make_float2(a, b);
    }
}

// one block per plane: fixed-order sum of that plane's block partials in double -> per-plane sums
__global__ void __launch_bounds__(256)
l1_ssim_finalize_kernel(int blocks_per_plane, const float2* __restrict__ partials, float* __restrict__ out) {
    __shared__ double sl[256], ssm[256];
    const float2* p = partials + (size_t)blockIdx.x * blocks_per_plane;
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < blocks_per_plane; i += 256) { float2 v = p[i]; a += (double)v.x; b += (double)v.y; }
    sl[threadIdx.x] = a; ssm[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; ssm[threadIdx.x] += ssm[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[2 * blockIdx.x] = (float)sl[0];
        out[2 * blockIdx.x + 1] = (float)ssm[0];
    }
}

// the training loss in one block: loss3 = (loss, mean |a-b|, mean SSIM), loss = (1-lambda) l1 + lambda (1 - ssim)
__global__ void __launch_bounds__(256)
l1_ssim_loss_kernel(int n_partials, double inv_count, float lambda, const float2* __restrict__ partials, float* __restrict__ loss3) {
    __shared__ double sl[256], ssm[256];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < n_partials; i += 256) { float2 v = partials[i]; a += (double)v.x; b += (double)v.y; }
    sl[threadIdx.x] = a; ssm[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; ssm[threadIdx.x] += ssm[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float l1 = (float)(sl[0] * inv_count), ss = (float)(ssm[0] * inv_count);
        loss3[0] = (1.0f - lambda) * l1 + lambda * (1.0f - ss);
        loss3[1] = l1;
        loss3[2] = ss;
    }
}

// Backward: the same stencil over the three derivative maps, (d m/d mu1, d m/d E11) as packed pairs and d m/d E12
// as scalars.
__global__ void __launch_bounds__(LTHREADS, 5)
l1_ssim_bwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Taps tw,
                   const float* __restrict__ maps, size_t map_stride, const float* __restrict__ scale,
                   const float* __restrict__ upstream, float c_l1, float c_ssim, float* __restrict__ dimg) {
    __shared__ __align__(16) float2 s01[LIN][LPITCH];
    __shared__ __align__(16) float s2[LIN][LPITCH];
    __shared__ __align__(16) float2 h01[LIN][LT];
    __shared__ __align__(16) float h2[LIN][LT];

    const int plane = blockIdx.z;
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t pbase = (size_t)plane * H * W;

    {
        constexpr int NR = (LIN + LTHREADS / 32 - 1) / (LTHREADS / 32);      // 6 rows per warp
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        const int xa = x0 + lane - LR, xb = xa + 32;
        const bool xa_in = (xa >= 0) && (xa < W);
        const bool xb_in = (lane < LIN - 32) && (xb >= 0) && (xb < W);
        const float* m0 = maps + pbase;
        const float* m1 = m0 + map_stride;
        const float* m2 = m1 + map_stride;
        float v0[NR][2], v1[NR][2], v2[NR][2];
#pragma unroll
        for (int u = 0; u < NR; ++u) {
            const int r = wrp + (LTHREADS / 32) * u;
            const int y = y0 + r - LR;
            const bool yin = (r < LIN) && (y >= 0) && (y < H);
            const size_t ro = (size_t)(yin ? y : 0) * W;
            const bool ia = yin && xa_in, ib = yin && xb_in;
            v0[u][0] = ia ? __ldg(m0 + ro + xa) : 0.0f;
            v1[u][0] = ia ? __ldg(m1 + ro + xa) : 0.0f;
            v2[u][0] = ia ? __ldg(m2 + ro + xa) : 0.0f;
            v0[u][1] = ib ? __ldg(m0 + ro + xb) : 0.0f;
            v1[u][1] = ib ? __ldg(m1 + ro + xb) : 0.0f;
            v2[u][1] = ib ? __ldg(m2 + ro + xb) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < NR; ++u) {
            const int r = wrp + (LTHREADS / 32) * u;
            if (r < LIN) {
                s01[r][lane] = make_float2(v0[u][0], v1[u][0]);
                s2[r][lane] = v2[u][0];
                if (lane < LIN - 32) {
                    s01[r][lane + 32] = make_float2(v0[u][1], v1[u][1]);
                    s2[r][lane + 32] = v2[u][1];
                }
            }
        }
    }
    __syncthreads();
    // this thread's four pixels of img / gt for the epilogue: requested now, consumed after the two passes
    float pa[LROWS], pb[LROWS];
    {
        const int ex = x0 + (threadIdx.x & 31), ey = y0 + (threadIdx.x >> 5) * LROWS;
#pragma unroll
        for (int j = 0; j < LROWS; ++j) {
            bool in = (ex < W) && (ey + j < H);
            size_t o = in ? (pbase + (size_t)(ey + j) * W + ex) : 0;
            pa[j] = in ? __ldg(img + o) : 0.0f;
            pb[j] = in ? __ldg(gt + o) : 0.0f;
        }
    }
    const Taps2 t2 = pair_taps(tw);
    for (int t = threadIdx.x; t < XTASKS; t += LTHREADS) {
        const int r = t >> 4, c = (t & 15) * 2;
        f2 ab[12];
        float q[12];
        const float4* src = reinterpret_cast<const float4*>(&s01[r][c]);
        const float2* src2 = reinterpret_cast<const float2*>(&s2[r][c]);
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            float4 v = src[j];
            ab[2 * j] = mk2(v.x, v.y);
            ab[2 * j + 1] = mk2(v.z, v.w);
            float2 u = src2[j];
            q[2 * j] = u.x;
            q[2 * j + 1] = u.y;
        }
        f2 m0 = mk2(0.f, 0.f), m1 = m0;
        float x0s = 0.f, x1s = 0.f;
#pragma unroll
        for (int k = 0; k < 2 * LR + 1; ++k) {
            m0 = fma2(t2.w[k], ab[k], m0);
            m1 = fma2(t2.w[k], ab[k + 1], m1);
            x0s = fmaf(tw.w[k], q[k], x0s);
            x1s = fmaf(tw.w[k], q[k + 1], x1s);
        }
        float a0, a1, a2, a3;
        un2(m0, a0, a1); un2(m1, a2, a3);
        *reinterpret_cast<float4*>(&h01[r][c]) = make_float4(a0, a1, a2, a3);
        *reinterpret_cast<float2*>(&h2[r][c]) = make_float2(x0s, x1s);
    }
    __syncthreads();

    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int r0 = ty * LROWS;
    f2 f01[LROWS];
    float f2s[LROWS];
#pragma unroll
    for (int j = 0; j < LROWS; ++j) { f01[j] = mk2(0.f, 0.f); f2s[j] = 0.f; }
#pragma unroll
    for (int i = 0; i < LROWS + 2 * LR; ++i) {
        f2 v01 = ld2(&h01[r0 + i][tx]);
        float v2 = h2[r0 + i][tx];
#pragma unroll
        for (int j = 0; j < LROWS; ++j) {
            const int k = i - j;
            if (k >= 0 && k < 2 * LR + 1) {
                f01[j] = fma2(t2.w[k], v01, f01[j]);
                f2s[j] = fmaf(tw.w[k], v2, f2s[j]);
            }
        }
    }
    // per-plane factors, or one upstream scalar times two host constants (the fused training loss)
    const float up = upstream ? __ldg(upstream) : 1.0f;
    const float sl = scale ? __ldg(scale + 2 * plane) : up * c_l1;
    const float ssc = scale ? __ldg(scale + 2 * plane + 1) : up * c_ssim;
    const int x = x0 + tx;
#pragma unroll
    for (int j = 0; j < LROWS; ++j) {
        int y = y0 + r0 + j;
        if (x < W && y < H) {
            size_t o = pbase + (size_t)y * W + x;
            float a = pa[j], b = pb[j];
            float d = a - b;
            float sg = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);
            float f0, f1;
            un2(f01[j], f0, f1);
            float g = f0 + 2.0f * a * f1 + b * f2s[j];
            dimg[o] = sl * sg + ssc * g;
        }
    }
}

// ---- Adam ------------------------------------------------------------------------------------------------
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_PER_BLOCK = ADAM_THREADS * 16;   // 4 x float4 per thread

struct AdamLaunch {
    float* p[B200SPLAT_ADAM_MAX_GROUPS];
    const float* g[B200SPLAT_ADAM_MAX_GROUPS];
    float* m[B200SPLAT_ADAM_MAX_GROUPS];
    float* v[B200SPLAT_ADAM_MAX_GROUPS];
    long long n[B200SPLAT_ADAM_MAX_GROUPS];
    float step_size[B200SPLAT_ADAM_MAX_GROUPS];
    float step_size_tail[B200SPLAT_ADAM_MAX_GROUPS];   // for elements whose column (index % row_width) >= row_split
    int row_width[B200SPLAT_ADAM_MAX_GROUPS];          // 0: one learning rate for the whole tensor
    int row_split[B200SPLAT_ADAM_MAX_GROUPS];
    float bc2_sqrt[B200SPLAT_ADAM_MAX_GROUPS];
    unsigned block_start[B200SPLAT_ADAM_MAX_GROUPS + 1];
    int vec[B200SPLAT_ADAM_MAX_GROUPS];
    int ng;
    float beta2, w1, w2, eps;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float step_size, float bc2_sqrt,
                                         float beta2, float w1, float w2, float eps) {
    float mi = __fmaf_rn(__fsub_rn(g, m), w1, m);                    // m + (g - m) * (1 - beta1)
    float vi = __fmaf_rn(__fmul_rn(w2, g), g, __fmul_rn(v, beta2));  // v * beta2 + ((1 - beta2) * g) * g
    float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), bc2_sqrt), eps);
    p = __fmaf_rn(-step_size, __fdiv_rn(mi, denom), p);
    m = mi;
    v = vi;
}

__global__ void __launch_bounds__(ADAM_THREADS) adam_kernel(const AdamLaunch L) {
    int gi = 0;
#pragma unroll
    for (int i = 1; i < B200SPLAT_ADAM_MAX_GROUPS; ++i)
        if (i < L.ng && blockIdx.x >= L.block_start[i]) gi = i;
    float* __restrict__ P = L.p[gi];
    const float* __restrict__ G = L.g[gi];
    float* __restrict__ M = L.m[gi];
    float* __restrict__ V = L.v[gi];
    const long long n = L.n[gi];
    const float step_head = L.step_size[gi], step_tail = L.step_size_tail[gi], bc2_sqrt = L.bc2_sqrt[gi];
    const int rw = L.row_width[gi], rs = L.row_split[gi];
    const long long base = (long long)(blockIdx.x - L.block_start[gi]) * ADAM_PER_BLOCK;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        long long e = base + ((long long)it * ADAM_THREADS + threadIdx.x) * 4;
        if (e >= n) continue;
        // learning rate by column, per element (a float4 may straddle the split or a row end)
        float ss[4];
        if (rw > 0) {
            int c = (int)(e % rw);
#pragma unroll
            for (int j = 0; j < 4; ++j) { ss[j] = (c >= rs) ? step_tail : step_head; if (++c == rw) c = 0; }
        } else {
            ss[0] = ss[1] = ss[2] = ss[3] = step_head;
        }
        if (L.vec[gi] && e + 4 <= n) {
            float4 p = *reinterpret_cast<const float4*>(P + e);
            float4 g = __ldg(reinterpret_cast<const float4*>(G + e));
            float4 m = *reinterpret_cast<const float4*>(M + e);
            float4 v = *reinterpret_cast<const float4*>(V + e);
            adam_one(p.x, g.x, m.x, v.x, ss[0], bc2_sqrt, L.beta2, L.w1, L.w2, L.eps);
            adam_one(p.y, g.y, m.y, v.y, ss[1], bc2_sqrt, L.beta2, L.w1, L.w2, L.eps);
            adam_one(p.z, g.z, m.z, v.z, ss[2], bc2_sqrt, L.beta2, L.w1, L.w2, L.eps);
            adam_one(p.w, g.w, m.w, v.w, ss[3], bc2_sqrt, L.beta2, L.w1, L.w2, L.eps);
            *reinterpret_cast<float4*>(P + e) = p;
            *reinterpret_cast<float4*>(M + e) = m;
            *reinterpret_cast<float4*>(V + e) = v;
        } else {
            for (int j = 0; j < 4 && e + j < n; ++j) {
                float p = P[e + j], m = M[e + j], v = V[e + j];
                adam_one(p, G[e + j], m, v, ss[j], bc2_sqrt, L.beta2, L.w1, L.w2, L.eps);
                P[e + j] = p; M[e + j] = m; V[e + j] = v;
            }
        }
    }
}

// ---- activations of the raw parameters -----------------------------------------------------------------------
// scales = exp(raw), rotations = raw / max(||raw||, 1e-12), opacities = sigmoid(raw): one pass over the 8 raw
// floats of a Gaussian instead of ~8 elementwise/reduction launches, and one pass back instead of ~15.
__global__ void __launch_bounds__(256)
activate_fwd_kernel(int P, const float* __restrict__ scaling, const float4* __restrict__ rotation,
                    const float* __restrict__ opacity, float* __restrict__ scales, float4* __restrict__ rotations,
                    float* __restrict__ opacities) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) scales[3 * (size_t)i + k] = expf(scaling[3 * (size_t)i + k]);
    float4 q = rotation[i];
    float nrm = sqrtf(__fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmul_rn(q.x, q.x)))));
    float d = fmaxf(nrm, 1e-12f);
    rotations[i] = make_float4(q.x / d, q.y / d, q.z / d, q.w / d);
    opacities[i] = 1.0f / (1.0f + expf(-opacity[i]));
}

// raw gradients from the activated values (saved by the forward) and their upstream gradients
__global__ void __launch_bounds__(256)
activate_bwd_kernel(int P, const float4* __restrict__ rotation, const float* __restrict__ scales,
                    const float* __restrict__ opacities, const float* __restrict__ g_scales,
                    const float4* __restrict__ g_rotations, const float* __restrict__ g_opacities,
                    float* __restrict__ d_scaling, float4* __restrict__ d_rotation, float* __restrict__ d_opacity) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (g_scales) {
#pragma unroll
        for (int k = 0; k < 3; ++k) d_scaling[3 * (size_t)i + k] = g_scales[3 * (size_t)i + k] * scales[3 * (size_t)i + k];
    }
    if (g_rotations) {
        float4 q = rotation[i], g = g_rotations[i];
        float nrm = sqrtf(__fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmul_rn(q.x, q.x)))));
        float4 r;
        if (nrm > 1e-12f) {
            // y = q / n:  dq = (g - y (y . g)) / n
            float inv = 1.0f / nrm;
            float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
            float dot = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
            r = make_float4((g.x - y.x * dot) * inv, (g.y - y.y * dot) * inv, (g.z - y.z * dot) * inv, (g.w - y.w * dot) * inv);
        } else {
            r = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);   // clamp_min branch: y = q / eps
        }
        d_rotation[i] = r;
    }
    if (g_opacities) {
        float o = opacities[i];
        d_opacity[i] = g_opacities[i] * ((1.0f - o) * o);
    }
}

// ---- 3 nearest neighbours ----------------------------------------------------------------------------------
constexpr int KBOX = 64;          // points per leaf box (sorted by Morton code)
constexpr int KSUPER = 64;        // leaf boxes per super box

struct Box { float lo[3]; float hi[3]; };

__device__ __forceinline__ unsigned expand10(unsigned x) {   // 10 bits -> every third bit
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

// ordered-int encoding of a float so that integer min/max are float min/max
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__global__ void knn_bounds_init_kernel(int* b6) {
    if (threadIdx.x < 3) b6[threadIdx.x] = 0x7FFFFFFF;             // min
    else if (threadIdx.x < 6) b6[threadIdx.x] = (int)0x80000000;   // max
}

__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ xyz, int* __restrict__ b6) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (long long)gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { float c = xyz[3 * i + k]; lo[k] = fminf(lo[k], c); hi[k] = fmaxf(hi[k], c); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(b6 + k, f2ord(lo[k])); atomicMax(b6 + 3 + k, f2ord(hi[k])); }
    }
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ xyz, const int* __restrict__ b6,
                                                         unsigned* __restrict__ codes, unsigned* __restrict__ ids) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    unsigned code = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float lo = ord2f(b6[k]), hi = ord2f(b6[3 + k]);
        float ext = hi - lo;
        float t = ext > 0.0f ? (xyz[3 * i + k] - lo) / ext : 0.0f;
        t = fminf(fmaxf(t, 0.0f), 1.0f);          // NaN -> 0
        unsigned q = (unsigned)(t * 1023.0f);
        code |= expand10(q > 1023u ? 1023u : q) << k;
    }
    codes[i] = code;
    ids[i] = (unsigned)i;
}

// points gathered into Morton order (float4: x, y, z, unused) + leaf boxes of KBOX consecutive points
__global__ void __launch_bounds__(KBOX) knn_gather_boxes_kernel(int P, const float* __restrict__ xyz,
                                                                const unsigned* __restrict__ ids_sorted,
                                                                float4* __restrict__ pts, Box* __restrict__ boxes) {
    int i = blockIdx.x * KBOX + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        unsigned id = ids_sorted[i];
        float x = xyz[3 * (size_t)id], y = xyz[3 * (size_t)id + 1], z = xyz[3 * (size_t)id + 2];
        pts[i] = make_float4(x, y, z, 0.0f);
        lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
    }
    __shared__ float slo[KBOX / 32][3], shi[KBOX / 32][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5][k] = lo[k]; shi[threadIdx.x >> 5][k] = hi[k]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        Box b;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            b.lo[k] = fminf(slo[0][k], slo[1][k]);
            b.hi[k] = fmaxf(shi[0][k], shi[1][k]);
        }
        boxes[blockIdx.x] = b;
    }
}

__global__ void __launch_bounds__(KSUPER) knn_super_boxes_kernel(int n_boxes, const Box* __restrict__ boxes, Box* __restrict__ supers) {
    int i = blockIdx.x * KSUPER + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n_boxes) {
        Box b = boxes[i];
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = b.lo[k]; hi[k] = b.hi[k]; }
    }
    __shared__ float slo[KSUPER / 32][3], shi[KSUPER / 32][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5][k] = lo[k]; shi[threadIdx.x >> 5][k] = hi[k]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        Box b;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            b.lo[k] = fminf(slo[0][k], slo[1][k]);
            b.hi[k] = fmaxf(shi[0][k], shi[1][k]);
        }
        supers[blockIdx.x] = b;
    }
}

// squared distance in the one canonical form every comparison uses -- the instruction sequence nvcc emits for the
// reference's d.x*d.x + d.y*d.y + d.z*d.z (FMUL dy*dy, FFMA dx*dx, FFMA dz*dz; SASS of its boxMeanDist); per-axis monotone, so a box's value never
// exceeds that of any point inside it and pruning on it is exact
__device__ __forceinline__ float dist2(float dx, float dy, float dz) {
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}
__device__ __forceinline__ float box_dist2(const Box& b, float x, float y, float z) {
    float dx = fmaxf(0.0f, fmaxf(__fsub_rn(b.lo[0], x), __fsub_rn(x, b.hi[0])));
    float dy = fmaxf(0.0f, fmaxf(__fsub_rn(b.lo[1], y), __fsub_rn(y, b.hi[1])));
    float dz = fmaxf(0.0f, fmaxf(__fsub_rn(b.lo[2], z), __fsub_rn(z, b.hi[2])));
    return dist2(dx, dy, dz);
}
__device__ __forceinline__ void push3(float* best, float d) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
        if (best[k] > d) { float t = best[k]; best[k] = d; d = t; }
}
__device__ __forceinline__ void scan_leaf(int P, const float4* __restrict__ pts, int box, int self, float x, float y,
                                          float z, float* best) {
    int lo = box * KBOX, hi = min(P, lo + KBOX);
    for (int j = lo; j < hi; ++j) {
        if (j == self) continue;
        float4 q = __ldg(pts + j);
        push3(best, dist2(__fsub_rn(q.x, x), __fsub_rn(q.y, y), __fsub_rn(q.z, z)));
    }
}

__global__ void __launch_bounds__(128) knn_search_kernel(int P, const float4* __restrict__ pts,
                                                         const unsigned* __restrict__ ids_sorted, int n_boxes,
                                                         const Box* __restrict__ boxes, int n_supers,
                                                         const Box* __restrict__ supers, float* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float4 me = pts[i];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    const int own = i / KBOX;
    scan_leaf(P, pts, own, i, me.x, me.y, me.z, best);
    for (int s = 0; s < n_supers; ++s) {
        Box sb = supers[s];
        if (box_dist2(sb, me.x, me.y, me.z) > best[2]) continue;
        int b1 = min(n_boxes, (s + 1) * KSUPER);
        for (int b = s * KSUPER; b < b1; ++b) {
            if (b == own) continue;
            Box lb = boxes[b];
            if (box_dist2(lb, me.x, me.y, me.z) > best[2]) continue;
            scan_leaf(P, pts, b, i, me.x, me.y, me.z, best);
        }
    }
    out[ids_sorted[i]] = __fdiv_rn(__fadd_rn(__fadd_rn(best[0], best[1]), best[2]), 3.0f);
}

// ---- b200splat_dist2_knn3_static: the same search with its scratch in the module's own device memory ----------
// Scratch for up to B200SPLAT_KNN_STATIC_MAX_POINTS points; every GPU that loads the library holds these 12 MiB until
// exit.  In host code `g_knn_static_scratch` is only a handle for cudaGetSymbolAddress -- never a device address.
constexpr size_t KNN_STATIC_BYTES = (size_t)12 << 20;
static __device__ __align__(256) unsigned char g_knn_static_scratch[KNN_STATIC_BYTES];

// device address of the array on the CURRENT device, looked up once per device
cudaError_t knn_static_scratch(char** base) {
    static std::mutex mu;
    static char* addr[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        if ((e = cudaGetSymbolAddress(&p, g_knn_static_scratch)) != cudaSuccess) return e;
        addr[dev] = static_cast<char*>(p);
    }
    *base = addr[dev];
    return cudaSuccess;
}

inline bool aligned16p(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

#define CUN(call, where)                                                                   \
    do {                                                                                   \
        cudaError_t _e = (call);                                                           \
        if (_e != cudaSuccess) {                                                           \
            std::string _m = std::string(where) + ": " + cudaGetErrorString(_e);           \
            return report_error(B200SPLAT_ERR_CUDA, _m.c_str());                           \
        }                                                                                  \
    } while (0)

inline int tiles_1d(int n) { return (n + LT - 1) / LT; }

}  // namespace

extern "C" {

size_t b200splat_l1_ssim_partials(int32_t planes, int32_t height, int32_t width) {
    if (planes <= 0 || height <= 0 || width <= 0) return 0;
    return (size_t)planes * tiles_1d(height) * tiles_1d(width);
}

int b200splat_l1_ssim_forward(int32_t planes, int32_t height, int32_t width, const float* img, const float* gt,
                              const float* window11, float* maps, float* partials, float* out, void* stream_) {
    if (planes <= 0 || height <= 0 || width <= 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: bad sizes");
    if (planes > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: more than 65535 planes");
    if (!img || !gt || !window11 || !partials || !out) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: null argument");
    cudaStream_t s = (cudaStream_t)stream_;
    Taps tw;
    for (int k = 0; k < 2 * LR + 1; ++k) tw.w[k] = window11[k];
    dim3 grid(tiles_1d(width), tiles_1d(height), planes);
    if (grid.y > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: image too tall");
    l1_ssim_fwd_kernel<<<grid, LTHREADS, 0, s>>>(height, width, img, gt, tw, maps, (size_t)planes * height * width,
                                                 reinterpret_cast<float2*>(partials));
    CUN(cudaGetLastError(), "l1_ssim forward");
    l1_ssim_finalize_kernel<<<planes, 256, 0, s>>>((int)(grid.x * grid.y), reinterpret_cast<const float2*>(partials), out);
    CUN(cudaGetLastError(), "l1_ssim finalize");
    return B200SPLAT_OK;
}

int b200splat_l1_ssim_backward(int32_t planes, int32_t height, int32_t width, const float* img, const float* gt,
                               const float* window11, const float* maps, const float* scale, float* dL_dimg,
                               void* stream_) {
    if (planes <= 0 || height <= 0 || width <= 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: bad sizes");
    if (planes > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: more than 65535 planes");
    if (!img || !gt || !window11 || !maps || !scale || !dL_dimg) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: null argument");
    Taps tw;
    for (int k = 0; k < 2 * LR + 1; ++k) tw.w[k] = window11[k];
    dim3 grid(tiles_1d(width), tiles_1d(height), planes);
    if (grid.y > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "l1_ssim: image too tall");
    l1_ssim_bwd_kernel<<<grid, LTHREADS, 0, (cudaStream_t)stream_>>>(height, width, img, gt, tw, maps,
                                                                     (size_t)planes * height * width, scale, nullptr, 0.f, 0.f, dL_dimg);
    CUN(cudaGetLastError(), "l1_ssim backward");
    return B200SPLAT_OK;
}

int b200splat_training_loss_forward(int32_t planes, int32_t height, int32_t width, const float* img, const float* gt,
                                    const float* window11, float lambda_dssim, float* maps, float* partials, float* loss3,
                                    void* stream_) {
    if (planes <= 0 || height <= 0 || width <= 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: bad sizes");
    if (planes > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: more than 65535 planes");
    if (!img || !gt || !window11 || !partials || !loss3) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: null argument");
    cudaStream_t s = (cudaStream_t)stream_;
    Taps tw;
    for (int k = 0; k < 2 * LR + 1; ++k) tw.w[k] = window11[k];
    dim3 grid(tiles_1d(width), tiles_1d(height), planes);
    if (grid.y > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: image too tall");
    l1_ssim_fwd_kernel<<<grid, LTHREADS, 0, s>>>(height, width, img, gt, tw, maps, (size_t)planes * height * width,
                                                 reinterpret_cast<float2*>(partials));
    CUN(cudaGetLastError(), "training_loss forward");
    l1_ssim_loss_kernel<<<1, 256, 0, s>>>((int)(grid.x * grid.y * grid.z), 1.0 / ((double)planes * height * width), lambda_dssim,
                                          reinterpret_cast<const float2*>(partials), loss3);
    CUN(cudaGetLastError(), "training_loss finalize");
    return B200SPLAT_OK;
}

int b200splat_training_loss_backward(int32_t planes, int32_t height, int32_t width, const float* img, const float* gt,
                                     const float* window11, float lambda_dssim, const float* maps, const float* upstream,
                                     float* dL_dimg, void* stream_) {
    if (planes <= 0 || height <= 0 || width <= 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: bad sizes");
    if (planes > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: more than 65535 planes");
    if (!img || !gt || !window11 || !maps || !dL_dimg) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: null argument");
    Taps tw;
    for (int k = 0; k < 2 * LR + 1; ++k) tw.w[k] = window11[k];
    dim3 grid(tiles_1d(width), tiles_1d(height), planes);
    if (grid.y > 65535) return report_error(B200SPLAT_ERR_INVALID_ARG, "training_loss: image too tall");
    const double n = (double)planes * height * width;
    l1_ssim_bwd_kernel<<<grid, LTHREADS, 0, (cudaStream_t)stream_>>>(height, width, img, gt, tw, maps, (size_t)planes * height * width,
                                                                     nullptr, upstream, (float)((1.0 - (double)lambda_dssim) / n),
                                                                     (float)(-(double)lambda_dssim / n), dL_dimg);
    CUN(cudaGetLastError(), "training_loss backward");
    return B200SPLAT_OK;
}

int b200splat_adam_step(int32_t n_groups, const b200splat_adam_group* groups, double beta1, double beta2, double eps,
                        void* stream_) {
    if (n_groups < 0 || n_groups > B200SPLAT_ADAM_MAX_GROUPS) return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: bad group count");
    if (n_groups > 0 && !groups) return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: null groups");
    AdamLaunch L;
    memset(&L, 0, sizeof(L));
    unsigned blocks = 0;
    int ng = 0;
    for (int i = 0; i < n_groups; ++i) {
        const b200splat_adam_group& g = groups[i];
        if (g.n < 0 || g.step < 1) return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: bad n or step");
        if (!g.grad || g.n == 0) continue;
        if (!g.param || !g.exp_avg || !g.exp_avg_sq) return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: null state");
        long long nb = (g.n + ADAM_PER_BLOCK - 1) / ADAM_PER_BLOCK;
        if (nb + blocks > 0x7FFFFFFFll) return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: too many elements");
        L.p[ng] = g.param; L.g[ng] = g.grad; L.m[ng] = g.exp_avg; L.v[ng] = g.exp_avg_sq; L.n[ng] = g.n;
        double bc1 = 1.0 - pow(beta1, (double)g.step), bc2 = 1.0 - pow(beta2, (double)g.step);
        L.step_size[ng] = (float)(g.lr / bc1);
        if (g.row_width < 0 || g.row_split < 0 || g.row_split > g.row_width || (g.row_width > 0 && g.n % g.row_width != 0))
            return report_error(B200SPLAT_ERR_INVALID_ARG, "adam: bad row_width / row_split");
        L.step_size_tail[ng] = (float)(g.lr_tail / bc1);
        L.row_width[ng] = g.row_width;
        L.row_split[ng] = g.row_split;
        L.bc2_sqrt[ng] = (float)sqrt(bc2);
        L.vec[ng] = aligned16p(g.param) && aligned16p(g.grad) && aligned16p(g.exp_avg) && aligned16p(g.exp_avg_sq);
        L.block_start[ng] = blocks;
        blocks += (unsigned)nb;
        ++ng;
    }
    if (ng == 0) return B200SPLAT_OK;
    for (int i = ng; i <= B200SPLAT_ADAM_MAX_GROUPS; ++i) L.block_start[i] = blocks;
    L.ng = ng;
    L.beta2 = (float)beta2;
    L.w1 = (float)(1.0 - beta1);
    L.w2 = (float)(1.0 - beta2);
    L.eps = (float)eps;
    adam_kernel<<<blocks, ADAM_THREADS, 0, (cudaStream_t)stream_>>>(L);
    CUN(cudaGetLastError(), "adam");
    return B200SPLAT_OK;
}

int b200splat_activate_forward(int32_t P, const float* scaling, const float* rotation, const float* opacity, float* scales,
                               float* rotations, float* opacities, void* stream_) {
    if (P < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "activate: bad P");
    if (P == 0) return B200SPLAT_OK;
    if (!scaling || !rotation || !opacity || !scales || !rotations || !opacities) return report_error(B200SPLAT_ERR_INVALID_ARG, "activate: null argument");
    if (!aligned16p(rotation) || !aligned16p(rotations)) return report_error(B200SPLAT_ERR_INVALID_ARG, "activate: rotations must be 16-byte aligned");
    activate_fwd_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, scaling, (const float4*)rotation, opacity, scales,
                                                                           (float4*)rotations, opacities);
    CUN(cudaGetLastError(), "activate forward");
    return B200SPLAT_OK;
}

int b200splat_activate_backward(int32_t P, const float* rotation, const float* scales, const float* opacities, const float* g_scales,
                                const float* g_rotations, const float* g_opacities, float* d_scaling, float* d_rotation,
                                float* d_opacity, void* stream_) {
    if (P < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "activate: bad P");
    if (P == 0) return B200SPLAT_OK;
    if ((g_scales && (!scales || !d_scaling)) || (g_rotations && (!rotation || !d_rotation)) || (g_opacities && (!opacities || !d_opacity)))
        return report_error(B200SPLAT_ERR_INVALID_ARG, "activate backward: null argument");
    if (g_rotations && (!aligned16p(rotation) || !aligned16p(g_rotations) || !aligned16p(d_rotation)))
        return report_error(B200SPLAT_ERR_INVALID_ARG, "activate backward: rotations must be 16-byte aligned");
    activate_bwd_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, (const float4*)rotation, scales, opacities, g_scales,
                                                                           (const float4*)g_rotations, g_opacities, d_scaling,
                                                                           (float4*)d_rotation, d_opacity);
    CUN(cudaGetLastError(), "activate backward");
    return B200SPLAT_OK;
}

int b200splat_dist2_knn3(int32_t P, const float* xyz, float* out, void* stream_) {
    if (P < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "knn: bad P");
    if (P == 0) return B200SPLAT_OK;
    if (!xyz || !out) return report_error(B200SPLAT_ERR_INVALID_ARG, "knn: null argument");
    cudaStream_t s = (cudaStream_t)stream_;
    const int n_boxes = (P + KBOX - 1) / KBOX, n_supers = (n_boxes + KSUPER - 1) / KSUPER;
    size_t sort_bytes = 0;
    CUN(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned*)nullptr, (unsigned*)nullptr,
                                        (const unsigned*)nullptr, (unsigned*)nullptr, P, 0, 30, s), "knn sort size");
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    size_t o_b6 = take(6 * sizeof(int)), o_codes = take(sizeof(unsigned) * P), o_codes2 = take(sizeof(unsigned) * P),
           o_ids = take(sizeof(unsigned) * P), o_ids2 = take(sizeof(unsigned) * P), o_pts = take(sizeof(float4) * (size_t)P),
           o_boxes = take(sizeof(Box) * (size_t)n_boxes), o_supers = take(sizeof(Box) * (size_t)n_supers),
           o_sort = take(sort_bytes);
    char* base = nullptr;
    CUN(cudaMallocAsync((void**)&base, off, s), "knn scratch");
    int* b6 = (int*)(base + o_b6);
    unsigned *codes = (unsigned*)(base + o_codes), *codes2 = (unsigned*)(base + o_codes2), *ids = (unsigned*)(base + o_ids),
             *ids2 = (unsigned*)(base + o_ids2);
    float4* pts = (float4*)(base + o_pts);
    Box *boxes = (Box*)(base + o_boxes), *supers = (Box*)(base + o_supers);
    cudaError_t e = cudaSuccess;
    do {
        knn_bounds_init_kernel<<<1, 32, 0, s>>>(b6);
        int nb = (P + 255) / 256;
        knn_bounds_kernel<<<nb < 1184 ? nb : 1184, 256, 0, s>>>(P, xyz, b6);
        knn_morton_kernel<<<nb, 256, 0, s>>>(P, xyz, b6, codes, ids);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cub::DeviceRadixSort::SortPairs(base + o_sort, sort_bytes, codes, codes2, ids, ids2, P, 0, 30, s)) != cudaSuccess) break;
        knn_gather_boxes_kernel<<<n_boxes, KBOX, 0, s>>>(P, xyz, ids2, pts, boxes);
        knn_super_boxes_kernel<<<n_supers, KSUPER, 0, s>>>(n_boxes, boxes, supers);
        knn_search_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, pts, ids2, n_boxes, boxes, n_supers, supers, out);
        e = cudaGetLastError();
    } while (0);
    cudaError_t e2 = cudaFreeAsync(base, s);
    CUN(e, "knn kernels");
    CUN(e2, "knn scratch free");
    return B200SPLAT_OK;
}

int b200splat_dist2_knn3_static(int32_t P, const float* xyz, float* out, void* stream_) {
    if (P < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "knn (static scratch): bad P");
    if (P > B200SPLAT_KNN_STATIC_MAX_POINTS) {
        std::string m = "knn (static scratch): " + std::to_string(P) + " points exceed the built-in scratch (maximum " +
                        std::to_string(B200SPLAT_KNN_STATIC_MAX_POINTS) + "); use b200splat_dist2_knn3";
        return report_error(B200SPLAT_ERR_INVALID_ARG, m.c_str());
    }
    if (P == 0) return B200SPLAT_OK;
    if (!xyz || !out) return report_error(B200SPLAT_ERR_INVALID_ARG, "knn (static scratch): null argument");
    cudaStream_t s = (cudaStream_t)stream_;
    const int n_boxes = (P + KBOX - 1) / KBOX, n_supers = (n_boxes + KSUPER - 1) / KSUPER;
    size_t sort_bytes = 0;      // size query only: with caller-provided temporary storage the sort allocates nothing
    CUN(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned*)nullptr, (unsigned*)nullptr,
                                        (const unsigned*)nullptr, (unsigned*)nullptr, P, 0, 30, s), "knn (static scratch) sort size");
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    size_t o_b6 = take(6 * sizeof(int)), o_codes = take(sizeof(unsigned) * P), o_codes2 = take(sizeof(unsigned) * P),
           o_ids = take(sizeof(unsigned) * P), o_ids2 = take(sizeof(unsigned) * P), o_pts = take(sizeof(float4) * (size_t)P),
           o_boxes = take(sizeof(Box) * (size_t)n_boxes), o_supers = take(sizeof(Box) * (size_t)n_supers),
           o_sort = take(sort_bytes);
    if (off > KNN_STATIC_BYTES) return report_error(B200SPLAT_ERR_INVALID_ARG, "knn (static scratch): scratch layout exceeds the built-in array");
    char* base = nullptr;
    CUN(knn_static_scratch(&base), "knn (static scratch) symbol address");
    int* b6 = (int*)(base + o_b6);
    unsigned *codes = (unsigned*)(base + o_codes), *codes2 = (unsigned*)(base + o_codes2), *ids = (unsigned*)(base + o_ids),
             *ids2 = (unsigned*)(base + o_ids2);
    float4* pts = (float4*)(base + o_pts);
    Box *boxes = (Box*)(base + o_boxes), *supers = (Box*)(base + o_supers);
    // every word of scratch the kernels read is written earlier in this same sequence (b6 by the init kernel): the previous
    // call's contents never matter
    knn_bounds_init_kernel<<<1, 32, 0, s>>>(b6);
    const int nb = (P + 255) / 256;
    knn_bounds_kernel<<<nb < 1184 ? nb : 1184, 256, 0, s>>>(P, xyz, b6);
    knn_morton_kernel<<<nb, 256, 0, s>>>(P, xyz, b6, codes, ids);
    CUN(cudaGetLastError(), "knn (static scratch) morton");
    CUN(cub::DeviceRadixSort::SortPairs(base + o_sort, sort_bytes, codes, codes2, ids, ids2, P, 0, 30, s), "knn (static scratch) sort");
    knn_gather_boxes_kernel<<<n_boxes, KBOX, 0, s>>>(P, xyz, ids2, pts, boxes);
    knn_super_boxes_kernel<<<n_supers, KSUPER, 0, s>>>(n_boxes, boxes, supers);
    knn_search_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, pts, ids2, n_boxes, boxes, n_supers, supers, out);
    CUN(cudaGetLastError(), "knn (static scratch) kernels");
    return B200SPLAT_OK;
}

}  // extern "C"
