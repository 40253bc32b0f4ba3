// render_tiles.cu -- tiled alpha compositing, forward and backward, sm_100a: NW WARPS PER 16x16 TILE.
//
// Stands in for the renderCUDA forward / backward stages of _C.rasterize_gaussians(_backward) named
// by BASELINE.json's north_star: front-to-back blending over 16x16 tiles with the 1/255 alpha cutoff,
// the 0.99 alpha clamp and the 1e-4 transmittance exit; reverse traversal with warp-level reduction
// of the per-Gaussian sums before any global atomic.
//
// Mapping (DESIGN.md "Render kernels").  One block of NW warps per tile (NW = 2: 64 threads); warp w owns the
// 16 x (16/NW) pixel strip of rows (16/NW) w .., and lane l the pixel (l & 7, l >> 3) of each 8x4 cell of it
// (8/NW pixels per lane).
//   * A tile's depth-sorted list is streamed through shared memory NW*32 packed 48-byte records at a
//     time (one record per thread, three 16-byte loads, the next batch prefetched into registers).
//   * Every record carries the bounding box of its {alpha >= 1/255} ellipse; the staging thread
//     reduces it to an NW-bit strip mask, a ballot turns the masks into one "reaches my strip"
//     bitmap per warp, and each warp walks ONLY the set bits of its bitmap (find-first-set), so
//     a splat that cannot touch a strip costs that warp nothing.  Results are identical: a skipped
//     pixel would have failed the exact alpha test.
//   * The pixels of a lane are evaluated in straight-line predicated code (independent chains).
//   * Backward: the per-Gaussian sums of a strip are finished with ONE transposed butterfly
//     (8 values over 32 lanes in 4+2+1+1+1 shuffles) and nine red.global.add per (strip, Gaussian)
//     instead of nine atomics per pixel.
// Tiles are issued longest list first (tile_order) so the heavy ones do not start last.  Measured and
// rejected (round 1): wider blocks for the heaviest tiles only (two launches side by side: slower, the
// stage is throughput- not critical-path-bound), per-cell skipping in the backward.
#include "splat_internal.cuh"

namespace b200splat {

namespace {

#ifndef FWD_NW
#define FWD_NW 2                       // warps per tile, forward   (strip = 16 x 16/NW pixels)
#endif
#ifndef BWD_NW
#define BWD_NW 2                       // warps per tile, backward
#endif
#ifndef FWD_PACKED
#define FWD_PACKED 1                   // forward power on packed fp32 pairs
#endif
#ifndef BWD_MINB
#define BWD_MINB 14                    // resident blocks per SM the backward is compiled for
#endif
#ifndef BWD_PACKED
#define BWD_PACKED 1                   // backward pixel arithmetic on packed fp32 pairs (FFMA2)
#endif
constexpr unsigned FULL = 0xffffffffu;

// NW-bit mask of the 16 x (16/NW) strips the box (rx +- ex, ry +- ey), tile-relative, can reach
template <int NW>
__device__ __forceinline__ unsigned strip_mask(float rx, float ry, float ex, float ey) {
    if (!(ex >= 0.0f)) return 0u;
    const float lox = rx - ex, hix = rx + ex, loy = ry - ey, hiy = ry + ey;
    if (!(lox <= 15.0f && hix >= 0.0f)) return 0u;
    constexpr int RH = 16 / NW;
    unsigned m = 0;
#pragma unroll
    for (int g = 0; g < NW; g++)
        m |= (loy <= (float)(g * RH + RH - 1) && hiy >= (float)(g * RH)) ? (1u << g) : 0u;
    return m;
}

// Pixel slots of a lane.  NW <= 4: lane -> (lane & 7, lane >> 3) inside an 8x4 cell, slot k -> cell (k & 1, k >> 1)
// of the strip; NW == 8: the strip is 16x2 and a lane owns the single pixel (lane & 15, lane >> 4).
template <int NW> struct Layout {
    static constexpr int NP = 8 / NW;            // pixels per thread
    static constexpr int NY = NP > 1 ? NP / 2 : 1;   // distinct rows per thread
    static constexpr int NX = NP > 1 ? 2 : 1;
    static constexpr int RH = 16 / NW;
    __device__ static __forceinline__ int x(int lane, int kx) { return NW == 8 ? (lane & 15) : (lane & 7) + 8 * kx; }
    __device__ static __forceinline__ int y(int lane, int warp, int ky) {
        return NW == 8 ? 2 * warp + (lane >> 4) : RH * warp + (lane >> 3) + 4 * ky;
    }
};

__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ float fast_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Canonical exponential (DESIGN.md "Canonical arithmetic"; oracle/splat_oracle.c implements the same sequence):
// the alpha >= 1/255 cutoff and the 1e-4 transmittance exit are discontinuous, so the forward's exp must agree
// with the oracle's to the bit -- libm-style expf implementations do not (a few flipped pixels per full-size view).
// Cephes-style: n = round(x log2 e) through the 1.5 * 2^23 trick, r = x - n ln 2 in two fma, degree-7 Horner,
// 2^n by adding n to the exponent field.  x must lie in [-87, 0]: callers mask lanes whose power is below the
// record's threshold (>= -80) instead of clamping.
__device__ __forceinline__ float canon_exp(float x) {
    const float t = __fmaf_rn(x, 1.4426950408889634f, 12582912.0f);
    const float n = __fsub_rn(t, 12582912.0f);
    float r = __fmaf_rn(n, -0.693359375f, x);
    r = __fmaf_rn(n, 2.12194440e-4f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    p = __fmaf_rn(p, r, 1.0f);
    p = __fmaf_rn(p, r, 1.0f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// Packed pairs of fp32 (sm_100 FFMA2 / FMUL2 / FADD2: two IEEE fp32 results per lane per issue slot).  The render
// kernels are bound by the issue rate, and the pixels of a lane run the same straight-line arithmetic, so the two
// pixels of a row go through one instruction stream.
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 mk2(float lo, float hi) {
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f2 bc2(float s) { return mk2(s, s); }
__device__ __forceinline__ float lo2(f2 a) {
    float l, h;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v));
    return l;
}
__device__ __forceinline__ float hi2(f2 a) {
    float l, h;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v));
    return h;
}
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
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    f2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
    v += __shfl_xor_sync(FULL, v, 16);
    v += __shfl_xor_sync(FULL, v, 8);
    v += __shfl_xor_sync(FULL, v, 4);
    v += __shfl_xor_sync(FULL, v, 2);
    v += __shfl_xor_sync(FULL, v, 1);
    return v;
}


template <int NW, bool DEPTH>
__global__ void __launch_bounds__(NW * 32) render_fwd_kernel(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
    const float* __restrict__ depths, int W, int H, int gx, const uint32_t* __restrict__ tile_order,
    const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ out_invdepth,
    float* __restrict__ final_T, uint32_t* __restrict__ n_contrib) {
    using L = Layout<NW>;
    constexpr int TB = NW * 32, NX = L::NX, NY = L::NY, NP = L::NP;
    __shared__ float4 s_rec[TB * 3];
    __shared__ float s_invd[DEPTH ? TB : 1];
    __shared__ unsigned s_bits[NW][NW];          // [consumer strip][staging warp]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order ? (int)__ldg(tile_order + blockIdx.x) : (int)blockIdx.x;
    const int2 range = __ldg(ranges + tile);
    const int n = range.y - range.x;
    const int ox = (tile % gx) * TILE, oy = (tile / gx) * TILE;
    const float fox = (float)ox, foy = (float)oy;
    int pxi[NX], pyi[NY];
    float pxf[NX], pyf[NY];
#pragma unroll
    for (int k = 0; k < NX; k++) { pxi[k] = ox + L::x(lane, k); pxf[k] = (float)pxi[k]; }
#pragma unroll
    for (int k = 0; k < NY; k++) { pyi[k] = oy + L::y(lane, warp, k); pyf[k] = (float)pyi[k]; }
    float T[NP], C0[NP], C1[NP], C2[NP], Dp[NP];
    uint32_t last[NP];
    bool any_in = false;
#pragma unroll
    for (int s = 0; s < NP; s++) {
        const bool inside = pxi[s % NX] < W && pyi[s / NX] < H;
        T[s] = inside ? 1.0f : -1.0f;                 // T < 0: pixel finished, |T| is final
        C0[s] = 0.f; C1[s] = 0.f; C2[s] = 0.f; Dp[s] = 0.f; last[s] = 0;
        any_in |= inside;
    }
    bool warp_live = __any_sync(FULL, any_in);
    const f2 npx2 = mk2(-pxf[0], -pxf[NX - 1]), npy2 = mk2(-pyf[0], -pyf[NY - 1]);

    float4 ra = make_float4(0.f, 0.f, -1.f, -1.f), rb = ra, rc = ra;
    float rinvd = 0.f;
    if (tid < n) {
        const uint32_t id = __ldg(point_list + range.x + tid);
        ra = __ldg(rec + 3 * (size_t)id); rb = __ldg(rec + 3 * (size_t)id + 1); rc = __ldg(rec + 3 * (size_t)id + 2);
        if (DEPTH) rinvd = __frcp_rn(__ldg(depths + id));
    }
    for (int base = 0; base < n; base += TB) {
        // all strips finished -> the rest of the list cannot change anything
        if (__syncthreads_and(!warp_live)) break;
        s_rec[tid * 3 + 0] = ra; s_rec[tid * 3 + 1] = rb; s_rec[tid * 3 + 2] = rc;
        if (DEPTH) s_invd[tid] = rinvd;
        const unsigned m = (base + tid < n) ? strip_mask<NW>(ra.x - fox, ra.y - foy, ra.z, ra.w) : 0u;
#pragma unroll
        for (int g = 0; g < NW; g++) {
            const unsigned bits = __ballot_sync(FULL, (m >> g) & 1u);
            if (lane == 0) s_bits[g][warp] = bits;
        }
        __syncthreads();
        {   // prefetch the next batch while this one is consumed
            const int nidx = base + TB + tid;
            if (nidx < n) {
                const uint32_t id = __ldg(point_list + range.x + nidx);
                ra = __ldg(rec + 3 * (size_t)id); rb = __ldg(rec + 3 * (size_t)id + 1); rc = __ldg(rec + 3 * (size_t)id + 2);
                if (DEPTH) rinvd = __frcp_rn(__ldg(depths + id));
            }
        }
        if (warp_live) {
#pragma unroll 1
            for (int sw = 0; sw < NW; sw++) {
                unsigned bits = s_bits[warp][sw];
                while (bits) {
                    const int j = sw * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float4 a = s_rec[3 * j], b = s_rec[3 * j + 1], c = s_rec[3 * j + 2];
                    float pw[NP];
                    float pmax = -3.0e38f;
                    if constexpr (FWD_PACKED && NP == 4) {
                        // The canonical power, two pixels per instruction (FFMA2): every packed operation is the same
                        // IEEE operation per half as its scalar twin below (a - p as a + (-p), -(b dy) as (-b) dy),
                        // and no packed multiply feeds a packed add, so nothing can be contracted differently.
                        const f2 dx2 = add2(bc2(a.x), npx2), dy2 = add2(bc2(a.y), npy2);
                        const f2 nbdy = mul2(bc2(-b.y), dy2);
                        const f2 t22 = mul2(mul2(bc2(b.z), dy2), dy2);
                        const f2 p0 = fma2(dx2, fma2(bc2(b.x), dx2, bc2(lo2(nbdy))), bc2(lo2(t22)));
                        const f2 p1 = fma2(dx2, fma2(bc2(b.x), dx2, bc2(hi2(nbdy))), bc2(hi2(t22)));
                        pw[0] = lo2(p0); pw[1] = hi2(p0); pw[NP - 2] = lo2(p1); pw[NP - 1] = hi2(p1);
                        pmax = fmaxf(fmaxf(pw[0], pw[1]), fmaxf(pw[NP - 2], pw[NP - 1]));
                    } else {
                        float dx[NX];
#pragma unroll
                        for (int k = 0; k < NX; k++) dx[k] = csub(a.x, pxf[k]);
#pragma unroll
                        for (int ky = 0; ky < NY; ky++) {
                            const float dy = csub(a.y, pyf[ky]);
                            const float bdy = cmul(b.y, dy);
                            const float t2 = cmul(cmul(b.z, dy), dy);
#pragma unroll
                            for (int kx = 0; kx < NX; kx++) {
                                const float p = cfma(dx[kx], cfma(b.x, dx[kx], -bdy), t2);
                                pw[ky * NX + kx] = p;
                                pmax = fmaxf(pmax, p);
                            }
                        }
                    }
                    // below thr no pixel of the strip can reach alpha >= 1/255: skip the exps
                    if (!__any_sync(FULL, pmax >= b.w)) continue;
                    const uint32_t contributor = (uint32_t)(base + j + 1);
                    const float invd = DEPTH ? s_invd[j] : 0.f;
#pragma unroll
                    for (int s = 0; s < NP; s++) {
                        // A slot is one 8x4 cell of the strip; a cell none of whose pixels can change (power outside
                        // [thr, 0], or pixel finished) is skipped by the whole warp: most splats cover one or two of
                        // a strip's cells.  (Measured: -9 % forward; the backward, tried the same way, gains nothing.)
                        if (NP > 1 && !__any_sync(FULL, pw[s] >= b.w && pw[s] <= 0.0f && T[s] > 0.0f)) continue;
                        // exact tests: power <= 0, alpha >= 1/255, T (1 - alpha) >= 1e-4; a finished pixel has T < 0
                        const float alpha = fminf(ALPHA_MAX, cmul(c.x, canon_exp(pw[s])));
                        const float test_T = cmul(T[s], csub(1.0f, alpha));
                        // power below the threshold: alpha < 1/255 for certain, and canon_exp's argument range ends there
                        const bool valid = alpha >= ALPHA_MIN && pw[s] <= 0.0f && pw[s] >= b.w;
                        const bool blend = valid && !(test_T < T_EPS);
                        const float w = blend ? cmul(alpha, T[s]) : 0.0f;     // adding c * 0 leaves the sums unchanged
                        C0[s] = cfma(c.y, w, C0[s]);
                        C1[s] = cfma(c.z, w, C1[s]);
                        C2[s] = cfma(c.w, w, C2[s]);
                        if (DEPTH) Dp[s] = cfma(invd, w, Dp[s]);
                        T[s] = blend ? test_T : (valid ? -fabsf(T[s]) : T[s]);
                        last[s] = blend ? contributor : last[s];
                    }
                }
            }
            bool live = false;
#pragma unroll
            for (int s = 0; s < NP; s++) live |= T[s] > 0.0f;
            warp_live = __any_sync(FULL, live);
        }
    }
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    const size_t HW = (size_t)H * W;
#pragma unroll
    for (int s = 0; s < NP; s++) {
        const int px = pxi[s % NX], py = pyi[s / NX];
        if (px < W && py < H) {
            const size_t pix = (size_t)py * W + px;
            const float Tf = fabsf(T[s]);
            final_T[pix] = Tf;
            n_contrib[pix] = last[s];
            out_color[pix] = cfma(Tf, bg0, C0[s]);
            out_color[HW + pix] = cfma(Tf, bg1, C1[s]);
            out_color[2 * HW + pix] = cfma(Tf, bg2, C2[s]);
            if (DEPTH) out_invdepth[pix] = Dp[s];
        }
    }
}

// Backward: reverse traversal.  Per pixel the running state is (T_i, S_i) with
//   S_i = sum_{j behind i} T_j alpha_j (c_j . g) + T_final (bg . g),   g = dL/dC(pixel)
//   dL/dalpha_i = T_i (c_i . g) - S_i / (1 - alpha_i)
// and per (strip, Gaussian) the pixel sums are the moments of z = G * o * dL/dalpha:
//   m0 = sum z, mx = sum z dx, my = sum z dy, mxx = sum z dx^2, mxy = sum z dx dy, myy = sum z dy^2
//   dL/dpx = -(A mx + B my), dL/dpy = -(C my + B mx), dL/dA = -mxx/2, dL/dB = -mxy, dL/dC = -myy/2,
//   dL/do = m0 / o, dL/dcolour = sum alpha T g.
// DEPTH: the inverse-depth image is a fourth blended channel (value 1/depth_i, no background term) with its
// own upstream gradient; it adds invd_i * gD to c_i . g and the sum of alpha T gD to acc[9].
template <int NW, bool DEPTH>
__global__ void __launch_bounds__(NW * 32, BWD_MINB) render_bwd_kernel(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
    const float* __restrict__ depths, int W, int H, int gx, const uint32_t* __restrict__ tile_order,
    const float* __restrict__ bg, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvd, float* __restrict__ acc,
    uint8_t* __restrict__ touched) {
    using L = Layout<NW>;
    constexpr int TB = NW * 32, NX = L::NX, NY = L::NY, NP = L::NP;
    __shared__ float4 s_rec[TB * 3];
    __shared__ uint32_t s_id[TB];
    __shared__ float s_invd[DEPTH ? TB : 1];
    __shared__ unsigned s_bits[NW][NW];
    __shared__ int s_wmax[NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = tile_order ? (int)__ldg(tile_order + blockIdx.x) : (int)blockIdx.x;
    const int2 range = __ldg(ranges + tile);
    const int ox = (tile % gx) * TILE, oy = (tile / gx) * TILE;
    const float fox = (float)ox, foy = (float)oy;
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    const size_t HW = (size_t)H * W;
    float pxf[NX], pyf[NY];
#pragma unroll
    for (int k = 0; k < NX; k++) pxf[k] = (float)(ox + L::x(lane, k));
#pragma unroll
    for (int k = 0; k < NY; k++) pyf[k] = (float)(oy + L::y(lane, warp, k));
    float T[NP], S[NP], g0[NP], g1[NP], g2[NP], gd[DEPTH ? NP : 1];
    int last[NP];
    int lmax = 0;
#pragma unroll
    for (int s = 0; s < NP; s++) {
        const int px = ox + L::x(lane, s % NX), py = oy + L::y(lane, warp, s / NX);
        T[s] = 1.f; S[s] = 0.f; g0[s] = 0.f; g1[s] = 0.f; g2[s] = 0.f; last[s] = 0;
        if (DEPTH) gd[s] = 0.f;
        if (px < W && py < H) {
            const size_t pix = (size_t)py * W + px;
            T[s] = __ldg(final_T + pix);
            last[s] = (int)__ldg(n_contrib + pix);
            g0[s] = __ldg(dL_dpix + pix); g1[s] = __ldg(dL_dpix + HW + pix); g2[s] = __ldg(dL_dpix + 2 * HW + pix);
            S[s] = T[s] * (bg0 * g0[s] + bg1 * g1[s] + bg2 * g2[s]);
            if (DEPTH) gd[s] = __ldg(dL_dinvd + pix);
        }
        lmax = max(lmax, last[s]);
    }
    const int wmax = __reduce_max_sync(FULL, lmax);   // deepest contributor of this strip
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    int maxc = 0;
#pragma unroll
    for (int g = 0; g < NW; g++) maxc = max(maxc, s_wmax[g]);
    const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4;
    // packed state (PK: NW == 2, pair ky = the two pixels (kx = 0, 1) of row ky): T, -S, the upstream gradients
    constexpr bool PK = BWD_PACKED && NP == 4;
    f2 T2[2], Sn2[2], g0p[2], g1p[2], g2p[2], gdp[2], npx2, npy2;
    if (PK) {
#pragma unroll
        for (int ky = 0; ky < 2; ky++) {
            T2[ky] = mk2(T[2 * ky], T[2 * ky + 1]); Sn2[ky] = mk2(-S[2 * ky], -S[2 * ky + 1]);
            g0p[ky] = mk2(g0[2 * ky], g0[2 * ky + 1]); g1p[ky] = mk2(g1[2 * ky], g1[2 * ky + 1]);
            g2p[ky] = mk2(g2[2 * ky], g2[2 * ky + 1]);
            gdp[ky] = DEPTH ? mk2(gd[2 * ky], gd[2 * ky + 1]) : bc2(0.f);
        }
        npx2 = mk2(-pxf[0], -pxf[NX - 1]); npy2 = mk2(-pyf[0], -pyf[NY - 1]);
    }

    float4 ra = make_float4(0.f, 0.f, -1.f, -1.f), rb = ra, rc = ra;
    uint32_t rid = 0;
    float rinvd = 0.f;
    {
        const int idx = max(0, maxc - TB) + tid;
        if (idx < maxc) {
            rid = __ldg(point_list + range.x + idx);
            ra = __ldg(rec + 3 * (size_t)rid); rb = __ldg(rec + 3 * (size_t)rid + 1); rc = __ldg(rec + 3 * (size_t)rid + 2);
            if (DEPTH) rinvd = __frcp_rn(__ldg(depths + rid));
        }
    }
    for (int hi = maxc; hi > 0; hi -= TB) {
        const int lo = max(0, hi - TB);
        const int cnt = hi - lo;
        __syncthreads();                         // previous batch fully consumed
        s_rec[tid * 3 + 0] = ra; s_rec[tid * 3 + 1] = rb; s_rec[tid * 3 + 2] = rc;
        s_id[tid] = rid;
        if (DEPTH) s_invd[tid] = rinvd;
        const unsigned m = tid < cnt ? strip_mask<NW>(ra.x - fox, ra.y - foy, ra.z, ra.w) : 0u;
#pragma unroll
        for (int g = 0; g < NW; g++) {
            const unsigned bits = __ballot_sync(FULL, (m >> g) & 1u);
            if (lane == 0) s_bits[g][warp] = bits;
        }
        __syncthreads();
        {   // prefetch the next (shallower) batch
            const int nhi = lo, nlo = max(0, nhi - TB);
            const int idx = nlo + tid;
            if (idx < nhi) {
                rid = __ldg(point_list + range.x + idx);
                ra = __ldg(rec + 3 * (size_t)rid); rb = __ldg(rec + 3 * (size_t)rid + 1); rc = __ldg(rec + 3 * (size_t)rid + 2);
                if (DEPTH) rinvd = __frcp_rn(__ldg(depths + rid));
            }
        }
        if (lo >= wmax) continue;                // this batch lies entirely behind this strip's deepest contributor
#pragma unroll 1
        for (int sw = NW - 1; sw >= 0; sw--) {
            unsigned bits = s_bits[warp][sw];
            const int wlo = lo + sw * 32;
            if (wlo >= wmax) continue;
            if (wmax - wlo < 32) bits &= (1u << (wmax - wlo)) - 1u;
            while (bits) {
                const int jb = 31 - __clz(bits);
                bits ^= 1u << jb;
                const int j = sw * 32 + jb;
                const int pos = lo + j;
                const float4 a = s_rec[3 * j], b = s_rec[3 * j + 1], c = s_rec[3 * j + 2];
                float m0, mx, my, mxx, mxy, myy, c0, c1, c2, cd = 0.f;
                bool any_valid = false;
                if constexpr (PK) {
                    // the same arithmetic as the scalar path below, two pixels per instruction; alpha, S and every sum
                    // are carried NEGATED so that no separate negation is ever needed (the signs return in the
                    // per-lane maps).  An invalid pixel has -alpha = 0: 1 / (1 + 0) = 1, T and S unchanged, z = 0.
                    const f2 dx2 = add2(bc2(a.x), npx2), dy2 = add2(bc2(a.y), npy2);
                    const f2 nbdy = mul2(bc2(-b.y), dy2);
                    const f2 t22 = mul2(mul2(bc2(b.z), dy2), dy2);
                    f2 p2[2];
                    p2[0] = fma2(dx2, fma2(bc2(b.x), dx2, bc2(lo2(nbdy))), bc2(lo2(t22)));
                    p2[1] = fma2(dx2, fma2(bc2(b.x), dx2, bc2(hi2(nbdy))), bc2(hi2(t22)));
                    const float pw[4] = {lo2(p2[0]), hi2(p2[0]), lo2(p2[1]), hi2(p2[1])};
                    bool ok[4], any_ok = false;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        ok[q] = pw[q] <= 0.0f && pw[q] >= b.w && pos < last[q];
                        any_ok |= ok[q];
                    }
                    if (!__any_sync(FULL, any_ok)) continue;
                    const float invd = DEPTH ? s_invd[j] : 0.f;
                    f2 nm0, nmx, nmy, nmxx, nmxy, nmyy, nc0, nc1, nc2, ncd = bc2(0.f);
#pragma unroll
                    for (int ky = 0; ky < 2; ky++) {
                        const f2 e2 = mul2(p2[ky], bc2(1.4426950408889634f));
                        const f2 nau0 = mul2(bc2(-c.x), mk2(fast_ex2(lo2(e2)), fast_ex2(hi2(e2))));
                        const bool v0 = ok[2 * ky] && lo2(nau0) <= -ALPHA_MIN, v1 = ok[2 * ky + 1] && hi2(nau0) <= -ALPHA_MIN;
                        const float na0 = v0 ? lo2(nau0) : 0.0f, na1 = v1 ? hi2(nau0) : 0.0f;
                        const f2 nau = mk2(na0, na1);
                        const f2 nal = mk2(fmaxf(-ALPHA_MAX, na0), fmaxf(-ALPHA_MAX, na1));
                        const f2 oma = add2(bc2(1.0f), nal);
                        const f2 r = mk2(fast_rcp(lo2(oma)), fast_rcp(hi2(oma)));
                        const f2 Ti = mul2(T2[ky], r);
                        f2 cg = fma2(bc2(c.w), g2p[ky], fma2(bc2(c.z), g1p[ky], mul2(bc2(c.y), g0p[ky])));
                        if (DEPTH) cg = fma2(bc2(invd), gdp[ky], cg);
                        const f2 dLda = fma2(Sn2[ky], r, mul2(Ti, cg));
                        const f2 nw = mul2(nal, Ti);
                        Sn2[ky] = fma2(nw, cg, Sn2[ky]);
                        T2[ky] = Ti;
                        const f2 nz = mul2(nau, dLda);
                        const f2 dyb = bc2(ky == 0 ? lo2(dy2) : hi2(dy2));
                        const f2 nzx = mul2(nz, dx2), nzy = mul2(nz, dyb);
                        if (ky == 0) {
                            nc0 = mul2(nw, g0p[0]); nc1 = mul2(nw, g1p[0]); nc2 = mul2(nw, g2p[0]);
                            if (DEPTH) ncd = mul2(nw, gdp[0]);
                            nm0 = nz; nmx = nzx; nmy = nzy;
                            nmxx = mul2(nzx, dx2); nmxy = mul2(nzx, dyb); nmyy = mul2(nzy, dyb);
                        } else {
                            nc0 = fma2(nw, g0p[1], nc0); nc1 = fma2(nw, g1p[1], nc1); nc2 = fma2(nw, g2p[1], nc2);
                            if (DEPTH) ncd = fma2(nw, gdp[1], ncd);
                            nm0 = add2(nm0, nz); nmx = add2(nmx, nzx); nmy = add2(nmy, nzy);
                            nmxx = fma2(nzx, dx2, nmxx); nmxy = fma2(nzx, dyb, nmxy); nmyy = fma2(nzy, dyb, nmyy);
                        }
                        any_valid |= v0 | v1;
                    }
                    m0 = -lo2(nm0) - hi2(nm0); mx = -lo2(nmx) - hi2(nmx); my = -lo2(nmy) - hi2(nmy);
                    mxx = -lo2(nmxx) - hi2(nmxx); mxy = -lo2(nmxy) - hi2(nmxy); myy = -lo2(nmyy) - hi2(nmyy);
                    c0 = -lo2(nc0) - hi2(nc0); c1 = -lo2(nc1) - hi2(nc1); c2 = -lo2(nc2) - hi2(nc2);
                    if (DEPTH) cd = -lo2(ncd) - hi2(ncd);
                } else {
                float dx[NX], dyv[NY], pw[NP];
                bool ok[NP], any_ok = false;
#pragma unroll
                for (int k = 0; k < NX; k++) dx[k] = csub(a.x, pxf[k]);
#pragma unroll
                for (int ky = 0; ky < NY; ky++) {
                    const float dy = csub(a.y, pyf[ky]);
                    dyv[ky] = dy;
                    const float bdy = cmul(b.y, dy);
                    const float t2 = cmul(cmul(b.z, dy), dy);
#pragma unroll
                    for (int kx = 0; kx < NX; kx++) {
                        const int s = ky * NX + kx;
                        const float p = cfma(dx[kx], cfma(b.x, dx[kx], -bdy), t2);
                        pw[s] = p;
                        ok[s] = p <= 0.0f && p >= b.w && pos < last[s];
                        any_ok |= ok[s];
                    }
                }
                if (!__any_sync(FULL, any_ok)) continue;
                m0 = 0.f; mx = 0.f; my = 0.f; mxx = 0.f; mxy = 0.f; myy = 0.f; c0 = 0.f; c1 = 0.f; c2 = 0.f;
                const float invd = DEPTH ? s_invd[j] : 0.f;
#pragma unroll
                for (int s = 0; s < NP; s++) {
                    // au = o G (unclamped alpha), gated ONCE: with au = 0 every term below is an exact no-op
                    // (alpha = 0, 1 / (1 - 0) = 1, T unchanged, w = 0, z = 0)
                    // fast exp (ex2.approx, <= ~6 ulp for power in [-5.6, 0]): the backward has no bit-exact gate, and
                    // the rare pixel whose alpha straddles 1/255 differently from the forward moves its gradient
                    // terms by at most alpha = 0.4 %
                    const float au0 = cmul(c.x, __expf(pw[s]));
                    const bool valid = ok[s] && fminf(ALPHA_MAX, au0) >= ALPHA_MIN;
                    const float au = valid ? au0 : 0.0f;
                    const float alpha = fminf(ALPHA_MAX, au);
                    const float r = fast_rcp(1.0f - alpha);
                    const float Ti = T[s] * r;
                    float cg = c.y * g0[s] + c.z * g1[s] + c.w * g2[s];
                    if (DEPTH) cg = fmaf(invd, gd[s], cg);
                    const float w = alpha * Ti;
                    const float dLda = Ti * cg - S[s] * r;
                    const float z = au * dLda;
                    T[s] = Ti;
                    S[s] = fmaf(w, cg, S[s]);
                    c0 = fmaf(w, g0[s], c0); c1 = fmaf(w, g1[s], c1); c2 = fmaf(w, g2[s], c2);
                    if (DEPTH) cd = fmaf(w, gd[s], cd);
                    const float ddx = dx[s % NX], ddy = dyv[s / NX];
                    const float zx = z * ddx, zy = z * ddy;
                    m0 += z; mx += zx; my += zy;
                    mxx = fmaf(zx, ddx, mxx); mxy = fmaf(zx, ddy, mxy); myy = fmaf(zy, ddy, myy);
                    any_valid |= valid;
                }
                }
                if (!__any_sync(FULL, any_valid)) continue;
                // per-lane linear maps first (they commute with the sum over lanes), then ONE transposed
                // butterfly; lane L ends up holding the strip total of value L >> 2, the ninth value takes
                // a plain butterfly.
                const float A = -2.0f * b.x, B = b.y, C = -2.0f * b.z;
                const float q0 = -(A * mx + B * my), q1 = -(C * my + B * mx);
                const float q2 = -0.5f * mxx, q3 = -mxy, q4 = -0.5f * myy, q5 = m0 * fast_rcp(c.x);
                float w0 = (up16 ? q4 : q0) + __shfl_xor_sync(FULL, up16 ? q0 : q4, 16);
                float w1 = (up16 ? q5 : q1) + __shfl_xor_sync(FULL, up16 ? q1 : q5, 16);
                float w2 = (up16 ? c0 : q2) + __shfl_xor_sync(FULL, up16 ? q2 : c0, 16);
                float w3 = (up16 ? c1 : q3) + __shfl_xor_sync(FULL, up16 ? q3 : c1, 16);
                float u0 = (up8 ? w2 : w0) + __shfl_xor_sync(FULL, up8 ? w0 : w2, 8);
                float u1 = (up8 ? w3 : w1) + __shfl_xor_sync(FULL, up8 ? w1 : w3, 8);
                float t0 = (up4 ? u1 : u0) + __shfl_xor_sync(FULL, up4 ? u0 : u1, 4);
                t0 += __shfl_xor_sync(FULL, t0, 2);
                t0 += __shfl_xor_sync(FULL, t0, 1);
                c2 = warp_sum(c2);
                if (DEPTH) cd = warp_sum(cd);
                // component slot of acc held by lane L (L & 3 == 0): L >> 2 ->
                //   0 dpx  1 dpy  2 dA  3 dB  4 dC  5 do  6 dr  7 dg ; slot 8 (db) from lane 1
                float* dst = acc + 12 * (size_t)s_id[j];
                if ((lane & 3) == 0) atomicAdd(dst + (lane >> 2), t0);
                else if (lane == 1) atomicAdd(dst + 8, c2);
                else if (DEPTH && lane == 2) atomicAdd(dst + 9, cd);
                else if (lane == 3) touched[s_id[j]] = 1;      // this Gaussian takes part in the per-Gaussian backward
            }
        }
    }
}

// Inspection only (b200splat_render_stats): workload counters of the render stage of one view, used for
// the FP32-issue roofline -- not on the hot path.  One block per tile, one thread per pixel.
//   out[0..4]  pairs by number of 16x4 strips their {alpha >= 1/255} box reaches (0 = skipped outright)
//   out[9]     pairs visited (list positions below the tile's deepest contributor)
//   out[10]    pixel evaluations (pixels of reached strips, position below the pixel's last contributor)
//   out[11]    of those, power inside [thr, 0] (an exp is executed)
//   out[12]    of those, alpha >= 1/255 (a blend / a gradient contribution)
//   out[13]    (strip, Gaussian) pairs with at least one contribution (one reduction + nine red.global)
//   out[14]    of the out[11] evaluations, those where the backward's alpha test (ex2.approx of power * log2 e, as
//              render_bwd_kernel computes it) decides differently from the forward's (expf): a count, not a gate
__global__ void __launch_bounds__(256) render_stats_kernel(const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                           const float4* __restrict__ rec, int W, int H, int gx,
                                                           const uint32_t* __restrict__ n_contrib,
                                                           unsigned long long* __restrict__ out) {
    __shared__ int s_max;
    __shared__ unsigned long long s_cnt[16];
    __shared__ int s_strip[4];
    const int tile = blockIdx.x, t = threadIdx.x;
    const int ox = (tile % gx) * TILE, oy = (tile / gx) * TILE;
    const int px = ox + (t & 15), py = oy + (t >> 4);
    const bool inside = px < W && py < H;
    const int last = inside ? (int)n_contrib[(size_t)py * W + px] : 0;
    if (t == 0) s_max = 0;
    if (t < 16) s_cnt[t] = 0ull;
    __syncthreads();
    atomicMax(&s_max, last);
    __syncthreads();
    const int2 range = ranges[tile];
    const int n = min(range.y - range.x, s_max);
    const int strip = t >> 6;                     // rows 4*strip .. 4*strip+3
    unsigned long long ev = 0, ex = 0, co = 0, dis = 0;
    for (int j = 0; j < n; j++) {
        const uint32_t id = point_list[range.x + j];
        const float4 a = rec[3 * (size_t)id], b = rec[3 * (size_t)id + 1], c = rec[3 * (size_t)id + 2];
        const unsigned m = strip_mask<4>(a.x - (float)ox, a.y - (float)oy, a.z, a.w);
        bool contrib = false;
        if (t < 4) s_strip[t] = 0;
        __syncthreads();
        if (((m >> strip) & 1u) && j < last) {
            ev++;
            const float dx = csub(a.x, (float)px), dy = csub(a.y, (float)py);
            const float power = cfma(dx, cfma(b.x, dx, -cmul(b.y, dy)), cmul(cmul(b.z, dy), dy));
            if (power <= 0.0f && power >= b.w) {
                ex++;
                const bool fwd_valid = fminf(ALPHA_MAX, cmul(c.x, canon_exp(power))) >= ALPHA_MIN;
                const bool bwd_valid = -c.x * fast_ex2(power * 1.4426950408889634f) <= -ALPHA_MIN;
                if (fwd_valid) { co++; contrib = true; }
                if (fwd_valid != bwd_valid) dis++;
            }
        }
        if (contrib) s_strip[strip] = 1;
        __syncthreads();
        if (t == 0) { s_cnt[__popc(m)]++; s_cnt[9]++; s_cnt[13] += s_strip[0] + s_strip[1] + s_strip[2] + s_strip[3]; }
        __syncthreads();
    }
    atomicAdd(&s_cnt[10], ev); atomicAdd(&s_cnt[11], ex); atomicAdd(&s_cnt[12], co); atomicAdd(&s_cnt[14], dis);
    __syncthreads();
    if (t < 16 && s_cnt[t]) atomicAdd(out + t, s_cnt[t]);
}

}  // namespace

cudaError_t launch_render_fwd(const b200splat_view& v, GeomView gv, BinView bv, ImgView iv, float* out_color,
                              float* out_invdepth, cudaStream_t s) {
    const int W = v.image_width, H = v.image_height;
    const int gx = (W + TILE - 1) / TILE;
    const int T = tiles_of(W, H);
    if (T == 0) return cudaSuccess;
    if (out_invdepth)
        render_fwd_kernel<FWD_NW, true><<<T, FWD_NW * 32, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, gx, iv.tile_order, v.bg,
                                                 out_color, out_invdepth, iv.final_T, iv.n_contrib);
    else
        render_fwd_kernel<FWD_NW, false><<<T, FWD_NW * 32, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, gx, iv.tile_order, v.bg,
                                                  out_color, nullptr, iv.final_T, iv.n_contrib);
    return cudaGetLastError();
}

cudaError_t launch_render_bwd(const b200splat_view& v, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                              const float* dL_dpix, const float* dL_dinvdepth, float* acc, uint8_t* touched, cudaStream_t s) {
    const int W = v.image_width, H = v.image_height;
    const int gx = (W + TILE - 1) / TILE;
    const int T = tiles_of(W, H);
    if (T == 0) return cudaSuccess;
    if (dL_dinvdepth)
        render_bwd_kernel<BWD_NW, true><<<T, BWD_NW * 32, 0, s>>>(iv.ranges, vals_sorted, gv.rec, gv.depths, W, H, gx, iv.tile_order,
                                                                 v.bg, iv.final_T, iv.n_contrib, dL_dpix, dL_dinvdepth, acc, touched);
    else
        render_bwd_kernel<BWD_NW, false><<<T, BWD_NW * 32, 0, s>>>(iv.ranges, vals_sorted, gv.rec, gv.depths, W, H, gx, iv.tile_order,
                                                                  v.bg, iv.final_T, iv.n_contrib, dL_dpix, nullptr, acc, touched);
    return cudaGetLastError();
}

cudaError_t launch_render_stats(int W, int H, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                                unsigned long long* out16, cudaStream_t s) {
    const int gx = (W + TILE - 1) / TILE;
    const int T = tiles_of(W, H);
    cudaError_t e = cudaMemsetAsync(out16, 0, 16 * sizeof(unsigned long long), s);
    if (e != cudaSuccess || T == 0) return e;
    render_stats_kernel<<<T, 256, 0, s>>>(iv.ranges, vals_sorted, gv.rec, W, H, gx, iv.n_contrib, out16);
    return cudaGetLastError();
}

}  // namespace b200splat
