// render.cu -- tiled alpha compositing, forward and backward, sm_100a.
//
// Stands in for the renderCUDA forward / backward stages of _C.rasterize_gaussians(_backward) named
// by BASELINE.json's north_star: front-to-back blending over 16x16 tiles with the 1/255 alpha cutoff,
// the 0.99 alpha clamp and the 1e-4 transmittance exit; reverse traversal with warp-level reduction
// of the per-Gaussian sums before any global atomic.
//
// B200 mapping (DESIGN.md "Render kernels"): ONE WARP OWNS ONE 16x16 TILE, eight pixels per thread.
// Thread `lane` holds the pixel at (lane & 7, lane >> 3) of each of the eight 8x4 sub-blocks of the
// tile.  Consequences:
//   * no block-wide barrier anywhere: a warp streams its tile's depth-sorted list through a private
//     1.5 KB shared-memory stage (32 packed 48-byte records per step, three 16-byte loads per lane,
//     next step prefetched into registers while the current one is consumed);
//   * every record carries the bounding box of its {alpha >= 1/255} ellipse, so the staging lane
//     reduces it to an 8-bit sub-block mask and the warp only evaluates the sub-blocks the splat can
//     reach (results are identical: a skipped pixel would have failed the exact alpha test);
//   * the exp is skipped when power < thr (same guarantee);
//   * in the backward pass the per-Gaussian sums of a whole tile are finished with ONE butterfly
//     reduction and ONE 9-lane red.global per (tile, Gaussian), instead of 9 atomics per pixel.
// Tiles are handed out dynamically (atomic counter) to persistent warps.
#include "splat_internal.cuh"

namespace b200splat {

namespace {

constexpr int RW = 4;                 // warps per block
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned subblock_mask(float rx, float ry, float ex, float ey) {
    if (!(ex >= 0.0f)) return 0u;
    const float lox = rx - ex, hix = rx + ex, loy = ry - ey, hiy = ry + ey;
    const unsigned mx = ((lox <= 7.0f && hix >= 0.0f) ? 1u : 0u) | ((lox <= 15.0f && hix >= 8.0f) ? 2u : 0u);
    const unsigned my = ((loy <= 3.0f && hiy >= 0.0f) ? 1u : 0u) | ((loy <= 7.0f && hiy >= 4.0f) ? 4u : 0u) |
                        ((loy <= 11.0f && hiy >= 8.0f) ? 16u : 0u) | ((loy <= 15.0f && hiy >= 12.0f) ? 64u : 0u);
    return mx * my;   // bit (2*sy + sx)
}

template <bool DEPTH>
__global__ void __launch_bounds__(RW * 32) render_fwd_kernel(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
    const float* __restrict__ depths, int W, int H, int gx, int num_tiles, const float* __restrict__ bg,
    float* __restrict__ out_color, float* __restrict__ out_invdepth, float* __restrict__ final_T,
    uint32_t* __restrict__ n_contrib, unsigned int* __restrict__ tile_counter) {
    __shared__ float4 s_rec[RW][32 * 3];
    __shared__ float s_invd[RW][32];
    __shared__ unsigned s_mask[RW][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lx = lane & 7, ly = lane >> 3;
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    const size_t HW = (size_t)H * W;

    while (true) {
        int tile = 0;
        if (lane == 0) tile = (int)atomicAdd(tile_counter, 1u);
        tile = __shfl_sync(FULL, tile, 0);
        if (tile >= num_tiles) break;

        const int2 range = __ldg(ranges + tile);
        const int n = range.y - range.x;
        const int ox = (tile % gx) * TILE, oy = (tile / gx) * TILE;
        const float fox = (float)ox, foy = (float)oy;
        const float pxf[2] = {(float)(ox + lx), (float)(ox + 8 + lx)};
        const float pyf[4] = {(float)(oy + ly), (float)(oy + 4 + ly), (float)(oy + 8 + ly), (float)(oy + 12 + ly)};
        float T[8], C0[8], C1[8], C2[8], Dp[8];
        uint32_t last[8];
        unsigned live = 0;
#pragma unroll
        for (int s = 0; s < 8; s++) {
            const bool inside = (ox + (s & 1) * 8 + lx < W) && (oy + (s >> 1) * 4 + ly < H);
            T[s] = inside ? 1.0f : -1.0f;
            C0[s] = 0.f; C1[s] = 0.f; C2[s] = 0.f; Dp[s] = 0.f; last[s] = 0;
            if (__any_sync(FULL, inside)) live |= 1u << s;
        }

        float4 ra = make_float4(0.f, 0.f, -1.f, -1.f), rb = ra, rc = ra;
        float rinvd = 0.f;
        if (lane < n) {
            const uint32_t id = __ldg(point_list + range.x + lane);
            ra = __ldg(rec + 3 * (size_t)id); rb = __ldg(rec + 3 * (size_t)id + 1); rc = __ldg(rec + 3 * (size_t)id + 2);
            if (DEPTH) rinvd = __frcp_rn(__ldg(depths + id));
        }
        for (int base = 0; base < n && live != 0; base += 32) {
            const int cnt = min(32, n - base);
            __syncwarp();
            if (lane < cnt) {
                s_rec[warp][lane * 3 + 0] = ra; s_rec[warp][lane * 3 + 1] = rb; s_rec[warp][lane * 3 + 2] = rc;
                s_mask[warp][lane] = subblock_mask(ra.x - fox, ra.y - foy, ra.z, ra.w);
                if (DEPTH) s_invd[warp][lane] = rinvd;
            }
            __syncwarp();
            {   // prefetch the next step while this one is consumed
                const int nidx = base + 32 + lane;
                if (nidx < n) {
                    const uint32_t id = __ldg(point_list + range.x + nidx);
                    ra = __ldg(rec + 3 * (size_t)id); rb = __ldg(rec + 3 * (size_t)id + 1); rc = __ldg(rec + 3 * (size_t)id + 2);
                    if (DEPTH) rinvd = __frcp_rn(__ldg(depths + id));
                }
            }
            for (int j = 0; j < cnt; j++) {
                const unsigned m = s_mask[warp][j] & live;
                if (m == 0) continue;
                const float4 a = s_rec[warp][3 * j], b = s_rec[warp][3 * j + 1], c = s_rec[warp][3 * j + 2];
                const float invd = DEPTH ? s_invd[warp][j] : 0.f;
                const uint32_t contributor = (uint32_t)(base + j + 1);
#pragma unroll
                for (int sy = 0; sy < 4; sy++) {
                    if (m & (3u << (2 * sy))) {
                        const float dy = csub(a.y, pyf[sy]);
                        const float bdy = cmul(b.y, dy);
                        const float t2 = cmul(cmul(b.z, dy), dy);
#pragma unroll
                        for (int sx = 0; sx < 2; sx++) {
                            const int s = sy * 2 + sx;
                            if (m & (1u << s)) {
                                const float dx = csub(a.x, pxf[sx]);
                                const float u = cfma(b.x, dx, -bdy);
                                const float power = cfma(dx, u, t2);
                                if (power <= 0.0f && power >= b.w) {
                                    const float alpha = fminf(ALPHA_MAX, cmul(c.x, expf(power)));
                                    if (alpha >= ALPHA_MIN) {
                                        const float test_T = cmul(T[s], csub(1.0f, alpha));
                                        if (test_T < T_EPS) {
                                            T[s] = -fabsf(T[s]);       // done; |T| is the final transmittance
                                        } else {
                                            const float w = cmul(alpha, T[s]);
                                            C0[s] = cfma(c.y, w, C0[s]);
                                            C1[s] = cfma(c.z, w, C1[s]);
                                            C2[s] = cfma(c.w, w, C2[s]);
                                            if (DEPTH) Dp[s] = cfma(invd, w, Dp[s]);
                                            T[s] = test_T;
                                            last[s] = contributor;
                                        }
                                    }
                                }
                            }
                        }
                    }
                }
            }
            unsigned nl = 0;
#pragma unroll
            for (int s = 0; s < 8; s++)
                if ((live >> s) & 1u) { if (__any_sync(FULL, T[s] > 0.0f)) nl |= 1u << s; }
            live = nl;
        }
#pragma unroll
        for (int s = 0; s < 8; s++) {
            const int px = ox + (s & 1) * 8 + lx, py = oy + (s >> 1) * 4 + ly;
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
}

__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
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

// Backward: reverse traversal.  Per pixel the running state is (T_i, S_i) with
//   S_i = sum_{j behind i} T_j alpha_j (c_j . g) + T_final (bg . g),   g = dL/dC(pixel)
//   dL/dalpha_i = T_i (c_i . g) - S_i / (1 - alpha_i)
// and per (tile, Gaussian) the pixel sums are the moments of z = G * o * dL/dalpha:
//   m0 = sum z, mx = sum z dx, my = sum z dy, mxx = sum z dx^2, mxy = sum z dx dy, myy = sum z dy^2
//   dL/dpx = -(A mx + B my), dL/dpy = -(C my + B mx), dL/dA = -mxx/2, dL/dB = -mxy, dL/dC = -myy/2,
//   dL/do = m0 / o, dL/dcolour = sum alpha T g.
__global__ void __launch_bounds__(RW * 32) render_bwd_kernel(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
    int W, int H, int gx, int num_tiles, const float* __restrict__ bg, const float* __restrict__ final_T,
    const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix, float* __restrict__ acc,
    unsigned int* __restrict__ tile_counter) {
    __shared__ float4 s_rec[RW][32 * 3];
    __shared__ unsigned s_mask[RW][32];
    __shared__ uint32_t s_id[RW][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lx = lane & 7, ly = lane >> 3;
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    const size_t HW = (size_t)H * W;

    while (true) {
        int tile = 0;
        if (lane == 0) tile = (int)atomicAdd(tile_counter, 1u);
        tile = __shfl_sync(FULL, tile, 0);
        if (tile >= num_tiles) break;

        const int2 range = __ldg(ranges + tile);
        const int ox = (tile % gx) * TILE, oy = (tile / gx) * TILE;
        const float fox = (float)ox, foy = (float)oy;
        const float pxf[2] = {(float)(ox + lx), (float)(ox + 8 + lx)};
        const float pyf[4] = {(float)(oy + ly), (float)(oy + 4 + ly), (float)(oy + 8 + ly), (float)(oy + 12 + ly)};
        float T[8], S[8], g0[8], g1[8], g2[8];
        int last[8], smax[8];
        int maxc = 0;
#pragma unroll
        for (int s = 0; s < 8; s++) {
            const int px = ox + (s & 1) * 8 + lx, py = oy + (s >> 1) * 4 + ly;
            T[s] = 1.f; S[s] = 0.f; g0[s] = 0.f; g1[s] = 0.f; g2[s] = 0.f; last[s] = 0;
            if (px < W && py < H) {
                const size_t pix = (size_t)py * W + px;
                T[s] = __ldg(final_T + pix);
                last[s] = (int)__ldg(n_contrib + pix);
                g0[s] = __ldg(dL_dpix + pix); g1[s] = __ldg(dL_dpix + HW + pix); g2[s] = __ldg(dL_dpix + 2 * HW + pix);
                S[s] = T[s] * (bg0 * g0[s] + bg1 * g1[s] + bg2 * g2[s]);
            }
            smax[s] = __reduce_max_sync(FULL, last[s]);
            maxc = max(maxc, smax[s]);
        }

        for (int hi = maxc; hi > 0; hi -= 32) {
            const int lo = max(0, hi - 32);
            const int cnt = hi - lo;
            unsigned live = 0;
#pragma unroll
            for (int s = 0; s < 8; s++) if (smax[s] > lo) live |= 1u << s;
            __syncwarp();
            if (lane < cnt) {
                const uint32_t id = __ldg(point_list + range.x + lo + lane);
                const float4 ra = __ldg(rec + 3 * (size_t)id);
                s_rec[warp][lane * 3 + 0] = ra;
                s_rec[warp][lane * 3 + 1] = __ldg(rec + 3 * (size_t)id + 1);
                s_rec[warp][lane * 3 + 2] = __ldg(rec + 3 * (size_t)id + 2);
                s_mask[warp][lane] = subblock_mask(ra.x - fox, ra.y - foy, ra.z, ra.w);
                s_id[warp][lane] = id;
            }
            __syncwarp();
            for (int j = cnt - 1; j >= 0; j--) {
                const unsigned m = s_mask[warp][j] & live;
                if (m == 0) continue;
                const float4 a = s_rec[warp][3 * j], b = s_rec[warp][3 * j + 1], c = s_rec[warp][3 * j + 2];
                const int pos = lo + j;
                float m0 = 0.f, mx = 0.f, my = 0.f, mxx = 0.f, mxy = 0.f, myy = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f;
                bool any = false;
#pragma unroll
                for (int sy = 0; sy < 4; sy++) {
                    if (m & (3u << (2 * sy))) {
                        const float dy = csub(a.y, pyf[sy]);
                        const float bdy = cmul(b.y, dy);
                        const float t2 = cmul(cmul(b.z, dy), dy);
#pragma unroll
                        for (int sx = 0; sx < 2; sx++) {
                            const int s = sy * 2 + sx;
                            if ((m & (1u << s)) && pos < last[s]) {
                                const float dx = csub(a.x, pxf[sx]);
                                const float u = cfma(b.x, dx, -bdy);
                                const float power = cfma(dx, u, t2);
                                if (power <= 0.0f && power >= b.w) {
                                    const float G = expf(power);
                                    const float alpha = fminf(ALPHA_MAX, cmul(c.x, G));
                                    if (alpha >= ALPHA_MIN) {
                                        const float ra = fast_rcp(1.0f - alpha);
                                        const float Ti = T[s] * ra;
                                        T[s] = Ti;
                                        const float cg = c.y * g0[s] + c.z * g1[s] + c.w * g2[s];
                                        const float w = alpha * Ti;
                                        c0 = fmaf(w, g0[s], c0); c1 = fmaf(w, g1[s], c1); c2 = fmaf(w, g2[s], c2);
                                        const float dLda = Ti * cg - S[s] * ra;
                                        S[s] = fmaf(w, cg, S[s]);
                                        const float z = G * c.x * dLda;
                                        const float zx = z * dx, zy = z * dy;
                                        m0 += z; mx += zx; my += zy;
                                        mxx = fmaf(zx, dx, mxx); mxy = fmaf(zx, dy, mxy); myy = fmaf(zy, dy, myy);
                                        any = true;
                                    }
                                }
                            }
                        }
                    }
                }
                if (__any_sync(FULL, any)) {
                    // per-lane linear maps first (they commute with the sum over lanes), then ONE transposed
                    // butterfly: 8 values over 32 lanes in 4+2+1+1+1 shuffles; lane L ends up holding the
                    // warp total of value L >> 2, the ninth value takes a plain butterfly.
                    const float A = -2.0f * b.x, B = b.y, C = -2.0f * b.z;
                    const float v0 = -(A * mx + B * my), v1 = -(C * my + B * mx);
                    const float v2 = -0.5f * mxx, v3 = -mxy, v4 = -0.5f * myy, v5 = m0 * fast_rcp(c.x);
                    const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4;
                    float w0 = (up16 ? v4 : v0) + __shfl_xor_sync(FULL, up16 ? v0 : v4, 16);
                    float w1 = (up16 ? v5 : v1) + __shfl_xor_sync(FULL, up16 ? v1 : v5, 16);
                    float w2 = (up16 ? c0 : v2) + __shfl_xor_sync(FULL, up16 ? v2 : c0, 16);
                    float w3 = (up16 ? c1 : v3) + __shfl_xor_sync(FULL, up16 ? v3 : c1, 16);
                    float u0 = (up8 ? w2 : w0) + __shfl_xor_sync(FULL, up8 ? w0 : w2, 8);
                    float u1 = (up8 ? w3 : w1) + __shfl_xor_sync(FULL, up8 ? w1 : w3, 8);
                    float t0 = (up4 ? u1 : u0) + __shfl_xor_sync(FULL, up4 ? u0 : u1, 4);
                    t0 += __shfl_xor_sync(FULL, t0, 2);
                    t0 += __shfl_xor_sync(FULL, t0, 1);
                    c2 = warp_sum(c2);
                    // value index held by lane L: 4*bit4 + 2*bit3 + bit2 -> component slot of acc
                    //   0: v0 (dpx)  1: v1 (dpy)  2: v2 (dA)  3: v3 (dB)  4: v4 (dC)  5: v5 (do)  6: c0  7: c1
                    float* dst = acc + 12 * (size_t)s_id[warp][j];
                    if ((lane & 3) == 0) atomicAdd(dst + (lane >> 2), t0);
                    else if (lane == 1) atomicAdd(dst + 8, c2);
                }
            }
        }
    }
}

}  // namespace

cudaError_t launch_render_fwd(const b200splat_view& v, GeomView gv, BinView bv, ImgView iv, float* out_color,
                              float* out_invdepth, cudaStream_t s) {
    const int W = v.image_width, H = v.image_height;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const int T = gx * gy;
    if (T == 0) return cudaSuccess;
    // iv.counters was zeroed at the start of the forward call
    const int blocks = min((T + RW - 1) / RW, 148 * 8);
    if (out_invdepth)
        render_fwd_kernel<true><<<blocks, RW * 32, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, gx, T, v.bg,
                                                           out_color, out_invdepth, iv.final_T, iv.n_contrib, iv.counters + CTR_RENDER_FWD);
    else
        render_fwd_kernel<false><<<blocks, RW * 32, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, gx, T, v.bg,
                                                            out_color, nullptr, iv.final_T, iv.n_contrib, iv.counters + CTR_RENDER_FWD);
    return cudaGetLastError();
}

cudaError_t launch_render_bwd(const b200splat_view& v, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                              const float* dL_dpix, const float* dL_dinvdepth, float* acc, cudaStream_t s) {
    (void)dL_dinvdepth;
    const int W = v.image_width, H = v.image_height;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const int T = gx * gy;
    if (T == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(iv.counters + CTR_RENDER_BWD, 0, sizeof(unsigned int), s);
    if (e != cudaSuccess) return e;
    const int blocks = min((T + RW - 1) / RW, 148 * 8);
    render_bwd_kernel<<<blocks, RW * 32, 0, s>>>(iv.ranges, vals_sorted, gv.rec, W, H, gx, T, v.bg, iv.final_T,
                                                 iv.n_contrib, dL_dpix, acc, iv.counters + CTR_RENDER_BWD);
    return cudaGetLastError();
}

}  // namespace b200splat
