// render_classic.cu -- COMPARATOR ONLY (see binning_classic.cu): the textbook tile renderer.
// One 16x16 block per tile, one pixel per thread, 256 list entries staged per batch; the backward pass issues
// nine global float atomics per (pixel, Gaussian).  Same arithmetic as the product, so the images agree.
#include "../splat_internal.cuh"

namespace b200splat {

namespace {

constexpr int BS = TILE * TILE;

template <bool DEPTH>
__global__ void __launch_bounds__(BS) render_fwd_classic(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec,
    const float* __restrict__ depths, int W, int H, const float* __restrict__ bg, float* __restrict__ out_color,
    float* __restrict__ out_invdepth, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib) {
    __shared__ uint32_t s_id[BS];
    __shared__ float2 s_xy[BS];
    __shared__ float4 s_con[BS];     // hA, B, hC, opacity
    const int tx = threadIdx.x, ty = threadIdx.y, t = ty * TILE + tx;
    const int px = blockIdx.x * TILE + tx, py = blockIdx.y * TILE + ty;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const int2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int rounds = (range.y - range.x + BS - 1) / BS;
    int todo = range.y - range.x;
    bool done = !inside;
    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, Dp = 0.f;
    uint32_t contributor = 0, last = 0;
    for (int i = 0; i < rounds; i++, todo -= BS) {
        if (__syncthreads_count(done) == BS) break;
        const int idx = range.x + i * BS + t;
        if (idx < range.y) {
            const uint32_t id = point_list[idx];
            const float4 a = rec[3 * (size_t)id], b = rec[3 * (size_t)id + 1], c = rec[3 * (size_t)id + 2];
            s_id[t] = id; s_xy[t] = make_float2(a.x, a.y); s_con[t] = make_float4(b.x, b.y, b.z, c.x);
        }
        __syncthreads();
        for (int j = 0; !done && j < min(BS, todo); j++) {
            contributor++;
            const float2 xy = s_xy[j];
            const float4 con = s_con[j];
            const float dx = csub(xy.x, pxf), dy = csub(xy.y, pyf);
            const float power = cfma(dx, cfma(con.x, dx, -cmul(con.y, dy)), cmul(cmul(con.z, dy), dy));
            if (power > 0.0f) continue;
            const float alpha = fminf(ALPHA_MAX, cmul(con.w, expf(power)));
            if (alpha < ALPHA_MIN) continue;
            const float test_T = cmul(T, csub(1.0f, alpha));
            if (test_T < T_EPS) { done = true; continue; }
            const float4 c = rec[3 * (size_t)s_id[j] + 2];        // colour fetched from global memory per contribution
            const float w = cmul(alpha, T);
            C0 = cfma(c.y, w, C0); C1 = cfma(c.z, w, C1); C2 = cfma(c.w, w, C2);
            if (DEPTH) Dp = cfma(__frcp_rn(depths[s_id[j]]), w, Dp);
            T = test_T;
            last = contributor;
        }
    }
    if (inside) {
        const size_t pix = (size_t)py * W + px, HW = (size_t)H * W;
        final_T[pix] = T; n_contrib[pix] = last;
        out_color[pix] = cfma(T, bg[0], C0); out_color[HW + pix] = cfma(T, bg[1], C1); out_color[2 * HW + pix] = cfma(T, bg[2], C2);
        if (DEPTH) out_invdepth[pix] = Dp;
    }
}

__global__ void __launch_bounds__(BS) render_bwd_classic(
    const int2* __restrict__ ranges, const uint32_t* __restrict__ point_list, const float4* __restrict__ rec, int W, int H,
    const float* __restrict__ bg, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, float* __restrict__ acc, uint8_t* __restrict__ touched) {
    __shared__ uint32_t s_id[BS];
    __shared__ float2 s_xy[BS];
    __shared__ float4 s_con[BS];
    __shared__ float s_col[3 * BS];
    const int tx = threadIdx.x, ty = threadIdx.y, t = ty * TILE + tx;
    const int px = blockIdx.x * TILE + tx, py = blockIdx.y * TILE + ty;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const int2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int rounds = (range.y - range.x + BS - 1) / BS;
    int todo = range.y - range.x;
    const size_t pix = (size_t)py * W + px, HW = (size_t)H * W;
    const float Tf = inside ? final_T[pix] : 0.f;
    float T = Tf;
    uint32_t contributor = todo;
    const int last = inside ? (int)n_contrib[pix] : 0;
    float g[3] = {0.f, 0.f, 0.f};
    if (inside) { g[0] = dL_dpix[pix]; g[1] = dL_dpix[HW + pix]; g[2] = dL_dpix[2 * HW + pix]; }
    float accum[3] = {0.f, 0.f, 0.f}, last_alpha = 0.f, last_col[3] = {0.f, 0.f, 0.f};
    const float bgdot = bg[0] * g[0] + bg[1] * g[1] + bg[2] * g[2];
    for (int i = 0; i < rounds; i++, todo -= BS) {
        __syncthreads();
        const int idx = range.y - 1 - (i * BS + t);              // back to front
        if (idx >= range.x) {
            const uint32_t id = point_list[idx];
            const float4 a = rec[3 * (size_t)id], b = rec[3 * (size_t)id + 1], c = rec[3 * (size_t)id + 2];
            s_id[t] = id; s_xy[t] = make_float2(a.x, a.y); s_con[t] = make_float4(b.x, b.y, b.z, c.x);
            s_col[3 * t] = c.y; s_col[3 * t + 1] = c.z; s_col[3 * t + 2] = c.w;
        }
        __syncthreads();
        for (int j = 0; inside && j < min(BS, todo); j++) {
            contributor--;
            if ((int)contributor >= last) continue;
            const float2 xy = s_xy[j];
            const float4 con = s_con[j];
            const float dx = csub(xy.x, pxf), dy = csub(xy.y, pyf);
            const float power = cfma(dx, cfma(con.x, dx, -cmul(con.y, dy)), cmul(cmul(con.z, dy), dy));
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(ALPHA_MAX, cmul(con.w, G));
            if (alpha < ALPHA_MIN) continue;
            T = T / (1.0f - alpha);
            const float w = alpha * T;
            float dL_dalpha = 0.f;
            float* dst = acc + 12 * (size_t)s_id[j];
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                const float c = s_col[3 * j + ch];
                accum[ch] = last_alpha * last_col[ch] + (1.f - last_alpha) * accum[ch];
                last_col[ch] = c;
                dL_dalpha += (c - accum[ch]) * g[ch];
                atomicAdd(dst + 6 + ch, w * g[ch]);
            }
            dL_dalpha *= T;
            last_alpha = alpha;
            dL_dalpha += (-Tf / (1.f - alpha)) * bgdot;
            const float z = con.w * G * dL_dalpha;             // dL/dG * G
            const float A = -2.0f * con.x, B = con.y, C = -2.0f * con.z;
            atomicAdd(dst + 0, -(A * z * dx + B * z * dy));
            atomicAdd(dst + 1, -(C * z * dy + B * z * dx));
            atomicAdd(dst + 2, -0.5f * z * dx * dx);
            atomicAdd(dst + 3, -z * dx * dy);
            atomicAdd(dst + 4, -0.5f * z * dy * dy);
            atomicAdd(dst + 5, G * dL_dalpha);
            touched[id] = 1;
        }
    }
}

}  // namespace

cudaError_t launch_render_fwd(const b200splat_view& v, GeomView gv, BinView bv, ImgView iv, float* out_color,
                              float* out_invdepth, cudaStream_t s) {
    const int W = v.image_width, H = v.image_height;
    const dim3 grid((W + TILE - 1) / TILE, (H + TILE - 1) / TILE), block(TILE, TILE);
    if (grid.x * grid.y == 0) return cudaSuccess;
    if (out_invdepth)
        render_fwd_classic<true><<<grid, block, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, v.bg, out_color,
                                                        out_invdepth, iv.final_T, iv.n_contrib);
    else
        render_fwd_classic<false><<<grid, block, 0, s>>>(iv.ranges, bv.vals_sorted, gv.rec, gv.depths, W, H, v.bg, out_color,
                                                         nullptr, iv.final_T, iv.n_contrib);
    return cudaGetLastError();
}

cudaError_t launch_render_bwd(const b200splat_view& v, GeomView gv, const uint32_t* vals_sorted, ImgView iv,
                              const float* dL_dpix, const float* dL_dinvdepth, float* acc, uint8_t* touched, cudaStream_t s) {
    (void)dL_dinvdepth;
    const int W = v.image_width, H = v.image_height;
    const dim3 grid((W + TILE - 1) / TILE, (H + TILE - 1) / TILE), block(TILE, TILE);
    if (grid.x * grid.y == 0) return cudaSuccess;
    render_bwd_classic<<<grid, block, 0, s>>>(iv.ranges, vals_sorted, gv.rec, W, H, v.bg, iv.final_T, iv.n_contrib, dL_dpix, acc, touched);
    return cudaGetLastError();
}

cudaError_t launch_render_stats(int, int, GeomView, const uint32_t*, ImgView, unsigned long long* out16, cudaStream_t s) {
    return cudaMemsetAsync(out16, 0, 16 * sizeof(unsigned long long), s);
}

}  // namespace b200splat
