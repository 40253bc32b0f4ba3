// api.cu -- the C ABI of libb200splat.so (declared in include/b200splat.h): argument validation,
// buffer carving, stage orchestration, optional per-stage timing, error strings.
#include <cstdio>
#include <string>

#include "splat_internal.cuh"

using namespace b200splat;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

int cuda_fail(cudaError_t e, const char* where) {
    return fail(B200SPLAT_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}

#define CU(call, where)                                         \
    do {                                                        \
        cudaError_t _e = (call);                                \
        if (_e != cudaSuccess) return cuda_fail(_e, where);     \
    } while (0)

// In debug mode every stage is followed by a synchronisation + error check, like the `debug` flag
// of the call being replaced (/root/reference/README.md:159-162).
#define STAGE(call, where)                                                          \
    do {                                                                            \
        cudaError_t _e = (call);                                                    \
        if (_e != cudaSuccess) return cuda_fail(_e, where);                         \
        if (debug) {                                                                \
            _e = cudaStreamSynchronize(stream);                                     \
            if (_e != cudaSuccess) return cuda_fail(_e, where " (debug sync)");     \
        }                                                                           \
    } while (0)

struct StageTimer {
    bool on;
    cudaStream_t s;
    cudaEvent_t ev[10];
    int n = 0;
    StageTimer(bool on_, cudaStream_t s_) : on(on_), s(s_) {
        if (on) for (auto& e : ev) cudaEventCreate(&e);
    }
    ~StageTimer() {
        if (on) for (auto& e : ev) cudaEventDestroy(e);
    }
    void mark() { if (on && n < 10) cudaEventRecord(ev[n++], s); }
    float ms(int i) {
        float t = 0.f;
        if (on && i + 1 < n) cudaEventElapsedTime(&t, ev[i], ev[i + 1]);
        return t;
    }
};

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int validate(const b200splat_view* v, const b200splat_gaussians* g) {
    if (!v || !g) return fail(B200SPLAT_ERR_INVALID_ARG, "null view/gaussians");
    if (g->P < 0 || v->image_width <= 0 || v->image_height <= 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad sizes");
    if (v->sh_degree < 0 || v->sh_degree > 3) return fail(B200SPLAT_ERR_INVALID_ARG, "sh_degree must be in [0,3]");
    if (!(v->tanfovx > 0.f) || !(v->tanfovy > 0.f)) return fail(B200SPLAT_ERR_INVALID_ARG, "tanfov must be positive");
    if (!v->bg || !v->viewmatrix || !v->projmatrix) return fail(B200SPLAT_ERR_INVALID_ARG, "null bg/view/proj");
    if (g->P > 0) {
        if (!g->means3D || !g->opacities) return fail(B200SPLAT_ERR_INVALID_ARG, "null means3D/opacities");
        if ((g->shs == nullptr) == (g->colors_precomp == nullptr))
            return fail(B200SPLAT_ERR_INVALID_ARG, "provide exactly one of shs / colors_precomp");
        const bool sr = g->scales != nullptr && g->rotations != nullptr;
        if ((g->scales != nullptr) != (g->rotations != nullptr) || sr == (g->cov3D_precomp != nullptr))
            return fail(B200SPLAT_ERR_INVALID_ARG, "provide exactly one of (scales, rotations) / cov3D_precomp");
        if (g->shs) {
            if (!v->campos) return fail(B200SPLAT_ERR_INVALID_ARG, "campos required with shs");
            if (g->M < (v->sh_degree + 1) * (v->sh_degree + 1))
                return fail(B200SPLAT_ERR_INVALID_ARG, "shs holds fewer coefficients than sh_degree needs");
            if (!aligned16(g->shs)) return fail(B200SPLAT_ERR_INVALID_ARG, "shs must be 16-byte aligned");
        }
        if (g->rotations && !aligned16(g->rotations)) return fail(B200SPLAT_ERR_INVALID_ARG, "rotations must be 16-byte aligned");
    }
    return B200SPLAT_OK;
}

}  // namespace

// error string shared with the other translation units (next_steps.cu, ply_io.cu)
namespace b200splat {
int report_error(int code, const char* msg) { return fail(code, msg); }
}

extern "C" {

const char* b200splat_last_error(void) { return g_last_error.c_str(); }
int b200splat_version(void) { return B200SPLAT_VERSION; }

size_t b200splat_geom_bytes(int32_t P) { return carve_geom(nullptr, P).total; }
size_t b200splat_image_bytes(int32_t width, int32_t height) { return carve_image(nullptr, width, height).total; }
size_t b200splat_binning_bytes(int32_t num_rendered, int32_t width, int32_t height) {
    return carve_binning(nullptr, num_rendered, tiles_of(width, height)).total;
}

int b200splat_forward(const b200splat_view* view, const b200splat_gaussians* g, const b200splat_forward_out* out,
                      const b200splat_allocators* alloc, int32_t* num_rendered, b200splat_stage_times* times,
                      void* stream_) {
    int rc = validate(view, g);
    if (rc) return rc;
    if (!out || !out->color || (g->P > 0 && !out->radii)) return fail(B200SPLAT_ERR_INVALID_ARG, "null outputs");
    if (!alloc || !alloc->alloc || !num_rendered) return fail(B200SPLAT_ERR_INVALID_ARG, "null allocator / num_rendered");
    if (alloc->bin_capacity < 0 || alloc->bin_capacity > 0x7fffffffll) return fail(B200SPLAT_ERR_INVALID_ARG, "bin_capacity out of range");
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool debug = view->debug != 0;
    const int P = g->P, W = view->image_width, H = view->image_height;
    if ((W + TILE - 1) / TILE > max_tile_bins() || (H + TILE - 1) / TILE > max_tile_bins())
        return fail(B200SPLAT_ERR_INVALID_ARG, "image larger than 8192 x 8192 pixels");

    void* geom_mem = alloc->alloc(alloc->geom_ctx, b200splat_geom_bytes(P));
    void* img_mem = alloc->alloc(alloc->image_ctx, b200splat_image_bytes(W, H));
    if (!geom_mem || !img_mem) return fail(B200SPLAT_ERR_ALLOC, "allocator returned NULL (geometry / image buffer)");
    GeomView gv = carve_geom(geom_mem, P);
    ImgView iv = carve_image(img_mem, W, H);
    const int T = tiles_of(W, H);

    StageTimer tm(times != nullptr, stream);
    tm.mark();
    CU(cudaMemsetAsync(iv.ranges, 0, iv.clear_bytes, stream), "clear ranges + control block");   // adjacent: one node
    STAGE(launch_preprocess_fwd(*view, *g, gv, out->radii, iv.counters, stream), "preprocess");
    tm.mark();
    STAGE(launch_depth_order(gv, P, iv.counters, stream), "depth order");
    tm.mark();

    // Pair capacity of the binning buffer (b200splat.h "pair capacity").  bin_capacity > 0: the caller's bound, and
    // this call never waits for the device -- R stays in the control block, where every later kernel reads it.
    // bin_capacity == 0: exact -- one host round trip for R, after the depth order has been queued.
    long long cap = alloc->bin_capacity;
    if (cap == 0 && P > 0) {
        unsigned long long R = 0;
        CU(cudaMemcpyAsync(&R, iv.counters + CTR_PAIRS, sizeof(R), cudaMemcpyDeviceToHost, stream), "read num_rendered");
        CU(cudaStreamSynchronize(stream), "sync num_rendered");
        if (R > 0x7fffffffull) return fail(B200SPLAT_ERR_INVALID_ARG, "more than 2^31-1 tile/Gaussian pairs");
        cap = (long long)R;
    }
    *num_rendered = (int32_t)cap;

    void* bin_mem = alloc->alloc(alloc->binning_ctx, b200splat_binning_bytes((int32_t)cap, W, H));
    if (!bin_mem) return fail(B200SPLAT_ERR_ALLOC, "allocator returned NULL (binning buffer)");
    BinView bv = carve_binning(bin_mem, (int)cap, T);

    if (cap > 0) STAGE(launch_emit(gv, P, W, H, bv, iv.counters, stream), "duplicateWithKeys");
    tm.mark();
    STAGE(launch_tile_sort(bv, W, H, iv, stream), "sort");
    tm.mark();
    STAGE(launch_tile_order(iv, W, H, stream), "identifyTileRanges");
    tm.mark();
    STAGE(launch_render_fwd(*view, gv, bv, iv, out->color, out->invdepth, stream), "render");
    tm.mark();
    if (times) {
        CU(cudaStreamSynchronize(stream), "timing sync");
        times->preprocess_ms = tm.ms(0); times->scan_ms = tm.ms(1); times->duplicate_ms = tm.ms(2);
        times->sort_ms = tm.ms(3); times->ranges_ms = tm.ms(4); times->render_ms = tm.ms(5);
    }
    return B200SPLAT_OK;
}

size_t b200splat_image_status_offset(int32_t width, int32_t height) {
    ImgView iv = carve_image(nullptr, width, height);
    return (size_t)((char*)(iv.counters + CTR_STATUS) - (char*)nullptr);
}

int b200splat_forward_status(const void* image, int32_t width, int32_t height, int64_t* num_pairs, int64_t* capacity,
                             int32_t* overflow, void* stream_) {
    if (!image || width <= 0 || height <= 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad arguments");
    ImgView iv = carve_image(const_cast<void*>(image), width, height);
    unsigned int w[4] = {0, 0, 0, 0};
    unsigned long long R = 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    CU(cudaMemcpyAsync(&R, iv.counters + CTR_PAIRS, sizeof(R), cudaMemcpyDeviceToHost, stream), "read pair count");
    CU(cudaMemcpyAsync(w, iv.counters + CTR_STATUS, sizeof(w), cudaMemcpyDeviceToHost, stream), "read status");
    CU(cudaStreamSynchronize(stream), "status sync");
    if (num_pairs) *num_pairs = (int64_t)R;
    if (capacity) *capacity = (int64_t)w[2];
    if (overflow) *overflow = (int32_t)w[3];
    return B200SPLAT_OK;
}

// Internal side stream (per thread and device) for work that runs beside the caller's stream inside one call;
// always forked from and joined back into the caller's stream, so the call stays stream-ordered for the caller.
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
static thread_local SideStream g_side_streams[64];
static cudaError_t side_stream(SideStream** out) {
    SideStream* per_device = g_side_streams;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    SideStream& ss = per_device[dev & 63];
    if (!ss.stream) {
        if ((e = cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    *out = &ss;
    return cudaSuccess;
}

int b200splat_zero_fill(int32_t n, void* const* ptrs, const size_t* bytes, void* stream_) {
    if (n < 0 || n > 8 || (n > 0 && (!ptrs || !bytes))) return fail(B200SPLAT_ERR_INVALID_ARG, "zero_fill takes 0..8 ranges");
    ZeroRanges zr;
    zr.n = 0;
    for (int i = 0; i < n; i++)
        if (ptrs[i] && bytes[i]) { zr.ptr[zr.n] = ptrs[i]; zr.bytes[zr.n] = bytes[i]; zr.n++; }
    CU(launch_zero_fill(zr, (cudaStream_t)stream_), "zero_fill");
    return B200SPLAT_OK;
}

// Releases what the calling host thread holds inside the library: its helper streams and events.
void b200splat_shutdown(void) {
    int cur = 0;
    const bool have = cudaGetDevice(&cur) == cudaSuccess;
    for (int d = 0; d < 64; d++) {
        SideStream& ss = g_side_streams[d];
        if (!ss.stream) continue;
        if (have) cudaSetDevice(d);
        cudaStreamSynchronize(ss.stream);
        cudaStreamDestroy(ss.stream);
        if (ss.fork) cudaEventDestroy(ss.fork);
        if (ss.join) cudaEventDestroy(ss.join);
        ss = SideStream();
    }
    if (have) cudaSetDevice(cur);
}

int b200splat_backward(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                       const void* geom, const void* binning, const void* image, int32_t num_rendered,
                       const float* dL_dout_color, const float* dL_dout_invdepth, const b200splat_grads* grads,
                       b200splat_alloc_fn scratch_alloc, void* scratch_ctx, b200splat_stage_times* times,
                       void* stream_) {
    return b200splat_backward_ex(view, g, radii, geom, binning, image, num_rendered, dL_dout_color, dL_dout_invdepth, grads,
                                 nullptr, scratch_alloc, scratch_ctx, times, stream_);
}

static int backward_impl(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                         const void* geom, const void* binning, const void* image, int32_t num_rendered,
                         const float* dL_dout_color, const float* dL_dout_invdepth, const b200splat_grads* grads,
                         const b200splat_backward_hooks* hooks, float* packed_rows, uint32_t* packed_count,
                         b200splat_alloc_fn scratch_alloc, void* scratch_ctx, b200splat_stage_times* times, void* stream_,
                         void* clean_scratch = nullptr, float* norm_accum = nullptr);

int b200splat_backward_ex(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                          const void* geom, const void* binning, const void* image, int32_t num_rendered,
                          const float* dL_dout_color, const float* dL_dout_invdepth, const b200splat_grads* grads,
                          const b200splat_backward_hooks* hooks, b200splat_alloc_fn scratch_alloc, void* scratch_ctx,
                          b200splat_stage_times* times, void* stream_) {
    if (!grads || !grads->dL_dmeans3D) return fail(B200SPLAT_ERR_INVALID_ARG, "grads->dL_dmeans3D is required");
    return backward_impl(view, g, radii, geom, binning, image, num_rendered, dL_dout_color, dL_dout_invdepth, grads, hooks,
                         nullptr, nullptr, scratch_alloc, scratch_ctx, times, stream_);
}

int b200splat_backward_packed(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                              const void* geom, const void* binning, const void* image, int32_t num_rendered,
                              const float* dL_dout_color, const float* dL_dout_invdepth, float* dL_dmeans2D,
                              float* packed_rows, uint32_t* packed_count, b200splat_alloc_fn scratch_alloc,
                              void* scratch_ctx, b200splat_stage_times* times, void* stream_) {
    if (!packed_rows || !packed_count) return fail(B200SPLAT_ERR_INVALID_ARG, "null packed_rows / packed_count");
    if (!aligned16(packed_rows)) return fail(B200SPLAT_ERR_INVALID_ARG, "packed_rows must be 16-byte aligned");
    if (g && (!g->shs || g->cov3D_precomp || g->colors_precomp))
        return fail(B200SPLAT_ERR_INVALID_ARG, "packed rows carry SH-mode, scale/rotation-mode gradients only");
    b200splat_grads only2d;
    memset(&only2d, 0, sizeof(only2d));
    only2d.dL_dmeans2D = dL_dmeans2D;
    return backward_impl(view, g, radii, geom, binning, image, num_rendered, dL_dout_color, dL_dout_invdepth, &only2d, nullptr,
                         packed_rows, packed_count, scratch_alloc, scratch_ctx, times, stream_);
}

size_t b200splat_backward_scratch_bytes(int32_t P) {
    if (P <= 0) return 0;
    return (size_t)P * 12 * sizeof(float) + (((size_t)P + 255) / 256 * 256 + 256) + (size_t)P * sizeof(uint32_t);
}

int b200splat_backward_packed_clean(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                                    const void* geom, const void* binning, const void* image, int32_t num_rendered,
                                    const float* dL_dout_color, const float* dL_dout_invdepth, float* dL_dmeans2D,
                                    float* xyz_gradient_accum, float* packed_rows, uint32_t* packed_count, void* scratch,
                                    size_t scratch_bytes, b200splat_stage_times* times, void* stream_) {
    if (!packed_rows || !packed_count) return fail(B200SPLAT_ERR_INVALID_ARG, "null packed_rows / packed_count");
    if (!aligned16(packed_rows)) return fail(B200SPLAT_ERR_INVALID_ARG, "packed_rows must be 16-byte aligned");
    if (g && (!g->shs || g->cov3D_precomp || g->colors_precomp))
        return fail(B200SPLAT_ERR_INVALID_ARG, "packed rows carry SH-mode, scale/rotation-mode gradients only");
    if (g && g->P > 0 && (!scratch || !aligned16(scratch) || scratch_bytes < b200splat_backward_scratch_bytes(g->P)))
        return fail(B200SPLAT_ERR_INVALID_ARG, "clean scratch: null, not 16-byte aligned or smaller than b200splat_backward_scratch_bytes(P)");
    b200splat_grads only2d;
    memset(&only2d, 0, sizeof(only2d));
    only2d.dL_dmeans2D = dL_dmeans2D;
    return backward_impl(view, g, radii, geom, binning, image, num_rendered, dL_dout_color, dL_dout_invdepth, &only2d, nullptr,
                         packed_rows, packed_count, nullptr, nullptr, times, stream_, scratch, xyz_gradient_accum);
}

int b200splat_accumulate_packed(int32_t P, int32_t M, int32_t sh_degree, const float* means3D, int32_t n_src,
                                const float* const* rows, const uint32_t* const* counts, const float* campos,
                                float* dL_dsh, float* dL_dgeom, uint8_t* marks, int32_t store, void* stream_) {
    if (P < 0 || M <= 0 || M > 16 || sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1))
        return fail(B200SPLAT_ERR_INVALID_ARG, "bad P / M / sh_degree");
    if (n_src < 1 || n_src > B200SPLAT_MAX_VIEWS) return fail(B200SPLAT_ERR_INVALID_ARG, "n_src must be in [1, B200SPLAT_MAX_VIEWS]");
    if (store && n_src != 1) return fail(B200SPLAT_ERR_INVALID_ARG, "store mode takes exactly one source");
    if (P == 0) return B200SPLAT_OK;
    if (!means3D || !rows || !counts || !dL_dsh || !dL_dgeom) return fail(B200SPLAT_ERR_INVALID_ARG, "null argument");
    if (!aligned16(dL_dgeom)) return fail(B200SPLAT_ERR_INVALID_ARG, "dL_dgeom must be 16-byte aligned");
    for (int v = 0; v < n_src; v++)
        if (!rows[v] || !counts[v]) return fail(B200SPLAT_ERR_INVALID_ARG, "null source");
    CU(launch_accumulate_packed(P, M, sh_degree, means3D, n_src, rows, reinterpret_cast<const unsigned* const*>(counts), campos,
                                dL_dsh, dL_dgeom, marks, store != 0, (cudaStream_t)stream_), "accumulate_packed");
    return B200SPLAT_OK;
}

int b200splat_clear_accumulated(int32_t P, int32_t M, float* dL_dsh, float* dL_dgeom, uint8_t* marks, void* stream_) {
    if (P < 0 || M <= 0 || M > 16) return fail(B200SPLAT_ERR_INVALID_ARG, "bad P / M");
    if (P == 0) return B200SPLAT_OK;
    if (!dL_dsh || !dL_dgeom || !marks) return fail(B200SPLAT_ERR_INVALID_ARG, "null argument");
    if (!aligned16(dL_dgeom)) return fail(B200SPLAT_ERR_INVALID_ARG, "dL_dgeom must be 16-byte aligned");
    CU(launch_clear_accumulated(P, M, dL_dsh, dL_dgeom, marks, (cudaStream_t)stream_), "clear_accumulated");
    return B200SPLAT_OK;
}

static int backward_impl(const b200splat_view* view, const b200splat_gaussians* g, const int32_t* radii,
                         const void* geom, const void* binning, const void* image, int32_t num_rendered,
                         const float* dL_dout_color, const float* dL_dout_invdepth, const b200splat_grads* grads,
                         const b200splat_backward_hooks* hooks, float* packed_rows, uint32_t* packed_count,
                         b200splat_alloc_fn scratch_alloc, void* scratch_ctx, b200splat_stage_times* times, void* stream_,
                         void* clean_scratch, float* norm_accum) {
    int rc = validate(view, g);
    if (rc) return rc;
    const bool packed = packed_rows != nullptr;
    float* early_colors = hooks ? hooks->masked_colors : nullptr;
    if (early_colors && !g->shs) return fail(B200SPLAT_ERR_INVALID_ARG, "hooks->masked_colors needs shs");
    if (!packed && g->shs && !grads->dL_dsh && !grads->dL_dcolors && !early_colors)
        return fail(B200SPLAT_ERR_INVALID_ARG, "with shs, grads->dL_dsh (or dL_dcolors for the masked colour gradient) is required");
    if (grads->dL_dsh && !aligned16(grads->dL_dsh)) return fail(B200SPLAT_ERR_INVALID_ARG, "dL_dsh must be 16-byte aligned");
    if (grads->dL_drotations && !aligned16(grads->dL_drotations))
        return fail(B200SPLAT_ERR_INVALID_ARG, "dL_drotations must be 16-byte aligned");
    if (!dL_dout_color || (!scratch_alloc && !clean_scratch)) return fail(B200SPLAT_ERR_INVALID_ARG, "null dL_dout_color / scratch allocator");
    if (g->P > 0 && (!radii || !geom || !binning || !image)) return fail(B200SPLAT_ERR_INVALID_ARG, "null forward state");
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool debug = view->debug != 0;
    const int P = g->P, W = view->image_width, H = view->image_height;
    if (P == 0) {
        if (early_colors && hooks->after_masked_colors) hooks->after_masked_colors(hooks->ctx);
        if (packed) CU(cudaMemsetAsync(packed_count, 0, sizeof(uint32_t), stream), "clear packed count");
        return B200SPLAT_OK;
    }

    GeomView gv = carve_geom(const_cast<void*>(geom), P);
    BinView bv = carve_binning(const_cast<void*>(binning), num_rendered, tiles_of(W, H));
    ImgView iv = carve_image(const_cast<void*>(image), W, H);

    // scratch: accumulator rows [P,12] f32 | touched flags [P] u8 + counter (cleared together) | touched indices [P] u32
    const size_t acc_bytes = (size_t)P * 12 * sizeof(float);
    const size_t flag_bytes = ((size_t)P + 255) / 256 * 256 + 256;
    const size_t idx_bytes = (size_t)P * sizeof(uint32_t);
    // clean scratch (b200splat_backward_packed_clean): caller-owned, all zero on entry, all zero again on return -- the
    // 150 MB clear in front of the render backward disappears (the consumers re-zero what they read)
    const bool clean = clean_scratch != nullptr;
    char* scratch = clean ? (char*)clean_scratch : (char*)scratch_alloc(scratch_ctx, acc_bytes + flag_bytes + idx_bytes);
    if (!scratch) return fail(B200SPLAT_ERR_ALLOC, "allocator returned NULL (backward scratch)");
    float* acc = (float*)scratch;
    uint8_t* touched = (uint8_t*)(scratch + acc_bytes);
    unsigned* active_count = (unsigned*)(scratch + acc_bytes + flag_bytes - 256);
    uint32_t* active_idx = (uint32_t*)(scratch + acc_bytes + flag_bytes);

    // Every gradient of a Gaussian the render backward never touches is exactly zero, and on a real view that is
    // most of them: the outputs are zero-filled on a side stream WHILE the render backward runs (it leaves HBM
    // nearly idle), and the per-Gaussian backward then visits the touched ones only.
    SideStream* ss = nullptr;
    CU(side_stream(&ss), "side stream");
    CU(cudaEventRecord(ss->fork, stream), "fork");
    CU(cudaStreamWaitEvent(ss->stream, ss->fork, 0), "fork wait");
    {
        const size_t n = (size_t)P * sizeof(float);
        const b200splat_grads& o = *grads;
        ZeroRanges zr;
        zr.n = 0;
        auto add = [&](void* p, size_t bytes) { if (p) { zr.ptr[zr.n] = p; zr.bytes[zr.n] = bytes; zr.n++; } };
        add(o.dL_dmeans2D, 3 * n); add(o.dL_dcolors, 3 * n); add(o.dL_dopacity, n); add(o.dL_dmeans3D, 3 * n);
        add(o.dL_dcov3D, 6 * n); add(o.dL_dsh, 3 * (size_t)g->M * n); add(o.dL_dscales, 3 * n); add(o.dL_drotations, 4 * n);
        CU(launch_zero_fill(zr, ss->stream), "zero the gradient outputs");
    }
    CU(cudaEventRecord(ss->join, ss->stream), "join");

    StageTimer tm(times != nullptr, stream);
    tm.mark();
    if (!clean) CU(cudaMemsetAsync(scratch, 0, acc_bytes + flag_bytes, stream), "clear backward scratch");
    if (num_rendered > 0)
        STAGE(launch_render_bwd(*view, gv, bv.vals_sorted, iv, dL_dout_color, dL_dout_invdepth, acc, touched, stream), "render backward");
    tm.mark();
    if (early_colors) {
        STAGE(launch_masked_colour_grad(P, radii, gv, acc, early_colors, stream), "masked colour gradient");
        if (hooks->after_masked_colors) hooks->after_masked_colors(hooks->ctx);
    }
    STAGE(launch_compact_active(P, touched, clean, active_idx, active_count, stream), "compact touched Gaussians");
    CU(cudaStreamWaitEvent(stream, ss->join, 0), "join wait");
    STAGE(launch_preprocess_bwd(*view, *g, radii, gv, acc, clean, *grads, active_idx, active_count, packed_rows, packed_count,
                                norm_accum, clean && packed, stream), "preprocess backward");
    if (clean) CU(cudaMemsetAsync(active_count, 0, sizeof(unsigned), stream), "reset the touched count");
    tm.mark();
    if (times) {
        CU(cudaStreamSynchronize(stream), "timing sync");
        times->render_bwd_ms = tm.ms(0); times->preprocess_bwd_ms = tm.ms(1);
    }
    return B200SPLAT_OK;
}

int b200splat_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                           uint8_t* present, void* stream_) {
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return fail(B200SPLAT_ERR_INVALID_ARG, "bad arguments");
    CU(launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream_), "mark_visible");
    return B200SPLAT_OK;
}

int b200splat_sh_grad_from_views(int32_t P, int32_t M, int32_t sh_degree, const float* means3D, int32_t n_views,
                                 const float* const* dcol, const float* campos, float* dL_dsh, void* stream_) {
    if (P < 0 || M <= 0 || sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1))
        return fail(B200SPLAT_ERR_INVALID_ARG, "bad P / M / sh_degree");
    if (n_views < 1 || n_views > B200SPLAT_MAX_VIEWS) return fail(B200SPLAT_ERR_INVALID_ARG, "n_views must be in [1, B200SPLAT_MAX_VIEWS]");
    if (P == 0) return B200SPLAT_OK;
    if (!means3D || !dcol || !campos || !dL_dsh) return fail(B200SPLAT_ERR_INVALID_ARG, "null argument");
    for (int v = 0; v < n_views; v++)
        if (!dcol[v]) return fail(B200SPLAT_ERR_INVALID_ARG, "null dcol pointer");
    if (!aligned16(dL_dsh)) return fail(B200SPLAT_ERR_INVALID_ARG, "dL_dsh must be 16-byte aligned");
    CU(launch_sh_grad_from_views(P, M, sh_degree, means3D, n_views, dcol, campos, dL_dsh, (cudaStream_t)stream_), "sh_grad_from_views");
    return B200SPLAT_OK;
}

int b200splat_densify_stats(int32_t P, const float* dL_dmeans2D, const int32_t* radii, float* xyz_gradient_accum,
                            float* denom, float* max_radii2D, void* stream_) {
    if (P < 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad P");
    if (P == 0) return B200SPLAT_OK;
    if (!dL_dmeans2D || !radii || !xyz_gradient_accum || !denom || !max_radii2D) return fail(B200SPLAT_ERR_INVALID_ARG, "null argument");
    CU(launch_densify_stats(P, dL_dmeans2D, radii, xyz_gradient_accum, denom, max_radii2D, (cudaStream_t)stream_), "densify_stats");
    return B200SPLAT_OK;
}

int b200splat_densify_stats_visible(int32_t P, const int32_t* radii, float* denom, float* max_radii2D, void* stream_) {
    if (P < 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad P");
    if (P == 0) return B200SPLAT_OK;
    if (!radii || !denom || !max_radii2D) return fail(B200SPLAT_ERR_INVALID_ARG, "null argument");
    CU(launch_densify_stats_visible(P, radii, denom, max_radii2D, (cudaStream_t)stream_), "densify_stats_visible");
    return B200SPLAT_OK;
}

int b200splat_render_stats(int32_t P, int32_t width, int32_t height, int32_t num_rendered, const void* geom,
                           const void* binning, const void* image, uint64_t* out16, void* stream_) {
    if (!out16 || P < 0 || width <= 0 || height <= 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad arguments");
    if (P == 0 || num_rendered <= 0) {
        CU(cudaMemsetAsync(out16, 0, 16 * sizeof(uint64_t), (cudaStream_t)stream_), "clear stats");
        return B200SPLAT_OK;
    }
    if (!geom || !binning || !image) return fail(B200SPLAT_ERR_INVALID_ARG, "null forward state");
    GeomView gv = carve_geom(const_cast<void*>(geom), P);
    BinView bv = carve_binning(const_cast<void*>(binning), num_rendered, tiles_of(width, height));
    ImgView iv = carve_image(const_cast<void*>(image), width, height);
    CU(launch_render_stats(width, height, gv, bv.vals_sorted, iv, reinterpret_cast<unsigned long long*>(out16), (cudaStream_t)stream_),
       "render stats");
    return B200SPLAT_OK;
}

int b200splat_export_state(int32_t P, int32_t width, int32_t height, int32_t num_rendered, const int32_t* radii,
                           const void* geom, const void* binning, const void* image, const b200splat_state_out* out,
                           void* stream_) {
    if (!out || P < 0) return fail(B200SPLAT_ERR_INVALID_ARG, "bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (geom && radii && P > 0) {
        GeomView gv = carve_geom(const_cast<void*>(geom), P);
        CU(launch_export_geom(P, radii, gv, *out, stream), "export geometry");
    }
    if (geom && image && P > 0) {
        GeomView gv = carve_geom(const_cast<void*>(geom), P);
        BinView bv = carve_binning(const_cast<void*>(binning), binning ? num_rendered : 0, tiles_of(width, height));
        ImgView iv = carve_image(const_cast<void*>(image), width, height);
        CU(launch_export_binning(P, width, height, gv, bv, iv, *out, stream), "export binning");
    }
    if (image) {
        ImgView iv = carve_image(const_cast<void*>(image), width, height);
        const size_t T = (size_t)((width + TILE - 1) / TILE) * ((height + TILE - 1) / TILE);
        const size_t N = (size_t)width * height;
        if (out->ranges) CU(cudaMemcpyAsync(out->ranges, iv.ranges, sizeof(int2) * T, cudaMemcpyDeviceToDevice, stream), "export ranges");
        if (out->final_T) CU(cudaMemcpyAsync(out->final_T, iv.final_T, sizeof(float) * N, cudaMemcpyDeviceToDevice, stream), "export final_T");
        if (out->n_contrib) CU(cudaMemcpyAsync(out->n_contrib, iv.n_contrib, sizeof(uint32_t) * N, cudaMemcpyDeviceToDevice, stream), "export n_contrib");
    }
    return B200SPLAT_OK;
}

}  // extern "C"
