/*
 * b200splat.h -- C ABI of libb200splat.so, the B200 (sm_100a) differentiable Gaussian-splat
 * rasterizer.  Plain pointers and sizes only: no C++ types, no torch types, no exceptions.
 *
 * What each entry point replaces in the reference (PoplarPoplar/3dgs_study):
 *
 *   The reference reaches this path through the Python package `diff_gaussian_rasterization`
 *   (import: /root/reference/gaussian_renderer/__init__.py:14; settings: :47-60; module: :62;
 *   call: :98-106), whose native half is the pybind module `_C` of the git submodule
 *   submodules/diff-gaussian-rasterization (pinned @ 2cf59156, /root/reference/.gitmodules:4-6;
 *   ABSENT from the snapshot, /root/reference/.SUBMODULES.json:16-22).  BASELINE.json's north_star
 *   names the two native calls this library stands in for:
 *
 *     b200splat_forward   <->  _C.rasterize_gaussians            (forward of the call at :98-106)
 *     b200splat_backward  <->  _C.rasterize_gaussians_backward   (autograd backward of the same call;
 *                                                                 its means2D gradient is consumed at
 *                                                                 /root/reference/scene/gaussian_model.py:576-580)
 *     b200splat_mark_visible <-> _C.mark_visible                 (frustum test; not called by the reference)
 *
 *   The binding a maintainer adds on the reference side is in INTEGRATION.md; the one this repo
 *   ships is 3dgs_study_b200/diff_gaussian_rasterization/_C.py (ctypes).
 *
 * Memory: every pointer in b200splat_view / b200splat_gaussians / outputs / gradients is a DEVICE
 * pointer owned by the caller, valid on `stream`'s device.  Scratch that must survive until the
 * backward pass is obtained through the caller's allocator callbacks (one per opaque buffer), so
 * the caller's caching allocator owns it.
 *
 * State: the library keeps no state between calls except, per host thread, an error string and (once a backward
 * pass has run) one helper stream + two events per device; b200splat_shutdown() releases those.
 * b200splat_accumulate_packed stages the row counts of its sources in 256 bytes of the module's own device memory
 * and is therefore not to be called concurrently from several host threads on the same device.
 *
 * Threading / streams: re-entrant; all work is enqueued on `stream` (a cudaStream_t passed as void*).
 *
 * Pair capacity.  The binning buffer holds one entry per (tile, Gaussian) pair; their number R is only known on
 * the device, after the preprocess kernel.  The call being replaced reads R back and blocks the host once per
 * view.  Here the caller chooses (b200splat_allocators.bin_capacity):
 *   bin_capacity == 0   exact: ONE host round trip for R (after the preprocess kernel and the depth sort have been
 *                       queued), the binning buffer is allocated for exactly R pairs, *num_rendered = R.
 *   bin_capacity  > 0   the binning buffer is allocated for that many pairs before R is known and the call never
 *                       waits for the device: R stays in the image buffer's control block, where the binning and
 *                       render kernels read it.  *num_rendered = bin_capacity (the value the backward needs to find
 *                       its way around the binning buffer).  If a view produces more pairs than the capacity, no
 *                       kernel writes out of bounds, the view's outputs are unspecified, and the overflow is
 *                       recorded: b200splat_forward_status reports R, the capacity and the flag (the same four
 *                       words can be copied asynchronously from the image buffer at b200splat_image_status_offset).
 *                       A caller that renders a sequence of similar views sizes the capacity from the R of the
 *                       previous views with head room and checks the flag one call late -- the policy of the
 *                       Python operator in this repo (diff_gaussian_rasterization/_C.py).
 */
#ifndef B200SPLAT_H
#define B200SPLAT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SPLAT_VERSION 4

#if defined(__GNUC__)
#define B200SPLAT_API __attribute__((visibility("default")))
#else
#define B200SPLAT_API
#endif

/* status codes */
#define B200SPLAT_OK 0
#define B200SPLAT_ERR_INVALID_ARG 1 /* bad sizes, missing/ambiguous inputs                     */
#define B200SPLAT_ERR_CUDA 2        /* a CUDA call or kernel failed; see b200splat_last_error  */
#define B200SPLAT_ERR_ALLOC 3       /* an allocator callback returned NULL                     */

/* Allocator callback: return a device pointer to at least `bytes` bytes, aligned to 256 B, that
 * stays valid until the matching backward call has completed (or NULL on failure). */
typedef void* (*b200splat_alloc_fn)(void* ctx, size_t bytes);

/* Per-view constants == the fields of GaussianRasterizationSettings
 * (/root/reference/gaussian_renderer/__init__.py:47-60). */
typedef struct b200splat_view {
    int32_t image_height;
    int32_t image_width;
    float tanfovx;
    float tanfovy;
    float scale_modifier;
    int32_t sh_degree;      /* active degree D (0..3), may be < the degree implied by M        */
    int32_t prefiltered;    /* accepted for API parity; culled Gaussians are simply skipped     */
    int32_t debug;          /* !=0: synchronise and check after every kernel                    */
    int32_t antialiasing;   /* !=0: opacity *= sqrt(max(2.5e-5, det(cov2D)/det(cov2D+0.3I)))    */
    const float* bg;         /* [3]                                                             */
    const float* viewmatrix; /* [16] world_view_transform, row-vector convention (p' = [p,1] @ M) */
    const float* projmatrix; /* [16] full_proj_transform, same convention                       */
    const float* campos;     /* [3]                                                             */
} b200splat_view;

/* Per-Gaussian inputs == the tensor arguments of GaussianRasterizer.forward (:98-106).
 * Exactly one of {shs, colors_precomp} and exactly one of {(scales, rotations), cov3D_precomp}
 * must be non-NULL. */
typedef struct b200splat_gaussians {
    int32_t P;                   /* number of Gaussians                                          */
    int32_t M;                   /* SH coefficients per Gaussian (tensor dim 1 of shs); 0 if none */
    const float* means3D;        /* [P,3]                                                        */
    const float* shs;            /* [P,M,3] coefficient-major, RGB innermost, or NULL            */
    const float* colors_precomp; /* [P,3] or NULL                                                */
    const float* opacities;      /* [P]                                                          */
    const float* scales;         /* [P,3] or NULL                                                */
    const float* rotations;      /* [P,4] unit quaternion (w,x,y,z) or NULL                      */
    const float* cov3D_precomp;  /* [P,6] (xx,xy,xz,yy,yz,zz) or NULL                            */
} b200splat_gaussians;

typedef struct b200splat_forward_out {
    float* color;     /* [3,H,W]                       */
    int32_t* radii;   /* [P]; 0 = culled               */
    float* invdepth;  /* [1,H,W] or NULL               */
} b200splat_forward_out;

/* The three opaque state buffers of the call being replaced (geometry / binning / image). */
typedef struct b200splat_allocators {
    b200splat_alloc_fn alloc;
    void* geom_ctx;
    void* binning_ctx;
    void* image_ctx;
    int64_t bin_capacity;   /* 0: exact (one host round trip for R); > 0: pairs to allocate for, no host wait */
} b200splat_allocators;

/* Optional per-stage device times in milliseconds (NULL to skip).  Requesting them adds one
 * synchronisation at the end of the call. */
typedef struct b200splat_stage_times {
    float preprocess_ms;
    float scan_ms;       /* depth order: stable sort of the visible Gaussians by depth + scan of tiles touched */
    float duplicate_ms;  /* duplicateWithKeys (in exact mode: includes the host round trip for the pair count)  */
    float sort_ms;       /* stable sort of the (tile, Gaussian) pairs by tile id, tile ranges                   */
    float ranges_ms;     /* tile order                                                                          */
    float render_ms;
    float render_bwd_ms;
    float preprocess_bwd_ms;
} b200splat_stage_times;

/* Forward.  On success *num_rendered = the pair capacity of the binning buffer: the number of (tile, Gaussian)
 * pairs R in exact mode, alloc->bin_capacity otherwise ("Pair capacity" above).  Pass it unchanged to the backward.
 * Images up to 8192 x 8192 pixels. */
B200SPLAT_API int b200splat_forward(const b200splat_view* view, const b200splat_gaussians* g,
                      const b200splat_forward_out* out, const b200splat_allocators* alloc,
                      int32_t* num_rendered, b200splat_stage_times* times /* nullable */,
                      void* stream);

/* Pair count R, pair capacity and overflow flag of the view whose image buffer is given.  Synchronises `stream`.
 * Any output pointer may be NULL. */
B200SPLAT_API int b200splat_forward_status(const void* image, int32_t width, int32_t height, int64_t* num_pairs,
                      int64_t* capacity, int32_t* overflow, void* stream);
/* Byte offset, inside the image buffer, of the four u32 status words {R low, R high, capacity, overflow} the forward
 * leaves there: a caller may copy them to pinned host memory on its own stream instead of synchronising. */
B200SPLAT_API size_t b200splat_image_status_offset(int32_t width, int32_t height);

/* Background clear: zeroes up to 8 device ranges with ONE block per SM, so that the clear can run beside a
 * compute-bound kernel on another stream (a full-grid memset takes every thread slot of the GPU while it runs).
 * Used for the dense gradient outputs, which are cleared while the render backward runs. */
B200SPLAT_API int b200splat_zero_fill(int32_t n, void* const* ptrs, const size_t* bytes, void* stream);

/* Releases what the calling host thread holds inside the library (helper streams and events). */
B200SPLAT_API void b200splat_shutdown(void);

/* Gradient outputs; each pointer may be NULL if the caller does not need it, except
 * dL_dmeans3D.  Every non-NULL output is fully written (zeros for culled Gaussians): the caller
 * need not clear them. */
typedef struct b200splat_grads {
    float* dL_dmeans2D;   /* [P,3]  xy = dL/d(NDC xy) (pixel gradient x 0.5W, 0.5H), z = 0      */
    float* dL_dcolors;    /* [P,3]  gradient of colors_precomp; with shs and dL_dsh == NULL: the colour
                           *        gradient times the SH clamp mask (see b200splat_sh_grad_from_views)   */
    float* dL_dopacity;   /* [P]                                                                  */
    float* dL_dmeans3D;   /* [P,3]                                                                */
    float* dL_dcov3D;     /* [P,6]  gradient of cov3D_precomp                                    */
    float* dL_dsh;        /* [P,M,3] (may be NULL with shs if dL_dcolors is given)                        */
    float* dL_dscales;    /* [P,3]                                                                */
    float* dL_drotations; /* [P,4]                                                                */
} b200splat_grads;

/* Backward.  geom / binning / image are the buffers the forward obtained through the allocator;
 * radii is the forward's output.  scratch_alloc provides transient scratch for this call. */
B200SPLAT_API int b200splat_backward(const b200splat_view* view, const b200splat_gaussians* g,
                       const int32_t* radii, const void* geom, const void* binning,
                       const void* image, int32_t num_rendered,
                       const float* dL_dout_color /* [3,H,W] */,
                       const float* dL_dout_invdepth /* [1,H,W] or NULL */,
                       const b200splat_grads* grads, b200splat_alloc_fn scratch_alloc,
                       void* scratch_ctx, b200splat_stage_times* times /* nullable */, void* stream);

/* Backward with an early hand-off for the view-parallel exchange.  If hooks->masked_colors is non-NULL (SH mode
 * only), the colour gradient times the SH clamp mask ([P,3], zeros for culled Gaussians) is written right after
 * the render backward, BEFORE the per-Gaussian backward is enqueued, and hooks->after_masked_colors (if non-NULL)
 * is then called on the host: the caller records an event on `stream` there and starts exchanging that buffer
 * on another stream while this call goes on to enqueue the per-Gaussian backward.  With masked_colors given,
 * grads->dL_dsh and grads->dL_dcolors may both be NULL. */
typedef struct b200splat_backward_hooks {
    float* masked_colors;
    void (*after_masked_colors)(void* ctx);
    void* ctx;
} b200splat_backward_hooks;

B200SPLAT_API int b200splat_backward_ex(const b200splat_view* view, const b200splat_gaussians* g,
                       const int32_t* radii, const void* geom, const void* binning,
                       const void* image, int32_t num_rendered,
                       const float* dL_dout_color, const float* dL_dout_invdepth,
                       const b200splat_grads* grads, const b200splat_backward_hooks* hooks /* nullable */,
                       b200splat_alloc_fn scratch_alloc, void* scratch_ctx,
                       b200splat_stage_times* times /* nullable */, void* stream);

/* present[i] = 1 if Gaussian i passes the near-plane test (view z > 0.2), else 0. */
B200SPLAT_API int b200splat_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                           const float* projmatrix, uint8_t* present, void* stream);

/* Parity / inspection: copy pieces of the opaque forward state into caller device buffers (any
 * pointer may be NULL).  Sorted keys are `(tile << 32) | float_bits(depth)`. */
typedef struct b200splat_state_out {
    float* means2D;           /* [P,2] pixel coordinates                  */
    float* depths;            /* [P]                                      */
    float* conic_opacity;     /* [P,4] (A,B,C,effective opacity)          */
    float* rgb;               /* [P,3]                                    */
    uint32_t* tiles_touched;  /* [P]                                      */
    uint32_t* point_offsets;  /* [P] inclusive scan of tiles_touched      */
    uint64_t* keys_sorted;    /* [R]                                      */
    uint32_t* vals_sorted;    /* [R]                                      */
    int32_t* ranges;          /* [T,2]                                    */
    float* final_T;           /* [H,W]                                    */
    uint32_t* n_contrib;      /* [H,W]                                    */
} b200splat_state_out;

/* num_rendered: the value the forward returned (the pair capacity); keys_sorted / vals_sorted need room for R
 * entries (b200splat_forward_status). */
B200SPLAT_API int b200splat_export_state(int32_t P, int32_t width, int32_t height, int32_t num_rendered,
                           const int32_t* radii, const void* geom, const void* binning,
                           const void* image, const b200splat_state_out* out, void* stream);

/* View-parallel gradient exchange (one view per GPU, Gaussians replicated; the reference itself is single-GPU,
 * /root/reference/train.py:86-98).  The SH gradient of a view is an outer product,
 *   dL/dsh[i][k][c] = Y_k(normalize(mean_i - campos_v)) * dcol_v[i][c],
 * with dcol_v the colour gradient of view v already multiplied by the SH clamp mask -- what b200splat_backward
 * writes to grads->dL_dcolors when called with shs != NULL and grads->dL_dsh == NULL.  Ranks therefore exchange
 * the [P,3] colour gradients (12 B per Gaussian and view instead of 12*M B) and each rebuilds
 *   dL_dsh[P,M,3] = sum_v Y(dir_v) (x) dcol_v
 * with this call.  dcol[v] are n_views DEVICE pointers (host array), which may address peer GPUs' memory
 * mapped into this process (NVLink loads); campos is a device array [n_views,3].  dL_dsh is fully written. */
#define B200SPLAT_MAX_VIEWS 16
B200SPLAT_API int b200splat_sh_grad_from_views(int32_t P, int32_t M, int32_t sh_degree, const float* means3D,
                           int32_t n_views, const float* const* dcol, const float* campos,
                           float* dL_dsh, void* stream);

/* Sparse view-parallel exchange (what the multi-GPU training step uses; the reference is single-GPU,
 * /root/reference/train.py:86-98).  One view touches a small fraction of the Gaussians (bench scene: 7 %), and
 * every gradient of an untouched Gaussian is exactly zero, so a rank publishes "packed gradient rows": one
 * 64-byte row of 16 f32 words per Gaussian its view touched,
 *   [0] index i (u32 bits)   [1..3] colour gradient x SH clamp mask   [4..7] dL/drotation
 *   [8..10] dL/dmean3D       [11..13] dL/dscale                       [14] dL/dopacity      [15] 0
 * b200splat_backward_packed is b200splat_backward producing these rows (packed_rows: room for P rows, 16-byte
 * aligned; *packed_count: their number) instead of the dense gradients; SH + scale/rotation inputs only.
 * dL_dmeans2D ([P,3], nullable) is still written densely: the densification statistics are rank-local.
 * b200splat_accumulate_packed ADDS the rows of n_src views into dense gradients the caller has cleared:
 * dL_dsh [P,M,3] (M <= 16), rebuilt on the spot as Y(normalize(mean_i - campos_v)) (x) colour gradient, and
 * dL_dgeom [P,12] = words 4..15 of the rows, i.e. per Gaussian (dL/drotation[4], dL/dmean3D[3], dL/dscale[3],
 * dL/dopacity, 0) interleaved -- one 48-byte run per Gaussian instead of four arrays, because the L2 retires
 * a bounded number of atomic sectors per second and the exchange is bound by exactly that.  rows[v] / counts[v] are DEVICE
 * pointers (host arrays of n_src entries) and may address a peer GPU's memory mapped into this process: the
 * kernel then reads the peer's rows over NVLink, i.e. the exchange is this kernel -- there is no collective.
 * campos: device array [n_src,3], the camera centre of each source view; or NULL: each counts[v] then points at a 64-byte
 * header {row count, 0, 0, 0, cx, cy, cz, ...} and the kernel takes the camera centre of view v from words 4..6 of it
 * (the owner of the rows writes it there, so the ranks need no collective to learn each other's cameras).
 * store != 0 (n_src must be 1): the rows are WRITTEN instead of added -- for the first source into freshly
 * cleared gradients, e.g. a rank's own rows while it waits for its peers; nothing else may write the
 * gradients while that launch runs. */
B200SPLAT_API int b200splat_backward_packed(const b200splat_view* view, const b200splat_gaussians* g,
                       const int32_t* radii, const void* geom, const void* binning, const void* image,
                       int32_t num_rendered, const float* dL_dout_color,
                       const float* dL_dout_invdepth /* [1,H,W] or NULL */, float* dL_dmeans2D,
                       float* packed_rows, uint32_t* packed_count, b200splat_alloc_fn scratch_alloc,
                       void* scratch_ctx, b200splat_stage_times* times /* nullable */, void* stream);
/* b200splat_backward_packed with a caller-owned CLEAN scratch instead of an allocator (a training step calls the
 * backward thousands of times with the same P; the reference allocates and clears its backward buffers per call,
 * /root/reference/gaussian_renderer/__init__.py:98 -> rasterize_gaussians_backward).  `scratch`: device memory of at least
 * b200splat_backward_scratch_bytes(P) bytes, 16-byte aligned, ALL ZERO before its first use.  The call leaves it clean
 * again -- the per-Gaussian backward re-zeroes exactly the accumulator rows it consumed, the compaction the flags it
 * read -- so the next call needs no clear: the 150 MB memset in front of the render backward (3 M Gaussians) is gone.
 * If the call returns an error the scratch must be zero-filled by the caller before it is used again.  One scratch per
 * stream: calls that may overlap on the device need their own.
 * packed_count must point at the 64-byte header b200splat_accumulate_packed reads with campos = NULL: the call writes
 * the row count to word 0 and the view's camera centre to words 4..6.
 * xyz_gradient_accum != NULL: the per-Gaussian backward also adds ||dL/dmean2D[i]|| of every Gaussian it visits into
 * xyz_gradient_accum[i] -- the gradient-norm line of b200splat_densify_stats (a Gaussian it does not visit has a zero
 * gradient); together with b200splat_densify_stats_visible after the forward that is the whole statistics update, and
 * dL_dmeans2D can be NULL (no dense [P,3] output to clear and write). */
B200SPLAT_API size_t b200splat_backward_scratch_bytes(int32_t P);
B200SPLAT_API int b200splat_backward_packed_clean(const b200splat_view* view, const b200splat_gaussians* g,
                       const int32_t* radii, const void* geom, const void* binning, const void* image,
                       int32_t num_rendered, const float* dL_dout_color,
                       const float* dL_dout_invdepth /* [1,H,W] or NULL */, float* dL_dmeans2D /* nullable */,
                       float* xyz_gradient_accum /* [P] or NULL */, float* packed_rows, uint32_t* packed_count,
                       void* scratch, size_t scratch_bytes, b200splat_stage_times* times /* nullable */, void* stream);
B200SPLAT_API int b200splat_accumulate_packed(int32_t P, int32_t M, int32_t sh_degree, const float* means3D,
                       int32_t n_src, const float* const* rows, const uint32_t* const* counts,
                       const float* campos, float* dL_dsh, float* dL_dgeom, uint8_t* marks /* [P] or NULL */,
                       int32_t store, void* stream);

/* Sparse clear of gradients that b200splat_accumulate_packed filled.  With marks != NULL the accumulate calls set
 * marks[i] = 1 for every Gaussian they add (or store) a row into; this call zeroes dL_dsh[i], dL_dgeom[i] and marks[i]
 * of exactly those Gaussians.  For buffers (and marks) that were all zero before the accumulate calls and that nobody
 * else wrote into, everything is all zero again afterwards -- at the cost of the rows a step touched (7 % of 3 M
 * Gaussians on the bench scene: 3 MB of marks read, 53 MB of rows written) instead of a clear of everything (0.85 GB,
 * which also slows whatever runs next to it: measured 0.1 ms per step beside the render backward). */
B200SPLAT_API int b200splat_clear_accumulated(int32_t P, int32_t M, float* dL_dsh, float* dL_dgeom, uint8_t* marks,
                       void* stream);

/* Densification statistics of one rendered view, fused (consumers of the operator's outputs in the reference:
 * add_densification_stats, /root/reference/scene/gaussian_model.py:576-581, and the max_radii2D update,
 * /root/reference/train.py:127).  For every Gaussian with radii > 0:
 *   xyz_gradient_accum[i] += ||dL_dmeans2D[i, :2]||, denom[i] += 1, max_radii2D[i] = max(max_radii2D[i], radii[i]).
 * All arrays are device pointers of P (dL_dmeans2D: [P,3]) elements. */
B200SPLAT_API int b200splat_densify_stats(int32_t P, const float* dL_dmeans2D, const int32_t* radii,
                           float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream);

/* The lines of b200splat_densify_stats that need the forward's radii only: denom[i] += 1 and
 * max_radii2D[i] = max(max_radii2D[i], radii[i]) for radii[i] > 0.  Can run as soon as the forward has been queued. */
B200SPLAT_API int b200splat_densify_stats_visible(int32_t P, const int32_t* radii, float* denom, float* max_radii2D,
                           void* stream);

/* Inspection: workload counters of the render stage of the view whose forward state is given, written
 * to out16 (device, 16 x u64): [0..8] (tile, Gaussian) pairs by number of 8x4-pixel sub-blocks the splat's
 * {alpha >= 1/255} box reaches, [9] pairs visited, [10] pixel evaluations, [11] of those with an exp,
 * [12] of those contributing (alpha >= 1/255), [13] pairs with at least one contribution, [14] evaluations whose
 * alpha >= 1/255 test comes out differently with the backward's fast exponential than with the forward's expf (a
 * count for DESIGN.md, not a gate), [15] reserved.  These are the flop / atomic counts behind the issue-rate
 * roofline of the render kernels; not on the hot path. */
B200SPLAT_API int b200splat_render_stats(int32_t P, int32_t width, int32_t height, int32_t num_rendered,
                           const void* geom, const void* binning, const void* image,
                           uint64_t* out16, void* stream);

/* Sizes of the geometry and image buffers for given dimensions (the binning buffer depends on R and,
 * through the width of the tile key, on the number of tiles). */
B200SPLAT_API size_t b200splat_geom_bytes(int32_t P);
B200SPLAT_API size_t b200splat_image_bytes(int32_t width, int32_t height);
B200SPLAT_API size_t b200splat_binning_bytes(int32_t num_rendered, int32_t width, int32_t height);

B200SPLAT_API const char* b200splat_last_error(void);
B200SPLAT_API int b200splat_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200SPLAT_H */
