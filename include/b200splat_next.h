/*
 * b200splat_next.h -- C ABI of the steps either side of the rasterizer ("next"),
 * implemented by the same libb200splat.so.  Plain pointers and sizes only.
 *
 * What each entry point replaces in the reference (PoplarPoplar/3dgs_study):
 *
 *   b200splat_l1_ssim_forward / _backward
 *       <-> utils/loss_utils.py  l1_loss (:17-24), ssim / _ssim (:59-108; window :35-57), as combined by the
 *           training loss at train.py:103-104 and differentiated by loss.backward() (train.py:105).
 *           One forward kernel + one backward kernel instead of five grouped conv2d and ~25 elementwise
 *           kernels each way.
 *   b200splat_adam_step
 *       <-> gaussians.optimizer.step() (train.py:137-141) on the optimizer built at
 *           scene/gaussian_model.py:185-197: torch.optim.Adam(lr=0, eps=1e-15) over six parameter groups
 *           with per-group learning rates (xyz rescheduled every iteration, :207-214).  All groups in ONE launch.
 *   b200splat_dist2_knn3
 *       <-> simple_knn._C.distCUDA2 (submodules/simple-knn/spatial.cu:15-26 -> simple_knn.cu:186-221), called
 *           at scene/gaussian_model.py:153 to initialise the scales.
 *   b200splat_dist2_knn3_static
 *       no counterpart in the reference: the same search with built-in scratch, no allocation.
 *   b200splat_ply_write / _read_header / _read   (HOST memory; no GPU involved)
 *       <-> GaussianModel.save_ply / load_ply (scene/gaussian_model.py:240-259, 267-320), which go through
 *           the third-party `plyfile` package (absent from this image).
 *
 * Device pointers are owned by the caller and valid on `stream`'s device; all device work is enqueued on
 * `stream` (a cudaStream_t passed as void*) and nothing synchronises with the host.
 * Status codes and b200splat_last_error() are those of b200splat.h.
 */
#ifndef B200SPLAT_NEXT_H
#define B200SPLAT_NEXT_H

#include "b200splat.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- L1 + SSIM ------------------------------------------------------------------------------------
 * `planes` = B*C images of height x width, contiguous (the [C,H,W] or [B,C,H,W] tensor of the reference).
 * window11: HOST pointer to the 11 taps of the 1-D Gaussian window (loss_utils.py:35-45); a shorter odd
 * window is passed centred and zero padded.  The reference's 2-D window is the outer product of these taps
 * (:54) and its convolutions are zero padded (:83-93): the kernels apply the taps separably, x then y.
 *
 * forward:  out[2p] = sum of |img-gt| over plane p, out[2p+1] = sum of the SSIM map over plane p (device,
 *           2*planes floats, accumulated in double in a fixed order; the reference's global / per-batch means
 *           are sums of these divided by the element count).  maps (device, 3*planes*H*W floats, or NULL when no gradient is wanted) receives
 *           d m/d mu1, d m/d conv(img^2), d m/d conv(img*gt) of the SSIM map m.  partials: device scratch of
 *           b200splat_l1_ssim_partials(planes, H, W) float pairs.
 * backward: dL_dimg[p] = scale[2p] * sign(img-gt) + scale[2p+1] * d(sum of the SSIM map of plane p)/d img,
 *           scale = device array of 2*planes floats = the upstream gradients of out (1/count for a plain mean);
 *           a device array so that no host synchronisation is needed to read them.
 */
B200SPLAT_API size_t b200splat_l1_ssim_partials(int32_t planes, int32_t height, int32_t width);
B200SPLAT_API int b200splat_l1_ssim_forward(int32_t planes, int32_t height, int32_t width, const float* img,
                           const float* gt, const float* window11, float* maps, float* partials, float* out,
                           void* stream);
B200SPLAT_API int b200splat_l1_ssim_backward(int32_t planes, int32_t height, int32_t width, const float* img,
                           const float* gt, const float* window11, const float* maps, const float* scale,
                           float* dL_dimg, void* stream);

/* The training loss of train.py:103-104 in one go: loss3 (device, 3 floats) = (loss, Ll1, ssim) with
 * loss = (1 - lambda_dssim) * Ll1 + lambda_dssim * (1 - ssim), means over all planes*H*W elements.  The
 * backward multiplies by `upstream` (device scalar = d L / d loss, NULL = 1), so loss.backward() needs no
 * elementwise glue kernels at all.  maps / partials as above. */
B200SPLAT_API int b200splat_training_loss_forward(int32_t planes, int32_t height, int32_t width, const float* img,
                           const float* gt, const float* window11, float lambda_dssim, float* maps, float* partials,
                           float* loss3, void* stream);
B200SPLAT_API int b200splat_training_loss_backward(int32_t planes, int32_t height, int32_t width, const float* img,
                           const float* gt, const float* window11, float lambda_dssim, const float* maps,
                           const float* upstream, float* dL_dimg, void* stream);

/* ---- Activations of the raw parameters --------------------------------------------------------------
 * GaussianModel.get_scaling / get_rotation / get_opacity (scene/gaussian_model.py:106-127 over the
 * functions set at :34-42): scales = exp(scaling), rotations = normalize(rotation) = rotation /
 * max(||rotation||_2, 1e-12), opacities = sigmoid(opacity); [P,3], [P,4], [P,1] device arrays.  The backward
 * takes the activated values the forward produced and the gradients the rasterizer returned for them; any of
 * the three gradient inputs may be NULL (that output is then left untouched). */
B200SPLAT_API int b200splat_activate_forward(int32_t P, const float* scaling, const float* rotation,
                           const float* opacity, float* scales, float* rotations, float* opacities, void* stream);
B200SPLAT_API int b200splat_activate_backward(int32_t P, const float* rotation, const float* scales,
                           const float* opacities, const float* g_scales, const float* g_rotations,
                           const float* g_opacities, float* d_scaling, float* d_rotation, float* d_opacity,
                           void* stream);

/* ---- Adam -----------------------------------------------------------------------------------------
 * One optimizer step over up to B200SPLAT_ADAM_MAX_GROUPS tensors in one launch (torch.optim.Adam
 * formulas, no weight decay, no amsgrad):
 *   m += (g - m)(1 - beta1);  v = beta2 v + (1 - beta2) g g;
 *   p -= (lr / (1 - beta1^step)) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps)
 * with the roundings of torch's CUDA kernels as nvcc compiles them (the platform the reference runs on):
 *   m = fma(g - m, 1-beta1, m);  v = fma((1-beta2) g, g, rn(v beta2));  p = fma(-step_size, m / denom, p).
 * `step` (per group) is the 1-based count of this step.  grad == NULL skips a group (a parameter without .grad, as
 * torch does).  The scalar factors are evaluated in double and rounded to fp32 once, like torch's.
 */
#define B200SPLAT_ADAM_MAX_GROUPS 8
typedef struct {
    float* param;        /* device, n elements, updated in place */
    const float* grad;   /* device, n elements                   */
    float* exp_avg;      /* device, n elements, updated in place */
    float* exp_avg_sq;   /* device, n elements, updated in place */
    int64_t n;
    double lr;
    int32_t step;        /* per-group step count (groups re-created by densification restart at 1) */
    int32_t row_width;   /* 0: `lr` for every element.  > 0: the tensor is [n / row_width, row_width] and   */
    int32_t row_split;   /*    columns >= row_split use lr_tail -- one [P, M*3] SH tensor trained with the   */
    int32_t reserved;    /*    f_dc / f_rest learning rates of gaussian_model.py:187-188 without splitting it */
    double lr_tail;
} b200splat_adam_group;
B200SPLAT_API int b200splat_adam_step(int32_t n_groups, const b200splat_adam_group* groups, double beta1,
                           double beta2, double eps, void* stream);

/* ---- distCUDA2 ------------------------------------------------------------------------------------
 * out[i] = mean of the squared distances from point i to its 3 nearest other points (device, P floats);
 * xyz = device [P,3].  Exact search (Morton order, two-level bounding boxes), distances evaluated as
 * fma(dz,dz, fma(dx,dx, dy*dy)) -- the instruction sequence nvcc emits for the reference's expression -- so
 * results are bit-identical to the reference's own build and to an exhaustive search.  Scratch comes from the stream-ordered CUDA allocator. */
B200SPLAT_API int b200splat_dist2_knn3(int32_t P, const float* xyz, float* out, void* stream);

/* ---- distCUDA2 without any allocation --------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot reach an
 * allocator at all (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments, same kernels and bit-identical results as b200splat_dist2_knn3, but the scratch (Morton codes, sorted
 * ids, gathered points, boxes, sort workspace) lives in a 12 MiB array inside the library's own device image, which
 * every GPU that loads the library holds until exit; the call performs no cudaMalloc / cudaFree / stream-ordered
 * allocation.  P > B200SPLAT_KNN_STATIC_MAX_POINTS is rejected with B200SPLAT_ERR_INVALID_ARG.  All calls on one GPU
 * share that array: calls that overlap on the device (different streams, no ordering between them) race -- order them.
 * Nothing has to be cleared between calls (every scratch word read is written earlier in the same call). */
#define B200SPLAT_KNN_STATIC_MAX_POINTS 262144
B200SPLAT_API int b200splat_dist2_knn3_static(int32_t P, const float* xyz, float* out, void* stream);

/* ---- Gaussian PLY file (host memory) --------------------------------------------------------------
 * binary_little_endian 1.0, one `vertex` element, float properties x y z nx ny nz f_dc_0..2
 * f_rest_0..(n_rest-1) opacity scale_0..2 rot_0..3 (scene/gaussian_model.py:218-238), raw pre-activation
 * values.  f_dc is [P,1,3] and f_rest [P,n_rest/3,3] in memory; the file holds them channel-major
 * (transpose(1,2).flatten, :245-246), and the reader transposes back (:291-293, 310-313).  Normals are written as
 * zeros and ignored on read.  The reader finds properties by NAME (any order, extra properties skipped).
 */
B200SPLAT_API int b200splat_ply_write(const char* path, int64_t P, int32_t n_rest, const float* xyz,
                           const float* f_dc, const float* f_rest, const float* opacity, const float* scale,
                           const float* rot);
B200SPLAT_API int b200splat_ply_read_header(const char* path, int64_t* P, int32_t* n_rest);
B200SPLAT_API int b200splat_ply_read(const char* path, int64_t P, int32_t n_rest, float* xyz, float* f_dc,
                           float* f_rest, float* opacity, float* scale, float* rot);

#ifdef __cplusplus
}
#endif
#endif /* B200SPLAT_NEXT_H */
