/* rain_rast.h — C ABI of librainrast.so, the sm_100a replacement for the two CUDA extensions of sharonal10/rain:
 *   submodules/diff_gaussian_rasterization  (pybind module `_C`, ext.cpp:4-8)
 *   submodules/simple-knn                   (pybind module `simple_knn._C`, ext.cpp:4-6)
 *
 * Conventions
 *  - every pointer is a raw DEVICE pointer (tensor.data_ptr()) unless its name ends in `_host`;
 *  - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream); the reference always
 *    used the legacy default stream (forward.cu:386,428; rasterizer_impl.cu:281,306);
 *  - the library never allocates: the caller owns outputs and the three opaque scratch blobs (geometry / binning /
 *    image) that the reference grew through std::function<char*(size_t)> callbacks (rasterizer.h:21-23,
 *    rasterize_points.cu:16-22); their sizes come from the rr_*_bytes functions;
 *  - functions return 0 on success and a negative value on error; rr_last_error() gives the message
 *    (the reference threw std::runtime_error / AT_ERROR through pybind, rasterize_points.cu:47-49,
 *    rasterizer_impl.cu:233-236, auxiliary.h:155-162);
 *  - null `shs` / `colors_precomp` and null `scales`+`rotations` / `cov3D_precomp` select the modes exactly like the
 *    nullptr tests of the reference (forward.cu:195,231; backward.cu:380,384);
 *  - `debug != 0` makes each entry synchronise the stream and report asynchronous CUDA errors (CHECK_CUDA semantics,
 *    auxiliary.h:155-162).
 * The library has one kernel per stage and no run-time kernel selection; its single environment switch is the opt-in
 * RAIN_TILE_CULL=1 (read once per process), which shortens the internal tile lists and therefore is off by default.
 */
#ifndef RAIN_RAST_H
#define RAIN_RAST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RR_BWD_ACCUMULATE 1
#define RR_BWD_LEAVE_ROOM 4  /* rr_backward_views: cap the kernel at 3 CTAs per SM so that a CTA of a concurrent rr_allreduce fits beside it */
#define RR_BWD_LEAF 2        /* rr_backward_views: inputs are the raw leaf tensors, gradients w.r.t. them */
#define RR_GEOM_LEAF 2       /* rr_forward_geom_views: inputs are the raw leaf tensors (activations applied in the kernel) */
#define RR_GEOM_KEEP_COV3D 1 /* rr_forward_geom_views: also store the 3D covariance in every view's blob (the per-view rr_backward reads it) */

const char* rr_last_error(void);
int rr_version(void);

/* number of CUDA kernels this library has launched in this process so far */
int64_t rr_launch_count(void);
/* 1 if the opt-in exact tile culling is on in this process (RAIN_TILE_CULL=1, read once; see the header comment) */
int rr_tile_cull_enabled(void);

/* Optional per-stage device timing: when enabled every entry point brackets each of its stages with CUDA events on the
 * caller's stream (no reference counterpart; the reference only has CHECK_CUDA syncs, auxiliary.h:155-162).
 * rr_prof_read synchronises the recorded events and returns accumulated milliseconds and call counts per stage
 * (arrays of rr_prof_num_stages() entries; either may be null), optionally resetting the accumulators. */
int rr_prof_enable(int on);
int rr_prof_num_stages(void);
const char* rr_prof_stage_name(int i);
int rr_prof_read(double* ms, int64_t* calls, int reset);

/* scratch sizes: replace required<GeometryState>(P), required<ImageState>(W*H), required<BinningState>(N)
 * (rasterizer_impl.h:55-62, rasterizer_impl.cu:144-183) */
size_t rr_geom_bytes(int P);
size_t rr_image_bytes(int W, int H);
size_t rr_binning_bytes(int64_t N, int W, int H);
size_t rr_backward_ws_bytes(int P);
size_t rr_geom_acc_offset(int P);      /* byte offset of the backward accumulator region inside the geometry blob (rr_backward_render) */

/* Forward, first half — replaces FORWARD::preprocess + cub InclusiveSum + the blocking cudaMemcpy of num_rendered
 * (Rasterizer::forward, rasterizer_impl.cu:187-273).  Writes radii[P] and the geometry blob, and enqueues an async
 * copy of num_rendered into *num_rendered_host (pinned host memory); the caller synchronises the stream before
 * reading it and sizing the binning blob.  The image blob (rr_image_bytes) already receives the tile ranges here: the "tile
 * plan" derives them, and the digit bases of the later tile sort, from the tile rects alone (replacing identifyTileRanges and
 * the sort's histogram pass, rasterizer_impl.cu:302-318). */
int rr_forward_geom(void* stream, int P, int D, int M, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass, int prefiltered, int debug,
	int* radii, void* geom, size_t geom_bytes, void* image, size_t image_bytes, int64_t* num_rendered_host);

/* rr_forward_geom for ALL nviews views of a step in one set of launches (no reference counterpart: the reference runs its
 * preprocess, scan and sort once per view): one pass over the Gaussians that reads the parameters once and projects them
 * through every camera, the nviews depth sorts as one segmented onesweep, the nviews prefix sums as one launch.  Every array
 * argument is a HOST array of nviews entries (viewmatrix / projmatrix / campos / radii / geom: device pointers; tan_fovx /
 * tan_fovy: floats; num_rendered_host[v]: pinned host int64 receiving view v's count).  Each view's outputs (radii[v], its
 * geometry blob) are bit-identical to rr_forward_geom's for that camera.  SH colours + scales / rotations only, M = 16,
 * nviews <= 16; flags: RR_GEOM_KEEP_COV3D, RR_GEOM_LEAF.
 * With RR_GEOM_LEAF the inputs are the six RAW leaf tensors of scene/gaussian_model.py:131-136 — means3D = _xyz, shs =
 * _features_dc [P,1,3], f_rest = _features_rest [P,15,3], opacities / scales / rotations = _opacity / _scaling / _rotation before
 * their activations — and sigmoid / exp / normalise (:85-105) are applied inside (f_rest is ignored otherwise; pass null).
 * The outputs then equal rr_forward_geom's on the activated tensors rr_activate produces, bit for bit. */
int rr_forward_geom_views(void* stream, int P, int D, int M, int W, int H, int nviews,
	const float* const* viewmatrix, const float* const* projmatrix, const float* const* campos, const float* tan_fovx, const float* tan_fovy,
	const float* means3D, const float* shs, const float* f_rest, const float* opacities, const float* scales, float scale_modifier, const float* rotations, float low_pass,
	int* const* radii, void* const* geom, size_t geom_bytes, void* const* image, size_t image_bytes, int64_t* const* num_rendered_host, int flags, int debug);

/* Forward, second half — replaces duplicateWithKeys + cub SortPairs + identifyTileRanges + FORWARD::render
 * (rasterizer_impl.cu:275-328).  out_color[3,H,W], out_depth[1,H,W]. */
int rr_forward_raster(void* stream, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	float* out_color, float* out_depth, int debug);

/* The two parts of rr_forward_raster as separate calls, so that a caller rendering a batch of views can run the binning of
 * the next view (HBM bound) on a second stream beside the render kernels of the current one (issue bound):
 *   rr_forward_bin     duplicateWithKeys + SortPairs + identifyTileRanges (rasterizer_impl.cu:275-318)
 *   rr_forward_render  FORWARD::render (rasterizer_impl.cu:320-328); must be ordered after rr_forward_bin of the same blobs
 * rr_forward_raster(...) == rr_forward_bin(...) then rr_forward_render(...) on one stream. */
int rr_forward_bin(void* stream, int P, int W, int H, int64_t N,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes, int debug);
int rr_forward_render(void* stream, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	float* out_color, float* out_depth, int debug);

/* Backward — replaces Rasterizer::backward (rasterizer_impl.cu:334-430) and the torch::zeros fills of
 * RasterizeGaussiansBackwardCUDA (rasterize_points.cu:145-153): every element of every gradient buffer is written.
 * flags & RR_BWD_ACCUMULATE (view-sharded batches, no reference counterpart): the parameter gradients (dL_dmean3D,
 * dL_dsh / dL_dcolor, dL_dopacity, dL_dscale, dL_drot / dL_dcov3D) are ADDED to the buffers' contents and Gaussians
 * with radii == 0 are left untouched; the per-view outputs dL_dmean2D and dL_dconic are still overwritten.
 * dL_dconic may be null (the reference keeps it internal).  dL_dsh is [P,M,3] (unused when shs is null),
 * dL_dscale / dL_drot may be null when cov3D_precomp is given; dL_dcolor may be null when shs is used and dL_dcov3D when
 * cov3D_precomp is not given (gradients of inputs the caller did not pass: intermediates, skipped).  ws: rr_backward_ws_bytes(P) of scratch, or NULL for the geometry blob's own accumulator region (first backward after a forward only, see rr_backward_render)
 * (the per-Gaussian gradient accumulators the backward render adds into; zero-filled inside). */
int rr_backward(void* stream, int P, int D, int M, int W, int H, int64_t N, const float* background,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass, const int* radii,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	const float* dL_dpix, void* ws, size_t ws_bytes,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
	int debug, int flags);

/* ---- view-batched backward (multi-view steps; no reference counterpart: the reference runs its whole backward once per view) ----
 * rr_backward_render is the first half of rr_backward alone — BACKWARD::render (rasterizer_impl.cu:386-404, backward.cu:389-547):
 * it zero-fills ws (rr_backward_ws_bytes(P)) and leaves there, per Gaussian, the 48-byte accumulator record of this view
 * (colour gradient and the moments of G * dL_dalpha over the pixels that blended it).  ws == NULL (ws_bytes 0): the records go
 * to the accumulator region of the view's geometry blob (byte offset rr_geom_acc_offset(P), rr_backward_ws_bytes(P) - 256 bytes),
 * which rr_forward_render / rr_forward_raster of the same view cleared while it rendered (its stores ride on idle HBM
 * bandwidth; a separate fill costs 13 us per million Gaussians): valid for ONE rr_backward_render per forward of the view.
 * rr_backward_views is the second half — computeCov2DCUDA + preprocessCUDA backward (backward.cu:133-386) — for ALL nviews
 * views of a step in ONE pass over the Gaussians: every array argument up to dL_dmean2D is a HOST array of nviews entries
 * (device pointers, resp. floats); geom[v] / radii[v] / grad_acc[v] are view v's geometry blob, radii and the ws its
 * rr_backward_render filled (all views share P, W, H, D, scale_modifier, low_pass).  Parameters are read once, the bucket row
 * (59 floats per Gaussian with M = 16, the only supported M) is written once — overwritten, or added with RR_BWD_ACCUMULATE
 * (rows no view touched are then left alone) — and dL_dmean2D[v] ([P,3] per view; the array or single entries may be null)
 * receives each view's viewspace gradient.  stat_*: optional (all or none) f32[P] densification statistics updated in place
 * for every view that saw the Gaussian (train.py:132-134, scene/gaussian_model.py:419-421).  [first, first + count) is the
 * range of Gaussians this call covers (0, P for all): a caller may run the pass chunk by chunk and all-reduce the finished
 * chunk's bucket rows (rr_allreduce_sum_f32_ranges) on another stream while the next chunk is computed.
 *   flags without RR_BWD_LEAF: shs_or_f_dc = shs[P,16,3], f_rest = null, scales / rotations activated; gradients w.r.t. them:
 *       dL_dmean3D[P,3], dL_dsh_or_f_dc = dL_dsh[P,16,3], dL_dopacity[P], dL_dscale[P,3], dL_drot[P,4]
 *   with RR_BWD_LEAF: means3D = _xyz, shs_or_f_dc = _features_dc[P,1,3], f_rest = _features_rest[P,15,3], scales = _scaling (log),
 *       rotations = _rotation (un-normalised) — the raw leaves of scene/gaussian_model.py:131-136; the activations of :85-105 and
 *       their chain rule are applied inside; outputs dL_dxyz, dL_df_dc[P,3], dL_df_rest[P,45], dL_dopacity (raw), dL_dscaling, dL_drotation */
int rr_backward_render(void* stream, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	const float* dL_dpix, void* ws, size_t ws_bytes, int debug);
int rr_backward_views(void* stream, int P, int D, int M, int W, int H, int nviews,
	const float* const* viewmatrix, const float* const* projmatrix, const float* const* campos, const float* tan_fovx, const float* tan_fovy,
	const int* const* radii, void* const* geom, const float* const* grad_acc, float* const* dL_dmean2D,
	const float* means3D, const float* shs_or_f_dc, const float* f_rest, const float* scales, float scale_modifier, const float* rotations, float low_pass,
	float* dL_dmean3D, float* dL_dsh_or_f_dc, float* dL_df_rest, float* dL_dopacity, float* dL_dscale, float* dL_drot,
	float* stat_max_radii2D, float* stat_grad_accum, float* stat_denom, int first, int count, int flags, int debug);

/* Self test of the packed exponential the render kernels use (exp2x, csrc/render_common.cuh) against libdevice expf —
 * the function the reference's forward.cu:336 / backward.cu:488 compile to — over the float bit patterns
 * [first_bits, first_bits + count).  Adds the number of results that differ in any bit to *mismatches_dev (device, u64,
 * zeroed by the caller).  The forward's bit-exact n_contrib / colour parity rests on this being 0 for all x <= 0. */
int rr_selftest_exp(void* stream, uint32_t first_bits, uint32_t count, unsigned long long* mismatches_dev);

/* replaces Rasterizer::markVisible / checkFrustum (rasterizer_impl.cu:43-55,130-142); present is uint8/bool [P] */
int rr_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present);

/* ---- image loss right after the render -------------------------------------------------------
 * rr_loss_l1_ssim replaces, for the rendered image, utils/loss_utils.py:6-7 (l1_loss) and :23-51 (ssim / _ssim, window 11,
 * sigma 1.5, zero padding, size_average=True) plus train.py:112-113 and their autograd backward:
 *   out3 = { Ll1, ssim, (1 - lambda) * Ll1 + lambda * (1 - ssim) }            (device, 3 floats; may be null)
 *   dL_dimage[C,H,W] = d loss / d image                                          (device; may be null: forward only)
 * image and gt are [C,H,W] f32 device arrays; ws is rr_loss_ws_bytes(C,H,W) bytes of device scratch.
 * rr_loss_l2: dL_dimage = image - gt and *loss_accum += 0.5 * sum (image - gt)^2 (loss_accum may be null) in one pass. */
size_t rr_loss_ws_bytes(int C, int H, int W);
int rr_loss_l1_ssim(void* stream, int C, int H, int W, const float* image, const float* gt, float lambda_dssim,
	void* ws, size_t ws_bytes, float* out3, float* dL_dimage);
int rr_loss_l2(void* stream, int64_t n, const float* image, const float* gt, float* dL_dimage, float* loss_accum);
/* same with the target kept as 8-bit samples (gt = gt_u8 * scale, e.g. scale = 1/255): a quarter of the host->device bytes per image */
int rr_loss_l2_u8(void* stream, int64_t n, const float* image, const uint8_t* gt_u8, float scale, float* dL_dimage, float* loss_accum);

/* ---- parameter step after the backward ------------------------------------------------
 * One pass per Gaussian: chain rule through the caller-side activations of scene/gaussian_model.py:85-105 (exp, sigmoid,
 * F.normalize, cat), torch.optim.Adam(..., eps=1e-15).step() over the six parameter groups of training_setup
 * (scene/gaussian_model.py:139-157, train.py:146), and the activated inputs of the next render
 * (gaussian_renderer/__init__.py:37-60).
 *   xyz[P,3] f_dc[P,1,3] f_rest[P,M-1,3] opacity[P,1] scaling[P,3] rotation[P,4]: raw leaf tensors, updated in place
 *   exp_avg, exp_avg_sq, grads: flat [P*(11+3M)] f32 in the gradient-bucket layout
 *       means3D [P,3] | sh [P,M,3] | opacity [P] | scales [P,3] | rotations [P,4]
 *     grads are w.r.t. the ACTIVATED values (what rr_backward writes / accumulates); the moments are updated in place
 *   hyper: HOST array { lr_xyz, lr_f_dc, lr_f_rest, lr_opacity, lr_scaling, lr_rotation, beta1, beta2, eps }
 *   step: 1-based Adam step count (bias correction 1 - beta^step)
 *   shs_out[P,M,3], opacities_out[P,1], scales_out[P,3], rotations_out[P,4]: activated values after the update (each may be null) */
int rr_param_step(void* stream, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* exp_avg, float* exp_avg_sq, const float* grads, const float* hyper, int step,
	float* shs_out, float* opacities_out, float* scales_out, float* rotations_out);

/* rr_param_step for LEAF gradients (the bucket the leaf mode of rr_backward_views fills: xyz | f_dc | f_rest | opacity | scaling |
 * rotation, moments in the same layout): plain torch.optim.Adam semantics over the six leaf tensors in one flat pass — the chain
 * rule already happened in the backward and the next render reads the leaves themselves, so nothing else is written. */
int rr_param_step_leaf(void* stream, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* exp_avg, float* exp_avg_sq, const float* grads, const float* hyper, int step);

/* The caller-side activations of scene/gaussian_model.py:85-105 alone — opacities_out[P] = sigmoid(opacity), scales_out[P,3] =
 * exp(scaling), rotations_out[P,4] = F.normalize(rotation) — computed with the same device functions as the leaf modes of
 * rr_forward_geom_views / rr_backward_views and the re-activation of rr_param_step (identical bits in all of them). */
int rr_activate(void* stream, int P, const float* opacity, const float* scaling, const float* rotation, float* opacities_out, float* scales_out, float* rotations_out);

/* Densification statistics of one rendered view, fused (train.py:132-134 + GaussianModel.add_densification_stats,
 * scene/gaussian_model.py:419-421): for every Gaussian with radii > 0 (the view's visibility_filter)
 *   max_radii2D[p] = max(max_radii2D[p], radii[p]); xyz_gradient_accum[p] += |dL_dmean2D[p, 0:2]|; denom[p] += 1
 * radii i32[P] and dL_dmean2D f32[P,3] are the view's outputs (viewspace_points.grad); the three f32[P] statistics are
 * updated in place.  Per view and non-linear, so they are accumulated locally and summed / maxed across ranks separately
 * from the gradient all-reduce. */
int rr_densify_stats(void* stream, int P, const int* radii, const float* dL_dmean2D, float* max_radii2D, float* xyz_gradient_accum, float* denom);

/* ---- gradient-bucket all-reduce over NVLink peer / multicast memory (no reference counterpart) -------------
 * In-place sum of an f32 buffer of n floats that every rank holds in SYMMETRIC memory: buffers[r] / signals[r] (HOST arrays of
 * `world` device pointers) are rank r's bucket and flag buffer as mapped into the calling rank's address space ([rank] = the
 * local ones; flag buffers are rr_allreduce_signal_bytes() bytes, zero before the first call); multicast is the multicast
 * address aliasing all buckets (in-switch reduction: multimem.ld_reduce + multimem.st), or null (peer loads / stores).  One
 * kernel per rank, `ctas` CTAs (0 = one per SM; the same value on every rank), launched by every rank of the job on its own
 * GPU; epoch must increase by 2 from call to call (same value on all ranks).  *error_flag (local device u32) is set if a
 * cross-rank wait timed out.  Replaces ncclAllReduce for the view-sharded step; world <= 8.
 * rr_allreduce_sum_f32_ranges reduces nranges (<= 8) sub-ranges [offsets[i], offsets[i] + lengths[i]) of the buffer (floats; HOST
 * arrays; all multiples of 4 when nranges > 1) in one kernel — e.g. the rows of one chunk of Gaussians in each of the bucket's
 * segments; threads = CTA size (0 = 512). */
size_t rr_allreduce_signal_bytes(void);
int rr_allreduce_sum_f32_ranges(void* stream, int rank, int world, int nranges, const int64_t* offsets, const int64_t* lengths, void* const* buffers, void* multicast,
	void* const* signals, uint32_t* error_flag, uint32_t epoch, int ctas, int threads);
int rr_allreduce_sum_f32(void* stream, int rank, int world, int64_t n, void* const* buffers, void* multicast, void* const* signals,
	uint32_t* error_flag, uint32_t epoch, int ctas);

/* replaces distCUDA2 / SimpleKNN::knn (simple-knn/spatial.cu:4-15, simple_knn.cu:174-210); points f32[P,3] -> out f32[P];
 * ws: rr_dist2_ws_bytes(P) bytes of device scratch */
size_t rr_dist2_ws_bytes(int P);
int rr_dist2(void* stream, int P, const float* points, float* out, void* ws, size_t ws_bytes);

/* ---- stage-level entry points used by the parity tests ---- */
/* stable ascending sort of (key,value) pairs on key bits [begin_bit,end_bit); result is written to keys_out/vals_out;
 * keys_in/vals_in are clobbered (they are the ping-pong partners; vals_in must be a real buffer). */
size_t rr_sort_ws_bytes(int64_t n, int key_bytes, int begin_bit, int end_bit);
int rr_sort_pairs_u32(void* stream, int64_t n, uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_in, uint32_t* vals_out,
	int begin_bit, int end_bit, void* ws);
int rr_sort_pairs_u64(void* stream, int64_t n, uint64_t* keys_in, uint64_t* keys_out, uint32_t* vals_in, uint32_t* vals_out,
	int begin_bit, int end_bit, void* ws);

/* rr_sort_pairs_u32 for callers that cannot allocate device memory (code under CUDA graph capture, serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  Same sort, same
 * arguments minus ws: the workspace is a 4.06 MiB array inside the library's device image (every GPU that loads the library
 * holds it until exit), so n is limited to rr_sort_noalloc_max_pairs() = 4 194 304 pairs — a larger n is rejected with -1 and
 * a message.  The call makes no cudaMalloc / cudaFree / allocating library call.  All calls on a device share the one array
 * (each call clears the words it needs on its stream first): calls that overlap on the same GPU — two streams, two host
 * threads — race; order them on one stream or with events.  vals_in must be a real buffer (it is a ping-pong partner). */
size_t rr_sort_noalloc_max_pairs(void);
int rr_sort_pairs_u32_noalloc(void* stream, int64_t n, uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_in, uint32_t* vals_out,
	int begin_bit, int end_bit);

/* byte offsets of the arrays inside the opaque blobs (for stage-wise diffing against the oracle / reference).
 * geom:    [0] rec f32[P,12]  [1] cov3D f32[P,6]  [2] clamped u8[P]  [3] tiles_touched u32[P]  [4] rect u32[P,2]
 *          [5] depth_order u32[P]  [6] offsets u32[P] (inclusive, in depth order)  [7] = [6]  [8] num_rendered u32
 * binning: [0] tile_keys_sorted u32[N]  [1] point_list u32[N]  [2] warp_flags u32[4*(N/32+T+2)]  [3] = [1]  [4] reserved
 * image:   [0] final_T f32[H*W]  [1] n_contrib u32[H*W]  [2] ranges u32[T,2]  [3] reserved u32[T] */
int rr_geom_layout(int P, int64_t offsets_host[9]);
int rr_binning_layout(int64_t N, int W, int H, int64_t offsets_host[5]);
int rr_image_layout(int W, int H, int64_t offsets_host[4]);
/* the 64-bit (tile<<32 | depth bits) key of every sorted duplicate, as the reference's point_list_keys holds them */
int rr_export_sorted_keys(void* stream, int P, int64_t N, int W, int H, void* geom, void* binning, void* image, uint64_t* keys_out);

#ifdef __cplusplus
}
#endif
#endif
