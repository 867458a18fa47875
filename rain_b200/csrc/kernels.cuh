// Launcher interfaces between the C-ABI layer (api.cu) and the kernel translation units.
#pragma once
#include "common.cuh"

namespace rr {

struct PreprocessArgs {
	int P, D, M, W, H, grid_x, grid_y;
	float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier, low_pass;
	const float *means3D, *scales, *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp, *view, *proj, *campos;
	int* radii;               // [P] user-visible output
	SplatRec* rec;            // [P] projected record (valid where radii > 0)
	float* cov3D;             // [P,6]
	uint8_t* clamped;         // [P] bit ch set if colour channel ch was clamped at 0
	uint32_t* tiles_touched;  // [P]
	uint32_t* depth_key;      // [P] float bits of view-space depth, 0xFFFFFFFF when culled
	uint2* rect;              // [P] x0 | x1<<16, y0 | y1<<16
	unsigned long long* tile_mask;   // [P] tile culling (opt-in): bit ry * w + rx of the rect is set if that tile can receive a contribution; ~0 = all
	int tile_cull;
};
int launch_preprocess_fwd(cudaStream_t, const PreprocessArgs&);

// view-batched forward preprocess (preprocess.cu): SH colour + scale / rotation mode only, M = 16
constexpr int PV_MAX_VIEWS = 16;
struct PreprocessViewsArgs {
	int P, D, W, H, grid_x, grid_y, nviews, write_cov3D;
	int leaf;                             // 1: raw leaf tensors in (shs unused; f_dc [P,1,3] + f_rest [P,15,3]), activations applied in the kernel
	float scale_modifier, low_pass;
	const float *means3D, *scales, *rotations, *opacities, *shs, *f_dc, *f_rest;
	const float* view[PV_MAX_VIEWS]; const float* proj[PV_MAX_VIEWS]; const float* campos[PV_MAX_VIEWS];
	float tan_fovx[PV_MAX_VIEWS], tan_fovy[PV_MAX_VIEWS];
	int* radii[PV_MAX_VIEWS];
	char* geom[PV_MAX_VIEWS];             // each view's geometry blob; the arrays below sit at the same byte offsets in every one
	size_t off_rec, off_cov3D, off_clamped, off_tiles, off_rect, off_key;
};
int launch_preprocess_views(cudaStream_t, const PreprocessViewsArgs&);
int launch_mark_visible(cudaStream_t, int P, const float* means3D, const float* view, uint8_t* present);

// binning.cu
size_t scan_ws_elems(int P);
int launch_scan_tiles(cudaStream_t, int P, const uint32_t* order, const uint32_t* tiles_touched, uint32_t* offsets /*inclusive, depth order*/,
	uint32_t* block_sums, uint32_t* num_rendered_dev, uint32_t* num_rendered_mapped /*device-visible alias of a pinned host word, or null*/);
// the prefix sums of up to 16 views as one launch (blockIdx.y = view)
struct ScanSegs {
	const uint32_t* order[16]; const uint32_t* tiles_touched[16]; uint32_t* status[16]; uint32_t* offsets[16]; uint32_t* total[16]; uint32_t* total_host[16];
};
int launch_scan_tiles_segments(cudaStream_t, int P, int nseg, const ScanSegs&);
int launch_duplicate(cudaStream_t, int P, const uint32_t* order, const uint32_t* offsets, const uint2* rect, const unsigned long long* tile_mask /*or null*/, int grid_x,
	uint32_t* tile_keys /*[N] emission order*/, uint32_t* gaussian_ids /*[N] emission order*/);
// tile plan (binning.cu): ranges and the tile sort's digit bases from the rects alone, for up to 16 views per launch pair
bool tile_plan_supported(int grid_x, int grid_y);
size_t tile_plan_ws_bytes(int grid_x, int grid_y);
struct PlanSegs { const uint32_t* tiles_touched[16]; const uint2* rect[16]; int* gdiff[16] /*tile_plan_ws_bytes scratch*/; uint2* ranges[16]; uint32_t* ghist[16] /*[2][256] out*/; };
int launch_tile_plan(cudaStream_t, int P, int grid_x, int grid_y, int nseg, const PlanSegs&, int npass, const int* shift, const uint32_t* mask);
int launch_duplicate16(cudaStream_t, int P, const uint32_t* order, const uint32_t* offsets, const uint2* rect, int grid_x, uint16_t* tile_keys, uint32_t* gaussian_ids);
int launch_tile_ranges(cudaStream_t, size_t N, const uint32_t* sorted_tile_keys, uint2* ranges, int num_tiles);
int launch_export_keys(cudaStream_t, int T, const uint2* ranges, const uint32_t* point_list, const uint32_t* rec_words, uint64_t* out);

// render_fwd.cu / render_bwd.cu
// Per-warp blend flags ("which list entries did warp w's 8x8 pixel block blend"), written by the forward and walked by
// the backward (both give warp w the block at ((w & 1) * 8, (w >> 1) * 8) of the tile).  One bit per (tile list entry,
// warp), stored as 32-bit words; the words of tile t start at FLAG_WARPS * (ranges[t].x / 32 + t) and warp w owns the
// ceil(n / 32) consecutive words from + w * ceil(n / 32) (n = list length of the tile; bit j % 32 of word j / 32 is list
// entry j).  A warp writes its words only up to the batch of its last contributor; the blob is not zero-filled.
constexpr int FLAG_WARPS = 4;
__host__ __device__ inline size_t warp_flag_words(size_t N, size_t tiles) { return FLAG_WARPS * (N / 32 + tiles + 2); }

// Per-Gaussian gradient accumulator filled by the backward render (one RED.ADD per value and (warp, list entry) visit)
// and consumed by the backward preprocess: 12 floats = 48 bytes,
//   0..2 dL_dcolor   3 Su.dx  4 Su.dy  5 Su.dx^2  6 Su.dx.dy  7 Su.dy^2  8 Su   (u = G * dL_dalpha per pixel; d = centre - pixel)
constexpr int ACC_FLOATS = 12;

struct RenderArgs {
	int W, H, grid_x, grid_y;
	const uint2* ranges;        // [tiles]
	const SplatRec* rec;        // [P] per-Gaussian records, gathered through point_list
	const uint32_t* point_list; // [N] sorted Gaussian ids
	const float* bg;            // [3]
	float* final_T;             // [H*W]
	uint32_t* n_contrib;        // [H*W]
	uint32_t* warp_flags;       // [warp_flag_words(N, tiles)] fwd writes, bwd reads
	float* out_color;           // [3,H,W]
	float* out_depth;           // [H,W]
	float4* zero;               // forward only: zero_n16 16-byte units the CTAs clear while they render (the geometry blob's accumulator
	uint32_t zero_n16;          // region of this view's backward; the kernel is issue bound, the stores ride on idle HBM bandwidth)
	// backward only
	const float* dL_dpix;       // [3,H,W]
	float* grad_acc;            // [P, ACC_FLOATS], zero on entry
	unsigned long long* xc;     // cross-check build only (-DRR_CROSSCHECK): [16] work counters, see render_common.cuh; null in the product library
};
int launch_render_fwd(cudaStream_t, const RenderArgs&);
int launch_render_bwd(cudaStream_t, const RenderArgs&);
int launch_exp_selftest(cudaStream_t, uint32_t first_bits, uint32_t count, unsigned long long* mismatches /*device, zeroed by the caller*/);

// preprocess_bwd.cu
struct PreprocessBwdArgs {
	int P, D, M, W, H;
	float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier, low_pass;
	const float *means3D, *scales, *rotations, *shs, *cov3D /*precomp or geom*/, *view, *proj, *campos;
	int has_colors_precomp;
	int accumulate;           // 1: += into the parameter-gradient buffers, untouched where radii == 0
	const int* radii;
	const uint8_t* clamped;
	const SplatRec* rec;      // [P] projected records (conic, opacity)
	const float* grad_acc;    // [P, ACC_FLOATS] from the backward render
	float *dL_dmean2D /*[P,3]*/, *dL_dconic /*[P,4] or null*/, *dL_dopacity /*[P]*/, *dL_dcolor /*[P,3]*/, *dL_dmean3D /*[P,3]*/,
		*dL_dcov3D /*[P,6]*/, *dL_dsh /*[P,M,3]*/, *dL_dscale /*[P,3]*/, *dL_drot /*[P,4]*/;
};
int launch_preprocess_bwd(cudaStream_t, const PreprocessBwdArgs&);

// view-batched backward (preprocess_bwd.cu): the per-Gaussian half of the backward for up to BV_MAX_VIEWS views in one pass
constexpr int BV_MAX_VIEWS = 16;
struct BackwardViewsArgs {
	int P, D, M, W, H, nviews;
	int first, count;         // the Gaussians [first, first + count) this launch covers
	int leave_room;           // 1: cap the kernel at 3 CTAs per SM so a co-running all-reduce CTA fits on every SM
	float scale_modifier, low_pass;
	int accumulate;           // 1: += into the bucket, rows no view of this batch touched are left alone
	int leaf;                 // 1: inputs are the raw leaf tensors, gradients w.r.t. them (see the kernel's header)
	const float *means3D, *scales, *rotations, *shs /*[P,16,3], leaf = 0*/, *f_dc /*[P,1,3]*/, *f_rest /*[P,15,3], leaf = 1*/;
	// per view
	const float* view[BV_MAX_VIEWS]; const float* proj[BV_MAX_VIEWS]; const float* campos[BV_MAX_VIEWS];
	float tan_fovx[BV_MAX_VIEWS], tan_fovy[BV_MAX_VIEWS];
	const int* radii[BV_MAX_VIEWS]; const SplatRec* rec[BV_MAX_VIEWS]; const uint8_t* clamped[BV_MAX_VIEWS]; const float* grad_acc[BV_MAX_VIEWS];
	float* dL_dmean2D[BV_MAX_VIEWS];        // [P,3] per view, or null
	// bucket segments (see the kernel's header for the two layouts); dL_dsh is the f_dc segment and dL_dfrest the f_rest one when leaf = 1
	float *dL_dmean3D, *dL_dsh, *dL_dfrest, *dL_dopacity, *dL_dscale, *dL_drot;
	float *stat_max_radii, *stat_grad_accum, *stat_denom;      // optional fused densification statistics ([P] each), or null
};
int launch_backward_views(cudaStream_t, const BackwardViewsArgs&);

// loss.cu
size_t loss_ws_bytes(int C, int H, int W);
int launch_loss_l1_ssim(cudaStream_t, int C, int H, int W, const float* img, const float* gt, float lambda, void* ws, float* out3, float* dL_dimg);
int launch_loss_l2(cudaStream_t, size_t n, const float* img, const void* gt, int gt_is_u8, float scale, float* dL_dimg, float* loss_acc);

// param_step.cu
int launch_param_step(cudaStream_t, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* m, float* v, const float* grads, const float* hyper, int step, float* shs_out, float* opacities_out, float* scales_out, float* rotations_out);

int launch_param_step_leaf(cudaStream_t, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* m, float* v, const float* grads, const float* hyper, int step);
int launch_activate(cudaStream_t, int P, const float* opacity, const float* scaling, const float* rotation, float* opacities_out, float* scales_out, float* rotations_out);
int launch_densify_stats(cudaStream_t, int P, const int* radii, const float* dL_dmean2D, float* max_radii2D, float* grad_accum, float* denom);

// allreduce.cu: one-kernel two-shot all-reduce of a symmetric f32 buffer over NVLink peer / multicast memory
constexpr int AR_MAX_WORLD = 8;
constexpr int AR_MAX_RANGES = 8;
constexpr int AR_MAX_CTAS = 256;          // CTAs per rank; the signal buffer holds AR_MAX_CTAS * AR_MAX_WORLD flag words
struct AllReduceArgs {
	float* multicast;                     // multicast address of the symmetric buffer, or null (peer-pointer variant)
	float* buffers[AR_MAX_WORLD];         // every rank's buffer as mapped into THIS rank's address space ([rank] = the local one)
	uint32_t* signal[AR_MAX_WORLD];       // every rank's flag buffer (u32[AR_MAX_CTAS * world], zero before the first call), peer-mapped
	uint32_t* error_flag;                 // local device word, set to 1 if a cross-rank wait timed out
	int nranges; int64_t off[AR_MAX_RANGES], len[AR_MAX_RANGES];      // sub-ranges of the buffer (floats) reduced by this launch
	int rank, world;
	uint32_t epoch;                       // flag value of this call's first barrier; the second uses epoch + 1; strictly increasing across calls
};
int launch_allreduce(cudaStream_t, const AllReduceArgs&, int ctas, int threads /*0 = default*/);

// knn.cu
size_t knn_ws_bytes(int P);
int launch_dist2(cudaStream_t, int P, const float* points, float* out, void* ws, size_t ws_bytes);

}  // namespace rr
