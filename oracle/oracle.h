/* TEST INFRASTRUCTURE — CPU oracle for the splat-rasterizer hot path.
 *
 * A scalar C++17/OpenMP restatement of the reference's algorithm
 * (/root/reference/submodules/diff_gaussian_rasterization/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu,
 *  auxiliary.h; /root/reference/submodules/simple-knn/simple_knn.cu).  Each function cites the lines it follows.
 *
 * PARITY PIN: the reference ships no tests, golden vectors or fixtures for this path, so this
 * restatement cannot be pinned against reference-owned vectors.  It is pinned instead against OUTPUTS OF THE
 * REFERENCE ITSELF: oracle/_ref/libref_rast.so is the unmodified reference CUDA compiled for sm_100a
 * (oracle/Makefile `ref`), and tests/test_gpu_parity.py compares oracle == reference on the B200 box; the
 * vectors produced there are committed under tests/golden/ with the generating script.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product (rain_b200/) never does.
 */
#ifndef RAIN_ORACLE_H
#define RAIN_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

orc_ctx* orc_create(void);
void orc_destroy(orc_ctx*);

/* Rasterizer::forward (rasterizer_impl.cu:187-330). All pointers are HOST pointers. Null shs/colors_precomp,
 * scales+rotations/cov3D_precomp select the modes exactly like the nullptr tests at forward.cu:195,231.
 * out_color[3*H*W], out_depth[H*W], radii[P] are written. Returns num_rendered (>=0) or -1. */
int64_t orc_forward(orc_ctx*, int P, int D, int M, const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass,
	float* out_color, float* out_depth, int* radii);

/* Rasterizer::backward (rasterizer_impl.cu:334-430) on the state left in ctx by the last orc_forward.
 * Gradient buffers are overwritten (the zero fill of rasterize_points.cu:145-153 is done inside).
 * dL_dconic is [P,4] (slots 0,1,3 used, backward.cu:539-541). */
int orc_backward(orc_ctx*, const float* dL_dpix,
	float* dL_dmean2D /*[P,3]*/, float* dL_dconic /*[P,4]*/, float* dL_dopacity /*[P]*/, float* dL_dcolor /*[P,3]*/,
	float* dL_dmean3D /*[P,3]*/, float* dL_dcov3D /*[P,6]*/, float* dL_dsh /*[P,M,3]*/,
	float* dL_dscale /*[P,3]*/, float* dL_drot /*[P,4]*/);

/* Intermediates of the last forward (pointers stay valid until the next orc_forward / orc_destroy).
 * names: "depths" f32[P], "means2D" f32[P,2], "cov3D" f32[P,6], "conic_opacity" f32[P,4], "rgb" f32[P,3],
 * "clamped" u8[P,3], "tiles_touched" u32[P], "point_offsets" u32[P], "keys_unsorted" u64[N],
 * "values_unsorted" u32[N], "keys" u64[N], "point_list" u32[N], "ranges" u32[T,2], "final_T" f32[H*W],
 * "n_contrib" u32[H*W].  Returns element count, or -1 for an unknown name. */
int64_t orc_get(orc_ctx*, const char* name, const void** ptr);

/* Work counters of the last forward/backward: [0]=E_f evaluated (pixel,entry) pairs fwd, [1]=contributing pairs fwd,
 * [2]=E_b evaluated pairs bwd, [3]=contributing pairs bwd. */
void orc_counters(orc_ctx*, int64_t out[4]);

/* checkFrustum (rasterizer_impl.cu:43-55) */
void orc_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present);

/* SimpleKNN::knn (simple_knn.cu:174-210): mean of the 3 smallest squared distances to other points.
 * orc_dist2 is exact via a uniform grid (any exact 3-NN gives the same value, see knn_oracle.cpp);
 * orc_dist2_brute is the O(P^2) definition for small P. */
void orc_dist2(int P, const float* points, float* out);
void orc_dist2_brute(int P, const float* points, float* out);
/* exact answer for a subset of query indices only (used at 5M points) */
void orc_dist2_queries(int P, const float* points, int Q, const int* query_idx, float* out);
/* coord2Morton (simple_knn.cu:34-50) for fixture checks */
uint32_t orc_morton(float x, float y, float z, const float* minn, const float* maxx);

int orc_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
