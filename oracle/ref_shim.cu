// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI shim around the UNMODIFIED reference CUDA rasterizer and simple-knn,
// compiled (by oracle/Makefile, target `ref`) straight from the sources under
// /root/reference into oracle/_ref/libref_rast.so.  No reference source is
// copied: this file only #includes the reference's public headers
//   submodules/diff_gaussian_rasterization/cuda_rasterizer/rasterizer.h      (Rasterizer::forward/backward/markVisible)
//   submodules/diff_gaussian_rasterization/cuda_rasterizer/rasterizer_impl.h (GeometryState/BinningState/ImageState::fromChunk)
//   submodules/simple-knn/simple_knn.h                                       (SimpleKNN::knn)
// and plays the role of rasterize_points.cu (the torch binding) without libtorch:
// the three `std::function<char*(size_t)>` resize callbacks hand out grow-only
// cudaMalloc'ed scratch blobs owned by this shim (the torch binding used
// tensor.resize_, rasterize_points.cu:16-22).
//
// Used by: tests/ (-m gpu) as the bit-exact gate, bench.py --impl reference as
// the "reference CUDA on the same B200" arm.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <cuda_runtime.h>
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"

namespace {
struct Blob {
	char* ptr = nullptr;
	size_t cap = 0;
	size_t used = 0;
	char* get(size_t n) {
		if (n > cap) {
			if (ptr) cudaFree(ptr);
			size_t want = n + n / 4 + 256;
			if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
			cap = want;
		}
		used = n;
		return ptr;
	}
};
Blob g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;
std::string g_err;
}

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

// rasterize_points.cu:24-108 minus the tensor allocation (caller passes zero-filled outputs).
int ref_forward(int P, int D, int M, const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, float* out_depth, float low_pass, int* radii, int debug)
{
	try {
		g_P = P; g_W = W; g_H = H;
		std::function<char*(size_t)> gf = [](size_t n) { return g_geom.get(n); };
		std::function<char*(size_t)> bf = [](size_t n) { return g_bin.get(n); };
		std::function<char*(size_t)> imf = [](size_t n) { return g_img.get(n); };
		g_R = CudaRasterizer::Rasterizer::forward(gf, bf, imf, P, D, M, background, W, H,
			means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp,
			viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, prefiltered != 0,
			out_color, out_depth, low_pass, radii, debug != 0);
		return g_R;
	} catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// rasterize_points.cu:110-191 minus torch::zeros (caller passes zero-filled gradient buffers).
int ref_backward(int P, int D, int M, int R, const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, const int* radii, const float* dL_dpix,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
	int debug, float low_pass)
{
	try {
		CudaRasterizer::Rasterizer::backward(P, D, M, R, background, W, H, means3D, shs, colors_precomp,
			scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos,
			tan_fovx, tan_fovy, radii, g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix,
			dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
			debug != 0, low_pass);
		return 0;
	} catch (const std::exception& e) { g_err = e.what(); return -1; }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, unsigned char* present)
{
	CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, (bool*)present);
}

// Pointers into the reference's private scratch of the LAST ref_forward call
// (layout recovered with the reference's own fromChunk, rasterizer_impl.cu:144-183).
int ref_state_ptrs(void** out /*[16]*/)
{
	if (!g_geom.ptr || !g_img.ptr) return -1;
	char* c = g_geom.ptr;
	CudaRasterizer::GeometryState geom = CudaRasterizer::GeometryState::fromChunk(c, g_P);
	c = g_img.ptr;
	CudaRasterizer::ImageState img = CudaRasterizer::ImageState::fromChunk(c, (size_t)g_W * g_H);
	out[0] = geom.depths; out[1] = geom.clamped; out[2] = geom.means2D; out[3] = geom.cov3D;
	out[4] = geom.conic_opacity; out[5] = geom.rgb; out[6] = geom.tiles_touched; out[7] = geom.point_offsets;
	out[8] = img.accum_alpha; out[9] = img.n_contrib; out[10] = img.ranges;
	out[11] = out[12] = out[13] = out[14] = nullptr;
	if (g_bin.ptr && g_R > 0) {
		c = g_bin.ptr;
		CudaRasterizer::BinningState bin = CudaRasterizer::BinningState::fromChunk(c, g_R);
		out[11] = bin.point_list_keys; out[12] = bin.point_list;
		out[13] = bin.point_list_keys_unsorted; out[14] = bin.point_list_unsorted;
	}
	out[15] = nullptr;
	return 0;
}

// KNN/spatial.cu:4-15 minus the tensor allocation.
void ref_dist2(int P, float* points, float* out)
{
	SimpleKNN::knn(P, (float3*)points, out);
}

}
