// Forward preprocess: per-Gaussian cull, 3D covariance, EWA projection to a 2D conic, SH colour, radius / tile rect.
// Replaces preprocessCUDA<3> (reference forward.cu:144-246, with computeCov3D :107-141, computeCov2D :63-102,
// computeColorFromSH :9-60, in_frustum auxiliary.h:128-153, ndc2Pix :30-33, getRect :35-45) and checkFrustum
// (rasterizer_impl.cu:43-55).
//
// Integer outputs (radii, tile rects, depth key bits) must be bit-identical to the reference build, so every
// rounding step on the path to them is written with explicit round-to-nearest intrinsics (__fmaf_rn / __fmul_rn /
// __fadd_rn / __fdiv_rn / __fsqrt_rn) in the order the reference's nvcc build performs them ("FMA
// contract"); the compiler can neither contract nor reassociate these.
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"
#include "splat_math.cuh"

namespace rr {

namespace {

__constant__ float c_SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
__constant__ float c_SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
constexpr float kSH_C0 = 0.28209479177387814f;
constexpr float kSH_C1 = 0.4886025119029199f;

// ---- per-Gaussian pieces, shared by the per-view kernel and the view-batched one ---------------------------------------
struct Projected {
	int radius; uint32_t tiles, depth_key; uint2 rect; bool visible;
	float conx, cony, conz, pix_x, pix_y;
};

// cull, EWA projection of the 3D covariance c3 to the 2D conic, radius and tile rect of one Gaussian whose view-space depth
// already passed the near-plane test (forward.cu:172-229; computeCov2D :63-102, ndc2Pix auxiliary.h:30-33, getRect :35-45)
__device__ __forceinline__ void project_splat(const float* V, const float* Pm, float px, float py, float pz, float view_z, const float* c3,
	float tan_fovx, float tan_fovy, float focal_x, float focal_y, float low_pass, int W, int H, int grid_x, int grid_y, Projected& o)
{
	const float hx = xform(Pm, 0, px, py, pz), hy = xform(Pm, 1, px, py, pz), hw = xform(Pm, 3, px, py, pz);
	const float p_w = DIV(1.0f, ADD(hw, 0.0000001f));
	const float projx = MUL(hx, p_w), projy = MUL(hy, p_w);
	// computeCov2D
	const float tx0 = xform(V, 0, px, py, pz), ty0 = xform(V, 1, px, py, pz), tz = view_z;
	const float limx = MUL(tan_fovx, 1.3f), limy = MUL(tan_fovy, 1.3f);
	const float cx = fminf(limx, fmaxf(-limx, DIV(tx0, tz)));
	const float cy = fminf(limy, fmaxf(-limy, DIV(ty0, tz)));
	const float tz2 = MUL(tz, tz);
	const float J00 = DIV(focal_x, tz);
	const float J02 = DIV(MUL(focal_x, MUL(cx, -tz)), tz2);
	const float J11 = DIV(focal_y, tz);
	const float J12 = DIV(MUL(focal_y, MUL(cy, -tz)), tz2);
	const float z0 = 0.0f;
	const float T00 = FMA(V[2], J02, FMA(V[0], J00, MUL(V[1], z0)));
	const float T01 = FMA(V[6], J02, FMA(V[4], J00, MUL(V[5], z0)));
	const float T02 = FMA(J02, V[10], FMA(V[8], J00, MUL(V[9], z0)));
	const float T10 = FMA(V[2], J12, FMA(V[0], z0, MUL(J11, V[1])));
	const float T11 = FMA(V[6], J12, FMA(V[4], z0, MUL(J11, V[5])));
	const float T12 = FMA(J12, V[10], FMA(V[8], z0, MUL(J11, V[9])));
	const float A00 = sop3(T00, c3[0], T01, c3[1], T02, c3[2]);
	const float A01 = sop3(T10, c3[0], T11, c3[1], T12, c3[2]);
	const float A10 = sop3(T00, c3[1], T01, c3[3], T02, c3[4]);
	const float A11 = sop3(T10, c3[1], T11, c3[3], T12, c3[4]);
	const float A20 = sop3(T00, c3[2], T01, c3[4], T02, c3[5]);
	const float A21 = sop3(T10, c3[2], T11, c3[4], T12, c3[5]);
	const float covx = ADD(low_pass, sop3(T00, A00, T01, A10, T02, A20));
	const float covy = sop3(T00, A01, T01, A11, T02, A21);
	const float covz = ADD(low_pass, sop3(T10, A01, T11, A11, T12, A21));

	const float det = FMA(covx, covz, -MUL(covy, covy));
	if (det != 0.0f) {
		const float det_inv = DIV(1.f, det);
		o.conx = MUL(covz, det_inv); o.cony = MUL(det_inv, -covy); o.conz = MUL(covx, det_inv);
		const float mid = MUL(ADD(covx, covz), 0.5f);
		const float sq = __fsqrt_rn(fmaxf(FMA(mid, mid, -det), 0.1f));
		const float lam = fmaxf(ADD(mid, sq), SUB(mid, sq));
		const float my_radius = ceilf(MUL(__fsqrt_rn(lam), 3.f));
		// ndc2Pix in double, as the reference's 1.0 / 0.5 literals force it (auxiliary.h:30-33)
		o.pix_x = (float)__dmul_rn(__fma_rn(__dadd_rn((double)projx, 1.0), (double)W, -1.0), 0.5);
		o.pix_y = (float)__dmul_rn(__fma_rn(__dadd_rn((double)projy, 1.0), (double)H, -1.0), 0.5);
		const int r_i = (int)my_radius;
		const float r_f = (float)r_i;
		// getRect
		const uint32_t x0 = min((uint32_t)grid_x, (uint32_t)max(0, (int)MUL(SUB(o.pix_x, r_f), 0.0625f)));
		const uint32_t y0 = min((uint32_t)grid_y, (uint32_t)max(0, (int)MUL(SUB(o.pix_y, r_f), 0.0625f)));
		const uint32_t x1 = min((uint32_t)grid_x, (uint32_t)max(0, (int)MUL(ADD(ADD(ADD(o.pix_x, r_f), 16.0f), -1.0f), 0.0625f)));
		const uint32_t y1 = min((uint32_t)grid_y, (uint32_t)max(0, (int)MUL(ADD(ADD(ADD(o.pix_y, r_f), 16.0f), -1.0f), 0.0625f)));
		const uint32_t area = (x1 - x0) * (y1 - y0);
		if (area != 0) {
			o.visible = true;
			o.radius = r_i;
			o.tiles = area;
			o.rect = make_uint2(x0 | (x1 << 16), y0 | (y1 << 16));
			o.depth_key = __float_as_uint(view_z);
		}
	}
}

// colour of one Gaussian from its SH row (coefficient-major, 3 channels each; only the (D+1)^2 coefficients of the active
// degree are read) — computeColorFromSH, forward.cu:9-60: clamped at 0 per channel, bit ch of `clamped` set where it was
__device__ __forceinline__ void sh_colour(int D, const float* sh, float px, float py, float pz, const float* campos, float& rgb0, float& rgb1, float& rgb2, uint8_t& clamped)
{
	const float dx = SUB(px, campos[0]), dy = SUB(py, campos[1]), dz = SUB(pz, campos[2]);
	const float len = __fsqrt_rn(sop3(dx, dx, dy, dy, dz, dz));
	const float x = DIV(dx, len), y = DIV(dy, len), z = DIV(dz, len);
	float res[3];
	#pragma unroll
	for (int ch = 0; ch < 3; ch++) res[ch] = MUL(sh[ch], kSH_C0);
	if (D > 0) {
		const float ca = MUL(y, kSH_C1), cb = MUL(z, kSH_C1), cc = MUL(x, kSH_C1);
		#pragma unroll
		for (int ch = 0; ch < 3; ch++) {
			float r = FMA(-ca, sh[3 + ch], res[ch]);
			r = FMA(cb, sh[6 + ch], r);
			res[ch] = FMA(-cc, sh[9 + ch], r);
		}
		if (D > 1) {
			const float xx = MUL(x, x), yy = MUL(y, y), zz = MUL(z, z), xy = MUL(x, y), yz = MUL(y, z), xz = MUL(x, z);
			const float zz2 = ADD(zz, zz);
			const float t4 = MUL(xy, c_SH_C2[0]), t5 = MUL(yz, c_SH_C2[1]), t6 = MUL(SUB(SUB(zz2, xx), yy), c_SH_C2[2]), t7 = MUL(xz, c_SH_C2[3]);
			const float xx_yy = SUB(xx, yy), t8 = MUL(xx_yy, c_SH_C2[4]);
			#pragma unroll
			for (int ch = 0; ch < 3; ch++) {
				float r = FMA(t4, sh[12 + ch], res[ch]);
				r = FMA(t5, sh[15 + ch], r);
				r = FMA(t6, sh[18 + ch], r);
				r = FMA(t7, sh[21 + ch], r);
				res[ch] = FMA(t8, sh[24 + ch], r);
			}
			if (D > 2) {
				const float t9 = MUL(MUL(y, c_SH_C3[0]), FMA(xx, 3.0f, -yy));
				const float t10 = MUL(z, MUL(xy, c_SH_C3[1]));
				const float q4 = SUB(FMA(zz, 4.0f, -xx), yy);
				const float t11 = MUL(MUL(y, c_SH_C3[2]), q4);
				const float t12 = MUL(MUL(z, c_SH_C3[3]), FMA(yy, -3.0f, FMA(xx, -3.0f, zz2)));
				const float t13 = MUL(MUL(x, c_SH_C3[4]), q4);
				const float t14 = MUL(MUL(z, c_SH_C3[5]), xx_yy);
				const float t15 = MUL(MUL(x, c_SH_C3[6]), FMA(yy, -3.0f, xx));
				#pragma unroll
				for (int ch = 0; ch < 3; ch++) {
					float r = FMA(t9, sh[27 + ch], res[ch]);
					r = FMA(t10, sh[30 + ch], r);
					r = FMA(t11, sh[33 + ch], r);
					r = FMA(t12, sh[36 + ch], r);
					r = FMA(t13, sh[39 + ch], r);
					r = FMA(t14, sh[42 + ch], r);
					res[ch] = FMA(t15, sh[45 + ch], r);
				}
			}
		}
	}
	const float v0 = ADD(res[0], 0.5f), v1 = ADD(res[1], 0.5f), v2 = ADD(res[2], 0.5f);
	clamped = (v0 < 0.0f ? 1 : 0) | (v1 < 0.0f ? 2 : 0) | (v2 < 0.0f ? 4 : 0);
	rgb0 = v0 < 0.0f ? 0.0f : v0; rgb1 = v1 < 0.0f ? 0.0f : v1; rgb2 = v2 < 0.0f ? 0.0f : v2;
}

// Conservative half-extents (pixels) of the region where a splat can reach alpha >= 1/255, for the render kernels' per-warp
// culling: alpha = min(0.99, o * exp(power)) >= 1/255 needs power >= -t with t = ln(255 o), and {d : 0.5 d^T Q d <= t} (Q =
// the stored conic) has x-extent sqrt(2 t Q^-1_xx) = sqrt(2 t C / (AC - B^2)).  Evaluated in double from the STORED conic; t
// is inflated for the fp32 rounding of `power` in the blend, which is relative to the magnitude of its terms, i.e. grows with
// the conditioning of Q.  +inf = never cull, -inf = the splat can never pass the alpha test (o < 1/255).
__device__ __forceinline__ void cull_extent(float conx, float cony, float conz, float opacity, float& ext_x, float& ext_y)
{
	ext_x = INFINITY; ext_y = INFINITY;
	const double A = conx, B = cony, C = conz;
	const double dd = A * C - B * B;
	const float t = __logf(255.0f * opacity) + 1e-3f;
	if (opacity == opacity && t == t && A > 0.0 && C > 0.0 && dd > 0.0) {
		if (!(t > 0.0f)) { ext_x = -INFINITY; ext_y = -INFINITY; }
		else {
			const double cond = (A + C) * (A + C) / dd;
			const double shrink = 1.0 - 4e-6 * cond;
			if (shrink > 0.5) {
				const double t2 = 2.0 * ((double)t + 1e-3) / shrink;
				ext_x = (float)(sqrt(t2 * C / dd) * 1.00001 + 1e-3);
				ext_y = (float)(sqrt(t2 * A / dd) * 1.00001 + 1e-3);
			}
		}
	}
}

#ifndef PF_MIN_CTAS
#define PF_MIN_CTAS 4      // 64 registers: 4 CTAs (32 warps) per SM instead of 3; 0.081 -> 0.078 ms
#endif
template <bool TILE_CULL>
__global__ void __launch_bounds__(256, PF_MIN_CTAS) preprocess_fwd_kernel(PreprocessArgs a)
{
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	const bool live = idx < a.P;
	if (!TILE_CULL && !live) return;            // the culling variant keeps whole warps for its cooperative pass
	const float* __restrict__ V = a.view;
	const float* __restrict__ Pm = a.proj;

	float px = 0, py = 0, pz = 0;
	if (live) { px = a.means3D[3 * (size_t)idx]; py = a.means3D[3 * (size_t)idx + 1]; pz = a.means3D[3 * (size_t)idx + 2]; }
	const float view_z = live ? xform(V, 2, px, py, pz) : 0.0f;

	Projected o;
	o.radius = 0; o.tiles = 0; o.depth_key = 0xFFFFFFFFu; o.rect = make_uint2(0u, 0u); o.visible = false;
	o.conx = o.cony = o.conz = o.pix_x = o.pix_y = 0.0f;

	if (live && !(view_z <= 0.2f)) {
		float c3[6];
		if (a.cov3D_precomp) {
			#pragma unroll
			for (int k = 0; k < 6; k++) c3[k] = a.cov3D_precomp[6 * (size_t)idx + k];
		} else {
			const float4 q = load_f4(a.rotations, (size_t)idx);
			cov3d_from_scale_rot(a.scales[3 * (size_t)idx], a.scales[3 * (size_t)idx + 1], a.scales[3 * (size_t)idx + 2], a.scale_modifier, q, c3);
			#pragma unroll
			for (int k = 0; k < 6; k++) a.cov3D[6 * (size_t)idx + k] = c3[k];     // written even if culled below (forward.cu:201)
		}
		project_splat(V, Pm, px, py, pz, view_z, c3, a.tan_fovx, a.tan_fovy, a.focal_x, a.focal_y, a.low_pass, a.W, a.H, a.grid_x, a.grid_y, o);
	}
	int radius = o.radius;
	uint32_t tiles = o.tiles;
	const uint32_t depth_key = o.depth_key;
	const uint2 rect = o.rect;
	const bool visible = o.visible;
	const float conx = o.conx, cony = o.cony, conz = o.conz, pix_x = o.pix_x, pix_y = o.pix_y;

	if (TILE_CULL) {
		// opt-in (RAIN_TILE_CULL=1): drop the tiles of the rect in which the reference's own alpha test (power > 0 or
		// alpha < 1/255 -> skip, forward.cu:330-340) rejects this Gaussian at every pixel — the same conservative test the
		// render kernel applies per 8x4 block.  Images and gradients do not change; the internal lists and num_rendered
		// get shorter.  Rects of more than 64 tiles are kept whole.  The warp tests the tiles of its 32 Gaussians
		// together, one tile per lane and step (a per-thread loop runs as long as the largest rect of the warp).
		__shared__ float s_par[8][32][8];
		__shared__ uint32_t s_rc[8][32][2], s_ex[8][33];
		__shared__ uint32_t s_mask[8][32][2];
		const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
		const uint32_t x0 = rect.x & 0xffffu, w = (rect.x >> 16) - x0, y0 = rect.y & 0xffffu;
		ContribTest ct = make_contrib_test(conx, cony, conz, visible ? a.opacities[idx] : 1.0f);
		const uint32_t mine = (visible && tiles <= 64u && !ct.keep_all) ? tiles : 0u;
		uint32_t incl = mine;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
		const uint32_t total = __shfl_sync(FULL, incl, 31);
		float* pr = s_par[warp][lane];
		pr[0] = pix_x; pr[1] = pix_y; pr[2] = ct.A; pr[3] = ct.B; pr[4] = ct.C; pr[5] = ct.ia; pr[6] = ct.ic; pr[7] = ct.t;
		s_rc[warp][lane][0] = x0 | (w << 16); s_rc[warp][lane][1] = y0;
		s_ex[warp][lane] = incl - mine;
		if (lane == 0) s_ex[warp][32] = total;
		s_mask[warp][lane][0] = 0u; s_mask[warp][lane][1] = 0u;
		__syncwarp();
		for (uint32_t s = lane; s < total; s += 32) {
			int lo = 0;                                      // last lane whose exclusive offset is <= s and that owns tiles
			#pragma unroll
			for (int step = 16; step >= 1; step >>= 1) if (s_ex[warp][lo + step] <= s) lo += step;
			const uint32_t local = s - s_ex[warp][lo];
			const float* q = s_par[warp][lo];
			ContribTest c; c.A = q[2]; c.B = q[3]; c.C = q[4]; c.ia = q[5]; c.ic = q[6]; c.t = q[7]; c.keep_all = false;
			const uint32_t xw = s_rc[warp][lo][0], ww = xw >> 16;
			const uint32_t ry = local / ww, rx = local - ry * ww;
			const float X0 = (float)(((xw & 0xffffu) + rx) * 16u), Y0 = (float)((s_rc[warp][lo][1] + ry) * 16u);
			if (may_contribute(c, q[0], q[1], X0, X0 + 15.0f, Y0, Y0 + 15.0f)) atomicOr(&s_mask[warp][lo][local >> 5], 1u << (local & 31u));
		}
		__syncwarp();
		if (!live) return;
		unsigned long long mask = ~0ull;
		if (mine) {
			const unsigned long long m = (unsigned long long)s_mask[warp][lane][0] | ((unsigned long long)s_mask[warp][lane][1] << 32);
			const uint32_t kept = (uint32_t)__popcll(m);
			if (kept != tiles) { mask = m; tiles = kept; }
		}
		if (visible) a.tile_mask[idx] = mask;
	}
	a.radii[idx] = radius;
	a.tiles_touched[idx] = tiles;
	a.depth_key[idx] = depth_key;
	if (!visible) return;
	a.rect[idx] = rect;

	// colour
	float rgb0, rgb1, rgb2;
	uint8_t clamped = 0;
	if (a.colors_precomp) {
		rgb0 = a.colors_precomp[3 * (size_t)idx]; rgb1 = a.colors_precomp[3 * (size_t)idx + 1]; rgb2 = a.colors_precomp[3 * (size_t)idx + 2];
	} else {
		// SH coefficients of the active degree only: (D+1)^2 x 3 floats, fetched as float4 when the row is the usual
		// 16 x 3 = 192-byte aligned row, scalar otherwise
		float sh[48];
		{
			const int nfl = (a.D + 1) * (a.D + 1) * 3;
			const float* __restrict__ row = a.shs + (size_t)idx * a.M * 3;
			if (a.M == 16 && (reinterpret_cast<uintptr_t>(a.shs) & 15) == 0) {
				const float4* __restrict__ row4 = reinterpret_cast<const float4*>(row);
				#pragma unroll
				for (int i = 0; i < 12; i++) {
					if (4 * i < nfl) { const float4 v = __ldg(row4 + i); sh[4 * i] = v.x; sh[4 * i + 1] = v.y; sh[4 * i + 2] = v.z; sh[4 * i + 3] = v.w; }
				}
			} else {
				#pragma unroll
				for (int i = 0; i < 48; i++) if (i < nfl && i < a.M * 3) sh[i] = row[i];
			}
		}
		sh_colour(a.D, sh, px, py, pz, a.campos, rgb0, rgb1, rgb2, clamped);
	}
	a.clamped[idx] = clamped;

	const float opacity = a.opacities[idx];
	float ext_x, ext_y;
	cull_extent(conx, cony, conz, opacity, ext_x, ext_y);

	SplatRec rec;
	rec.a = make_float4(pix_x, pix_y, conx, cony);
	rec.b = make_float4(conz, opacity, rgb0, rgb1);
	rec.c = make_float4(rgb2, view_z, ext_x, ext_y);
	a.rec[idx] = rec;
}

// ---- view-batched forward preprocess --------------------------------------------------------------------------------------
// preprocessCUDA (forward.cu:144-246) for ALL views of a step in one pass over the Gaussians (rr_forward_geom_views; no
// reference counterpart: the reference runs it once per view).  One thread owns one Gaussian: its mean, scale, rotation and
// opacity are read once, its 3D covariance is built once, the warp's 32 SH rows are staged once (coalesced float4 reads into
// shared memory: 192 B per Gaussian, the largest input), and the views' cameras sit in shared memory.  Per view the thread then
// runs exactly the per-view kernel's arithmetic (project_splat / sh_colour / cull_extent above: same functions, same roundings,
// so every per-view output is bit-identical to the per-view kernel's) and writes that view's radii, depth key, tile count,
// rect, clamp flags and 48-byte record into the view's own geometry blob.
constexpr int PV_THREADS = 128;
constexpr int PV_WARPS = PV_THREADS / 32;
constexpr int PV_SH_STRIDE = 49;          // 48 floats + 1: a lane walking its own row is bank-conflict free
constexpr int PV_CAM_FLOATS = 40;         // V[16] Pj[16] campos[3] tan_fovx tan_fovy focal_x focal_y pad

// LEAF: the inputs are the six raw leaf tensors of scene/gaussian_model.py:131-136 (xyz, _features_dc [P,1,3], _features_rest
// [P,15,3], raw opacity / scaling / rotation); sigmoid, exp and normalise (:85-105) are applied here (splat_math.cuh), so no
// activated copy of the parameters is ever written or re-read.
template <bool LEAF>
__global__ void __launch_bounds__(PV_THREADS, 4) preprocess_views_kernel(PreprocessViewsArgs a)
{
	__shared__ __align__(16) float s_sh[PV_WARPS][32 * PV_SH_STRIDE];
	__shared__ float s_cam[PV_MAX_VIEWS][PV_CAM_FLOATS];
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int nv = a.nviews;
	for (int i = tid; i < nv * 35; i += PV_THREADS) {
		const int v = i / 35, k = i - v * 35;
		s_cam[v][k] = k < 16 ? a.view[v][k] : k < 32 ? a.proj[v][k - 16] : a.campos[v][k - 32];
	}
	if (tid < nv) {
		s_cam[tid][35] = a.tan_fovx[tid]; s_cam[tid][36] = a.tan_fovy[tid];
		s_cam[tid][37] = a.W / (2.0f * a.tan_fovx[tid]); s_cam[tid][38] = a.H / (2.0f * a.tan_fovy[tid]);       // rasterizer_impl.cu:213-214
	}
	__syncthreads();
	const int warp_base = blockIdx.x * PV_THREADS + warp * 32;
	if (warp_base >= a.P) return;                               // warp-uniform
	const int idx = warp_base + lane;
	const bool live = idx < a.P;
	const int nvalid = min(32, a.P - warp_base);
	const size_t ii = (size_t)(live ? idx : 0);
	// the warp's SH rows (48 floats each), staged with coalesced loads
	if (LEAF) {
		const float* __restrict__ dc = a.f_dc + (size_t)warp_base * 3;
		const float* __restrict__ rest = a.f_rest + (size_t)warp_base * 45;
		for (int f = lane; f < nvalid * 3; f += 32) { const int row = f / 3; s_sh[warp][row * PV_SH_STRIDE + (f - row * 3)] = __ldg(dc + f); }
		for (int f = lane; f < nvalid * 45; f += 32) { const int row = f / 45; s_sh[warp][row * PV_SH_STRIDE + 3 + (f - row * 45)] = __ldg(rest + f); }
	} else {
		const float* __restrict__ src = a.shs + (size_t)warp_base * 48;
		if ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0) {
			const float4* __restrict__ src4 = reinterpret_cast<const float4*>(src);
			for (int f4 = lane; f4 < nvalid * 12; f4 += 32) {
				const int row = f4 / 12, col = (f4 - row * 12) * 4;
				const float4 v = __ldg(src4 + f4);
				float* d = s_sh[warp] + row * PV_SH_STRIDE + col;
				d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
			}
		} else {
			for (int f = lane; f < nvalid * 48; f += 32) { const int row = f / 48; s_sh[warp][row * PV_SH_STRIDE + (f - row * 48)] = __ldg(src + f); }
		}
	}
	__syncwarp();
	if (!live) return;
	const float* sh = s_sh[warp] + lane * PV_SH_STRIDE;
	const float px = a.means3D[3 * ii], py = a.means3D[3 * ii + 1], pz = a.means3D[3 * ii + 2];
	float opacity = a.opacities[ii];
	float s0 = a.scales[3 * ii], s1 = a.scales[3 * ii + 1], s2 = a.scales[3 * ii + 2];
	float4 q = load_f4(a.rotations, ii);
	if (LEAF) { opacity = act_opacity(opacity); s0 = act_scale(s0); s1 = act_scale(s1); s2 = act_scale(s2); q = act_rot(q); }
	float c3[6];
	cov3d_from_scale_rot(s0, s1, s2, a.scale_modifier, q, c3);

	for (int v = 0; v < nv; v++) {
		const float* cam = s_cam[v];
		char* __restrict__ g = a.geom[v];
		const float view_z = xform(cam, 2, px, py, pz);
		Projected o;
		o.radius = 0; o.tiles = 0; o.depth_key = 0xFFFFFFFFu; o.rect = make_uint2(0u, 0u); o.visible = false;
		o.conx = o.cony = o.conz = o.pix_x = o.pix_y = 0.0f;
		if (!(view_z <= 0.2f)) {
			if (a.write_cov3D) {                  // only the per-view backward reads it back (the view-batched one rebuilds Sigma)
				float* c = reinterpret_cast<float*>(g + a.off_cov3D) + 6 * ii;
				#pragma unroll
				for (int k = 0; k < 6; k++) c[k] = c3[k];
			}
			project_splat(cam, cam + 16, px, py, pz, view_z, c3, cam[35], cam[36], cam[37], cam[38], a.low_pass, a.W, a.H, a.grid_x, a.grid_y, o);
		}
		a.radii[v][ii] = o.radius;
		reinterpret_cast<uint32_t*>(g + a.off_tiles)[ii] = o.tiles;
		reinterpret_cast<uint32_t*>(g + a.off_key)[ii] = o.depth_key;
		if (!o.visible) continue;
		reinterpret_cast<uint2*>(g + a.off_rect)[ii] = o.rect;
		float rgb0, rgb1, rgb2, ext_x, ext_y;
		uint8_t clamped;
		sh_colour(a.D, sh, px, py, pz, cam + 32, rgb0, rgb1, rgb2, clamped);
		reinterpret_cast<uint8_t*>(g + a.off_clamped)[ii] = clamped;
		cull_extent(o.conx, o.cony, o.conz, opacity, ext_x, ext_y);
		SplatRec rec;
		rec.a = make_float4(o.pix_x, o.pix_y, o.conx, o.cony);
		rec.b = make_float4(o.conz, opacity, rgb0, rgb1);
		rec.c = make_float4(rgb2, view_z, ext_x, ext_y);
		reinterpret_cast<SplatRec*>(g + a.off_rec)[ii] = rec;
	}
}

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view, uint8_t* __restrict__ present)
{
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;
	const float z = xform(view, 2, means3D[3 * (size_t)idx], means3D[3 * (size_t)idx + 1], means3D[3 * (size_t)idx + 2]);
	present[idx] = z > 0.2f ? 1 : 0;
}

}  // namespace

int launch_preprocess_fwd(cudaStream_t stream, const PreprocessArgs& a)
{
	if (a.P == 0) return 0;
	if (a.tile_cull) count_launch(), preprocess_fwd_kernel<true><<<(a.P + 255) / 256, 256, 0, stream>>>(a);
	else count_launch(), preprocess_fwd_kernel<false><<<(a.P + 255) / 256, 256, 0, stream>>>(a);
	return check_launch("preprocess_fwd", stream, false);
}

int launch_preprocess_views(cudaStream_t stream, const PreprocessViewsArgs& a)
{
	if (a.P == 0 || a.nviews == 0) return 0;
	count_launch();
	if (a.leaf) preprocess_views_kernel<true><<<(a.P + PV_THREADS - 1) / PV_THREADS, PV_THREADS, 0, stream>>>(a);
	else preprocess_views_kernel<false><<<(a.P + PV_THREADS - 1) / PV_THREADS, PV_THREADS, 0, stream>>>(a);
	return check_launch("preprocess_views", stream, false);
}

int launch_mark_visible(cudaStream_t stream, int P, const float* means3D, const float* view, uint8_t* present)
{
	if (P == 0) return 0;
	count_launch(), mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
	return check_launch("mark_visible", stream, false);
}

}  // namespace rr
