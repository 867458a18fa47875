// TEST INFRASTRUCTURE — see oracle.h.  CPU restatement of the reference rasterizer.
//
// Floating-point contract.  The reference is only ever run as nvcc-compiled CUDA, whose default -fmad=true
// contracts specific multiply/add pairs into FMAs.  Integer outputs (radii, tile rects, keys) depend on those
// roundings, so the forward-preprocess restatement below spells out every rounding step explicitly
// (fmaf = one rounding, * and + = one rounding each; the file is compiled with -ffp-contract=off) in the
// order the reference build performs them (worked out from `nvcc -ptx` / `cuobjdump -sass` of the reference's
// forward.cu, see DESIGN.md "FMA contract"): for a glm sum of products a*b + c*d + e*f (type_mat3x3.inl:509-517,
// func_geometric.inl:52-53) the build evaluates fma(e,f, fma(a,b, c*d)); a product whose only consumers are
// additions/subtractions is folded into them.  IEEE division, sqrt and reciprocal are correctly rounded on both
// sides, so this half of the oracle is bit-comparable with the GPU.  The only library arithmetic that is not
// is expf in the blend (forward.cu:336, backward.cu:488): the GPU uses libdevice's ex2.approx-based expf,
// here it is glibc expf — hence tolerance-based comparison for colour/depth/gradients.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>
#include <omp.h>

namespace {

constexpr int BLOCK_X = 16, BLOCK_Y = 16;  // config.h:5-6

// auxiliary.h:11-28
constexpr float SH_C0 = 0.28209479177387814f;
constexpr float SH_C1 = 0.4886025119029199f;
constexpr float SH_C2[] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
constexpr float SH_C3[] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };

// a*b + c*d + e*f as the reference build rounds it
inline float sop3(float a, float b, float c, float d, float e, float f) { return fmaf(e, f, fmaf(a, b, c * d)); }

// transformPoint4x3 / 4x4 row `col` (auxiliary.h:47-66): m[col]*x + m[4+col]*y + m[8+col]*z + m[12+col]
inline float xform(const float* m, int col, float x, float y, float z) { return m[12 + col] + sop3(x, m[col], y, m[4 + col], z, m[8 + col]); }

struct Ctx {
	int P = 0, D = 0, M = 0, W = 0, H = 0, gx = 0, gy = 0;
	float scale_modifier = 1.f, tan_fovx = 0, tan_fovy = 0, low_pass = 0.3f;
	std::vector<float> bg, means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp, view, proj, campos;
	bool has_sh = false, has_colors = false, has_scales = false, has_cov = false;
	std::vector<int> radii;
	std::vector<float> depths, means2D, cov3D, conic_opacity, rgb, final_T;
	std::vector<uint8_t> clamped;
	std::vector<uint32_t> tiles_touched, point_offsets, values_unsorted, point_list, ranges, n_contrib;
	std::vector<uint64_t> keys_unsorted, keys;
	int64_t N = 0;
	int64_t counters[4] = { 0, 0, 0, 0 };
};

// getHigherMsb (rasterizer_impl.cu:24-39)
uint32_t higher_msb(uint32_t n) {
	uint32_t msb = sizeof(n) * 4, step = msb;
	while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
	if (n >> msb) msb++;
	return msb;
}

// getRect (auxiliary.h:35-45); division by BLOCK_X=16 is exact either way
inline void get_rect(float px, float py, int max_radius, int gx, int gy, uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
	float r = (float)max_radius;
	auto clampu = [](uint32_t g, int v) { return std::min(g, (uint32_t)std::max(0, v)); };
	x0 = clampu(gx, (int)((px - r) / BLOCK_X));
	y0 = clampu(gy, (int)((py - r) / BLOCK_Y));
	x1 = clampu(gx, (int)((((px + r) + (float)BLOCK_X) - 1.0f) / BLOCK_X));
	y1 = clampu(gy, (int)((((py + r) + (float)BLOCK_Y) - 1.0f) / BLOCK_Y));
}

// computeCov3D (forward.cu:107-141)
void cov3d_from_scale_rot(const float* sc, float mod, const float* q, float* c) {
	const float sx = mod * sc[0], sy = mod * sc[1], sz = mod * sc[2];
	const float r = q[0], x = q[1], y = q[2], z = q[3];
	const float xz = x * z, A = fmaf(r, y, xz), B = fmaf(-r, y, xz);
	const float rx = r * x, C = fmaf(y, z, -rx), Dd = fmaf(y, z, rx);
	const float rz = r * z, E = fmaf(x, y, -rz), F = fmaf(x, y, rz);
	const float yy = y * y, zz = z * z, G = fmaf(x, x, yy), Hh = yy + zz, I = fmaf(x, x, zz);
	// glm column-major R (forward.cu:124-128): R[c][r]
	const float R00 = 1.f - (Hh + Hh), R01 = E + E, R02 = A + A;
	const float R10 = F + F, R11 = 1.f - (I + I), R12 = C + C;
	const float R20 = B + B, R21 = Dd + Dd, R22 = 1.f - (G + G);
	// M = S * R with S = diag(s) (forward.cu:130); the zero products are kept as the build keeps them
	const float z0 = 0.0f;
	const float M00 = fmaf(R02, z0, fmaf(R01, z0, sx * R00));
	const float t0 = R00 * z0;
	const float M01 = fmaf(R02, z0, fmaf(sy, R01, t0));
	const float M02 = fmaf(sz, R02, fmaf(R01, z0, t0));
	const float t1 = R11 * z0;
	const float M10 = fmaf(R12, z0, fmaf(sx, R10, t1));
	const float M11 = fmaf(R12, z0, fmaf(R10, z0, sy * R11));
	const float M12 = fmaf(sz, R12, fmaf(R10, z0, t1));
	const float t2 = R21 * z0;
	const float M20 = fmaf(R22, z0, fmaf(sx, R20, t2));
	const float M21 = fmaf(R22, z0, fmaf(R20, z0, sy * R21));
	const float M22 = fmaf(sz, R22, fmaf(R20, z0, t2));
	// Sigma = M^T M (forward.cu:133), upper triangle (forward.cu:135-140)
	c[0] = sop3(M00, M00, M01, M01, M02, M02);
	c[1] = sop3(M10, M00, M11, M01, M12, M02);
	c[2] = sop3(M20, M00, M21, M01, M22, M02);
	c[3] = sop3(M10, M10, M11, M11, M12, M12);
	c[4] = sop3(M20, M10, M21, M11, M22, M12);
	c[5] = sop3(M20, M20, M21, M21, M22, M22);
}

// computeCov2D (forward.cu:63-102); returns (cov.x, cov.y, cov.z) and the T entries (reused by the backward check)
void cov2d(const float* p, float fx, float fy, float tan_fovx, float tan_fovy, const float* c, const float* V, float low_pass, float out[3]) {
	const float tx0 = xform(V, 0, p[0], p[1], p[2]), ty0 = xform(V, 1, p[0], p[1], p[2]), tz = xform(V, 2, p[0], p[1], p[2]);
	const float limx = tan_fovx * 1.3f, limy = tan_fovy * 1.3f;
	const float txtz = tx0 / tz, tytz = ty0 / tz;
	const float cx = std::min(limx, std::max(-limx, txtz));
	const float cy = std::min(limy, std::max(-limy, tytz));
	const float J00 = fx / tz;
	const float J02 = (fx * (cx * (-tz))) / (tz * tz);   // -(focal_x * t.x) / (t.z * t.z), t.x = cx * t.z
	const float J11 = fy / tz;
	const float J12 = (fy * (cy * (-tz))) / (tz * tz);
	const float z0 = 0.0f;
	// T = W * J (forward.cu:89), W columns = (V0,V4,V8),(V1,V5,V9),(V2,V6,V10)
	const float T00 = fmaf(V[2], J02, fmaf(V[0], J00, V[1] * z0));
	const float T01 = fmaf(V[6], J02, fmaf(V[4], J00, V[5] * z0));
	const float T02 = fmaf(J02, V[10], fmaf(V[8], J00, V[9] * z0));
	const float T10 = fmaf(V[2], J12, fmaf(V[0], z0, J11 * V[1]));
	const float T11 = fmaf(V[6], J12, fmaf(V[4], z0, J11 * V[5]));
	const float T12 = fmaf(J12, V[10], fmaf(V[8], z0, J11 * V[9]));
	// A = T^T * Vrk^T, cov = A * T (forward.cu:96)
	const float A00 = sop3(T00, c[0], T01, c[1], T02, c[2]);
	const float A01 = sop3(T10, c[0], T11, c[1], T12, c[2]);
	const float A10 = sop3(T00, c[1], T01, c[3], T02, c[4]);
	const float A11 = sop3(T10, c[1], T11, c[3], T12, c[4]);
	const float A20 = sop3(T00, c[2], T01, c[4], T02, c[5]);
	const float A21 = sop3(T10, c[2], T11, c[4], T12, c[5]);
	const float c00 = sop3(T00, A00, T01, A10, T02, A20);
	const float c01 = sop3(T00, A01, T01, A11, T02, A21);
	const float c11 = sop3(T10, A01, T11, A11, T12, A21);
	out[0] = low_pass + c00;   // forward.cu:99-100
	out[1] = c01;
	out[2] = low_pass + c11;
}

// computeColorFromSH (forward.cu:9-60)
void sh_to_rgb(int deg, const float* p, const float* campos, const float* sh /*[M][3]*/, float rgb[3], uint8_t clamped[3]) {
	const float dx = p[0] - campos[0], dy = p[1] - campos[1], dz = p[2] - campos[2];
	const float len = sqrtf(sop3(dx, dx, dy, dy, dz, dz));
	const float x = dx / len, y = dy / len, z = dz / len;
	float res[3];
	for (int ch = 0; ch < 3; ch++) res[ch] = sh[0 * 3 + ch] * SH_C0;
	if (deg > 0) {
		const float a = y * SH_C1, b = z * SH_C1, c = x * SH_C1;
		for (int ch = 0; ch < 3; ch++) {
			float r = fmaf(-a, sh[1 * 3 + ch], res[ch]);
			r = fmaf(b, sh[2 * 3 + ch], r);
			res[ch] = fmaf(-c, sh[3 * 3 + ch], r);
		}
		if (deg > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			const float zz2 = zz + zz;
			const float t4 = xy * SH_C2[0], t5 = yz * SH_C2[1], t6 = ((zz2 - xx) - yy) * SH_C2[2], t7 = xz * SH_C2[3];
			const float xx_yy = xx - yy, t8 = xx_yy * SH_C2[4];
			for (int ch = 0; ch < 3; ch++) {
				float r = fmaf(t4, sh[4 * 3 + ch], res[ch]);
				r = fmaf(t5, sh[5 * 3 + ch], r);
				r = fmaf(t6, sh[6 * 3 + ch], r);
				r = fmaf(t7, sh[7 * 3 + ch], r);
				res[ch] = fmaf(t8, sh[8 * 3 + ch], r);
			}
			if (deg > 2) {
				const float t9 = (y * SH_C3[0]) * fmaf(xx, 3.0f, -yy);
				const float t10 = z * (xy * SH_C3[1]);
				const float q4 = fmaf(zz, 4.0f, -xx) - yy;                 // 4zz - xx - yy
				const float t11 = (y * SH_C3[2]) * q4;
				const float t12 = (z * SH_C3[3]) * fmaf(yy, -3.0f, fmaf(xx, -3.0f, zz2)); // 2zz - 3xx - 3yy
				const float t13 = (x * SH_C3[4]) * q4;
				const float t14 = (z * SH_C3[5]) * xx_yy;
				const float t15 = (x * SH_C3[6]) * fmaf(yy, -3.0f, xx);
				for (int ch = 0; ch < 3; ch++) {
					float r = fmaf(t9, sh[9 * 3 + ch], res[ch]);
					r = fmaf(t10, sh[10 * 3 + ch], r);
					r = fmaf(t11, sh[11 * 3 + ch], r);
					r = fmaf(t12, sh[12 * 3 + ch], r);
					r = fmaf(t13, sh[13 * 3 + ch], r);
					r = fmaf(t14, sh[14 * 3 + ch], r);
					res[ch] = fmaf(t15, sh[15 * 3 + ch], r);
				}
			}
		}
	}
	for (int ch = 0; ch < 3; ch++) {
		const float v = res[ch] + 0.5f;
		clamped[ch] = v < 0.0f;                    // forward.cu:56-58
		rgb[ch] = (v < 0.0f) ? 0.0f : v;           // glm::max(result, 0.0f)
	}
}

// preprocessCUDA forward (forward.cu:144-246), one Gaussian
void preprocess_one(Ctx& s, int idx, float fx, float fy) {
	s.radii[idx] = 0; s.tiles_touched[idx] = 0;
	const float* p = &s.means3D[3 * (size_t)idx];
	const float* V = s.view.data(); const float* Pm = s.proj.data();
	const float view_z = xform(V, 2, p[0], p[1], p[2]);
	if (view_z <= 0.2f) return;                                    // in_frustum, auxiliary.h:143
	const float hx = xform(Pm, 0, p[0], p[1], p[2]), hy = xform(Pm, 1, p[0], p[1], p[2]), hw = xform(Pm, 3, p[0], p[1], p[2]);
	const float p_w = 1.0f / (hw + 0.0000001f);
	const float projx = hx * p_w, projy = hy * p_w;
	const float* c3;
	if (s.has_cov) c3 = &s.cov3D_precomp[6 * (size_t)idx];
	else { cov3d_from_scale_rot(&s.scales[3 * (size_t)idx], s.scale_modifier, &s.rotations[4 * (size_t)idx], &s.cov3D[6 * (size_t)idx]); c3 = &s.cov3D[6 * (size_t)idx]; }
	float cov[3];
	cov2d(p, fx, fy, s.tan_fovx, s.tan_fovy, c3, V, s.low_pass, cov);
	const float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));    // forward.cu:208
	if (det == 0.0f) return;
	const float det_inv = 1.f / det;
	const float conic[3] = { cov[2] * det_inv, det_inv * (-cov[1]), cov[0] * det_inv };
	const float mid = (cov[0] + cov[2]) * 0.5f;
	const float sq = sqrtf(std::max(fmaf(mid, mid, -det), 0.1f));
	const float lambda1 = mid + sq, lambda2 = mid - sq;
	const float my_radius = ceilf(sqrtf(std::max(lambda1, lambda2)) * 3.f);
	// ndc2Pix (auxiliary.h:30-33) is evaluated in double: ((v + 1.0) * S - 1.0) * 0.5
	const float pix_x = (float)(fma((double)projx + 1.0, (double)s.W, -1.0) * 0.5);
	const float pix_y = (float)(fma((double)projy + 1.0, (double)s.H, -1.0) * 0.5);
	uint32_t x0, y0, x1, y1;
	get_rect(pix_x, pix_y, (int)my_radius, s.gx, s.gy, x0, y0, x1, y1);
	if ((x1 - x0) * (y1 - y0) == 0) return;
	if (!s.has_colors)
		sh_to_rgb(s.D, p, s.campos.data(), &s.shs[(size_t)idx * s.M * 3], &s.rgb[3 * (size_t)idx], &s.clamped[3 * (size_t)idx]);
	s.depths[idx] = view_z;
	s.radii[idx] = (int)my_radius;
	s.means2D[2 * (size_t)idx] = pix_x; s.means2D[2 * (size_t)idx + 1] = pix_y;
	float* co = &s.conic_opacity[4 * (size_t)idx];
	co[0] = conic[0]; co[1] = conic[1]; co[2] = conic[2]; co[3] = s.opacities[idx];
	s.tiles_touched[idx] = (y1 - y0) * (x1 - x0);
}

inline uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

// renderCUDA forward (forward.cu:251-369) for one tile; per-pixel loop nest (block-cooperative fetching has no numerical effect)
void render_tile(Ctx& s, int tx, int ty, const float* feat, float* out_color, float* out_depth, int64_t& evals, int64_t& contribs) {
	const int W = s.W, H = s.H;
	const uint32_t r0 = s.ranges[2 * (size_t)(ty * s.gx + tx)], r1 = s.ranges[2 * (size_t)(ty * s.gx + tx) + 1];
	for (int ly = 0; ly < BLOCK_Y; ly++) for (int lx = 0; lx < BLOCK_X; lx++) {
		const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
		if (px >= W || py >= H) continue;
		const float pxf = (float)px, pyf = (float)py;
		float T = 1.0f, C[3] = { 0, 0, 0 }, Dp = 0;
		uint32_t contributor = 0, last_contributor = 0;
		for (uint32_t i = r0; i < r1; i++) {
			contributor++;
			const uint32_t id = s.point_list[i];
			const float dx = s.means2D[2 * (size_t)id] - pxf, dy = s.means2D[2 * (size_t)id + 1] - pyf;
			const float* co = &s.conic_opacity[4 * (size_t)id];
			// power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y  (forward.cu:328)
			const float q = fmaf(dx, dx * co[0], dy * (dy * co[2]));
			const float power = fmaf(q, -0.5f, -(dy * (dx * co[1])));
			if (power > 0.0f) continue;
			const float alpha = std::min(0.99f, co[3] * expf(power));
			if (alpha < 1.0f / 255.0f) continue;
			const float test_T = T * (1 - alpha);
			if (test_T < 0.0001f) break;                            // done = true
			for (int ch = 0; ch < 3; ch++) C[ch] = fmaf(T, alpha * feat[3 * (size_t)id + ch], C[ch]);
			Dp = fmaf(T, alpha * s.depths[id], Dp);
			T = test_T;
			last_contributor = contributor;
			contribs++;
		}
		evals += contributor;
		const size_t pid = (size_t)W * py + px;
		s.final_T[pid] = T; s.n_contrib[pid] = last_contributor;
		for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pid] = fmaf(T, s.bg[ch], C[ch]);
		out_depth[pid] = Dp;
	}
}

template <class T> void setv(std::vector<T>& v, const T* p, size_t n) { if (p) v.assign(p, p + n); else v.clear(); }

}  // namespace

extern "C" {

orc_ctx* orc_create(void) { return reinterpret_cast<orc_ctx*>(new Ctx()); }
void orc_destroy(orc_ctx* c) { delete reinterpret_cast<Ctx*>(c); }
int orc_num_threads(void) { return omp_get_max_threads(); }

int64_t orc_forward(orc_ctx* c, int P, int D, int M, const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass,
	float* out_color, float* out_depth, int* radii)
{
	Ctx& s = *reinterpret_cast<Ctx*>(c);
	s.P = P; s.D = D; s.M = M; s.W = W; s.H = H; s.scale_modifier = scale_modifier;
	s.tan_fovx = tan_fovx; s.tan_fovy = tan_fovy; s.low_pass = low_pass;
	s.gx = (W + BLOCK_X - 1) / BLOCK_X; s.gy = (H + BLOCK_Y - 1) / BLOCK_Y;
	setv(s.bg, background, 3); setv(s.means3D, means3D, (size_t)3 * P); setv(s.shs, shs, (size_t)P * M * 3);
	setv(s.colors_precomp, colors_precomp, (size_t)3 * P); setv(s.opacities, opacities, (size_t)P);
	setv(s.scales, scales, (size_t)3 * P); setv(s.rotations, rotations, (size_t)4 * P); setv(s.cov3D_precomp, cov3D_precomp, (size_t)6 * P);
	setv(s.view, viewmatrix, 16); setv(s.proj, projmatrix, 16); setv(s.campos, campos, 3);
	s.has_sh = shs != nullptr; s.has_colors = colors_precomp != nullptr; s.has_scales = scales != nullptr; s.has_cov = cov3D_precomp != nullptr;
	if (!s.has_sh && !s.has_colors) return -1;
	if (!s.has_cov && !(s.has_scales && rotations)) return -1;
	const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);   // rasterizer_impl.cu:213-214
	s.radii.assign(P, 0); s.depths.assign(P, 0.f); s.means2D.assign((size_t)2 * P, 0.f); s.cov3D.assign((size_t)6 * P, 0.f);
	s.conic_opacity.assign((size_t)4 * P, 0.f); s.rgb.assign((size_t)3 * P, 0.f); s.clamped.assign((size_t)3 * P, 0);
	s.tiles_touched.assign(P, 0); s.point_offsets.assign(P, 0);
	#pragma omp parallel for schedule(static)
	for (int i = 0; i < P; i++) preprocess_one(s, i, focal_x, focal_y);
	// InclusiveSum (rasterizer_impl.cu:269)
	uint64_t acc = 0;
	for (int i = 0; i < P; i++) { acc += s.tiles_touched[i]; s.point_offsets[i] = (uint32_t)acc; }
	const int64_t N = P > 0 ? (int64_t)s.point_offsets[P - 1] : 0;
	s.N = N;
	s.keys_unsorted.assign(N, 0); s.values_unsorted.assign(N, 0);
	// duplicateWithKeys (rasterizer_impl.cu:59-100)
	#pragma omp parallel for schedule(dynamic, 4096)
	for (int idx = 0; idx < P; idx++) {
		if (s.radii[idx] > 0) {
			uint32_t off = idx == 0 ? 0 : s.point_offsets[idx - 1];
			uint32_t x0, y0, x1, y1;
			get_rect(s.means2D[2 * (size_t)idx], s.means2D[2 * (size_t)idx + 1], s.radii[idx], s.gx, s.gy, x0, y0, x1, y1);
			for (uint32_t y = y0; y < y1; y++) for (uint32_t x = x0; x < x1; x++) {
				uint64_t key = (uint64_t)(y * s.gx + x);
				key <<= 32; key |= fbits(s.depths[idx]);
				s.keys_unsorted[off] = key; s.values_unsorted[off] = (uint32_t)idx; off++;
			}
		}
	}
	// SortPairs over bits [0, 32+bit) (rasterizer_impl.cu:292-300): stable, ascending
	const int bit = (int)higher_msb((uint32_t)(s.gx * s.gy));
	const uint64_t mask = (32 + bit >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
	std::vector<uint32_t> perm(N);
	std::iota(perm.begin(), perm.end(), 0u);
	std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return (s.keys_unsorted[a] & mask) < (s.keys_unsorted[b] & mask); });
	s.keys.resize(N); s.point_list.resize(N);
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < N; i++) { s.keys[i] = s.keys_unsorted[perm[i]]; s.point_list[i] = s.values_unsorted[perm[i]]; }
	// memset + identifyTileRanges (rasterizer_impl.cu:302-311,105-127)
	s.ranges.assign((size_t)2 * s.gx * s.gy, 0);
	for (int64_t i = 0; i < N; i++) {
		const uint32_t cur = (uint32_t)(s.keys[i] >> 32);
		if (i == 0) s.ranges[2 * (size_t)cur] = 0;
		else {
			const uint32_t prev = (uint32_t)(s.keys[i - 1] >> 32);
			if (cur != prev) { s.ranges[2 * (size_t)prev + 1] = (uint32_t)i; s.ranges[2 * (size_t)cur] = (uint32_t)i; }
		}
		if (i == N - 1) s.ranges[2 * (size_t)cur + 1] = (uint32_t)N;
	}
	// FORWARD::render (rasterizer_impl.cu:314-328)
	s.final_T.assign((size_t)W * H, 0.f); s.n_contrib.assign((size_t)W * H, 0);
	const float* feat = s.has_colors ? s.colors_precomp.data() : s.rgb.data();
	int64_t evals = 0, contribs = 0;
	#pragma omp parallel for collapse(2) schedule(dynamic, 1) reduction(+ : evals, contribs)
	for (int ty = 0; ty < s.gy; ty++) for (int tx = 0; tx < s.gx; tx++) render_tile(s, tx, ty, feat, out_color, out_depth, evals, contribs);
	s.counters[0] = evals; s.counters[1] = contribs;
	memcpy(radii, s.radii.data(), sizeof(int) * (size_t)P);
	return N;
}

int64_t orc_get(orc_ctx* c, const char* name, const void** ptr) {
	Ctx& s = *reinterpret_cast<Ctx*>(c);
	const std::string n(name);
#define G(nm, vec) if (n == nm) { *ptr = s.vec.data(); return (int64_t)s.vec.size(); }
	G("depths", depths) G("means2D", means2D) G("cov3D", cov3D) G("conic_opacity", conic_opacity) G("rgb", rgb)
	G("clamped", clamped) G("tiles_touched", tiles_touched) G("point_offsets", point_offsets)
	G("keys_unsorted", keys_unsorted) G("values_unsorted", values_unsorted) G("keys", keys) G("point_list", point_list)
	G("ranges", ranges) G("final_T", final_T) G("n_contrib", n_contrib) G("radii", radii)
#undef G
	return -1;
}

void orc_counters(orc_ctx* c, int64_t out[4]) { Ctx& s = *reinterpret_cast<Ctx*>(c); for (int i = 0; i < 4; i++) out[i] = s.counters[i]; }

void orc_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present) {
	(void)projmatrix;
	for (int i = 0; i < P; i++) present[i] = xform(viewmatrix, 2, means3D[3 * (size_t)i], means3D[3 * (size_t)i + 1], means3D[3 * (size_t)i + 2]) > 0.2f;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// Backward.  Gradients are compared at 1e-4 relative (the reference accumulates them with unordered float
// atomics, backward.cu:513-544), so this half is a plain-float restatement of the formulas; per-pair terms are
// formed in float exactly as written in the reference and SUMMED in double in a fixed order, i.e. the oracle is
// the exactly-rounded value both GPU implementations scatter around.
namespace {

struct V3 { float x, y, z; };

// dnormvdv (auxiliary.h:96-106)
V3 dnormvdv3(V3 v, V3 dv) {
	float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
	float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
	V3 o;
	o.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
	o.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
	o.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
	return o;
}

// BACKWARD renderCUDA (backward.cu:389-547), one tile; partial[(i - r0)*9 + k] accumulates, in double, the 9 sums
// of entry i: 0..2 dL_dcolor, 3..4 dL_dmean2D.xy, 5..7 dL_dconic.{x,y,w}, 8 dL_dopacity
void render_tile_bw(const Ctx& s, int tx, int ty, const float* colors, const float* dL_dpix, double* partial, int64_t& evals, int64_t& contribs) {
	const int W = s.W, H = s.H;
	const uint32_t r0 = s.ranges[2 * (size_t)(ty * s.gx + tx)], r1 = s.ranges[2 * (size_t)(ty * s.gx + tx) + 1];
	const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
	for (int ly = 0; ly < BLOCK_Y; ly++) for (int lx = 0; lx < BLOCK_X; lx++) {
		const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
		if (px >= W || py >= H) continue;
		const size_t pid = (size_t)W * py + px;
		const float pxf = (float)px, pyf = (float)py;
		const float T_final = s.final_T[pid];
		float T = T_final;
		const uint32_t last_contributor = s.n_contrib[pid];
		float accum_rec[3] = { 0, 0, 0 }, dL_dpixel[3], last_color[3] = { 0, 0, 0 }, last_alpha = 0;
		for (int ch = 0; ch < 3; ch++) dL_dpixel[ch] = dL_dpix[(size_t)ch * H * W + pid];
		uint32_t contributor = r1 - r0;
		for (uint32_t i = r1; i-- > r0;) {
			contributor--;
			if (contributor >= last_contributor) continue;
			evals++;
			const uint32_t id = s.point_list[i];
			const float dx = s.means2D[2 * (size_t)id] - pxf, dy = s.means2D[2 * (size_t)id + 1] - pyf;
			const float* co = &s.conic_opacity[4 * (size_t)id];
			const float q = fmaf(dx, dx * co[0], dy * (dy * co[2]));
			const float power = fmaf(q, -0.5f, -(dy * (dx * co[1])));    // same rounding as the forward so the same pairs pass
			if (power > 0.0f) continue;
			const float G = expf(power);
			const float alpha = std::min(0.99f, co[3] * G);
			if (alpha < 1.0f / 255.0f) continue;
			contribs++;
			T = T / (1.f - alpha);
			const float dchannel_dcolor = alpha * T;
			double* pr = partial + (size_t)(i - r0) * 9;
			float dL_dalpha = 0.0f;
			for (int ch = 0; ch < 3; ch++) {
				const float c = colors[3 * (size_t)id + ch];
				accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
				last_color[ch] = c;
				const float dL_dchannel = dL_dpixel[ch];
				dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
				pr[ch] += (double)(dchannel_dcolor * dL_dchannel);
			}
			dL_dalpha *= T;
			last_alpha = alpha;
			float bg_dot_dpixel = 0;
			for (int ch = 0; ch < 3; ch++) bg_dot_dpixel += s.bg[ch] * dL_dpixel[ch];
			dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
			const float dL_dG = co[3] * dL_dalpha;
			const float gdx = G * dx, gdy = G * dy;
			const float dG_ddelx = -gdx * co[0] - gdy * co[1];
			const float dG_ddely = -gdy * co[2] - gdx * co[1];
			pr[3] += (double)(dL_dG * dG_ddelx * ddelx_dx);
			pr[4] += (double)(dL_dG * dG_ddely * ddely_dy);
			pr[5] += (double)(-0.5f * gdx * dx * dL_dG);
			pr[6] += (double)(-0.5f * gdx * dy * dL_dG);
			pr[7] += (double)(-0.5f * gdy * dy * dL_dG);
			pr[8] += (double)(G * dL_dalpha);
		}
	}
}

// computeCov2DCUDA (backward.cu:133-264), one Gaussian
void cov2d_bw(const Ctx& s, int idx, float h_x, float h_y, const float* cov3D, const float* dL_dconics, float* dL_dmeans, float* dL_dcov) {
	const float* V = s.view.data();
	const float* mean = &s.means3D[3 * (size_t)idx];
	const float dcx = dL_dconics[4 * (size_t)idx], dcy = dL_dconics[4 * (size_t)idx + 1], dcz = dL_dconics[4 * (size_t)idx + 3];
	float tx = V[0] * mean[0] + V[4] * mean[1] + V[8] * mean[2] + V[12];
	float ty = V[1] * mean[0] + V[5] * mean[1] + V[9] * mean[2] + V[13];
	const float tz = V[2] * mean[0] + V[6] * mean[1] + V[10] * mean[2] + V[14];
	const float limx = 1.3f * s.tan_fovx, limy = 1.3f * s.tan_fovy;
	const float txtz = tx / tz, tytz = ty / tz;
	tx = std::min(limx, std::max(-limx, txtz)) * tz;
	ty = std::min(limy, std::max(-limy, tytz)) * tz;
	const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
	const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
	// J, W, T in the reference's glm [col][row] indexing
	float J[3][3] = { { h_x / tz, 0.0f, -(h_x * tx) / (tz * tz) }, { 0.0f, h_y / tz, -(h_y * ty) / (tz * tz) }, { 0, 0, 0 } };
	float Wm[3][3] = { { V[0], V[4], V[8] }, { V[1], V[5], V[9] }, { V[2], V[6], V[10] } };
	float Vrk[3][3] = { { cov3D[0], cov3D[1], cov3D[2] }, { cov3D[1], cov3D[3], cov3D[4] }, { cov3D[2], cov3D[4], cov3D[5] } };
	float T[3][3], A[3][3], C2[3][3];
	auto mul = [](const float a[3][3], const float b[3][3], float o[3][3]) {   // glm operator*: o[c][r] = sum_k a[k][r]*b[c][k]
		for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) o[c][r] = a[0][r] * b[c][0] + a[1][r] * b[c][1] + a[2][r] * b[c][2];
	};
	mul(Wm, J, T);
	float Tt[3][3], Vt[3][3];
	for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) { Tt[c][r] = T[r][c]; Vt[c][r] = Vrk[r][c]; }
	mul(Tt, Vt, A); mul(A, T, C2);
	const float a = C2[0][0] + s.low_pass, b = C2[0][1], c = C2[1][1] + s.low_pass;
	const float denom = a * c - b * b;
	float dL_da = 0, dL_db = 0, dL_dc = 0;
	const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
	if (denom2inv != 0) {
		dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
		dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
		dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
		dL_dcov[6 * (size_t)idx + 0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
		dL_dcov[6 * (size_t)idx + 3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
		dL_dcov[6 * (size_t)idx + 5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
		dL_dcov[6 * (size_t)idx + 1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
		dL_dcov[6 * (size_t)idx + 2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
		dL_dcov[6 * (size_t)idx + 4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
	} else {
		for (int i = 0; i < 6; i++) dL_dcov[6 * (size_t)idx + i] = 0;
	}
	const float dL_dT00 = 2 * (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_da + (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_db;
	const float dL_dT01 = 2 * (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_da + (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_db;
	const float dL_dT02 = 2 * (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_da + (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_db;
	const float dL_dT10 = 2 * (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc + (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_db;
	const float dL_dT11 = 2 * (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc + (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_db;
	const float dL_dT12 = 2 * (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc + (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_db;
	const float dL_dJ00 = Wm[0][0] * dL_dT00 + Wm[0][1] * dL_dT01 + Wm[0][2] * dL_dT02;
	const float dL_dJ02 = Wm[2][0] * dL_dT00 + Wm[2][1] * dL_dT01 + Wm[2][2] * dL_dT02;
	const float dL_dJ11 = Wm[1][0] * dL_dT10 + Wm[1][1] * dL_dT11 + Wm[1][2] * dL_dT12;
	const float dL_dJ12 = Wm[2][0] * dL_dT10 + Wm[2][1] * dL_dT11 + Wm[2][2] * dL_dT12;
	const float itz = 1.f / tz, tz2 = itz * itz, tz3 = tz2 * itz;
	const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
	const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
	const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tx) * tz3 * dL_dJ02 + (2 * h_y * ty) * tz3 * dL_dJ12;
	// transformVec4x3Transpose (auxiliary.h:86-94); ASSIGNMENT (backward.cu:263)
	dL_dmeans[3 * (size_t)idx + 0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
	dL_dmeans[3 * (size_t)idx + 1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
	dL_dmeans[3 * (size_t)idx + 2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;
}

// SH backward (backward.cu:9-128)
void sh_bw(const Ctx& s, int idx, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs) {
	const float* pos = &s.means3D[3 * (size_t)idx];
	V3 dir_orig = { pos[0] - s.campos[0], pos[1] - s.campos[1], pos[2] - s.campos[2] };
	const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
	const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
	const float* sh = &s.shs[(size_t)idx * s.M * 3];
	float dRGB[3];
	for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[3 * (size_t)idx + ch] * (s.clamped[3 * (size_t)idx + ch] ? 0.f : 1.f);
	float dx[3] = { 0, 0, 0 }, dy[3] = { 0, 0, 0 }, dz[3] = { 0, 0, 0 };
	float* dsh = dL_dshs + (size_t)idx * s.M * 3;
	auto put = [&](int k, float w) { for (int ch = 0; ch < 3; ch++) dsh[k * 3 + ch] = w * dRGB[ch]; };
	auto S = [&](int k, int ch) { return sh[k * 3 + ch]; };
	put(0, SH_C0);
	const int deg = s.D;
	if (deg > 0) {
		put(1, -SH_C1 * y); put(2, SH_C1 * z); put(3, -SH_C1 * x);
		for (int ch = 0; ch < 3; ch++) { dx[ch] = -SH_C1 * S(3, ch); dy[ch] = -SH_C1 * S(1, ch); dz[ch] = SH_C1 * S(2, ch); }
		if (deg > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			put(4, SH_C2[0] * xy); put(5, SH_C2[1] * yz); put(6, SH_C2[2] * (2.f * zz - xx - yy)); put(7, SH_C2[3] * xz); put(8, SH_C2[4] * (xx - yy));
			for (int ch = 0; ch < 3; ch++) {
				dx[ch] += SH_C2[0] * y * S(4, ch) + SH_C2[2] * 2.f * -x * S(6, ch) + SH_C2[3] * z * S(7, ch) + SH_C2[4] * 2.f * x * S(8, ch);
				dy[ch] += SH_C2[0] * x * S(4, ch) + SH_C2[1] * z * S(5, ch) + SH_C2[2] * 2.f * -y * S(6, ch) + SH_C2[4] * 2.f * -y * S(8, ch);
				dz[ch] += SH_C2[1] * y * S(5, ch) + SH_C2[2] * 2.f * 2.f * z * S(6, ch) + SH_C2[3] * x * S(7, ch);
			}
			if (deg > 2) {
				put(9, SH_C3[0] * y * (3.f * xx - yy)); put(10, SH_C3[1] * xy * z); put(11, SH_C3[2] * y * (4.f * zz - xx - yy));
				put(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)); put(13, SH_C3[4] * x * (4.f * zz - xx - yy));
				put(14, SH_C3[5] * z * (xx - yy)); put(15, SH_C3[6] * x * (xx - 3.f * yy));
				for (int ch = 0; ch < 3; ch++) {
					dx[ch] += (SH_C3[0] * S(9, ch) * 3.f * 2.f * xy + SH_C3[1] * S(10, ch) * yz + SH_C3[2] * S(11, ch) * -2.f * xy + SH_C3[3] * S(12, ch) * -3.f * 2.f * xz +
						SH_C3[4] * S(13, ch) * (-3.f * xx + 4.f * zz - yy) + SH_C3[5] * S(14, ch) * 2.f * xz + SH_C3[6] * S(15, ch) * 3.f * (xx - yy));
					dy[ch] += (SH_C3[0] * S(9, ch) * 3.f * (xx - yy) + SH_C3[1] * S(10, ch) * xz + SH_C3[2] * S(11, ch) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * S(12, ch) * -3.f * 2.f * yz +
						SH_C3[4] * S(13, ch) * -2.f * xy + SH_C3[5] * S(14, ch) * -2.f * yz + SH_C3[6] * S(15, ch) * -3.f * 2.f * xy);
					dz[ch] += (SH_C3[1] * S(10, ch) * xy + SH_C3[2] * S(11, ch) * 4.f * 2.f * yz + SH_C3[3] * S(12, ch) * 3.f * (2.f * zz - xx - yy) +
						SH_C3[4] * S(13, ch) * 4.f * 2.f * xz + SH_C3[5] * S(14, ch) * (xx - yy));
				}
			}
		}
	}
	V3 dL_ddir = { dx[0] * dRGB[0] + dx[1] * dRGB[1] + dx[2] * dRGB[2], dy[0] * dRGB[0] + dy[1] * dRGB[1] + dy[2] * dRGB[2], dz[0] * dRGB[0] + dz[1] * dRGB[1] + dz[2] * dRGB[2] };
	V3 dm = dnormvdv3(dir_orig, dL_ddir);
	dL_dmeans[3 * (size_t)idx + 0] += dm.x; dL_dmeans[3 * (size_t)idx + 1] += dm.y; dL_dmeans[3 * (size_t)idx + 2] += dm.z;
}

// computeCov3D backward (backward.cu:268-331)
void cov3d_bw(const Ctx& s, int idx, const float* dL_dcov3Ds, float* dL_dscales, float* dL_drots) {
	const float* q = &s.rotations[4 * (size_t)idx];
	const float r = q[0], x = q[1], y = q[2], z = q[3];
	// glm [col][row]
	float R[3][3] = { { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
		{ 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
		{ 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
	const float sv[3] = { s.scale_modifier * s.scales[3 * (size_t)idx], s.scale_modifier * s.scales[3 * (size_t)idx + 1], s.scale_modifier * s.scales[3 * (size_t)idx + 2] };
	float M[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) M[c][rr] = sv[rr] * R[c][rr];   // S*R with diagonal S
	const float* d = dL_dcov3Ds + 6 * (size_t)idx;
	float dS[3][3] = { { d[0], 0.5f * d[1], 0.5f * d[2] }, { 0.5f * d[1], d[3], 0.5f * d[4] }, { 0.5f * d[2], 0.5f * d[4], d[5] } };
	float dM[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) {   // (2*M) * dS
		dM[c][rr] = 2.0f * M[0][rr] * dS[c][0] + 2.0f * M[1][rr] * dS[c][1] + 2.0f * M[2][rr] * dS[c][2];
	}
	float Rt[3][3], dMt[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) { Rt[c][rr] = R[rr][c]; dMt[c][rr] = dM[rr][c]; }
	float* ds = dL_dscales + 3 * (size_t)idx;
	for (int k = 0; k < 3; k++) ds[k] = Rt[k][0] * dMt[k][0] + Rt[k][1] * dMt[k][1] + Rt[k][2] * dMt[k][2];
	for (int k = 0; k < 3; k++) for (int j = 0; j < 3; j++) dMt[k][j] *= sv[k];
	float* dq = dL_drots + 4 * (size_t)idx;
	dq[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
	dq[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
	dq[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
	dq[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

}  // namespace

extern "C" int orc_backward(orc_ctx* c, const float* dL_dpix,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot)
{
	Ctx& s = *reinterpret_cast<Ctx*>(c);
	const int P = s.P;
	const size_t Pz = (size_t)P;
	memset(dL_dmean2D, 0, sizeof(float) * 3 * Pz); memset(dL_dconic, 0, sizeof(float) * 4 * Pz); memset(dL_dopacity, 0, sizeof(float) * Pz);
	memset(dL_dcolor, 0, sizeof(float) * 3 * Pz); memset(dL_dmean3D, 0, sizeof(float) * 3 * Pz); memset(dL_dcov3D, 0, sizeof(float) * 6 * Pz);
	if (dL_dsh) memset(dL_dsh, 0, sizeof(float) * Pz * s.M * 3);
	if (dL_dscale) memset(dL_dscale, 0, sizeof(float) * 3 * Pz);
	if (dL_drot) memset(dL_drot, 0, sizeof(float) * 4 * Pz);
	const float* colors = s.has_colors ? s.colors_precomp.data() : s.rgb.data();
	std::vector<double> partial((size_t)s.N * 9, 0.0);
	int64_t evals = 0, contribs = 0;
	#pragma omp parallel for collapse(2) schedule(dynamic, 1) reduction(+ : evals, contribs)
	for (int ty = 0; ty < s.gy; ty++) for (int tx = 0; tx < s.gx; tx++) {
		const uint32_t r0 = s.ranges[2 * (size_t)(ty * s.gx + tx)];
		render_tile_bw(s, tx, ty, colors, dL_dpix, partial.data() + (size_t)r0 * 9, evals, contribs);
	}
	s.counters[2] = evals; s.counters[3] = contribs;
	// fixed-order (sorted-list order) accumulation per Gaussian, in double
	std::vector<double> acc(Pz * 9, 0.0);
	for (int64_t i = 0; i < s.N; i++) { const uint32_t id = s.point_list[i]; for (int k = 0; k < 9; k++) acc[(size_t)id * 9 + k] += partial[(size_t)i * 9 + k]; }
	#pragma omp parallel for schedule(static)
	for (int i = 0; i < P; i++) {
		const double* a = &acc[(size_t)i * 9];
		for (int ch = 0; ch < 3; ch++) dL_dcolor[3 * (size_t)i + ch] = (float)a[ch];
		dL_dmean2D[3 * (size_t)i] = (float)a[3]; dL_dmean2D[3 * (size_t)i + 1] = (float)a[4];
		dL_dconic[4 * (size_t)i] = (float)a[5]; dL_dconic[4 * (size_t)i + 1] = (float)a[6]; dL_dconic[4 * (size_t)i + 3] = (float)a[7];
		dL_dopacity[i] = (float)a[8];
	}
	const float focal_y = s.H / (2.0f * s.tan_fovy), focal_x = s.W / (2.0f * s.tan_fovx);
	const float* proj = s.proj.data();
	#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (!(s.radii[idx] > 0)) continue;
		const float* cov3D = s.has_cov ? &s.cov3D_precomp[6 * (size_t)idx] : &s.cov3D[6 * (size_t)idx];
		cov2d_bw(s, idx, focal_x, focal_y, cov3D, dL_dconic, dL_dmean3D, dL_dcov3D);
		// BACKWARD preprocessCUDA (backward.cu:336-386)
		const float* m = &s.means3D[3 * (size_t)idx];
		const float m_hw = proj[3] * m[0] + proj[7] * m[1] + proj[11] * m[2] + proj[15];
		const float m_w = 1.0f / (m_hw + 0.0000001f);
		const float mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
		const float mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
		const float gx2 = dL_dmean2D[3 * (size_t)idx], gy2 = dL_dmean2D[3 * (size_t)idx + 1];
		dL_dmean3D[3 * (size_t)idx + 0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
		dL_dmean3D[3 * (size_t)idx + 1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
		dL_dmean3D[3 * (size_t)idx + 2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;
		if (s.has_sh) sh_bw(s, idx, dL_dcolor, dL_dmean3D, dL_dsh);
		if (s.has_scales) cov3d_bw(s, idx, dL_dcov3D, dL_dscale, dL_drot);
	}
	return 0;
}
