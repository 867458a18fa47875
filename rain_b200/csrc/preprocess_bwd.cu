// Backward preprocess, fused: moments -> conic / screen-mean / opacity gradient, conic -> 2D covariance -> 3D
// covariance / mean gradient, projection gradient, SH gradient, scale / rotation gradient, and the zero fill of
// everything an untouched Gaussian leaves.  Replaces computeCov2DCUDA (reference backward.cu:133-264), preprocessCUDA
// backward (:336-386) with its SH (:9-128) and cov3D (:268-331) device functions, and the nine torch::zeros fills of
// rasterize_points.cu:145-153 (every output element is written exactly once here instead).
//
// One thread per Gaussian.  Gradients only need 1e-4 relative agreement with the reference (its own atomics are
// unordered), so this file uses plain float expressions.
#include "common.cuh"
#include "kernels.cuh"

namespace rr {

namespace {

__constant__ float b_SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
__constant__ float b_SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
constexpr float kC0 = 0.28209479177387814f;
constexpr float kC1 = 0.4886025119029199f;

constexpr int PB_THREADS = 128;
#ifndef PB_MIN_CTAS
#define PB_MIN_CTAS 8      // 64 registers: 8 CTAs (32 warps) per SM instead of 7; measured 0.153 -> 0.143 ms (the kernel waits on memory 58 % of the time)
#endif
constexpr int PB_WARPS = PB_THREADS / 32;
constexpr int SH_FLOATS = 48;            // 16 coefficients x 3 channels (the staged fast path needs M == 16)
constexpr int SH_STRIDE = SH_FLOATS + 1; // +1 float: a lane walking its own row is bank-conflict free

// conic / screen-mean gradient -> 2D covariance -> 3D covariance and mean gradient of one Gaussian that received gradient
// (reference computeCov2DCUDA, backward.cu:133-264, and the projection part of preprocessCUDA, :360-377)
__device__ __forceinline__ void cov2d_mean_grad(const PreprocessBwdArgs& a, int idx, float mx, float my, float mz, const float* g, float* dmean, float* dcov)
{
	const float* __restrict__ V = a.view;
	// ---- conic -> cov2D -> (cov3D, mean) : reference computeCov2DCUDA ----
	const float* __restrict__ c3 = a.cov3D + 6 * (size_t)idx;
	const float S[6] = { c3[0], c3[1], c3[2], c3[3], c3[4], c3[5] };
	float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
	float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
	const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
	const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
	const float txtz = tx / tz, tytz = ty / tz;
	tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
	ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
	const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
	const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
	const float J00 = a.focal_x / tz, J02 = -(a.focal_x * tx) / (tz * tz);
	const float J11 = a.focal_y / tz, J12 = -(a.focal_y * ty) / (tz * tz);
	// u = first row of the 2x3 projection T, w = second row
	float u[3], w[3];
	#pragma unroll
	for (int r = 0; r < 3; r++) { u[r] = V[4 * r] * J00 + V[4 * r + 2] * J02; w[r] = V[4 * r + 1] * J11 + V[4 * r + 2] * J12; }
	// Sigma u, Sigma w
	const float Su[3] = { S[0] * u[0] + S[1] * u[1] + S[2] * u[2], S[1] * u[0] + S[3] * u[1] + S[4] * u[2], S[2] * u[0] + S[4] * u[1] + S[5] * u[2] };
	const float Sw[3] = { S[0] * w[0] + S[1] * w[1] + S[2] * w[2], S[1] * w[0] + S[3] * w[1] + S[4] * w[2], S[2] * w[0] + S[4] * w[1] + S[5] * w[2] };
	const float ca = u[0] * Su[0] + u[1] * Su[1] + u[2] * Su[2] + a.low_pass;
	const float cb = u[0] * Sw[0] + u[1] * Sw[1] + u[2] * Sw[2];
	const float cc = w[0] * Sw[0] + w[1] * Sw[1] + w[2] * Sw[2] + a.low_pass;
	const float denom = ca * cc - cb * cb;
	const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
	float dL_da = 0, dL_db = 0, dL_dc = 0;
	if (denom2inv != 0) {
		const float dcx = g[5], dcy = g[6], dcz = g[7];
		dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
		dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
		dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
		dcov[0] = u[0] * u[0] * dL_da + u[0] * w[0] * dL_db + w[0] * w[0] * dL_dc;
		dcov[3] = u[1] * u[1] * dL_da + u[1] * w[1] * dL_db + w[1] * w[1] * dL_dc;
		dcov[5] = u[2] * u[2] * dL_da + u[2] * w[2] * dL_db + w[2] * w[2] * dL_dc;
		dcov[1] = 2 * u[0] * u[1] * dL_da + (u[0] * w[1] + u[1] * w[0]) * dL_db + 2 * w[0] * w[1] * dL_dc;
		dcov[2] = 2 * u[0] * u[2] * dL_da + (u[0] * w[2] + u[2] * w[0]) * dL_db + 2 * w[0] * w[2] * dL_dc;
		dcov[4] = 2 * u[2] * u[1] * dL_da + (u[1] * w[2] + u[2] * w[1]) * dL_db + 2 * w[1] * w[2] * dL_dc;
	}
	float dJ00 = 0, dJ02 = 0, dJ11 = 0, dJ12 = 0;
	#pragma unroll
	for (int r = 0; r < 3; r++) {
		const float du = 2 * Su[r] * dL_da + Sw[r] * dL_db;
		const float dw = 2 * Sw[r] * dL_dc + Su[r] * dL_db;
		dJ00 += V[4 * r] * du; dJ02 += V[4 * r + 2] * du;
		dJ11 += V[4 * r + 1] * dw; dJ12 += V[4 * r + 2] * dw;
	}
	const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
	const float dtx = x_grad_mul * -a.focal_x * itz2 * dJ02;
	const float dty = y_grad_mul * -a.focal_y * itz2 * dJ12;
	const float dtz = -a.focal_x * itz2 * dJ00 - a.focal_y * itz2 * dJ11 + (2 * a.focal_x * tx) * itz3 * dJ02 + (2 * a.focal_y * ty) * itz3 * dJ12;
	dmean[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
	dmean[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
	dmean[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;

	// ---- screen-space mean -> 3D mean through the projection (reference backward.cu:360-377) ----
	const float* __restrict__ Pj = a.proj;
	const float hw = Pj[3] * mx + Pj[7] * my + Pj[11] * mz + Pj[15];
	const float m_w = 1.0f / (hw + 0.0000001f);
	const float mul1 = (Pj[0] * mx + Pj[4] * my + Pj[8] * mz + Pj[12]) * m_w * m_w;
	const float mul2 = (Pj[1] * mx + Pj[5] * my + Pj[9] * mz + Pj[13]) * m_w * m_w;
	dmean[0] += (Pj[0] * m_w - Pj[3] * mul1) * g[3] + (Pj[1] * m_w - Pj[3] * mul2) * g[4];
	dmean[1] += (Pj[4] * m_w - Pj[7] * mul1) * g[3] + (Pj[5] * m_w - Pj[7] * mul2) * g[4];
	dmean[2] += (Pj[8] * m_w - Pj[11] * mul1) * g[3] + (Pj[9] * m_w - Pj[11] * mul2) * g[4];
}

// SH colour gradient of one Gaussian (reference computeColorFromSH backward, backward.cu:9-128): dL_dsh row through
// store(ptr, r, g, b), and the view-direction part of the mean gradient added to dmean.  sh and dsh may alias (a staged
// row is overwritten in place): every dot() of a band is taken before its put()s.
template <class Store>
__device__ __forceinline__ void sh_grad(const PreprocessBwdArgs& a, int idx, float mx, float my, float mz, const float* g, const float* __restrict__ sh,
	float* __restrict__ dsh, float* dmean, Store store)
{
	const uint8_t cl = a.clamped[idx];
	const float dR[3] = { (cl & 1) ? 0.f : g[0], (cl & 2) ? 0.f : g[1], (cl & 4) ? 0.f : g[2] };
	const float ox = mx - a.campos[0], oy = my - a.campos[1], oz = mz - a.campos[2];
	const float inv_len = 1.0f / sqrtf(ox * ox + oy * oy + oz * oz);
	const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
	float ddx = 0, ddy = 0, ddz = 0;   // dL/d(dir)
	auto dot = [&](int k) { return sh[3 * k] * dR[0] + sh[3 * k + 1] * dR[1] + sh[3 * k + 2] * dR[2]; };
	auto put = [&](int k, float wgt) { store(dsh + 3 * k, wgt * dR[0], wgt * dR[1], wgt * dR[2]); };
	put(0, kC0);
	if (a.D > 0) {
		const float s1 = dot(1), s2 = dot(2), s3 = dot(3);
		put(1, -kC1 * y); put(2, kC1 * z); put(3, -kC1 * x);
		ddx = -kC1 * s3; ddy = -kC1 * s1; ddz = kC1 * s2;
		if (a.D > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			const float s4 = dot(4), s5 = dot(5), s6 = dot(6), s7 = dot(7), s8 = dot(8);
			put(4, b_SH_C2[0] * xy); put(5, b_SH_C2[1] * yz); put(6, b_SH_C2[2] * (2.f * zz - xx - yy)); put(7, b_SH_C2[3] * xz); put(8, b_SH_C2[4] * (xx - yy));
			ddx += b_SH_C2[0] * y * s4 + b_SH_C2[2] * 2.f * -x * s6 + b_SH_C2[3] * z * s7 + b_SH_C2[4] * 2.f * x * s8;
			ddy += b_SH_C2[0] * x * s4 + b_SH_C2[1] * z * s5 + b_SH_C2[2] * 2.f * -y * s6 + b_SH_C2[4] * 2.f * -y * s8;
			ddz += b_SH_C2[1] * y * s5 + b_SH_C2[2] * 2.f * 2.f * z * s6 + b_SH_C2[3] * x * s7;
			if (a.D > 2) {
				const float s9 = dot(9), s10 = dot(10), s11 = dot(11), s12 = dot(12), s13 = dot(13), s14 = dot(14), s15 = dot(15);
				put(9, b_SH_C3[0] * y * (3.f * xx - yy)); put(10, b_SH_C3[1] * xy * z); put(11, b_SH_C3[2] * y * (4.f * zz - xx - yy));
				put(12, b_SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)); put(13, b_SH_C3[4] * x * (4.f * zz - xx - yy));
				put(14, b_SH_C3[5] * z * (xx - yy)); put(15, b_SH_C3[6] * x * (xx - 3.f * yy));
				ddx += b_SH_C3[0] * s9 * 3.f * 2.f * xy + b_SH_C3[1] * s10 * yz + b_SH_C3[2] * s11 * -2.f * xy + b_SH_C3[3] * s12 * -3.f * 2.f * xz +
					b_SH_C3[4] * s13 * (-3.f * xx + 4.f * zz - yy) + b_SH_C3[5] * s14 * 2.f * xz + b_SH_C3[6] * s15 * 3.f * (xx - yy);
				ddy += b_SH_C3[0] * s9 * 3.f * (xx - yy) + b_SH_C3[1] * s10 * xz + b_SH_C3[2] * s11 * (-3.f * yy + 4.f * zz - xx) + b_SH_C3[3] * s12 * -3.f * 2.f * yz +
					b_SH_C3[4] * s13 * -2.f * xy + b_SH_C3[5] * s14 * -2.f * yz + b_SH_C3[6] * s15 * -3.f * 2.f * xy;
				ddz += b_SH_C3[1] * s10 * xy + b_SH_C3[2] * s11 * 4.f * 2.f * yz + b_SH_C3[3] * s12 * 3.f * (2.f * zz - xx - yy) +
					b_SH_C3[4] * s13 * 4.f * 2.f * xz + b_SH_C3[5] * s14 * (xx - yy);
			}
		}
	}
	// through the normalisation of the view direction (dnormvdv, auxiliary.h:96-106)
	const float sum2 = ox * ox + oy * oy + oz * oz;
	const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
	dmean[0] += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
	dmean[1] += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
	dmean[2] += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
}

// 3D covariance gradient -> scale and rotation gradient (reference computeCov3D backward, backward.cu:268-331)
__device__ __forceinline__ void scale_rot_grad(const PreprocessBwdArgs& a, int idx, size_t i3, const float* dcov, float* ds, float4& dq)
{
	const float4 q = load_f4(a.rotations, (size_t)idx);
	const float r = q.x, x = q.y, y = q.z, z = q.w;
	// R in glm [col][row] indexing
	const float R[3][3] = { { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
		{ 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
		{ 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
	const float sv[3] = { a.scale_modifier * a.scales[i3], a.scale_modifier * a.scales[i3 + 1], a.scale_modifier * a.scales[i3 + 2] };
	const float dS[3][3] = { { dcov[0], 0.5f * dcov[1], 0.5f * dcov[2] }, { 0.5f * dcov[1], dcov[3], 0.5f * dcov[4] }, { 0.5f * dcov[2], 0.5f * dcov[4], dcov[5] } };
	// dMt[k][j] = dL_dM[j][k] with M[c][r] = sv[r] * R[c][r], dL_dM = 2 M dSigma
	float dMt[3][3];
	#pragma unroll
	for (int c = 0; c < 3; c++)
		#pragma unroll
		for (int rr = 0; rr < 3; rr++)
			dMt[rr][c] = 2.0f * sv[rr] * (R[0][rr] * dS[c][0] + R[1][rr] * dS[c][1] + R[2][rr] * dS[c][2]);
	#pragma unroll
	for (int k = 0; k < 3; k++) ds[k] = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
	#pragma unroll
	for (int k = 0; k < 3; k++)
		#pragma unroll
		for (int j = 0; j < 3; j++) dMt[k][j] *= sv[k];
	dq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
	dq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
	dq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
	dq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

// One warp handles 32 consecutive Gaussians.  The render gradients arrive as one 48-byte accumulator record per
// Gaussian (kernels.cuh: colour sums and the moments of u = G * dL_dalpha), read as three float4; the conic / mean /
// opacity gradients of reference backward.cu:519-544 are those moments times per-Gaussian constants.  A Gaussian
// whose record is all zero received no gradient from any pixel: every output of it is zero (the reference computes
// the same zeros the long way), so it is treated like a culled one — no SH row is read for it, and in accumulate mode
// nothing is written.  The 32 SH rows (6 KB contiguous) are read as float4, each lane works on its row in shared
// memory, and the 32 dL_dsh rows go back as float4 (STAGED; the plain per-thread path remains for M != 16 or
// unaligned pointers).
template <bool STAGED>
__global__ void __launch_bounds__(PB_THREADS, PB_MIN_CTAS) preprocess_bwd_kernel(PreprocessBwdArgs a)
{
	__shared__ __align__(16) float s_buf[PB_WARPS][32 * SH_STRIDE];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int warp_base = blockIdx.x * PB_THREADS + warp * 32;
	if (warp_base >= a.P) return;                               // warp-uniform
	float* __restrict__ sb = s_buf[warp];
	const int idx = warp_base + lane;
	const bool valid = idx < a.P;
	const int nvalid = min(32, a.P - warp_base);
	const size_t i3 = 3 * (size_t)(valid ? idx : 0);
	// accumulate mode (multi-view batches, see rr_backward flags): parameter gradients are added to what the buffers
	// hold and a Gaussian this view does not touch is not touched at all; otherwise every element is overwritten.
	const bool acc = a.accumulate != 0;
	auto put1 = [acc](float* p, float v) { *p = acc ? *p + v : v; };

	// ---- the render gradients of this Gaussian ----
	float g[9];
	#pragma unroll
	for (int k = 0; k < 9; k++) g[k] = 0.0f;
	bool visible = valid && a.radii[idx] > 0;
	if (visible) {
		const float4* __restrict__ ap = reinterpret_cast<const float4*>(a.grad_acc) + (size_t)idx * (ACC_FLOATS / 4);
		const float4 m0 = __ldg(ap), m1 = __ldg(ap + 1), m2 = __ldg(ap + 2);
		visible = (m0.x != 0.0f) | (m0.y != 0.0f) | (m0.z != 0.0f) | (m0.w != 0.0f) | (m1.x != 0.0f) | (m1.y != 0.0f) | (m1.z != 0.0f) | (m1.w != 0.0f) | (m2.x != 0.0f);
		if (visible) {
			const float4* __restrict__ rp = reinterpret_cast<const float4*>(a.rec) + (size_t)idx * 3;
			const float4 ra = __ldg(rp), rb = __ldg(rp + 1);         // ra = { pix.xy, conic.x, conic.y }, rb = { conic.z, opacity, .. }
			const float A = ra.z, B = ra.w, C = rb.x, o = rb.y;
			const float Sx = m0.w, Sy = m1.x, Sxx = m1.y, Sxy = m1.z, Syy = m1.w, S0 = m2.x;
			g[0] = m0.x; g[1] = m0.y; g[2] = m0.z;
			g[3] = -o * (0.5f * a.W) * (A * Sx + B * Sy);           // dL_dmean2D.x (backward.cu:519-527, 535)
			g[4] = -o * (0.5f * a.H) * (C * Sy + B * Sx);           // dL_dmean2D.y
			g[5] = -0.5f * o * Sxx; g[6] = -0.5f * o * Sxy; g[7] = -0.5f * o * Syy;   // dL_dconic .x .y .w (backward.cu:539-541)
			g[8] = S0;                                               // dL_dopacity (backward.cu:544)
		}
	}
	const unsigned vismask = __ballot_sync(FULL, visible);
	if (valid) {
		a.dL_dmean2D[i3] = g[3]; a.dL_dmean2D[i3 + 1] = g[4]; a.dL_dmean2D[i3 + 2] = 0.0f;
		if (a.dL_dconic) reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(g[5], g[6], 0.0f, g[7]);
	}
	if (acc && vismask == 0) return;                            // warp-uniform: nothing to add anywhere
	const bool wr = valid && !(acc && !visible);                // does this lane write its per-Gaussian parameter gradients?
	if (wr) {
		put1(a.dL_dopacity + idx, g[8]);
		if (a.dL_dcolor) { put1(a.dL_dcolor + i3, g[0]); put1(a.dL_dcolor + i3 + 1, g[1]); put1(a.dL_dcolor + i3 + 2, g[2]); }
	}

	// ---- stage the warp's 32 SH rows (coalesced float4 reads) ----
	const bool use_sh = a.shs != nullptr;
	if (STAGED && use_sh && vismask != 0) {
		const float4* __restrict__ src = reinterpret_cast<const float4*>(a.shs) + (size_t)warp_base * (SH_FLOATS / 4);
		const int n4 = nvalid * (SH_FLOATS / 4);
		#pragma unroll
		for (int it = 0; it < SH_FLOATS / 4; it++) {
			const int f4 = it * 32 + lane;
			if (f4 < n4) {
				const int row = f4 / (SH_FLOATS / 4), col = (f4 - row * (SH_FLOATS / 4)) * 4;
				if ((vismask >> row) & 1u) {
					const float4 v = __ldg(src + f4);
					float* d = sb + row * SH_STRIDE + col;
					d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
				}
			}
		}
	}
	__syncwarp();

	float dmean[3] = { 0, 0, 0 }, dcov[6] = { 0, 0, 0, 0, 0, 0 };
	const float mx = a.means3D[i3], my = a.means3D[i3 + 1], mz = a.means3D[i3 + 2];
	const float* __restrict__ V = a.view;

	if (visible) {
		cov2d_mean_grad(a, idx, mx, my, mz, g, dmean, dcov);
	}
	#pragma unroll
	for (int k = 0; k < 6; k++) if (wr && a.dL_dcov3D) put1(a.dL_dcov3D + 6 * (size_t)idx + k, dcov[k]);

	// ---- SH colour gradient (reference backward.cu:9-128) ----
	if (use_sh) {
		// row pointers: shared-memory row (STAGED) or the Gaussian's own global rows
		const float* __restrict__ sh = STAGED ? sb + lane * SH_STRIDE : a.shs + (size_t)(valid ? idx : 0) * a.M * 3;
		float* __restrict__ dsh = STAGED ? sb + lane * SH_STRIDE : a.dL_dsh + (size_t)(valid ? idx : 0) * a.M * 3;
		const int nact = visible ? (a.D + 1) * (a.D + 1) : 0;
		if (visible) {
			sh_grad(a, idx, mx, my, mz, g, sh, dsh, dmean, [&](float* p, float r0, float r1, float r2) {
				if (STAGED) { p[0] = r0; p[1] = r1; p[2] = r2; }
				else { put1(p, r0); put1(p + 1, r1); put1(p + 2, r2); }
			});
		}
		if (STAGED) {
			// coefficients above the active degree (and whole rows of culled Gaussians) are zero; then the 32 rows go
			// out as coalesced float4 (added to the buffer in accumulate mode, rows of culled Gaussians skipped there)
			for (int k = nact; k < 16; k++) { dsh[3 * k] = 0.0f; dsh[3 * k + 1] = 0.0f; dsh[3 * k + 2] = 0.0f; }
			__syncwarp();
			float4* __restrict__ dst = reinterpret_cast<float4*>(a.dL_dsh) + (size_t)warp_base * (SH_FLOATS / 4);
			const int n4 = nvalid * (SH_FLOATS / 4);
			#pragma unroll
			for (int it = 0; it < SH_FLOATS / 4; it++) {
				const int f4 = it * 32 + lane;
				if (f4 < n4) {
					const int row = f4 / (SH_FLOATS / 4), col = (f4 - row * (SH_FLOATS / 4)) * 4;
					const float* srow = sb + row * SH_STRIDE + col;
					float4 v = make_float4(srow[0], srow[1], srow[2], srow[3]);
					if (acc) {
						if (!((vismask >> row) & 1u)) continue;
						const float4 o = dst[f4];
						v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
					}
					dst[f4] = v;
				}
			}
		} else if (!acc && valid) {
			for (int k = nact; k < a.M; k++) { dsh[3 * k] = 0.0f; dsh[3 * k + 1] = 0.0f; dsh[3 * k + 2] = 0.0f; }
		}
	}
	if (!wr) return;
	put1(a.dL_dmean3D + i3, dmean[0]); put1(a.dL_dmean3D + i3 + 1, dmean[1]); put1(a.dL_dmean3D + i3 + 2, dmean[2]);

	// ---- 3D covariance -> scale, rotation (reference backward.cu:268-331) ----
	float ds[3] = { 0, 0, 0 };
	float4 dq = make_float4(0, 0, 0, 0);
	if (visible && a.scales) {
		scale_rot_grad(a, idx, i3, dcov, ds, dq);
	}
	if (a.dL_dscale) { put1(a.dL_dscale + i3, ds[0]); put1(a.dL_dscale + i3 + 1, ds[1]); put1(a.dL_dscale + i3 + 2, ds[2]); }
	if (a.dL_drot) {
		// the rotations segment of a flat gradient bucket starts at float offset P * (7 + 3M): 16-byte aligned only when
		// P % 4 == 0, so the float4 path is taken only for an aligned base (uniform branch)
		if ((reinterpret_cast<uintptr_t>(a.dL_drot) & 15) == 0) {
			float4* o = reinterpret_cast<float4*>(a.dL_drot) + idx;
			if (acc) { const float4 c = *o; dq.x += c.x; dq.y += c.y; dq.z += c.z; dq.w += c.w; }
			*o = dq;
		} else {
			float* o = a.dL_drot + 4 * (size_t)idx;
			put1(o, dq.x); put1(o + 1, dq.y); put1(o + 2, dq.z); put1(o + 3, dq.w);
		}
	}
}


}  // namespace

int launch_preprocess_bwd(cudaStream_t stream, const PreprocessBwdArgs& a)
{
	if (a.P == 0) return 0;
	const bool staged = a.shs && a.M == 16 && (reinterpret_cast<uintptr_t>(a.shs) & 15) == 0 && (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0;
	const unsigned grid = (unsigned)((a.P + PB_THREADS - 1) / PB_THREADS);
	if (staged) count_launch(), preprocess_bwd_kernel<true><<<grid, PB_THREADS, 0, stream>>>(a);
	else count_launch(), preprocess_bwd_kernel<false><<<grid, PB_THREADS, 0, stream>>>(a);
	return check_launch("preprocess_bwd", stream, false);
}

}  // namespace rr
