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
#include "splat_math.cuh"

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
// camera constants of one view as the per-Gaussian gradient functions need them; V / Pj may point to global or shared memory
struct ViewCam { const float* V; const float* Pj; float cam[3]; float tan_fovx, tan_fovy, focal_x, focal_y, low_pass; };

__device__ __forceinline__ void cov2d_mean_grad(const ViewCam& a, const float* S /*cov3D[6]*/, float mx, float my, float mz, const float* g, float* dmean, float* dcov)
{
	const float* V = a.V;
	// ---- conic -> cov2D -> (cov3D, mean) : reference computeCov2DCUDA ----
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
	const float* Pj = a.Pj;
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
__device__ __forceinline__ void sh_grad(int D, uint8_t cl /*clamped bits*/, const float* campos, float mx, float my, float mz, const float* g, const float* sh,
	float* dsh, float* dmean, Store store)
{
	const float dR[3] = { (cl & 1) ? 0.f : g[0], (cl & 2) ? 0.f : g[1], (cl & 4) ? 0.f : g[2] };
	const float ox = mx - campos[0], oy = my - campos[1], oz = mz - campos[2];
	const float inv_len = 1.0f / sqrtf(ox * ox + oy * oy + oz * oz);
	const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
	float ddx = 0, ddy = 0, ddz = 0;   // dL/d(dir)
	auto dot = [&](int k) { return sh[3 * k] * dR[0] + sh[3 * k + 1] * dR[1] + sh[3 * k + 2] * dR[2]; };
	auto put = [&](int k, float wgt) { store(dsh + 3 * k, wgt * dR[0], wgt * dR[1], wgt * dR[2]); };
	put(0, kC0);
	if (D > 0) {
		const float s1 = dot(1), s2 = dot(2), s3 = dot(3);
		put(1, -kC1 * y); put(2, kC1 * z); put(3, -kC1 * x);
		ddx = -kC1 * s3; ddy = -kC1 * s1; ddz = kC1 * s2;
		if (D > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			const float s4 = dot(4), s5 = dot(5), s6 = dot(6), s7 = dot(7), s8 = dot(8);
			put(4, b_SH_C2[0] * xy); put(5, b_SH_C2[1] * yz); put(6, b_SH_C2[2] * (2.f * zz - xx - yy)); put(7, b_SH_C2[3] * xz); put(8, b_SH_C2[4] * (xx - yy));
			ddx += b_SH_C2[0] * y * s4 + b_SH_C2[2] * 2.f * -x * s6 + b_SH_C2[3] * z * s7 + b_SH_C2[4] * 2.f * x * s8;
			ddy += b_SH_C2[0] * x * s4 + b_SH_C2[1] * z * s5 + b_SH_C2[2] * 2.f * -y * s6 + b_SH_C2[4] * 2.f * -y * s8;
			ddz += b_SH_C2[1] * y * s5 + b_SH_C2[2] * 2.f * 2.f * z * s6 + b_SH_C2[3] * x * s7;
			if (D > 2) {
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
__device__ __forceinline__ void scale_rot_grad(float4 q, float s0, float s1, float s2, float scale_modifier, const float* dcov, float* ds, float4& dq)
{
	const float r = q.x, x = q.y, y = q.z, z = q.w;
	// R in glm [col][row] indexing
	const float R[3][3] = { { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
		{ 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
		{ 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
	const float sv[3] = { scale_modifier * s0, scale_modifier * s1, scale_modifier * s2 };
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
	ViewCam vc;
	vc.V = a.view; vc.Pj = a.proj; vc.cam[0] = a.campos[0]; vc.cam[1] = a.campos[1]; vc.cam[2] = a.campos[2];
	vc.tan_fovx = a.tan_fovx; vc.tan_fovy = a.tan_fovy; vc.focal_x = a.focal_x; vc.focal_y = a.focal_y; vc.low_pass = a.low_pass;
	if (visible) {
		const float* __restrict__ c3 = a.cov3D + 6 * (size_t)idx;
		const float S[6] = { c3[0], c3[1], c3[2], c3[3], c3[4], c3[5] };
		cov2d_mean_grad(vc, S, mx, my, mz, g, dmean, dcov);
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
			sh_grad(a.D, a.clamped[idx], vc.cam, mx, my, mz, g, sh, dsh, dmean, [&](float* p, float r0, float r1, float r2) {
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
		scale_rot_grad(load_f4(a.rotations, (size_t)idx), a.scales[i3], a.scales[i3 + 1], a.scales[i3 + 2], a.scale_modifier, dcov, ds, dq);
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


// ---- view-batched backward ------------------------------------------------------------------------------------------------
// The per-Gaussian half of the backward for ALL views of a step in one pass (rr_backward_views; no reference counterpart: the
// reference runs computeCov2DCUDA + preprocessCUDA once per view, backward.cu:133-386).  A multi-view step otherwise runs
// preprocess_bwd once per view, each time re-reading the mean / scale / rotation / SH row of every touched Gaussian and
// read-modify-writing its 59 gradient floats.  Here one thread owns one Gaussian: it reads its parameters ONCE, walks the
// views of the batch (phase A: radii and accumulator records of all views, loads in flight four views at a time, giving the
// set of views that touched it; phase B: only those views — projected record, clamp flags, the view's camera from shared
// memory), sums the 3D-covariance gradient over the views (Sigma -> (scale, rotation) is linear in it, so that step runs once),
// accumulates dL_dsh in a shared-memory row, and writes its bucket row ONCE.  Per-view outputs (dL_dmean2D = viewspace
// gradients) are still written per view, and the densification statistics of the batch (train.py:132-134) can be folded in.
//   LEAF = false: inputs are the activated tensors, gradients w.r.t. them, bucket layout  means3D | sh[P,16,3] | opacity | scales | rotations
//   LEAF = true : inputs are the six RAW leaf tensors of scene/gaussian_model.py:131-136 (activations of :85-105 applied
//                 here: splat_math.cuh), gradients w.r.t. the leaves (chain rule applied here), bucket layout
//                 xyz | f_dc[P,1,3] | f_rest[P,15,3] | opacity | scaling | rotation
#ifndef BV_PREFETCH
#define BV_PREFETCH 1
#endif
constexpr int BV_THREADS = 128;
constexpr int BV_WARPS = BV_THREADS / 32;
constexpr int CAM_FLOATS = 40;           // V[16] Pj[16] campos[3] tan_fovx tan_fovy focal_x focal_y pad
constexpr size_t BV_SMEM_BYTES = sizeof(float) * (2 * BV_WARPS * 32 * SH_STRIDE + BV_MAX_VIEWS * CAM_FLOATS);

template <bool LEAF>
__global__ void __launch_bounds__(BV_THREADS, 4) backward_views_kernel(BackwardViewsArgs a)
{
	// dynamic shared memory (BV_SMEM_BYTES > the 48 KB static limit): per warp 32 SH rows (coefficient-major, 3 channels each)
	// and their gradient rows summed over the views, then the cameras of the batch
	extern __shared__ __align__(16) float bv_smem[];
	float (*s_sh)[32 * SH_STRIDE] = reinterpret_cast<float (*)[32 * SH_STRIDE]>(bv_smem);
	float (*s_dsh)[32 * SH_STRIDE] = reinterpret_cast<float (*)[32 * SH_STRIDE]>(bv_smem + BV_WARPS * 32 * SH_STRIDE);
	float (*s_cam)[CAM_FLOATS] = reinterpret_cast<float (*)[CAM_FLOATS]>(bv_smem + 2 * BV_WARPS * 32 * SH_STRIDE);
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int nv = a.nviews;
	for (int i = tid; i < nv * 35; i += BV_THREADS) {
		const int v = i / 35, k = i - v * 35;
		s_cam[v][k] = k < 16 ? a.view[v][k] : k < 32 ? a.proj[v][k - 16] : a.campos[v][k - 32];
	}
	if (tid < nv) {
		s_cam[tid][35] = a.tan_fovx[tid]; s_cam[tid][36] = a.tan_fovy[tid];
		s_cam[tid][37] = a.W / (2.0f * a.tan_fovx[tid]); s_cam[tid][38] = a.H / (2.0f * a.tan_fovy[tid]);       // rasterizer_impl.cu:213-214
	}
	__syncthreads();
	// this launch covers the Gaussians [first, last) (the whole set, or one chunk of it when the caller overlaps the all-reduce of
	// finished chunks with the backward of the next one)
	const int last = a.first + a.count;
	const int warp_base = a.first + blockIdx.x * BV_THREADS + warp * 32;
	if (warp_base >= last) return;                              // warp-uniform
	const int idx = warp_base + lane;
	const bool valid = idx < last;
	const int nvalid = min(32, last - warp_base);
	const size_t ii = (size_t)(valid ? idx : a.first), i3 = 3 * ii;
	float* __restrict__ sh = s_sh[warp] + lane * SH_STRIDE;
	float* __restrict__ dsh = s_dsh[warp] + lane * SH_STRIDE;

	// ---- stage the warp's 32 SH rows (coalesced) and clear the gradient rows ----
	if (LEAF) {
		// f_dc [P,1,3] and f_rest [P,15,3]: row r of the warp = 3 floats of f_dc then 45 of f_rest
		const float* __restrict__ dc = a.f_dc + (size_t)warp_base * 3;
		const float* __restrict__ rest = a.f_rest + (size_t)warp_base * 45;
		for (int f = lane; f < nvalid * 3; f += 32) { const int row = f / 3; s_sh[warp][row * SH_STRIDE + (f - row * 3)] = __ldg(dc + f); }
		for (int f = lane; f < nvalid * 45; f += 32) { const int row = f / 45; s_sh[warp][row * SH_STRIDE + 3 + (f - row * 45)] = __ldg(rest + f); }
	} else {
		const float* __restrict__ src = a.shs + (size_t)warp_base * SH_FLOATS;
		if ((reinterpret_cast<uintptr_t>(a.shs) & 15) == 0) {
			const float4* __restrict__ src4 = reinterpret_cast<const float4*>(src);
			for (int f4 = lane; f4 < nvalid * (SH_FLOATS / 4); f4 += 32) {
				const int row = f4 / (SH_FLOATS / 4), col = (f4 - row * (SH_FLOATS / 4)) * 4;
				const float4 v = __ldg(src4 + f4);
				float* d = s_sh[warp] + row * SH_STRIDE + col;
				d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
			}
		} else {
			for (int f = lane; f < nvalid * SH_FLOATS; f += 32) { const int row = f / SH_FLOATS; s_sh[warp][row * SH_STRIDE + (f - row * SH_FLOATS)] = __ldg(src + f); }
		}
	}
	#pragma unroll
	for (int k = 0; k < SH_FLOATS; k++) dsh[k] = 0.0f;
	__syncwarp();

	// ---- phase A: which views touched this Gaussian (radii > 0 and a non-zero accumulator record) ----
	unsigned touched = 0;
	int nvis = 0, maxr = 0;
	for (int v0 = 0; v0 < nv; v0 += 4) {
		int r[4];
		#pragma unroll
		for (int j = 0; j < 4; j++) r[j] = (valid && v0 + j < nv) ? __ldg(a.radii[v0 + j] + ii) : 0;
		float4 m[4][3];
		#pragma unroll
		for (int j = 0; j < 4; j++) {
			if (r[j] > 0) {
				const float4* __restrict__ ap = reinterpret_cast<const float4*>(a.grad_acc[v0 + j]) + ii * (ACC_FLOATS / 4);
				m[j][0] = __ldg(ap); m[j][1] = __ldg(ap + 1); m[j][2] = __ldg(ap + 2);
			} else { m[j][0] = m[j][1] = m[j][2] = make_float4(0.f, 0.f, 0.f, 0.f); }
		}
		#pragma unroll
		for (int j = 0; j < 4; j++) {
			const bool nz = (m[j][0].x != 0.0f) | (m[j][0].y != 0.0f) | (m[j][0].z != 0.0f) | (m[j][0].w != 0.0f) | (m[j][1].x != 0.0f) | (m[j][1].y != 0.0f) |
				(m[j][1].z != 0.0f) | (m[j][1].w != 0.0f) | (m[j][2].x != 0.0f);
			if (r[j] > 0) { nvis++; maxr = max(maxr, r[j]); }
			if (nz) {
				touched |= 1u << (v0 + j);
#if BV_PREFETCH
				// phase B reads this view's projected record and clamp flags one view at a time, with its arithmetic in between:
				// start those lines on their way now
				const float4* rp = reinterpret_cast<const float4*>(a.rec[v0 + j]) + ii * 3;
				asm volatile("prefetch.global.L2 [%0];" ::"l"(rp));
				asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + 1));
				asm volatile("prefetch.global.L2 [%0];" ::"l"(a.clamped[v0 + j] + ii));
#endif
			}
			else if (valid && v0 + j < nv && a.dL_dmean2D[v0 + j]) {           // a view that left no gradient: its viewspace gradient row is zero
				float* o = a.dL_dmean2D[v0 + j] + i3;
				o[0] = 0.0f; o[1] = 0.0f; o[2] = 0.0f;
			}
		}
	}
	const unsigned anymask = __ballot_sync(FULL, touched != 0);
	const bool acc = a.accumulate != 0;
	if (acc && anymask == 0 && !a.stat_denom) return;          // warp-uniform: nothing to add anywhere

	// ---- phase B: the views that touched it ----
	float dmean[3] = { 0, 0, 0 }, dcov[6] = { 0, 0, 0, 0, 0, 0 }, dop = 0.0f, gnorm = 0.0f;
	const float mx = a.means3D[i3], my = a.means3D[i3 + 1], mz = a.means3D[i3 + 2];
	float S[6] = { 0, 0, 0, 0, 0, 0 };
	float4 q = make_float4(1.f, 0.f, 0.f, 0.f), qraw = q;
	float sc0 = 0, sc1 = 0, sc2 = 0, qinv = 1.0f;
	if (touched) {
		qraw = load_f4(a.rotations, ii);
		sc0 = a.scales[i3]; sc1 = a.scales[i3 + 1]; sc2 = a.scales[i3 + 2];
		if (LEAF) { qinv = rot_inv_norm(qraw); q = act_rot(qraw); sc0 = act_scale(sc0); sc1 = act_scale(sc1); sc2 = act_scale(sc2); }
		else q = qraw;
		cov3d_from_scale_rot(sc0, sc1, sc2, a.scale_modifier, q, S);         // the forward's Sigma, bit for bit (splat_math.cuh)
	}
	for (unsigned rest = touched; rest; rest &= rest - 1) {
		const int v = __ffs(rest) - 1;
		const float4* __restrict__ ap = reinterpret_cast<const float4*>(a.grad_acc[v]) + ii * (ACC_FLOATS / 4);
		const float4 m0 = __ldg(ap), m1 = __ldg(ap + 1), m2 = __ldg(ap + 2);
		const float4* __restrict__ rp = reinterpret_cast<const float4*>(a.rec[v]) + ii * 3;
		const float4 ra = __ldg(rp), rb = __ldg(rp + 1);             // ra = { pix.xy, conic.x, conic.y }, rb = { conic.z, opacity, .. }
		const float A = ra.z, B = ra.w, C = rb.x, o = rb.y;
		const float Sx = m0.w, Sy = m1.x, Sxx = m1.y, Sxy = m1.z, Syy = m1.w, S0 = m2.x;
		float g[9];
		g[0] = m0.x; g[1] = m0.y; g[2] = m0.z;
		g[3] = -o * (0.5f * a.W) * (A * Sx + B * Sy);               // dL_dmean2D.x (backward.cu:519-527, 535)
		g[4] = -o * (0.5f * a.H) * (C * Sy + B * Sx);               // dL_dmean2D.y
		g[5] = -0.5f * o * Sxx; g[6] = -0.5f * o * Sxy; g[7] = -0.5f * o * Syy;   // dL_dconic .x .y .w (backward.cu:539-541)
		g[8] = S0;                                                   // dL_dopacity (backward.cu:544)
		if (a.dL_dmean2D[v]) { float* out = a.dL_dmean2D[v] + i3; out[0] = g[3]; out[1] = g[4]; out[2] = 0.0f; }
		gnorm += sqrtf(g[3] * g[3] + g[4] * g[4]);
		dop += LEAF ? g[8] * o * (1.0f - o) : g[8];                 // d sigmoid: the forward's own opacity value
		const float* cam = s_cam[v];
		ViewCam vc;
		vc.V = cam; vc.Pj = cam + 16; vc.cam[0] = cam[32]; vc.cam[1] = cam[33]; vc.cam[2] = cam[34];
		vc.tan_fovx = cam[35]; vc.tan_fovy = cam[36]; vc.focal_x = cam[37]; vc.focal_y = cam[38]; vc.low_pass = a.low_pass;
		float dm[3], dc[6] = { 0, 0, 0, 0, 0, 0 };
		cov2d_mean_grad(vc, S, mx, my, mz, g, dm, dc);
		#pragma unroll
		for (int k = 0; k < 6; k++) dcov[k] += dc[k];
		sh_grad(a.D, __ldg(a.clamped[v] + ii), vc.cam, mx, my, mz, g, sh, dsh, dm, [](float* p, float r0, float r1, float r2) { p[0] += r0; p[1] += r1; p[2] += r2; });
		dmean[0] += dm[0]; dmean[1] += dm[1]; dmean[2] += dm[2];
	}

	// ---- densification statistics of the batch (scene/gaussian_model.py:419-421 over every view that saw the Gaussian) ----
	if (a.stat_denom && valid && nvis > 0) {
		a.stat_max_radii[ii] = fmaxf(a.stat_max_radii[ii], (float)maxr);
		a.stat_grad_accum[ii] += gnorm;
		a.stat_denom[ii] += (float)nvis;
	}
	if (acc && anymask == 0) return;

	// ---- the bucket row: written once (added in accumulate mode, where rows nothing touched are left alone) ----
	__syncwarp();
	if (LEAF) {
		float* __restrict__ o_dc = a.dL_dsh + (size_t)warp_base * 3;          // f_dc segment
		float* __restrict__ o_rest = a.dL_dfrest + (size_t)warp_base * 45;
		for (int f = lane; f < nvalid * 3; f += 32) {
			const int row = f / 3;
			const float v = s_dsh[warp][row * SH_STRIDE + (f - row * 3)];
			if (!acc) o_dc[f] = v; else if ((anymask >> row) & 1u) o_dc[f] += v;
		}
		for (int f = lane; f < nvalid * 45; f += 32) {
			const int row = f / 45;
			const float v = s_dsh[warp][row * SH_STRIDE + 3 + (f - row * 45)];
			if (!acc) o_rest[f] = v; else if ((anymask >> row) & 1u) o_rest[f] += v;
		}
	} else {
		float* __restrict__ dst = a.dL_dsh + (size_t)warp_base * SH_FLOATS;
		for (int f = lane; f < nvalid * SH_FLOATS; f += 32) {
			const int row = f / SH_FLOATS;
			const float v = s_dsh[warp][row * SH_STRIDE + (f - row * SH_FLOATS)];
			if (!acc) dst[f] = v; else if ((anymask >> row) & 1u) dst[f] += v;
		}
	}
	if (!valid || (acc && !touched)) return;
	auto put1 = [acc](float* p, float v) { *p = acc ? *p + v : v; };
	put1(a.dL_dmean3D + i3, dmean[0]); put1(a.dL_dmean3D + i3 + 1, dmean[1]); put1(a.dL_dmean3D + i3 + 2, dmean[2]);
	put1(a.dL_dopacity + ii, dop);
	float ds[3] = { 0, 0, 0 };
	float4 dq = make_float4(0, 0, 0, 0);
	if (touched) {
		scale_rot_grad(q, sc0, sc1, sc2, a.scale_modifier, dcov, ds, dq);
		if (LEAF) {
			ds[0] *= sc0; ds[1] *= sc1; ds[2] *= sc2;                            // d exp(s) = exp(s)
			const float dot = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;  // F.normalize backward: (g - qh (qh . g)) / |q|
			dq = make_float4((dq.x - q.x * dot) * qinv, (dq.y - q.y * dot) * qinv, (dq.z - q.z * dot) * qinv, (dq.w - q.w * dot) * qinv);
		}
	}
	put1(a.dL_dscale + i3, ds[0]); put1(a.dL_dscale + i3 + 1, ds[1]); put1(a.dL_dscale + i3 + 2, ds[2]);
	float* orot = a.dL_drot + 4 * ii;                                          // bucket segments need not be 16-byte aligned: scalar
	put1(orot, dq.x); put1(orot + 1, dq.y); put1(orot + 2, dq.z); put1(orot + 3, dq.w);
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

int launch_backward_views(cudaStream_t stream, const BackwardViewsArgs& a)
{
	if (a.P == 0 || a.nviews == 0 || a.count == 0) return 0;
	const unsigned grid = (unsigned)((a.count + BV_THREADS - 1) / BV_THREADS);
	// leave_room: at most 3 CTAs per SM (by shared memory), so that one CTA of the all-reduce kernel of an earlier chunk fits beside
	const size_t smem = a.leave_room ? (BV_SMEM_BYTES > 59392 ? BV_SMEM_BYTES : (size_t)59392) : BV_SMEM_BYTES;
	auto launch = [&](auto kern) -> int {
		if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 59392) != cudaSuccess) {
			set_error("backward_views: cannot reserve %d bytes of dynamic shared memory", 59392); return -1;
		}
		count_launch();
		kern<<<grid, BV_THREADS, smem, stream>>>(a);
		return 0;
	};
	if (a.leaf ? launch(backward_views_kernel<true>) : launch(backward_views_kernel<false>)) return -1;
	return check_launch("backward_views", stream, false);
}

}  // namespace rr
