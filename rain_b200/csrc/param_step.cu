// Parameter step after the (all-reduced) backward: the chain rule through the caller-side
// activations, the Adam update of the six leaf tensors, and the re-activation for the next render, in one pass over the
// 59 floats of a Gaussian.  Replaces, per training iteration of the reference:
//   * autograd backward of scene/gaussian_model.py:85-105 — exp (get_scaling), sigmoid (get_opacity), F.normalize
//     (get_rotation), torch.cat (get_features) — ~10 elementwise kernels and a 192 B/Gaussian cat copy each way;
//   * torch.optim.Adam(l, lr=0.0, eps=1e-15).step() over the six parameter groups of training_setup
//     (scene/gaussian_model.py:139-157; train.py:146): betas (0.9, 0.999), no weight decay, no amsgrad;
//   * the forward activations of the next render() call (gaussian_renderer/__init__.py:37-60).
// Gradients arrive in the flat bucket the backward accumulates into (multiview.py: means3D | sh | opacity | scales |
// rotations, w.r.t. the ACTIVATED values); Adam moments use the same layout.  Every rank applies the same step to its
// replica after the all-reduce, so replicas stay bit-identical without a parameter broadcast.
// Traffic: 4 reads + 4 writes of 59 floats per Gaussian (1.9 KB): HBM bound, ~0.29 ms at 1M Gaussians.
#include <cmath>
#include "common.cuh"
#include "kernels.cuh"
#include "splat_math.cuh"

namespace rr {

namespace {

struct AdamScalars { float lr[6]; float beta1, beta2, eps, inv_bc1, inv_sqrt_bc2; };   // groups: xyz, f_dc, f_rest, opacity, scaling, rotation

// torch.optim.Adam single-tensor step (torch/optim/adam.py _single_tensor_adam, default flags):
//   m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
__device__ __forceinline__ float adam(float p, float g, float& m, float& v, float lr, const AdamScalars& h)
{
	m = fmaf(h.beta1, m, (1.0f - h.beta1) * g);
	v = fmaf(h.beta2, v, (1.0f - h.beta2) * g * g);
	const float denom = sqrtf(v) * h.inv_sqrt_bc2 + h.eps;
	return p - lr * h.inv_bc1 * (m / denom);
}

// SH coefficients: element i of the [P, M, 3] gradient segment belongs to f_dc (coefficient 0) or f_rest; the updated
// value also goes to the concatenated [P, M, 3] render input (torch.cat((f_dc, f_rest), dim=1)).
__global__ void __launch_bounds__(256) param_step_sh_kernel(size_t n /*P*M*3*/, int M3 /*M*3*/, float* __restrict__ f_dc, float* __restrict__ f_rest,
	float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g, AdamScalars h, float* __restrict__ shs_out)
{
	for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
		const size_t p = i / (size_t)M3;
		const int j = (int)(i - p * (size_t)M3);
		float* dst = j < 3 ? f_dc + p * 3 + j : f_rest + p * (size_t)(M3 - 3) + (j - 3);
		float mm = m[i], vv = v[i];
		const float out = adam(*dst, __ldg(g + i), mm, vv, j < 3 ? h.lr[1] : h.lr[2], h);
		*dst = out; m[i] = mm; v[i] = vv;
		if (shs_out) shs_out[i] = out;
	}
}

// one thread per Gaussian: xyz (3), opacity (1), scaling (3), rotation (4)
__global__ void __launch_bounds__(256) param_step_geom_kernel(int P, float* __restrict__ xyz, float* __restrict__ opacity, float* __restrict__ scaling,
	float* __restrict__ rotation, float* __restrict__ m_xyz, float* __restrict__ v_xyz, float* __restrict__ m_op, float* __restrict__ v_op,
	float* __restrict__ m_sc, float* __restrict__ v_sc, float* __restrict__ m_rot, float* __restrict__ v_rot,
	const float* __restrict__ g_xyz, const float* __restrict__ g_op, const float* __restrict__ g_sc, const float* __restrict__ g_rot, AdamScalars h,
	float* __restrict__ opacities_out, float* __restrict__ scales_out, float* __restrict__ rotations_out)
{
	const int p = blockIdx.x * 256 + threadIdx.x;
	if (p >= P) return;
	#pragma unroll
	for (int c = 0; c < 3; c++) {                                   // xyz: identity activation
		const size_t i = (size_t)p * 3 + c;
		float mm = m_xyz[i], vv = v_xyz[i];
		xyz[i] = adam(xyz[i], g_xyz[i], mm, vv, h.lr[0], h);
		m_xyz[i] = mm; v_xyz[i] = vv;
	}
	{                                                               // opacity = sigmoid(o): d/do = s (1 - s)
		const float o = opacity[p], s = act_opacity(o);
		float mm = m_op[p], vv = v_op[p];
		const float on = adam(o, g_op[p] * s * (1.0f - s), mm, vv, h.lr[3], h);
		opacity[p] = on; m_op[p] = mm; v_op[p] = vv;
		if (opacities_out) opacities_out[p] = act_opacity(on);
	}
	#pragma unroll
	for (int c = 0; c < 3; c++) {                                   // scales = exp(s): d/ds = exp(s)
		const size_t i = (size_t)p * 3 + c;
		const float s = scaling[i];
		float mm = m_sc[i], vv = v_sc[i];
		const float sn = adam(s, g_sc[i] * expf(s), mm, vv, h.lr[4], h);
		scaling[i] = sn; m_sc[i] = mm; v_sc[i] = vv;
		if (scales_out) scales_out[i] = act_scale(sn);
	}
	{                                                               // rotations = q / max(|q|, 1e-12) (F.normalize): d/dq = (g - qh (qh.g)) / |q|
		// the bucket segments start at 55 P floats: only the leaf tensor itself is known to be 16-byte aligned
		const size_t r4 = (size_t)p * 4;
		const float4 q = reinterpret_cast<const float4*>(rotation)[p], g = make_float4(g_rot[r4], g_rot[r4 + 1], g_rot[r4 + 2], g_rot[r4 + 3]);
		const float nrm = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f), inv = 1.0f / nrm;
		const float4 qh = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
		const float dot = qh.x * g.x + qh.y * g.y + qh.z * g.z + qh.w * g.w;
		const float4 d = make_float4((g.x - qh.x * dot) * inv, (g.y - qh.y * dot) * inv, (g.z - qh.z * dot) * inv, (g.w - qh.w * dot) * inv);
		float4 mm = make_float4(m_rot[r4], m_rot[r4 + 1], m_rot[r4 + 2], m_rot[r4 + 3]), vv = make_float4(v_rot[r4], v_rot[r4 + 1], v_rot[r4 + 2], v_rot[r4 + 3]);
		float4 qn;
		qn.x = adam(q.x, d.x, mm.x, vv.x, h.lr[5], h); qn.y = adam(q.y, d.y, mm.y, vv.y, h.lr[5], h);
		qn.z = adam(q.z, d.z, mm.z, vv.z, h.lr[5], h); qn.w = adam(q.w, d.w, mm.w, vv.w, h.lr[5], h);
		reinterpret_cast<float4*>(rotation)[p] = qn;
		m_rot[r4] = mm.x; m_rot[r4 + 1] = mm.y; m_rot[r4 + 2] = mm.z; m_rot[r4 + 3] = mm.w;
		v_rot[r4] = vv.x; v_rot[r4 + 1] = vv.y; v_rot[r4 + 2] = vv.z; v_rot[r4 + 3] = vv.w;
		if (rotations_out) reinterpret_cast<float4*>(rotations_out)[p] = act_rot(qn);
	}
}

// Adam over the six raw leaf tensors with LEAF gradients (bucket layout xyz | f_dc | f_rest | opacity | scaling | rotation, what
// the leaf mode of rr_backward_views writes): no chain rule, no re-activation, no activated copies — one flat pass; the segment
// of a flat index selects the leaf tensor and its learning rate.
struct LeafSegs { float* p[6]; size_t end[6]; };
__global__ void __launch_bounds__(256) param_step_leaf_kernel(size_t n, LeafSegs sg, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g, AdamScalars h)
{
	for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
		int k = 0;
		#pragma unroll
		for (int j = 0; j < 5; j++) k += i >= sg.end[j] ? 1 : 0;
		float* dst = sg.p[k] + (i - (k ? sg.end[k - 1] : 0));
		float mm = m[i], vv = v[i];
		*dst = adam(*dst, __ldg(g + i), mm, vv, h.lr[k], h);
		m[i] = mm; v[i] = vv;
	}
}

// The caller-side activations alone (scene/gaussian_model.py:85-105: sigmoid, exp, F.normalize), with the same device functions
// the leaf-mode rasterizer kernels and the parameter step use (splat_math.cuh): one thread per Gaussian.
__global__ void __launch_bounds__(256) activate_kernel(int P, const float* __restrict__ opacity, const float* __restrict__ scaling, const float* __restrict__ rotation,
	float* __restrict__ opacities_out, float* __restrict__ scales_out, float* __restrict__ rotations_out)
{
	const int p = blockIdx.x * 256 + threadIdx.x;
	if (p >= P) return;
	opacities_out[p] = act_opacity(opacity[p]);
	#pragma unroll
	for (int c = 0; c < 3; c++) scales_out[(size_t)p * 3 + c] = act_scale(scaling[(size_t)p * 3 + c]);
	const float4 q = act_rot(load_f4(rotation, (size_t)p));
	float* o = rotations_out + (size_t)p * 4;
	o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
}

// Densification statistics of one rendered view (train.py:132-134, scene/gaussian_model.py:419-421), one thread per
// Gaussian, for the Gaussians the view saw (visibility_filter = radii > 0):
//   max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |dL/dmean2D.xy|; denom += 1
__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dL_dmean2D /*[P,3]*/,
	float* __restrict__ max_radii2D, float* __restrict__ grad_accum, float* __restrict__ denom)
{
	const int p = blockIdx.x * 256 + threadIdx.x;
	if (p >= P) return;
	const int r = radii[p];
	if (r <= 0) return;
	max_radii2D[p] = fmaxf(max_radii2D[p], (float)r);
	const float gx = dL_dmean2D[(size_t)p * 3], gy = dL_dmean2D[(size_t)p * 3 + 1];
	grad_accum[p] += sqrtf(gx * gx + gy * gy);          // torch.norm(grad[:, :2], dim=-1)
	denom[p] += 1.0f;
}

}  // namespace

int launch_activate(cudaStream_t stream, int P, const float* opacity, const float* scaling, const float* rotation, float* opacities_out, float* scales_out, float* rotations_out)
{
	count_launch(), activate_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, opacity, scaling, rotation, opacities_out, scales_out, rotations_out);
	return check_launch("activate", stream, false);
}

int launch_densify_stats(cudaStream_t stream, int P, const int* radii, const float* dL_dmean2D, float* max_radii2D, float* grad_accum, float* denom)
{
	count_launch(), densify_stats_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, radii, dL_dmean2D, max_radii2D, grad_accum, denom);
	return check_launch("densify_stats", stream, false);
}

static AdamScalars adam_scalars(const float* hyper, int step)
{
	AdamScalars h;
	for (int i = 0; i < 6; i++) h.lr[i] = hyper[i];
	h.beta1 = hyper[6]; h.beta2 = hyper[7]; h.eps = hyper[8];
	// bias corrections as torch computes them (python floats, i.e. double): 1 - beta ** step
	h.inv_bc1 = (float)(1.0 / (1.0 - std::pow((double)hyper[6], (double)step)));
	h.inv_sqrt_bc2 = (float)(1.0 / std::sqrt(1.0 - std::pow((double)hyper[7], (double)step)));
	return h;
}

int launch_param_step_leaf(cudaStream_t stream, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* m, float* v, const float* grads, const float* hyper, int step)
{
	const AdamScalars h = adam_scalars(hyper, step);
	const size_t Pz = (size_t)P;
	LeafSegs sg;
	sg.p[0] = xyz; sg.p[1] = f_dc; sg.p[2] = f_rest; sg.p[3] = opacity; sg.p[4] = scaling; sg.p[5] = rotation;
	const size_t len[6] = { 3 * Pz, 3 * Pz, (size_t)(M - 1) * 3 * Pz, Pz, 3 * Pz, 4 * Pz };
	size_t acc = 0;
	for (int k = 0; k < 6; k++) { acc += len[k]; sg.end[k] = acc; }
	const size_t blocks = (acc + 255) / 256;
	count_launch(), param_step_leaf_kernel<<<(unsigned)(blocks < 148u * 16u ? blocks : 148u * 16u), 256, 0, stream>>>(acc, sg, m, v, grads, h);
	return check_launch("param_step_leaf", stream, false);
}

int launch_param_step(cudaStream_t stream, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* m, float* v, const float* grads, const float* hyper /*host: lr[6], beta1, beta2, eps*/, int step,
	float* shs_out, float* opacities_out, float* scales_out, float* rotations_out)
{
	const AdamScalars h = adam_scalars(hyper, step);
	const size_t Pz = (size_t)P, nsh = Pz * M * 3;
	// bucket layout: means3D [P,3] | sh [P,M,3] | opacity [P] | scales [P,3] | rotations [P,4]
	const size_t o_sh = 3 * Pz, o_op = o_sh + nsh, o_sc = o_op + Pz, o_rot = o_sc + 3 * Pz;
	const size_t blocks = (nsh + 255) / 256;
	count_launch(), param_step_sh_kernel<<<(unsigned)(blocks < 148u * 16u ? blocks : 148u * 16u), 256, 0, stream>>>(nsh, M * 3, f_dc, f_rest, m + o_sh, v + o_sh,
		grads + o_sh, h, shs_out);
	count_launch(), param_step_geom_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, xyz, opacity, scaling, rotation, m, v, m + o_op, v + o_op, m + o_sc, v + o_sc,
		m + o_rot, v + o_rot, grads, grads + o_op, grads + o_sc, grads + o_rot, h, opacities_out, scales_out, rotations_out);
	return check_launch("param_step", stream, false);
}

}  // namespace rr
