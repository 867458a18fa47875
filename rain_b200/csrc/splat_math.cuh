// Explicit-rounding arithmetic shared by the forward preprocess and the view-batched backward: every step is a round-to-
// nearest intrinsic in the order the reference's nvcc build performs it ("FMA contract"), so the 3D covariance
// is the reference's bit for bit wherever it is (re)computed.
#pragma once
#include "common.cuh"

namespace rr {

#define FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))
#define DIV(a, b) __fdiv_rn((a), (b))

__device__ __forceinline__ float sop3(float a, float b, float c, float d, float e, float f) { return FMA(e, f, FMA(a, b, MUL(c, d))); }
// column `col` of the row-vector transform: m[col]*x + m[4+col]*y + m[8+col]*z + m[12+col]
__device__ __forceinline__ float xform(const float* __restrict__ m, int col, float x, float y, float z) {
	return ADD(m[12 + col], sop3(x, m[col], y, m[4 + col], z, m[8 + col]));
}

__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* c)
{
	const float sx = MUL(mod, s0), sy = MUL(mod, s1), sz = MUL(mod, s2);
	const float r = q.x, x = q.y, y = q.z, z = q.w;
	const float xz = MUL(x, z), A = FMA(r, y, xz), B = FMA(-r, y, xz);
	const float rx = MUL(r, x), C = FMA(y, z, -rx), Dd = FMA(y, z, rx);
	const float rz = MUL(r, z), E = FMA(x, y, -rz), F = FMA(x, y, rz);
	const float yy = MUL(y, y), zz = MUL(z, z), G = FMA(x, x, yy), Hh = ADD(yy, zz), I = FMA(x, x, zz);
	const float R00 = SUB(1.f, ADD(Hh, Hh)), R01 = ADD(E, E), R02 = ADD(A, A);
	const float R10 = ADD(F, F), R11 = SUB(1.f, ADD(I, I)), R12 = ADD(C, C);
	const float R20 = ADD(B, B), R21 = ADD(Dd, Dd), R22 = SUB(1.f, ADD(G, G));
	const float z0 = 0.0f;
	const float M00 = FMA(R02, z0, FMA(R01, z0, MUL(sx, R00)));
	const float t0 = MUL(R00, z0);
	const float M01 = FMA(R02, z0, FMA(sy, R01, t0));
	const float M02 = FMA(sz, R02, FMA(R01, z0, t0));
	const float t1 = MUL(R11, z0);
	const float M10 = FMA(R12, z0, FMA(sx, R10, t1));
	const float M11 = FMA(R12, z0, FMA(R10, z0, MUL(sy, R11)));
	const float M12 = FMA(sz, R12, FMA(R10, z0, t1));
	const float t2 = MUL(R21, z0);
	const float M20 = FMA(R22, z0, FMA(sx, R20, t2));
	const float M21 = FMA(R22, z0, FMA(R20, z0, MUL(sy, R21)));
	const float M22 = FMA(sz, R22, FMA(R20, z0, t2));
	c[0] = sop3(M00, M00, M01, M01, M02, M02);
	c[1] = sop3(M10, M00, M11, M01, M12, M02);
	c[2] = sop3(M20, M00, M21, M01, M22, M02);
	c[3] = sop3(M10, M10, M11, M11, M12, M12);
	c[4] = sop3(M20, M10, M21, M11, M22, M12);
	c[5] = sop3(M20, M20, M21, M21, M22, M22);
}

// ---- caller-side activations of the raw leaf tensors (scene/gaussian_model.py:85-105), for the opt-in leaf mode ----------
// opacity = torch.sigmoid(o), scales = torch.exp(s), rotations = F.normalize(q) = q / max(|q|, 1e-12).  One definition, used by
// the forward preprocess (leaf mode), the view-batched backward (leaf mode) and the parameter step's re-activation, so the
// three agree bit for bit with each other (and with torch to within an ulp or two: tests/test_param_step.py, golden vectors).
// Every step is an explicit intrinsic, so the compiler cannot contract it differently at different call sites.
__device__ __forceinline__ float act_opacity(float o) { return DIV(1.0f, ADD(1.0f, expf(-o))); }
__device__ __forceinline__ float act_scale(float s) { return expf(s); }
__device__ __forceinline__ float rot_inv_norm(float4 q) { return DIV(1.0f, fmaxf(__fsqrt_rn(FMA(q.w, q.w, FMA(q.z, q.z, FMA(q.y, q.y, MUL(q.x, q.x))))), 1e-12f)); }
__device__ __forceinline__ float4 act_rot(float4 q) { const float i = rot_inv_norm(q); return make_float4(MUL(q.x, i), MUL(q.y, i), MUL(q.z, i), MUL(q.w, i)); }

}  // namespace rr
