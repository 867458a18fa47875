// Image loss right after the render: the training loss of the reference,
//   loss = (1 - lambda) * mean|x - y| + lambda * (1 - mean SSIM(x, y))            (train.py:112-113)
// with the 11x11 Gaussian-window SSIM of utils/loss_utils.py:15-51 (sigma 1.5, zero padding, per channel), forward AND
// gradient with respect to the rendered image x in two kernels — the reference runs five grouped conv2d, ~15
// elementwise kernels and their autograd twins on [3,H,W] tensors for the same result.
//
//   1. ssim_stats_kernel: one CTA per 16x16 pixel tile and channel.  x and y with a 5-pixel halo go to shared memory
//      once; the window is separable, so the five windowed sums (x, y, xx, yy, xy) are a horizontal 11-tap pass into
//      shared memory and a vertical 11-tap pass into registers.  Per pixel it forms the SSIM value and the three
//      partial derivatives dS/dmu1, dS/dE[xx], dS/dE[xy] (12 B per pixel and channel written), and block-reduces
//      sum|x - y| and sum S into two double accumulators (one atomicAdd each per CTA).
//   2. ssim_grad_kernel: dL/dx(q) = g * [ (w * M_mu)(q) + 2 x(q) (w * M_xx)(q) + y(q) (w * M_xy)(q) ] + (1-lambda)/n * sign(x - y),
//      g = -lambda / n: the same separable two-pass window over the three derivative maps (the window is symmetric).
//   3. loss_finish_kernel: the three scalars { L1, SSIM, loss } from the accumulators.
// HBM traffic: 2 reads + 3 map writes, then 3 map reads + 2 reads + 1 write of [C,H,W] floats = 11 x 4 B per pixel and
// channel (274 MB at 3x1080x1920 -> ~42 us at the measured 6.5 TB/s); arithmetic is ~260 FMA per pixel and channel.
//
// Also here: the L2 loss the benchmark's end-to-end arm uses (dL/dx = x - y, loss += 0.5 sum (x-y)^2) as one fused pass.
#include <cmath>
#include "common.cuh"
#include "kernels.cuh"

namespace rr {

namespace {

constexpr int LT = 16;                 // output tile edge
constexpr int HALO = 5;                // window_size // 2 (loss_utils.py:34)
constexpr int LIN = LT + 2 * HALO;     // 26
constexpr int WIN = 2 * HALO + 1;      // 11
constexpr float kC1 = 0.01f * 0.01f, kC2 = 0.03f * 0.03f;     // loss_utils.py:45-46

struct Window { float w[WIN]; };

// normalised 1-D Gaussian, sigma 1.5 (loss_utils.py:12-14: float32 taps, float32 normalisation)
Window make_window()
{
	Window g; float sum = 0.0f;
	for (int i = 0; i < WIN; i++) { g.w[i] = (float)std::exp(-(double)((i - HALO) * (i - HALO)) / (2.0 * 1.5 * 1.5)); sum += g.w[i]; }
	for (int i = 0; i < WIN; i++) g.w[i] /= sum;
	return g;
}

__device__ __forceinline__ float block_sum_256(float v, float* s_red /*[8]*/)
{
	#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
	if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
	__syncthreads();
	float t = 0.0f;
	if (threadIdx.x < 8) t = s_red[threadIdx.x];
	if (threadIdx.x < 32) {
		#pragma unroll
		for (int o = 4; o >= 1; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
	}
	__syncthreads();
	return t;      // valid in thread 0
}

__global__ void __launch_bounds__(LT * LT) ssim_stats_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Window g,
	float* __restrict__ m_mu, float* __restrict__ m_xx, float* __restrict__ m_xy, double* __restrict__ acc /*[2]: sum|x-y|, sum S*/)
{
	__shared__ float sx[LIN][LIN + 1], sy[LIN][LIN + 1];
	__shared__ float sh[5][LIN][LT + 1];
	__shared__ float s_red[8];
	const int tid = threadIdx.x, tx = tid & (LT - 1), ty = tid / LT;
	const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
	const size_t plane = (size_t)blockIdx.z * H * W;
	for (int i = tid; i < LIN * LIN; i += LT * LT) {
		const int r = i / LIN, c = i - r * LIN, gy = y0 + r - HALO, gx = x0 + c - HALO;
		const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;            // zero padding (conv2d padding=5, loss_utils.py:34)
		sx[r][c] = in ? __ldg(img + plane + (size_t)gy * W + gx) : 0.0f;
		sy[r][c] = in ? __ldg(gt + plane + (size_t)gy * W + gx) : 0.0f;
	}
	__syncthreads();
	for (int i = tid; i < LIN * LT; i += LT * LT) {                        // horizontal pass
		const int r = i / LT, c = i - r * LT;
		float a = 0, b = 0, aa = 0, bb = 0, ab = 0;
		#pragma unroll
		for (int k = 0; k < WIN; k++) {
			const float x = sx[r][c + k], y = sy[r][c + k], w = g.w[k];
			a = fmaf(w, x, a); b = fmaf(w, y, b); aa = fmaf(w, x * x, aa); bb = fmaf(w, y * y, bb); ab = fmaf(w, x * y, ab);
		}
		sh[0][r][c] = a; sh[1][r][c] = b; sh[2][r][c] = aa; sh[3][r][c] = bb; sh[4][r][c] = ab;
	}
	__syncthreads();
	float mu1 = 0, mu2 = 0, e11 = 0, e22 = 0, e12 = 0;                     // vertical pass
	#pragma unroll
	for (int k = 0; k < WIN; k++) {
		const float w = g.w[k];
		mu1 = fmaf(w, sh[0][ty + k][tx], mu1); mu2 = fmaf(w, sh[1][ty + k][tx], mu2);
		e11 = fmaf(w, sh[2][ty + k][tx], e11); e22 = fmaf(w, sh[3][ty + k][tx], e22); e12 = fmaf(w, sh[4][ty + k][tx], e12);
	}
	const int px = x0 + tx, py = y0 + ty;
	const bool inside = px < W && py < H;
	float l1 = 0.0f, S = 0.0f;
	if (inside) {
		// loss_utils.py:37-48
		const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
		const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
		const float A1 = 2.0f * mu12 + kC1, A2 = 2.0f * s12 + kC2, B1 = mu1_sq + mu2_sq + kC1, B2 = s1 + s2 + kC2;
		const float inv = 1.0f / (B1 * B2);
		S = A1 * A2 * inv;
		const size_t o = plane + (size_t)py * W + px;
		m_mu[o] = 2.0f * mu2 * (A2 - A1) * inv - S * 2.0f * mu1 * (B2 - B1) * inv;     // dS/dmu1 (through A1, A2, B1, B2)
		m_xx[o] = -S / B2;                                                              // dS/dE[xx]
		m_xy[o] = 2.0f * A1 * inv;                                                      // dS/dE[xy]
		l1 = fabsf(sx[ty + HALO][tx + HALO] - sy[ty + HALO][tx + HALO]);
	}
	const float bl1 = block_sum_256(l1, s_red), bS = block_sum_256(S, s_red);
	if (tid == 0) { atomicAdd(&acc[0], (double)bl1); atomicAdd(&acc[1], (double)bS); }
}

__global__ void __launch_bounds__(LT * LT) ssim_grad_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, Window g,
	const float* __restrict__ m_mu, const float* __restrict__ m_xx, const float* __restrict__ m_xy, float gS /*-lambda/n*/, float gL1 /*(1-lambda)/n*/,
	float* __restrict__ dL_dimg)
{
	__shared__ float sm[3][LIN][LIN + 1];
	__shared__ float sh[3][LIN][LT + 1];
	const int tid = threadIdx.x, tx = tid & (LT - 1), ty = tid / LT;
	const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
	const size_t plane = (size_t)blockIdx.z * H * W;
	for (int i = tid; i < LIN * LIN; i += LT * LT) {
		const int r = i / LIN, c = i - r * LIN, gy = y0 + r - HALO, gx = x0 + c - HALO;
		const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
		const size_t o = plane + (size_t)gy * W + gx;
		sm[0][r][c] = in ? __ldg(m_mu + o) : 0.0f; sm[1][r][c] = in ? __ldg(m_xx + o) : 0.0f; sm[2][r][c] = in ? __ldg(m_xy + o) : 0.0f;
	}
	__syncthreads();
	for (int i = tid; i < LIN * LT; i += LT * LT) {
		const int r = i / LT, c = i - r * LT;
		float a = 0, b = 0, d = 0;
		#pragma unroll
		for (int k = 0; k < WIN; k++) { const float w = g.w[k]; a = fmaf(w, sm[0][r][c + k], a); b = fmaf(w, sm[1][r][c + k], b); d = fmaf(w, sm[2][r][c + k], d); }
		sh[0][r][c] = a; sh[1][r][c] = b; sh[2][r][c] = d;
	}
	__syncthreads();
	float cmu = 0, cxx = 0, cxy = 0;
	#pragma unroll
	for (int k = 0; k < WIN; k++) { const float w = g.w[k]; cmu = fmaf(w, sh[0][ty + k][tx], cmu); cxx = fmaf(w, sh[1][ty + k][tx], cxx); cxy = fmaf(w, sh[2][ty + k][tx], cxy); }
	const int px = x0 + tx, py = y0 + ty;
	if (px < W && py < H) {
		const size_t o = plane + (size_t)py * W + px;
		const float x = __ldg(img + o), y = __ldg(gt + o), d = x - y;
		const float sgn = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);       // torch.abs backward: sign(0) = 0
		dL_dimg[o] = gS * (cmu + 2.0f * x * cxx + y * cxy) + gL1 * sgn;
	}
}

__global__ void loss_finish_kernel(const double* __restrict__ acc, double n, float lambda, float* __restrict__ out3)
{
	const double l1 = acc[0] / n, s = acc[1] / n;
	out3[0] = (float)l1; out3[1] = (float)s; out3[2] = (float)((1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - s));
}

// dL = x - y, loss += 0.5 * sum (x - y)^2 in one pass (4 elements per thread and iteration, scalar tail).  T = float, or
// uint8_t for a target image kept in its native 8-bit form (y = byte * scale): a quarter of the host->device bytes.
template <class T>
__global__ void __launch_bounds__(256) l2_loss_grad_kernel(size_t n, const float* __restrict__ x, const T* __restrict__ y, float scale, float* __restrict__ d_out,
	float* __restrict__ loss_acc)
{
	__shared__ float s_red[8];
	float acc = 0.0f;
	const size_t n4 = n / 4, stride = (size_t)gridDim.x * 256;
	const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(d_out)) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & (4 * sizeof(T) - 1)) == 0;
	if (vec) {
		for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
			const float4 a = __ldg(reinterpret_cast<const float4*>(x) + i);
			float4 b;
			if (sizeof(T) == 1) {
				const uchar4 u = __ldg(reinterpret_cast<const uchar4*>(y) + i);
				b = make_float4(__fmul_rn(u.x, scale), __fmul_rn(u.y, scale), __fmul_rn(u.z, scale), __fmul_rn(u.w, scale));   // rounded product, not fused into the subtraction
			} else {
				b = __ldg(reinterpret_cast<const float4*>(y) + i);
			}
			const float4 d = make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
			reinterpret_cast<float4*>(d_out)[i] = d;
			acc += d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w;
		}
	}
	for (size_t i = (vec ? n4 * 4 : 0) + (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
		const float d = x[i] - (sizeof(T) == 1 ? __fmul_rn((float)y[i], scale) : (float)y[i]);
		d_out[i] = d; acc += d * d;
	}
	const float t = block_sum_256(acc, s_red);
	if (threadIdx.x == 0 && loss_acc) atomicAdd(loss_acc, 0.5f * t);
}

}  // namespace

size_t loss_ws_bytes(int C, int H, int W) { return align_up((size_t)3 * C * H * W * sizeof(float), 256) + 256; }

int launch_loss_l1_ssim(cudaStream_t stream, int C, int H, int W, const float* img, const float* gt, float lambda, void* ws, float* out3, float* dL_dimg)
{
	static const Window g = make_window();
	const size_t n = (size_t)C * H * W;
	float* m_mu = reinterpret_cast<float*>(ws);
	float* m_xx = m_mu + n; float* m_xy = m_xx + n;
	double* acc = reinterpret_cast<double*>(reinterpret_cast<char*>(ws) + align_up(3 * n * sizeof(float), 256));
	cudaMemsetAsync(acc, 0, 2 * sizeof(double), stream);
	const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
	count_launch(), ssim_stats_kernel<<<grid, LT * LT, 0, stream>>>(H, W, img, gt, g, m_mu, m_xx, m_xy, acc);
	if (dL_dimg)
		count_launch(), ssim_grad_kernel<<<grid, LT * LT, 0, stream>>>(H, W, img, gt, g, m_mu, m_xx, m_xy, -lambda / (float)n, (1.0f - lambda) / (float)n, dL_dimg);
	if (out3) count_launch(), loss_finish_kernel<<<1, 1, 0, stream>>>(acc, (double)n, lambda, out3);
	return check_launch("loss_l1_ssim", stream, false);
}

int launch_loss_l2(cudaStream_t stream, size_t n, const float* img, const void* gt, int gt_is_u8, float scale, float* dL_dimg, float* loss_acc)
{
	const size_t blocks = (n / 4 + 255) / 256 + 1;
	const unsigned grid = (unsigned)(blocks < 148 * 8 ? blocks : 148 * 8);
	count_launch();
	if (gt_is_u8) l2_loss_grad_kernel<uint8_t><<<grid, 256, 0, stream>>>(n, img, reinterpret_cast<const uint8_t*>(gt), scale, dL_dimg, loss_acc);
	else l2_loss_grad_kernel<float><<<grid, 256, 0, stream>>>(n, img, reinterpret_cast<const float*>(gt), 1.0f, dL_dimg, loss_acc);
	return check_launch("loss_l2", stream, false);
}

}  // namespace rr
