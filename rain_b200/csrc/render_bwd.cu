// Backward tile renderer: back-to-front replay of each tile's list, producing one gradient partial per
// (tile, Gaussian) duplicate.  Replaces renderCUDA<3> backward (reference backward.cu:389-547), whose 9 global float
// atomics per contributing (pixel, Gaussian) pair serialise in L2 on a single address per Gaussian.
//
// B200 design:
//  * same TMA-staged list walk as the forward (render_fwd.cu), in reverse order, starting at the tile's deepest
//    contributor (tile_max_contrib, recorded by the forward); each warp visits exactly the pairs its 8x4 pixels blended
//    in the forward (one flag byte per pair written by the forward) — no culling arithmetic, no wasted pairs;
//  * each lane forms the 9 per-pair terms for its pixel exactly as the reference does; the warp then reduces them with
//    a halving butterfly (8 values in 4+2+1+1+1 shuffles, the 9th in 5) — skipped when no lane contributed;
//  * the 8 warps' sums meet in a shared-memory slab (plain stores, a per-warp bitmask says which rows are live) and
//    ONE 40-byte partial per (tile, Gaussian) is stored to global memory at the duplicate's emission index, so the
//    partials of one Gaussian are contiguous for the per-Gaussian gather in preprocess_bwd.cu.
//  No atomics anywhere: the backward is deterministic (bit-reproducible run to run), unlike the reference.
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace rr {

namespace {

constexpr int BWD_BATCH = 64;
constexpr int NV = 9;

// Sum v[0..8] over the 32 lanes.  On return lane L with (L & 3) == 0 holds in `y` the total of value
// k = 4*bit4(L) + 2*bit3(L) + bit2(L); every lane holds the total of v[8] in `z`.
__device__ __forceinline__ void warp_reduce9(const float (&v)[NV], int lane, float& y, float& z)
{
	float w[4], x[2];
	const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
	#pragma unroll
	for (int i = 0; i < 4; i++) {
		const float send = h4 ? v[i] : v[i + 4], keep = h4 ? v[i + 4] : v[i];
		w[i] = keep + __shfl_xor_sync(FULL, send, 16);
	}
	#pragma unroll
	for (int i = 0; i < 2; i++) {
		const float send = h3 ? w[i] : w[i + 2], keep = h3 ? w[i + 2] : w[i];
		x[i] = keep + __shfl_xor_sync(FULL, send, 8);
	}
	{
		const float send = h2 ? x[0] : x[1], keep = h2 ? x[1] : x[0];
		y = keep + __shfl_xor_sync(FULL, send, 4);
	}
	y += __shfl_xor_sync(FULL, y, 2);
	y += __shfl_xor_sync(FULL, y, 1);
	z = v[8];
	#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) z += __shfl_xor_sync(FULL, z, o);
}

template <int MODE>
__global__ void __launch_bounds__(RENDER_THREADS) render_bwd_kernel(RenderArgs a)
{
	constexpr bool USE_TMA = StageMode<MODE>::tma;
	__shared__ __align__(128) float4 s_rec[STAGES][BWD_BATCH * 3];
	__shared__ __align__(8) uint64_t s_full[STAGES];
	__shared__ float s_slab[RENDER_WARPS][BWD_BATCH * NV];
	__shared__ uint8_t s_flags[2][BWD_BATCH];     // [batch parity] the forward's "which warps blended this pair" byte

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	int wx, wy, lx, ly;
	warp_footprint(warp, lane, wx, wy, lx, ly);
	const int pix_x = blockIdx.x * TILE_X + wx + lx, pix_y = blockIdx.y * TILE_Y + wy + ly;
	const bool inside = pix_x < a.W && pix_y < a.H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;

	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const int m = min(n, (int)a.tile_max_contrib[tile]);        // entries [m, n) contributed to no pixel
	const int nbatches = (m + BWD_BATCH - 1) / BWD_BATCH;
	const float4* __restrict__ src = reinterpret_cast<const float4*>(a.packed) + (size_t)range.x * 3;
	const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
	const uint32_t* __restrict__ ids = a.point_list + range.x;
	const uint32_t* __restrict__ emit = a.sorted_emit + range.x;
	const uint8_t* __restrict__ flags = a.pair_flags + range.x;

	// batch b covers local indices [lo_b, hi_b), hi_b = m - b*BWD_BATCH (back to front)
	if (tid == 0 && USE_TMA) {
		for (int s = 0; s < STAGES; s++) mbar_init(&s_full[s], 1);
		mbar_fence_init();
	}
	if (tid < BWD_BATCH && nbatches > 0) {
		const int lo0 = max(0, m - BWD_BATCH);
		s_flags[0][tid] = (lo0 + tid < m) ? flags[lo0 + tid] : (uint8_t)0;
	}
	__syncthreads();
	if (USE_TMA && warp == 0) {
		for (int b = 0; b < STAGES && b < nbatches; b++) {
			const int hi = m - b * BWD_BATCH, lo = max(0, hi - BWD_BATCH);
			stage_fill_async<MODE>(&s_rec[b][0], &s_full[b], src, grec, ids, lo, hi - lo, lane);
		}
	}

	// duplicates behind the deepest contributor get zero partials (one thread per pair: one slot load, five 8-byte stores)
	for (int e = m + tid; e < n; e += RENDER_THREADS) {
		float2* __restrict__ p = reinterpret_cast<float2*>(a.partial + (size_t)__ldg(emit + e) * PARTIAL_FLOATS);
		#pragma unroll
		for (int k = 0; k < PARTIAL_FLOATS / 2; k++) p[k] = make_float2(0.0f, 0.0f);
	}

	const size_t pid = (size_t)a.W * pix_y + pix_x, HW = (size_t)a.H * a.W;
	const float T_final = inside ? a.final_T[pid] : 0.0f;
	float T = T_final;
	const int last_contributor = inside ? (int)a.n_contrib[pid] : 0;
	float dLp0 = 0, dLp1 = 0, dLp2 = 0;
	if (inside) { dLp0 = a.dL_dpix[pid]; dLp1 = a.dL_dpix[HW + pid]; dLp2 = a.dL_dpix[2 * HW + pid]; }
	float bg_dot_dpixel = 0;
	bg_dot_dpixel += a.bg[0] * dLp0; bg_dot_dpixel += a.bg[1] * dLp1; bg_dot_dpixel += a.bg[2] * dLp2;
	float ar0 = 0, ar1 = 0, ar2 = 0, lc0 = 0, lc1 = 0, lc2 = 0, last_alpha = 0;
	const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;

	for (int b = 0; b < nbatches; b++) {
		const int s = b % STAGES;
		const int hi = m - b * BWD_BATCH, lo = max(0, hi - BWD_BATCH), cnt = hi - lo;
		if (USE_TMA) {
			mbar_wait(&s_full[s], (uint32_t)(b / STAGES) & 1u);
		} else {
			stage_fill_sync<MODE>(&s_rec[s][0], src, grec, ids, lo, cnt, tid);
			__syncthreads();
		}
		const float4* __restrict__ rec = s_rec[s];
		// the pairs this warp's 8x4 pixels blended in the forward, as a 64-bit mask over the batch
		const uint8_t* __restrict__ fl = s_flags[b & 1];
		uint64_t surv = (uint64_t)__ballot_sync(FULL, (fl[lane] >> warp) & 1) | ((uint64_t)__ballot_sync(FULL, (fl[lane + 32] >> warp) & 1) << 32);
		while (surv) {
			const int e = 63 - __clzll((long long)surv);
			surv ^= 1ull << e;
			const float4 ra = rec[e * 3], rb = rec[e * 3 + 1], rc = rec[e * 3 + 2];
			float v[NV];
			#pragma unroll
			for (int i = 0; i < NV; i++) v[i] = 0.0f;
			if (lo + e < last_contributor) {
				const float dx = __fsub_rn(ra.x, pxf), dy = __fsub_rn(ra.y, pyf);
				const float q = __fmaf_rn(dx, __fmul_rn(dx, ra.z), __fmul_rn(dy, __fmul_rn(dy, rb.x)));
				const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, ra.w)));
				if (!(power > 0.0f)) {
					const float G = expf(power);
					const float alpha = fminf(0.99f, __fmul_rn(rb.y, G));
					if (!(alpha < 1.0f / 255.0f)) {
						float inv_1ma;      // 1/(1-alpha), 1-alpha in [0.01, 1): the approximate reciprocal (1 ulp) is ample for 1e-4 gradients
						asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_1ma) : "f"(1.f - alpha));
						T = T * inv_1ma;
						const float dchannel_dcolor = alpha * T;
						const float one_m_la = 1.f - last_alpha;
						ar0 = last_alpha * lc0 + one_m_la * ar0; lc0 = rb.z;
						ar1 = last_alpha * lc1 + one_m_la * ar1; lc1 = rb.w;
						ar2 = last_alpha * lc2 + one_m_la * ar2; lc2 = rc.x;
						float dL_dalpha = (rb.z - ar0) * dLp0;
						dL_dalpha += (rb.w - ar1) * dLp1;
						dL_dalpha += (rc.x - ar2) * dLp2;
						v[0] = dchannel_dcolor * dLp0; v[1] = dchannel_dcolor * dLp1; v[2] = dchannel_dcolor * dLp2;
						dL_dalpha *= T;
						last_alpha = alpha;
						dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;
						const float dL_dG = rb.y * dL_dalpha;
						const float gdx = G * dx, gdy = G * dy;
						const float dG_ddelx = -gdx * ra.z - gdy * ra.w;
						const float dG_ddely = -gdy * rb.x - gdx * ra.w;
						v[3] = dL_dG * dG_ddelx * ddelx_dx;
						v[4] = dL_dG * dG_ddely * ddely_dy;
						v[5] = -0.5f * gdx * dx * dL_dG;
						v[6] = -0.5f * gdx * dy * dL_dG;
						v[7] = -0.5f * gdy * dy * dL_dG;
						v[8] = G * dL_dalpha;
					}
				}
			}
			float y, z;
			warp_reduce9(v, lane, y, z);
			if ((lane & 3) == 0) s_slab[warp][e * NV + ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)] = y;
			if (lane == 1) s_slab[warp][e * NV + 8] = z;
		}
		__syncthreads();
		// combine the warps that blended each pair and store one partial per (tile, Gaussian) at its slot
		for (int t = tid; t < cnt * PARTIAL_FLOATS; t += RENDER_THREADS) {
			const int e = t / PARTIAL_FLOATS, k = t - e * PARTIAL_FLOATS;
			float sum = 0.0f;
			if (k < NV) {
				const uint32_t f = fl[e];
				#pragma unroll
				for (int w = 0; w < RENDER_WARPS; w++) if ((f >> w) & 1u) sum += s_slab[w][e * NV + k];
			}
			a.partial[(size_t)emit[lo + e] * PARTIAL_FLOATS + k] = sum;
		}
		// flags of the next batch
		if (b + 1 < nbatches && tid < BWD_BATCH) {
			const int nhi = m - (b + 1) * BWD_BATCH, nlo = max(0, nhi - BWD_BATCH);
			s_flags[(b + 1) & 1][tid] = (nlo + tid < nhi) ? flags[nlo + tid] : (uint8_t)0;
		}
		__syncthreads();
		if (USE_TMA && warp == 0 && b + STAGES < nbatches) {
			const int nb = b + STAGES;
			const int nhi = m - nb * BWD_BATCH, nlo = max(0, nhi - BWD_BATCH);
			stage_fill_async<MODE>(&s_rec[s][0], &s_full[s], src, grec, ids, nlo, nhi - nlo, lane);
		}
	}
}

}  // namespace

int launch_render_bwd(cudaStream_t stream, const RenderArgs& a)
{
	const dim3 grid(a.grid_x, a.grid_y);
	switch (a.variant) {
	case MODE_PACKED_LDG: count_launch(), render_bwd_kernel<MODE_PACKED_LDG><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	case MODE_PACKED_TMA: count_launch(), render_bwd_kernel<MODE_PACKED_TMA><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	case MODE_GATHER_LDG: count_launch(), render_bwd_kernel<MODE_GATHER_LDG><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	default: count_launch(), render_bwd_kernel<MODE_GATHER_TMA><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	}
	return check_launch("render_bwd", stream, false);
}

}  // namespace rr
