// Forward tile renderer: front-to-back alpha blending of each 16x16 tile's depth-sorted splat list.
// Replaces renderCUDA<3> forward (reference forward.cu:251-369).
//
// B200 design:
//  * the tile's list is a contiguous span of 48-byte records (packed by pack_kernel), so each batch of 128 splats
//    is staged into shared memory with ONE cp.async.bulk (TMA) issued by one thread and tracked by an mbarrier,
//    3 stages deep — no per-thread gathers, no register staging (variant 1; variant 0 uses plain cooperative
//    float4 loads and exists as a cross-check);
//  * each warp owns an 8x4 pixel block and first culls the staged batch against that block (one splat per lane,
//    exact concave-maximum test, see render_common.cuh) and compacts the survivors with ballot/popc, so the blend
//    loop only visits splats that can reach alpha >= 1/255 somewhere in the warp's pixels;
//  * warps stop individually once their 32 pixels are saturated; the CTA stops when all 8 warps have.
// The per-pair arithmetic (power, exp, alpha, T test, accumulation) is spelled with explicit-rounding intrinsics in
// exactly the order of the reference build so final_T, n_contrib and the colours are reproduced bit for bit.
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace rr {

namespace {

template <int MODE>
__global__ void __launch_bounds__(RENDER_THREADS) render_fwd_kernel(RenderArgs a)
{
	constexpr bool USE_TMA = StageMode<MODE>::tma;
	__shared__ __align__(128) float4 s_rec[STAGES][BATCH * 3];
	__shared__ __align__(8) uint64_t s_full[STAGES];
	__shared__ uint8_t s_list[RENDER_WARPS][BATCH];
	__shared__ __align__(4) uint8_t s_used[RENDER_WARPS][BATCH];      // 1 if the warp blended the entry into at least one of its pixels

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	int wx, wy, lx, ly;
	warp_footprint(warp, lane, wx, wy, lx, ly);
	const int pix_x = blockIdx.x * TILE_X + wx + lx, pix_y = blockIdx.y * TILE_Y + wy + ly;
	const bool inside = pix_x < a.W && pix_y < a.H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;
	// rectangle of the warp's pixel centres
	const float X0 = (float)(blockIdx.x * TILE_X + wx), X1 = X0 + 7.0f;
	const float Y0 = (float)(blockIdx.y * TILE_Y + wy), Y1 = Y0 + 3.0f;

	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const int nbatches = (n + BATCH - 1) / BATCH;
	const float4* __restrict__ src = reinterpret_cast<const float4*>(a.packed) + (size_t)range.x * 3;
	const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
	const uint32_t* __restrict__ ids = a.point_list + range.x;

	if (tid == 0 && USE_TMA) {
		for (int s = 0; s < STAGES; s++) mbar_init(&s_full[s], 1);
		mbar_fence_init();
	}
	// this warp's blend-flag words (kernels.cuh): bit j % 32 of word j / 32 = "list entry j was blended into one of my pixels"
	uint32_t* __restrict__ wflags = a.warp_flags + 8ull * ((size_t)(range.x >> 5) + (size_t)tile) + (size_t)warp * ((n + 31) >> 5);
	__syncthreads();
	if (USE_TMA && warp == 0) {
		for (int b = 0; b < STAGES && b < nbatches; b++)
			stage_fill_async<MODE>(&s_rec[b][0], &s_full[b], src, grec, ids, b * BATCH, min(BATCH, n - b * BATCH), lane);
	}

	float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dp = 0.0f;
	uint32_t last_contributor = 0;
	bool done = !inside;
	int b_exit = nbatches;   // first batch index whose stage was never consumed

	for (int b = 0; b < nbatches; b++) {
		const int s = b % STAGES;
		const int cnt = min(BATCH, n - b * BATCH);
		// this warp's "used" row for the batch starts from zero (one 4-byte store per lane covers the 128 entries)
		reinterpret_cast<uint32_t*>(s_used[warp])[lane] = 0u;
		__syncwarp();
		if (USE_TMA) {
			mbar_wait(&s_full[s], (uint32_t)(b / STAGES) & 1u);
		} else {
			__syncthreads();
			stage_fill_sync<MODE>(&s_rec[s][0], src, grec, ids, b * BATCH, cnt, tid);
			__syncthreads();
		}
		const float4* __restrict__ rec = s_rec[s];
		const bool warp_done = __all_sync(FULL, done);
		if (!warp_done) {
			// stage 1: cheap AABB cull against this warp's 8x4 pixels, one splat per lane, survivors compacted
			int n1 = 0;
			for (int c = 0; c < cnt; c += 32) {
				const int e = c + lane;
				bool pass = false;
				if (e < cnt) {
					const float4 ra = rec[e * 3], rc = rec[e * 3 + 2];
					pass = aabb_may_contribute(ra.x, ra.y, rc.z, rc.w, X0, X1, Y0, Y1);
				}
				const unsigned m = __ballot_sync(FULL, pass);
				if (pass) s_list[warp][n1 + __popc(m & lanemask_lt())] = (uint8_t)e;
				n1 += __popc(m);
			}
			__syncwarp();
			// stage 2: exact concave-maximum test on the (dense) survivors, compacted in place
			int nsurv = 0;
			for (int c = 0; c < n1; c += 32) {
				const int k = c + lane;
				bool pass = false;
				int e = 0;
				if (k < n1) {
					e = s_list[warp][k];
					const float4 ra = rec[e * 3], rb = rec[e * 3 + 1];
					pass = rect_may_contribute(ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, X0, X1, Y0, Y1);
				}
				const unsigned m = __ballot_sync(FULL, pass);     // all lanes have consumed their s_list[k] by now
				if (pass) s_list[warp][nsurv + __popc(m & lanemask_lt())] = (uint8_t)e;
				nsurv += __popc(m);
			}
			__syncwarp();
			// blend the survivors, 32 at a time; each lane keeps one "I blended survivor kk" bit per chunk, OR-reduced once per
			// chunk with redux.sync, and the saturation vote is taken every 8 survivors (a saturated pixel skips the math anyway)
			for (int kc = 0; kc < nsurv; kc += 32) {
				const int kn = min(32, nsurv - kc);
				uint32_t usedbits = 0;
				bool stop = false;
				for (int kk = 0; kk < kn; kk++) {
					const int e = s_list[warp][kc + kk];
					const float4 ra = rec[e * 3], rb = rec[e * 3 + 1], rc = rec[e * 3 + 2];
					if (!done) {
						const float dx = __fsub_rn(ra.x, pxf), dy = __fsub_rn(ra.y, pyf);
						const float q = __fmaf_rn(dx, __fmul_rn(dx, ra.z), __fmul_rn(dy, __fmul_rn(dy, rb.x)));
						const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, ra.w)));
						if (!(power > 0.0f)) {
							const float alpha = fminf(0.99f, __fmul_rn(rb.y, expf(power)));
							if (!(alpha < 1.0f / 255.0f)) {
								const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
								if (test_T < 0.0001f) {
									done = true;
								} else {
									C0 = __fmaf_rn(T, __fmul_rn(alpha, rb.z), C0);
									C1 = __fmaf_rn(T, __fmul_rn(alpha, rb.w), C1);
									C2 = __fmaf_rn(T, __fmul_rn(alpha, rc.x), C2);
									Dp = __fmaf_rn(T, __fmul_rn(alpha, rc.y), Dp);
									T = test_T;
									last_contributor = (uint32_t)(b * BATCH + e + 1);
									usedbits |= 1u << kk;
								}
							}
						}
					}
					if ((kk & 7) == 7 && __all_sync(FULL, done)) { stop = true; break; }
				}
				const unsigned wbits = __reduce_or_sync(FULL, usedbits);
				if ((wbits >> lane) & 1u) s_used[warp][s_list[warp][kc + lane]] = 1;
				if (stop || __all_sync(FULL, done)) break;
			}
			// the warp's flag words of this batch (4 words = 128 entries), one 16-byte store; the backward walks exactly these
			__syncwarp();
			uint32_t myword = 0;
			#pragma unroll
			for (int c = 0; c < BATCH / 32; c++) {
				const uint32_t wd = __ballot_sync(FULL, s_used[warp][c * 32 + lane] != 0);
				if (lane == c) myword = wd;
			}
			if (lane < BATCH / 32 && lane * 32 < cnt) wflags[b * (BATCH / 32) + lane] = myword;
		}
		// every warp is past this stage: it can be refilled; stop early when all 256 pixels are saturated
		const int all_done = __syncthreads_and(done);
		if (all_done) { b_exit = b + 1; break; }
		if (USE_TMA && warp == 0 && b + STAGES < nbatches) {
			const int nb = b + STAGES;
			stage_fill_async<MODE>(&s_rec[s][0], &s_full[s], src, grec, ids, nb * BATCH, min(BATCH, n - nb * BATCH), lane);
		}
	}

	if (USE_TMA && tid == 0) {
		// bulk copies already in flight for later stages must land before the CTA (and its shared memory) retires
		for (int bb = b_exit; bb < nbatches && bb < b_exit + STAGES - 1; bb++) mbar_wait(&s_full[bb % STAGES], (uint32_t)(bb / STAGES) & 1u);
	}

	if (inside) {
		const size_t pid = (size_t)a.W * pix_y + pix_x, HW = (size_t)a.H * a.W;
		a.final_T[pid] = T;
		a.n_contrib[pid] = last_contributor;
		a.out_color[pid] = __fmaf_rn(T, a.bg[0], C0);
		a.out_color[HW + pid] = __fmaf_rn(T, a.bg[1], C1);
		a.out_color[2 * HW + pid] = __fmaf_rn(T, a.bg[2], C2);
		a.out_depth[pid] = Dp;
	}
}

}  // namespace

int launch_render_fwd(cudaStream_t stream, const RenderArgs& a)
{
	const dim3 grid(a.grid_x, a.grid_y);
	switch (a.variant) {
	case MODE_PACKED_LDG: count_launch(), render_fwd_kernel<MODE_PACKED_LDG><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	case MODE_PACKED_TMA: count_launch(), render_fwd_kernel<MODE_PACKED_TMA><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	case MODE_GATHER_LDG: count_launch(), render_fwd_kernel<MODE_GATHER_LDG><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	default: count_launch(), render_fwd_kernel<MODE_GATHER_TMA><<<grid, RENDER_THREADS, 0, stream>>>(a); break;
	}
	return check_launch("render_fwd", stream, false);
}

}  // namespace rr
