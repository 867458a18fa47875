// Forward tile renderer: front-to-back alpha blending of each 16x16 tile's depth-sorted splat list.
// Replaces renderCUDA<3> forward (reference forward.cu:251-369).
//
// B200 design (one CTA per tile, 8 consumer warps + 1 producer warp, no __syncthreads after the prologue):
//  * the PRODUCER warp stages the tile's list into a ring of FWD_STAGES shared-memory stages of 128 records (48 B
//    each), gathered asynchronously straight from the per-Gaussian record array (L2 resident) through point_list, 4
//    records per lane and batch, all landing on the stage's `full` mbarrier (render_common.cuh: three 16-byte cp.async
//    per record with cp.async.mbarrier.arrive.noinc; the 48-byte TMA bulk copy per record is the cross-check build's mode).
//    A stage is refilled when every consumer warp has released it: one arrival per warp and batch on
//    the stage's `empty` mbarrier, on which the producer sleeps in hardware (an earlier version polled per-warp
//    progress words every 256 ns: ~9 % of the kernel's warp instructions, all on the producer's scheduler);
//  * each CONSUMER warp owns an 8x4 pixel block and runs through the ring at its own pace: it culls a staged batch
//    against its block (one splat per lane: AABB, then the exact concave-maximum test of render_common.cuh), compacts
//    the survivors with ballot/popc, and blends only those — two survivors per iteration, sharing the packed steps
//    of the bit-exact exponential (exp2x) — so warps that own easy pixels do not wait for the others;
//  * a warp whose 32 pixels are saturated stops working but keeps the handshake going as a "ghost" (wait full, arrive
//    empty: a few instructions per batch, asleep otherwise; its registers belong to the CTA until the CTA retires
//    anyway); when all 8 warps are ghosts the producer stops filling, completes the next `full` phase without data
//    and raises s_stop, on which the ghosts leave;
//  * each warp leaves one bit per list entry ("blended into one of my pixels", kernels.cuh) for the backward.
// The per-pair arithmetic (power, exp, alpha, T test, accumulation) is spelled with explicit-rounding intrinsics in
// exactly the order of the reference build so final_T, n_contrib and the colours are reproduced bit for bit.
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace rr {

namespace {

#ifndef RR_FWD_STAGES
#define RR_FWD_STAGES 4
#endif
constexpr int FWD_STAGES = RR_FWD_STAGES;
constexpr int FWD_THREADS = RENDER_THREADS + 32;      // 8 consumer warps + the producer warp

#ifndef FWD_MIN_CTAS
#define FWD_MIN_CTAS 4      // CTAs per SM the register budget is held to (56 registers at 288 threads)
#endif
template <int MODE>
__global__ void __launch_bounds__(FWD_THREADS, FWD_MIN_CTAS) render_fwd_kernel(RenderArgs a)
{
	__shared__ __align__(128) float4 s_rec[FWD_STAGES][BATCH * 3];
	__shared__ __align__(8) uint64_t s_full[FWD_STAGES];
	__shared__ __align__(8) uint64_t s_empty[FWD_STAGES];               // one arrival per consumer warp and batch: "I am past this stage"
	__shared__ int s_gone;                                              // consumer warps whose 32 pixels are all saturated (or outside the image)
	__shared__ int s_stop;                                              // the batch at which the producer stopped filling early (all warps gone); -1: it has not
	__shared__ uint8_t s_list[RENDER_WARPS][BATCH];
	__shared__ __align__(4) uint8_t s_used[RENDER_WARPS][BATCH];      // 1 if the warp blended the entry into at least one of its pixels

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const int nbatches = (n + BATCH - 1) / BATCH;

	if (tid == 0) {
		for (int s = 0; s < FWD_STAGES; s++) { mbar_init(&s_full[s], StageMode<MODE>::full_count); mbar_init(&s_empty[s], RENDER_WARPS); }
		s_gone = 0; s_stop = -1;
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == RENDER_WARPS) {
		// ------------------------------------------------ producer ------------------------------------------------
		const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
		const uint32_t* __restrict__ ids = a.point_list + range.x;
		for (int b = 0; b < nbatches; b++) {
			const int s = b % FWD_STAGES;
			if (b >= FWD_STAGES) {
				// sleep (in hardware, no polling) until every consumer warp is past batch b - FWD_STAGES
				mbar_wait(&s_empty[s], (uint32_t)(b / FWD_STAGES - 1) & 1u);
				if (*reinterpret_cast<volatile int*>(&s_gone) == RENDER_WARPS) {
					// every pixel of the tile is saturated: stop filling.  The warps are all ghosts by now and will reach
					// batch b (the earlier ones are filled); complete its `full` phase without data so that they wake, see
					// s_stop and leave.
					if (lane == 0) *reinterpret_cast<volatile int*>(&s_stop) = b;
					__syncwarp();
					if (lane < (int)StageMode<MODE>::full_count) mbar_arrive(&s_full[s]);
					return;
				}
				__threadfence_block();       // the consumers' reads of the stage are ordered before the refill below
			}
			stage_fill_async<MODE>(&s_rec[s][0], &s_full[s], grec, ids, b * BATCH, min(BATCH, n - b * BATCH), lane);
		}
		// every consumer warp waits on every filled batch before it leaves, so no fill is in flight when the CTA retires
		return;
	}

	// ---------------------------------------------------- consumers ----------------------------------------------------
	int wx, wy, lx, ly;
	warp_footprint(warp, lane, wx, wy, lx, ly);
	const int pix_x = blockIdx.x * TILE_X + wx + lx, pix_y = blockIdx.y * TILE_Y + wy + ly;
	const bool inside = pix_x < a.W && pix_y < a.H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;
	// rectangle of the warp's pixel centres
	const float X0 = (float)(blockIdx.x * TILE_X + wx), X1 = X0 + 7.0f;
	const float Y0 = (float)(blockIdx.y * TILE_Y + wy), Y1 = Y0 + 3.0f;
	// this warp's blend-flag words (kernels.cuh): bit j % 32 of word j / 32 = "list entry j was blended into one of my pixels"
	uint32_t* __restrict__ wflags = a.warp_flags + 8ull * ((size_t)(range.x >> 5) + (size_t)tile) + (size_t)warp * ((n + 31) >> 5);
	uint8_t* __restrict__ list = s_list[warp];
	uint8_t* __restrict__ used = s_used[warp];

	float T = 1.0f;
	p2 C01 = pk(0.0f, 0.0f), C2D = C01;      // (C0, C1) and (C2, depth) accumulators
	const p2 pxy2 = pk(pxf, pyf);
	uint32_t last_contributor = 0;
	bool done = !inside;
	bool warp_done = __all_sync(FULL, done);
	if (warp_done && lane == 0) atomicAdd(&s_gone, 1);
	XC(unsigned xc_batches = 0; unsigned xc_entries = 0; unsigned xc_aabb = 0; unsigned xc_blend = 0; unsigned xc_live = 0; unsigned xc_bl = 0;)

	// Every consumer warp walks every filled batch: wait for the stage, work on it (or not: a warp whose pixels are all
	// saturated is a "ghost" that only keeps the handshake going, asleep in the barrier wait), release the stage.
	for (int b = 0; b < nbatches; b++) {
		const int s = b % FWD_STAGES;
		mbar_wait(&s_full[s], (uint32_t)(b / FWD_STAGES) & 1u);
		if (warp_done) {
			// the producer stopped at this batch: nothing was filled.  (Only at this batch: a ghost leaves after it has
			// waited for every fill that was issued, so none is in flight when the CTA retires.)
			if (*reinterpret_cast<volatile int*>(&s_stop) == b) break;
		} else {
			const int cnt = min(BATCH, n - b * BATCH);
			const float4* __restrict__ rec = s_rec[s];
			// this warp's "used" row for the batch starts from zero (one 4-byte store per lane covers the 128 entries)
			if (lane < BATCH / 4) reinterpret_cast<uint32_t*>(used)[lane] = 0u;
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
				if (pass) list[n1 + __popc(m & lanemask_lt())] = (uint8_t)e;
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
					e = list[k];
					const float4 ra = rec[e * 3], rb = rec[e * 3 + 1];
					pass = rect_may_contribute(ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, X0, X1, Y0, Y1);
				}
				const unsigned m = __ballot_sync(FULL, pass);     // all lanes have consumed their list[k] by now
				if (pass) list[nsurv + __popc(m & lanemask_lt())] = (uint8_t)e;
				nsurv += __popc(m);
			}
			__syncwarp();
			XC(xc_batches++; xc_entries += cnt; xc_aabb += n1;)
			// blend the survivors, 32 at a time.  Each lane keeps one "I blended survivor kk" bit per chunk (OR-reduced once
			// per chunk with redux.sync); the saturation vote is taken every 8 survivors (a saturated pixel is predicated
			// off anyway).  The body is branch-free: done / culled lanes are masked by predicates.
			const uint32_t pos1 = (uint32_t)(b * BATCH) + 1u;
			for (int kc = 0; kc < nsurv && !warp_done; kc += 32) {
				const int kn = min(32, nsurv - kc);
				uint32_t usedbits = 0, bit = 1u;
				for (int k8 = 0; k8 < kn; k8 += 8) {
					const int n8 = min(8, kn - k8);
					auto pair_step = [&](int li, bool two, uint32_t ubit) {
						const int e0 = list[li], e1 = two ? list[li + 1] : e0;
						const float4 ra0 = rec[e0 * 3], rb0 = rec[e0 * 3 + 1], ra1 = rec[e1 * 3], rb1 = rec[e1 * 3 + 1];
						float dx0, dy0, dx1, dy1;
						up(sub2(pk(ra0.x, ra0.y), pxy2), dx0, dy0);
						up(sub2(pk(ra1.x, ra1.y), pxy2), dx1, dy1);
						const float q0 = __fmaf_rn(dx0, __fmul_rn(dx0, ra0.z), __fmul_rn(dy0, __fmul_rn(dy0, rb0.x)));
						const float q1 = __fmaf_rn(dx1, __fmul_rn(dx1, ra1.z), __fmul_rn(dy1, __fmul_rn(dy1, rb1.x)));
						const float power0 = __fmaf_rn(q0, -0.5f, -__fmul_rn(dy0, __fmul_rn(dx0, ra0.w)));
						const float power1 = __fmaf_rn(q1, -0.5f, -__fmul_rn(dy1, __fmul_rn(dx1, ra1.w)));
						float G0, G1;
						exp2x(power0, power1, G0, G1);
						const float alpha0 = fminf(0.99f, __fmul_rn(rb0.y, G0)), alpha1 = fminf(0.99f, __fmul_rn(rb1.y, G1));
						// forward.cu:330-351 for one list entry, branch free: a pixel that does not blend the entry runs the same
						// instructions with alpha = 0, which leaves the accumulators unchanged exactly (fma(T, +-0, C) = C) and keeps
						// T; C01 = (C0, C1), C2D = (C2, depth): same roundings as the reference, packed
						auto blend = [&](int e, float power, float alpha, const float4& rb, uint32_t b1) {
							const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
							const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);     // forward.cu:330-340
							const bool sat = ok && (test_T < 0.0001f);                               // forward.cu:341-346
							done = done || sat;
							const bool bl = ok && !sat;
							XC(xc_live += ok; xc_bl += bl;)
							const float4 rc = rec[e * 3 + 2];
							const p2 a2 = pk1(bl ? alpha : 0.0f), T2 = pk1(T);
							C01 = fma2(T2, mul2(a2, pk(rb.z, rb.w)), C01);
							C2D = fma2(T2, mul2(a2, pk(rc.x, rc.y)), C2D);
							T = bl ? test_T : T;
							usedbits |= bl ? b1 : 0u;
						};
						blend(e0, power0, alpha0, rb0, ubit);
						if (two) blend(e1, power1, alpha1, rb1, ubit << 1);
						XC(xc_blend += two ? 2 : 1;)
					};
					// two survivors per step: everything up to alpha is independent of the blend state, so the two exponentials
					// share the packed steps of exp2x; the blend itself stays in list order.  A full group of 8 runs unguarded, so
					// the four steps can be scheduled across each other.
					if (n8 == 8) {
						#pragma unroll
						for (int j = 0; j < 8; j += 2) pair_step(kc + k8 + j, true, bit << j);
						bit <<= 8;
					} else {
						for (int j = 0; j < n8; j += 2, bit <<= 2) pair_step(kc + k8 + j, j + 1 < n8, bit);
					}
					if (__all_sync(FULL, done)) { warp_done = true; break; }
				}
				// n_contrib (forward.cu:353-355): the last entry this pixel blended is the highest bit of usedbits — looked up
				// once per chunk instead of a select per survivor
				if (usedbits) last_contributor = pos1 + (uint32_t)list[kc + 31 - __clz(usedbits)];
				const unsigned wbits = __reduce_or_sync(FULL, usedbits);
				if ((wbits >> lane) & 1u) used[list[kc + lane]] = 1;
			}
			// the warp's flag words of this batch (4 words = 128 entries), one 16-byte store; the backward walks exactly these
			__syncwarp();
			uint32_t myword = 0;
			#pragma unroll
			for (int c = 0; c < BATCH / 32; c++) {
				const uint32_t wd = __ballot_sync(FULL, used[c * 32 + lane] != 0);
				if (lane == c) myword = wd;
			}
			if (lane < BATCH / 32 && lane * 32 < cnt) wflags[b * (BATCH / 32) + lane] = myword;
			if (warp_done && lane == 0) atomicAdd(&s_gone, 1);         // before the release below, so the producer sees it
		}
		// this warp is past the stage (mbarrier.arrive is a release: the stage reads above come first)
		__syncwarp();
		if (lane == 0) mbar_arrive(&s_empty[s]);
	}

	XC({
		const unsigned live = __reduce_add_sync(FULL, xc_live), bl = __reduce_add_sync(FULL, xc_bl);
		if (lane == 0) { xc_add(a.xc, 0, xc_batches); xc_add(a.xc, 1, xc_entries); xc_add(a.xc, 2, xc_aabb); xc_add(a.xc, 3, xc_blend); xc_add(a.xc, 4, live); xc_add(a.xc, 5, bl); }
	})
	if (inside) {
		const size_t pid = (size_t)a.W * pix_y + pix_x, HW = (size_t)a.H * a.W;
		a.final_T[pid] = T;
		a.n_contrib[pid] = last_contributor;
		float C0, C1, C2, Dp;
		up(C01, C0, C1); up(C2D, C2, Dp);
		a.out_color[pid] = __fmaf_rn(T, a.bg[0], C0);
		a.out_color[HW + pid] = __fmaf_rn(T, a.bg[1], C1);
		a.out_color[2 * HW + pid] = __fmaf_rn(T, a.bg[2], C2);
		a.out_depth[pid] = Dp;
	}
}

// exp2x (render_common.cuh) against libdevice expf over every float bit pattern in [first, first + count): counts lanes that differ
__global__ void __launch_bounds__(256) exp_selftest_kernel(uint32_t first, uint32_t count, unsigned long long* __restrict__ mismatches)
{
	unsigned long long bad = 0;
	for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < count; i += gridDim.x * 256u) {
		const float xa = __uint_as_float(first + i), xb = __uint_as_float(first + (count - 1u - i));
		float ea, eb;
		exp2x(xa, xb, ea, eb);
		bad += (__float_as_uint(ea) != __float_as_uint(expf(xa))) + (__float_as_uint(eb) != __float_as_uint(expf(xb)));
	}
	if (bad) atomicAdd(mismatches, bad);
}

}  // namespace

int launch_exp_selftest(cudaStream_t stream, uint32_t first_bits, uint32_t count, unsigned long long* mismatches)
{
	count_launch(), exp_selftest_kernel<<<148 * 8, 256, 0, stream>>>(first_bits, count, mismatches);
	return check_launch("exp_selftest", stream, false);
}

int launch_render_fwd(cudaStream_t stream, const RenderArgs& a)
{
	const dim3 grid(a.grid_x, a.grid_y);
#ifdef RR_CROSSCHECK
	if (crosscheck_staging() == STAGE_TMA) { count_launch(), render_fwd_kernel<STAGE_TMA><<<grid, FWD_THREADS, 0, stream>>>(a); return check_launch("render_fwd", stream, false); }
#endif
	count_launch(), render_fwd_kernel<STAGE_CP_ASYNC><<<grid, FWD_THREADS, 0, stream>>>(a);
	return check_launch("render_fwd", stream, false);
}

}  // namespace rr
