// Forward tile renderer: front-to-back alpha blending of each 16x16 tile's depth-sorted splat list.
// Replaces renderCUDA<3> forward (reference forward.cu:251-369).
//
// B200 design (one CTA per tile, 4 consumer warps + 1 producer warp, no __syncthreads after the prologue):
//  * the PRODUCER warp stages the tile's list into a ring of FWD_STAGES shared-memory stages of 128 records (48 B
//    each), gathered asynchronously straight from the per-Gaussian record array (L2 resident) through point_list, 4
//    records per lane and batch, all landing on the stage's `full` mbarrier (render_common.cuh: three 16-byte cp.async
//    per record with cp.async.mbarrier.arrive.noinc; the 48-byte TMA bulk copy per record is the cross-check build's mode).
//    A stage is refilled when every consumer warp has released it: one arrival per warp and batch on
//    the stage's `empty` mbarrier, on which the producer sleeps in hardware (an earlier version polled per-warp
//    progress words every 256 ns: ~9 % of the kernel's warp instructions, all on the producer's scheduler);
//  * each CONSUMER warp owns an 8x8 pixel block, TWO pixels per lane (rows y and y + 4 of one column), and runs through the
//    ring at its own pace: it culls a staged batch against its block (one splat per lane: AABB, then the exact
//    concave-maximum test of render_common.cuh), compacts the survivors with ballot/popc, and blends only those.  The two
//    pixels of a lane are evaluated with PACKED fp32 arithmetic (add/mul/fma.rn.f32x2 -> FADD2 / FMUL2 / FFMA2, one issue
//    slot for both; the two exponentials share the packed steps of the bit-exact exp2x), and record loads, list walking and
//    loop control are paid once per 64 pixels: 2.1 M survivor visits of ~55 instructions per view at C3 instead of 3.8 M of
//    ~60 with one pixel per lane (live lanes 51 % vs 56 %, profiles/live_lanes.json), so warps that own easy pixels still
//    do not wait for the others;
//  * a warp whose 64 pixels are saturated stops working but keeps the handshake going as a "ghost" (wait full, arrive
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
constexpr int FWD_WARPS = 4;                          // consumer warps: one 8x8 pixel block each, two pixels per lane
constexpr int FWD_THREADS = (FWD_WARPS + 1) * 32;     // + the producer warp

#ifndef FWD_MIN_CTAS
#define FWD_MIN_CTAS 5      // CTAs per SM the register budget is held to (80 registers at 160 threads)
#endif
template <int MODE>
__global__ void __launch_bounds__(FWD_THREADS, FWD_MIN_CTAS) render_fwd_kernel(RenderArgs a)
{
	__shared__ __align__(128) float4 s_rec[FWD_STAGES][BATCH * 3];
	__shared__ __align__(8) uint64_t s_full[FWD_STAGES];
	__shared__ __align__(8) uint64_t s_empty[FWD_STAGES];               // one arrival per consumer warp and batch: "I am past this stage"
	__shared__ int s_gone;                                              // consumer warps whose 64 pixels are all saturated (or outside the image)
	__shared__ int s_stop;                                              // the batch at which the producer stopped filling early (all warps gone); -1: it has not
	__shared__ uint8_t s_list[FWD_WARPS][BATCH];
	__shared__ __align__(4) uint8_t s_used[FWD_WARPS][BATCH];         // 1 if the warp blended the entry into at least one of its pixels

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	// this CTA's share of the zero fill of the view's backward accumulators (api.cu: GeomLayout::acc): a few 16-byte stores per thread
	if (a.zero) {
		const uint32_t stride = (uint32_t)(a.grid_x * a.grid_y) * FWD_THREADS;
		for (uint32_t i = (uint32_t)tile * FWD_THREADS + tid; i < a.zero_n16; i += stride) a.zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
	}
	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const int nbatches = (n + BATCH - 1) / BATCH;

	if (tid == 0) {
		for (int s = 0; s < FWD_STAGES; s++) { mbar_init(&s_full[s], StageMode<MODE>::full_count); mbar_init(&s_empty[s], FWD_WARPS); }
		s_gone = 0; s_stop = -1;
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == FWD_WARPS) {
		// ------------------------------------------------ producer ------------------------------------------------
		const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
		const uint32_t* __restrict__ ids = a.point_list + range.x;
		for (int b = 0; b < nbatches; b++) {
			const int s = b % FWD_STAGES;
			if (b >= FWD_STAGES) {
				// sleep (in hardware, no polling) until every consumer warp is past batch b - FWD_STAGES
				mbar_wait(&s_empty[s], (uint32_t)(b / FWD_STAGES - 1) & 1u);
				if (*reinterpret_cast<volatile int*>(&s_gone) == FWD_WARPS) {
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
	// warp w owns the 8x8 block at ((w & 1) * 8, (w >> 1) * 8) of the tile (the backward's footprint, render_bwd.cu); lane ->
	// column lx, rows ly (pixel A) and ly + 4 (pixel B)
	const int bx = (warp & 1) * 8, by = (warp >> 1) * 8, lx = lane & 7, ly = lane >> 3;
	const int pix_x = blockIdx.x * TILE_X + bx + lx, pyA = blockIdx.y * TILE_Y + by + ly, pyB = pyA + 4;
	const bool inA = pix_x < a.W && pyA < a.H, inB = pix_x < a.W && pyB < a.H;
	const float pxf = (float)pix_x;
	const p2 py2 = pk((float)pyA, (float)pyB);
	// rectangle of the warp's pixel centres
	const float X0 = (float)(blockIdx.x * TILE_X + bx), X1 = X0 + 7.0f;
	const float Y0 = (float)(blockIdx.y * TILE_Y + by), Y1 = Y0 + 7.0f;
	// this warp's blend-flag words (kernels.cuh): bit j % 32 of word j / 32 = "list entry j was blended into one of my pixels"
	uint32_t* __restrict__ wflags = a.warp_flags + (size_t)FLAG_WARPS * ((size_t)(range.x >> 5) + (size_t)tile) + (size_t)warp * ((n + 31) >> 5);
	uint8_t* __restrict__ list = s_list[warp];
	uint8_t* __restrict__ used = s_used[warp];

	p2 T2 = pk(1.0f, 1.0f);
	p2 C0 = pk(0.0f, 0.0f), C1 = C0, C2 = C0, Dp = C0;      // per-channel accumulators of the lane's two pixels (A, B)
	const p2 ONE2 = pk(1.0f, 1.0f), NHALF2 = pk(-0.5f, -0.5f);
	uint32_t lastA = 0, lastB = 0;
	bool doneA = !inA, doneB = !inB;
	bool warp_done = __all_sync(FULL, doneA && doneB);
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
			// stage 1: cheap AABB cull against this warp's 8x8 pixels, one splat per lane, survivors compacted
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
			// blend the survivors, 32 at a time.  Each lane keeps one "pixel A / B blended survivor kk" bit per chunk (OR-reduced
			// once per chunk with redux.sync); the saturation vote is taken every 8 survivors (a saturated pixel is predicated
			// off anyway).  The body is branch-free: done / culled pixels are masked by predicates.
			const uint32_t pos1 = (uint32_t)(b * BATCH) + 1u;
			for (int kc = 0; kc < nsurv && !warp_done; kc += 32) {
				const int kn = min(32, nsurv - kc);
				uint32_t usedA = 0, usedB = 0, bit = 1u;
				for (int k8 = 0; k8 < kn; k8 += 8) {
					const int n8 = min(8, kn - k8);
					// forward.cu:325-351 for one list entry and this lane's two pixels, branch free, with the reference build's
					// roundings (f32x2 operations are IEEE round-to-nearest per element): a pixel that does not blend the entry
					// runs the same instructions with alpha = 0, which leaves its accumulators unchanged exactly
					// (fma(T, +-0, C) = C) and keeps its T
					auto step = [&](int li, uint32_t ubit) {
						const int e = list[li];
						const float4 ra = rec[e * 3], rb = rec[e * 3 + 1], rc = rec[e * 3 + 2];
						const float dx = __fsub_rn(ra.x, pxf);
						const p2 dy = sub2(pk1(ra.y), py2);
						// power = fma(fma(dx, dx*A, dy*(dy*C)), -0.5, -(dy*(dx*B)))
						const p2 qd = fma2(pk1(dx), pk1(__fmul_rn(dx, ra.z)), mul2(dy, mul2(dy, pk1(rb.x))));
						const p2 power = fma2(qd, NHALF2, mul2(dy, pk1(-__fmul_rn(dx, ra.w))));
						float pA, pB, GA, GB;
						up(power, pA, pB);
						exp2x(pA, pB, GA, GB);
						const float aA = fminf(0.99f, __fmul_rn(rb.y, GA)), aB = fminf(0.99f, __fmul_rn(rb.y, GB));
						float tA, tB, TA, TB;
						up(mul2(T2, sub2(ONE2, pk(aA, aB))), tA, tB);                               // test_T = T * (1 - alpha)
						up(T2, TA, TB);
						const bool okA = !doneA && !(pA > 0.0f) && !(aA < 1.0f / 255.0f);           // forward.cu:330-340
						const bool okB = !doneB && !(pB > 0.0f) && !(aB < 1.0f / 255.0f);
						const bool satA = okA && (tA < 0.0001f), satB = okB && (tB < 0.0001f);      // forward.cu:341-346
						doneA = doneA || satA; doneB = doneB || satB;
						const bool blA = okA && !satA, blB = okB && !satB;
						XC(xc_live += (unsigned)okA + (unsigned)okB; xc_bl += (unsigned)blA + (unsigned)blB;)
						const p2 a2 = pk(blA ? aA : 0.0f, blB ? aB : 0.0f);
						C0 = fma2(T2, mul2(a2, pk1(rb.z)), C0);
						C1 = fma2(T2, mul2(a2, pk1(rb.w)), C1);
						C2 = fma2(T2, mul2(a2, pk1(rc.x)), C2);
						Dp = fma2(T2, mul2(a2, pk1(rc.y)), Dp);
						T2 = pk(blA ? tA : TA, blB ? tB : TB);
						usedA |= blA ? ubit : 0u; usedB |= blB ? ubit : 0u;
					};
					// a full group of 8 runs unguarded, so the independent fronts (power, exp, alpha) of consecutive entries
					// can be scheduled across each other; the blend itself stays in list order
					if (n8 == 8) {
						#pragma unroll
						for (int j = 0; j < 8; j++) step(kc + k8 + j, bit << j);
						bit <<= 8;
					} else {
						for (int j = 0; j < n8; j++, bit <<= 1) step(kc + k8 + j, bit);
					}
					XC(xc_blend += n8;)
					if (__all_sync(FULL, doneA && doneB)) { warp_done = true; break; }
				}
				// n_contrib (forward.cu:353-355): the last entry a pixel blended is the highest bit of its used mask — looked up
				// once per chunk instead of a select per survivor
				if (usedA) lastA = pos1 + (uint32_t)list[kc + 31 - __clz(usedA)];
				if (usedB) lastB = pos1 + (uint32_t)list[kc + 31 - __clz(usedB)];
				const unsigned wbits = __reduce_or_sync(FULL, usedA | usedB);
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
	const size_t HW = (size_t)a.H * a.W;
	float TA, TB, c0A, c0B, c1A, c1B, c2A, c2B, dA, dB;
	up(T2, TA, TB); up(C0, c0A, c0B); up(C1, c1A, c1B); up(C2, c2A, c2B); up(Dp, dA, dB);
	const float bg0 = a.bg[0], bg1 = a.bg[1], bg2 = a.bg[2];
	if (inA) {
		const size_t pid = (size_t)a.W * pyA + pix_x;
		a.final_T[pid] = TA; a.n_contrib[pid] = lastA;
		a.out_color[pid] = __fmaf_rn(TA, bg0, c0A); a.out_color[HW + pid] = __fmaf_rn(TA, bg1, c1A); a.out_color[2 * HW + pid] = __fmaf_rn(TA, bg2, c2A);
		a.out_depth[pid] = dA;
	}
	if (inB) {
		const size_t pid = (size_t)a.W * pyB + pix_x;
		a.final_T[pid] = TB; a.n_contrib[pid] = lastB;
		a.out_color[pid] = __fmaf_rn(TB, bg0, c0B); a.out_color[HW + pid] = __fmaf_rn(TB, bg1, c1B); a.out_color[2 * HW + pid] = __fmaf_rn(TB, bg2, c2B);
		a.out_depth[pid] = dB;
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
