// Backward tile renderer, two pixels per lane: each WARP owns an 8x8 pixel block of the tile (lane = column x row
// pair {y, y+4}) and replays, back to front, exactly the list entries its block blended in the forward.
// Replaces renderCUDA<3> backward (reference backward.cu:389-547).  Per-warp visit list from the forward's blend
// flags, asynchronous per-lane gather of the visited 48-byte records, shared-memory transpose reduction, one RED.ADD.F32
// per value and GROUP visits; and, to cut issue slots per pixel:
//  * the two pixels of a lane are processed with PACKED fp32 arithmetic (PTX add/mul/fma.rn.f32x2, SASS FADD2 / FMUL2 /
//    FFMA2): one issue slot for both pixels.  Measured on B200 (tools/micro/f32x2_bench.cu): FFMA2 sustains the same
//    73 TFLOP/s as FFMA with half the issue slots, and these kernels are issue bound with the FMA pipe at ~34 %;
//  * record loads, loop control, the 9 transpose stores and the row reduction are paid once per 64 pixels;
//  * no divergent "is this pixel live" branch: a pixel that did not blend the entry runs the same arithmetic with
//    alpha and G forced to 0, which leaves its T, its accumulated colour and every sum unchanged exactly
//    (T/(1-0) = T; accum' = 0*c + 1*accum), so the pending-update recurrence of backward.cu:493-507 is preserved.
// The pair test (power > 0, alpha < 1/255, entry index < n_contrib) is evaluated with the forward's exact roundings
// (f32x2 operations are IEEE round-to-nearest per element), so the set of live (pixel, entry) pairs is the forward's.
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace rr {

namespace {

constexpr int NV = 9;
constexpr int B2_WARPS = 4;                     // 4 warps x (8x8 pixels) = one 16x16 tile
constexpr int B2_THREADS = B2_WARPS * 32;
#ifndef B2_EXPF_CALL
#define B2_EXPF_CALL 0
#endif
constexpr bool EXACT_EXP_CALL = B2_EXPF_CALL != 0;   // 1: call libdevice expf per pixel (cross-check of exp2x)
#ifndef B2_MIN_CTAS
#define B2_MIN_CTAS 5
#endif
constexpr int LIST_WORDS = 8;                   // flag words (of 32 list entries) turned into a visit list per round
constexpr int LIST_CAP = LIST_WORDS * 32;
constexpr int CHUNK = 16;                       // visits staged per buffer
constexpr int GROUP = 3;                        // visits reduced together through shared memory
constexpr int RED_ROWS = GROUP * NV;

__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// STAGE: STAGE_CP_ASYNC (per-lane cp.async, the shipped mode) or STAGE_TMA (per-lane TMA bulk copies, cross-check build); render_common.cuh
template <int STAGE>
__global__ void __launch_bounds__(B2_THREADS, B2_MIN_CTAS) render_bwd_kernel(RenderArgs a)
{
	constexpr bool USE_TMA = STAGE == STAGE_TMA;
	__shared__ __align__(128) float4 s_rec[B2_WARPS][2][CHUNK * 3];    // per warp, double buffered: CHUNK records of 48 B
	__shared__ __align__(16) float s_red[B2_WARPS][RED_ROWS * 32];     // transpose rows [visit in group][value][lane]
	__shared__ uint32_t s_gid[B2_WARPS][2][CHUNK];                      // Gaussian id of each staged visit
	__shared__ uint16_t s_pos[B2_WARPS][LIST_CAP];                      // visit list of the round (positions relative to the round's first entry, descending)
	__shared__ __align__(8) uint64_t s_bar[B2_WARPS][2];

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	// 8x8 block of this warp; lane -> column lx, rows ly (pixel A) and ly + 4 (pixel B)
	const int bx = (warp & 1) * 8, by = (warp >> 1) * 8, lx = lane & 7, ly = lane >> 3;
	const int pix_x = blockIdx.x * TILE_X + bx + lx, pyA = blockIdx.y * TILE_Y + by + ly, pyB = pyA + 4;
	const bool inA = pix_x < a.W && pyA < a.H, inB = pix_x < a.W && pyB < a.H;
	const float pxf = (float)pix_x;
	const p2 py2 = pk((float)pyA, (float)pyB);

	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const size_t HW = (size_t)a.H * a.W;
	const size_t pidA = (size_t)a.W * pyA + pix_x, pidB = (size_t)a.W * pyB + pix_x;
	const uint32_t lastA = inA ? a.n_contrib[pidA] : 0u, lastB = inB ? a.n_contrib[pidB] : 0u;
	const uint32_t wmax = __reduce_max_sync(FULL, max(lastA, lastB));
	if (wmax == 0 || n == 0) return;             // warp-uniform: nothing was blended into this warp's pixels

	const float TfA = inA ? a.final_T[pidA] : 0.0f, TfB = inB ? a.final_T[pidB] : 0.0f;
	p2 T2 = pk(TfA, TfB);
	float dA0 = 0, dA1 = 0, dA2 = 0, dB0 = 0, dB1 = 0, dB2 = 0;
	if (inA) { dA0 = a.dL_dpix[pidA]; dA1 = a.dL_dpix[HW + pidA]; dA2 = a.dL_dpix[2 * HW + pidA]; }
	if (inB) { dB0 = a.dL_dpix[pidB]; dB1 = a.dL_dpix[HW + pidB]; dB2 = a.dL_dpix[2 * HW + pidB]; }
	const p2 dLp0 = pk(dA0, dB0), dLp1 = pk(dA1, dB1), dLp2 = pk(dA2, dB2);
	const float bg0 = a.bg[0], bg1 = a.bg[1], bg2 = a.bg[2];
	float bdA = 0, bdB = 0;
	bdA += bg0 * dA0; bdA += bg1 * dA1; bdA += bg2 * dA2;
	bdB += bg0 * dB0; bdB += bg1 * dB1; bdB += bg2 * dB2;
	const p2 neg_Tf_bg = pk(-TfA * bdA, -TfB * bdB);
	p2 ar0 = pk(0.f, 0.f), ar1 = ar0, ar2 = ar0, la2 = ar0;     // accumulated colour behind the pixel, alpha of the previous visit
	float lc0 = 0, lc1 = 0, lc2 = 0;                               // colour of the previous visit (same for every pixel)
	const p2 ONE2 = pk(1.f, 1.f), NHALF2 = pk(-0.5f, -0.5f);

	uint64_t* bar = s_bar[warp];
	if (lane == 0) { mbar_init(&bar[0], StageMode<STAGE>::full_count); mbar_init(&bar[1], StageMode<STAGE>::full_count); mbar_fence_init(); }
	__syncwarp();
	// shared-memory transpose: lane r < 9 * GROUP owns row r = visit * 9 + value of the group
	const int row_visit = lane / NV;
	float* __restrict__ acc_row = a.grad_acc + (lane - row_visit * NV);
	float* __restrict__ red = s_red[warp];
	const float4* __restrict__ red_row = reinterpret_cast<const float4*>(red) + lane * 8;

	// the forward's flag words of this 8x8 block (same warp footprint in both kernels, kernels.cuh)
	const int ngroups = (n + 31) >> 5;
	const uint32_t* __restrict__ wf = a.warp_flags + (size_t)FLAG_WARPS * ((size_t)(range.x >> 5) + (size_t)tile) + (size_t)warp * ngroups;
	const uint32_t* __restrict__ ids = a.point_list + range.x;
	const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
	// The forward warp of this block wrote its flag words only for the batches it worked on, i.e. up to the batch of its last
	// contributor (the blob is not zero-filled): words are read only below ceil(wmax / 32).
	const int nwords = (int)((wmax + 31u) >> 5);
	XC(unsigned xc_visits = 0; unsigned xc_live = 0;)
	uint32_t q = 0;                               // chunks staged so far by this warp: buffer = q & 1, mbarrier parity = (q >> 1) & 1

	for (int wbase = ((nwords - 1) / LIST_WORDS) * LIST_WORDS; wbase >= 0; wbase -= LIST_WORDS) {
		// ---- visit list of this round: set bits of words [wbase, wbase + LIST_WORDS), highest list entry first ----
		const int wi = wbase + lane;
		uint32_t word = (lane < LIST_WORDS && wi < nwords) ? __ldg(wf + wi) : 0u;
		const int cnt = __popc(word);
		int inc = cnt;                             // inclusive sum over lanes >= this one
		#pragma unroll
		for (int o = 1; o < LIST_WORDS; o <<= 1) { const int t = __shfl_down_sync(FULL, inc, o); if (lane + o < 32) inc += t; }
		const int total = __shfl_sync(FULL, inc, 0);
		if (total == 0) continue;
		int at = inc - cnt;
		while (word) { const int bit = 31 - __clz(word); word ^= 1u << bit; s_pos[warp][at++] = (uint16_t)(lane * 32 + bit); }
		__syncwarp();
		const uint32_t ebase = (uint32_t)wbase * 32u;
		const int nch = (total + CHUNK - 1) / CHUNK;

		auto load_gid = [&](int k) -> uint32_t {
			const int i = k * CHUNK + lane;
			return (lane < CHUNK && i < total) ? __ldg(ids + ebase + s_pos[warp][i]) : 0u;
		};
		// stage chunk k of the round into buffer qq & 1
		auto issue = [&](int k, uint32_t qq, uint32_t gid) {
			const int buf = (int)(qq & 1u);
			const int cc = min(CHUNK, total - k * CHUNK);
			if (lane < cc) s_gid[warp][buf][lane] = gid;
			if (USE_TMA) {
				if (lane == 0) mbar_arrive_expect_tx(&bar[buf], (uint32_t)cc * (uint32_t)sizeof(SplatRec));
				__syncwarp();
				if (lane < cc) bulk_g2s(&s_rec[warp][buf][lane * 3], grec + (size_t)gid * 3, (uint32_t)sizeof(SplatRec), &bar[buf]);
			} else {
				if (lane < cc) {
					const float4* __restrict__ src = grec + (size_t)gid * 3;
					float4* dst = &s_rec[warp][buf][lane * 3];
					cp_async16(dst, src); cp_async16(dst + 1, src + 1); cp_async16(dst + 2, src + 2);
				}
				cp_async_arrive_noinc(&bar[buf]);
			}
		};

		uint32_t gid_next = load_gid(0);
		issue(0, q, gid_next);
		gid_next = nch > 1 ? load_gid(1) : 0u;
		for (int k = 0; k < nch; k++, q++) {
			if (k + 1 < nch) {
				issue(k + 1, q + 1, gid_next);
				if (k + 2 < nch) gid_next = load_gid(k + 2);
			}
			const int buf = (int)(q & 1u);
			mbar_wait(&bar[buf], (q >> 1) & 1u);
			__syncwarp();
			const float4* __restrict__ rec = s_rec[warp][buf];
			const int cc = min(CHUNK, total - k * CHUNK);
			const uint16_t* __restrict__ cpos = s_pos[warp] + k * CHUNK;
			XC(xc_visits += cc;)

			// the 9 sums of one visit over this lane's two pixels (reference backward.cu:481-544, moments of u = G * dL_dalpha
			// instead of the conic / mean products), stored as one transpose column
			auto visit = [&](int kk, float* __restrict__ col) {
				const uint32_t pos = ebase + cpos[kk];
				const float4 ra = rec[kk * 3], rb = rec[kk * 3 + 1];
				const float c2 = rec[kk * 3 + 2].x;
				const float dx = __fsub_rn(ra.x, pxf);
				const p2 dy = sub2(pk1(ra.y), py2);
				// power = fma(fma(dx, dx*A, dy*(dy*C)), -0.5, -(dy*(dx*B))), every step rounded as in the forward
				const p2 qd = fma2(pk1(dx), pk1(__fmul_rn(dx, ra.z)), mul2(dy, mul2(dy, pk1(rb.x))));
				const p2 power = fma2(qd, NHALF2, mul2(dy, pk1(-__fmul_rn(dx, ra.w))));
				float pA, pB;
				up(power, pA, pB);
				float GA, GB;
				if (EXACT_EXP_CALL) { GA = expf(fminf(pA, 0.0f)); GB = expf(fminf(pB, 0.0f)); }
				else exp2x(fminf(pA, 0.0f), fminf(pB, 0.0f), GA, GB);
				const float aA = fminf(0.99f, __fmul_rn(rb.y, GA)), aB = fminf(0.99f, __fmul_rn(rb.y, GB));
				const bool liveA = !(pA > 0.0f) && !(aA < 1.0f / 255.0f) && pos < lastA;
				const bool liveB = !(pB > 0.0f) && !(aB < 1.0f / 255.0f) && pos < lastB;
				XC(xc_live += (unsigned)liveA + (unsigned)liveB;)
				const p2 G = pk(liveA ? GA : 0.0f, liveB ? GB : 0.0f);
				const p2 al = pk(liveA ? aA : 0.0f, liveB ? aB : 0.0f);
				float oA, oB;      // 1 - alpha in [0.01, 1]: the approximate reciprocal (1 ulp) is ample for 1e-4 gradients; 1/1 is exact
				up(sub2(ONE2, al), oA, oB);
				const p2 inv = pk(rcp_approx(oA), rcp_approx(oB));
				T2 = mul2(T2, inv);
				const p2 dcdc = mul2(al, T2);
				const p2 oml = sub2(ONE2, la2);
				ar0 = fma2(la2, pk1(lc0), mul2(oml, ar0)); lc0 = rb.z;
				ar1 = fma2(la2, pk1(lc1), mul2(oml, ar1)); lc1 = rb.w;
				ar2 = fma2(la2, pk1(lc2), mul2(oml, ar2)); lc2 = c2;
				p2 dLda = mul2(sub2(pk1(rb.z), ar0), dLp0);
				dLda = fma2(sub2(pk1(rb.w), ar1), dLp1, dLda);
				dLda = fma2(sub2(pk1(c2), ar2), dLp2, dLda);
				dLda = fma2(dLda, T2, mul2(neg_Tf_bg, inv));
				la2 = al;
				const p2 u = mul2(G, dLda);
				const p2 w = mul2(u, dy);
				const float su = hsum(u), sw = hsum(w);
				const float v3 = su * dx;
				col[0 * 32] = hsum(mul2(dcdc, dLp0));
				col[1 * 32] = hsum(mul2(dcdc, dLp1));
				col[2 * 32] = hsum(mul2(dcdc, dLp2));
				col[3 * 32] = v3;
				col[4 * 32] = sw;
				col[5 * 32] = v3 * dx;
				col[6 * 32] = sw * dx;
				col[7 * 32] = hsum(mul2(w, dy));
				col[8 * 32] = su;
			};
			for (int g0 = 0; g0 < cc; g0 += GROUP) {
				const int gn = min(GROUP, cc - g0);               // warp-uniform
				if (gn == GROUP) {
					// full group: no guards between the visits, so the independent front halves (power, exp, alpha) of the three
					// visits can be scheduled across each other
					#pragma unroll
					for (int j = 0; j < GROUP; j++) visit(g0 + j, red + j * NV * 32 + lane);
				} else {
					#pragma unroll
					for (int j = 0; j < GROUP; j++)
						if (j < gn) visit(g0 + j, red + j * NV * 32 + lane);
				}
				__syncwarp();
				if (lane < gn * NV) {
					p2 s0 = pk(0.f, 0.f), s1 = s0;
					#pragma unroll
					for (int i = 0; i < 8; i++) {
						const float4 t = red_row[(i + lane) & 7];
						s0 = add2(s0, pk(t.x, t.y)); s1 = add2(s1, pk(t.z, t.w));
					}
					atomicAdd(acc_row + (size_t)s_gid[warp][buf][g0 + row_visit] * ACC_FLOATS, hsum(add2(s0, s1)));
				}
				__syncwarp();      // rows are free again
			}
			__syncwarp();      // every lane is done with this buffer before the next iteration refills it
		}
	}
	XC({
		const unsigned live = __reduce_add_sync(FULL, xc_live);
		if (lane == 0) { xc_add(a.xc, 8, xc_visits); xc_add(a.xc, 9, live); }
	})
}

}  // namespace

int launch_render_bwd(cudaStream_t stream, const RenderArgs& a)
{
	const dim3 grid(a.grid_x, a.grid_y);
	count_launch();
#ifdef RR_CROSSCHECK
	if (crosscheck_staging() == STAGE_TMA) { render_bwd_kernel<STAGE_TMA><<<grid, B2_THREADS, 0, stream>>>(a); return check_launch("render_bwd", stream, false); }
#endif
	render_bwd_kernel<STAGE_CP_ASYNC><<<grid, B2_THREADS, 0, stream>>>(a);
	return check_launch("render_bwd", stream, false);
}

}  // namespace rr
