// Backward tile renderer: each WARP independently replays, back to front, exactly the list entries its 8x4 pixel
// block blended in the forward, and adds one warp-reduced gradient record per visit to the Gaussian's accumulator.
// Replaces renderCUDA<3> backward (reference backward.cu:389-547), whose 9 global float atomics per contributing
// (pixel, Gaussian) pair serialise in L2 on a single address per Gaussian.
//
// B200 design:
//  * no CTA-wide coupling at all: the forward leaves one bit per (list entry, warp) ("this warp blended the entry",
//    kernels.cuh); a warp reads its own bit words, turns them into a descending visit list, and walks it.  Warps that
//    own saturated or empty pixel blocks finish at once; there is no __syncthreads in the kernel.
//  * the 48-byte records of the next CHUNK (16) visits are gathered straight from the per-Gaussian record array (L2
//    resident: 48 MB at 1M Gaussians) by per-lane 1-D TMA bulk copies onto a per-warp mbarrier, double buffered, while
//    the current chunk is processed; Gaussian ids are prefetched one chunk further ahead.  Shared memory per CTA is
//    17.5 KB static + 27 KB of transpose rows, so 4 CTAs (32 warps) stay resident per SM — measured: 3 CTAs/SM with
//    32-visit chunks 0.534 ms, 4 CTAs/SM with 16-visit chunks 0.494 ms (C3, 8.4M pairs).
//  * each lane forms the terms of its pixel as the reference does (backward.cu:481-544), but accumulates the MOMENTS
//    u, u.dx, u.dy, u.dx^2, u.dx.dy, u.dy^2 of u = G * dL_dalpha instead of the five conic / mean products — they are
//    linear in the per-Gaussian constants (conic, opacity), which preprocess_bwd.cu applies once per Gaussian;
//  * the 9 values of GROUP consecutive visits are reduced across the warp by a TRANSPOSE through shared memory instead
//    of shuffles: every lane stores its 9 terms of a visit as 9 conflict-free STS into rows [visit][value][lane];
//    after GROUP visits lane r < 9*GROUP owns row r, reads its 32 floats as 8 LDS.128 (rotated by r so the 8 lanes of a
//    quarter-warp hit 8 different bank groups) and sums them with packed add.rn.f32x2 (SASS FADD2), then issues its
//    share of ONE predicated RED.ADD.F32 per group (9*GROUP active lanes).  Per visit this is 9 STS + ~10 issue
//    slots of row work against 14 SHFL + 14 FADD + 14 FSEL for the halving butterfly it replaces (REDUCE_SHFL keeps
//    the butterfly as the cross-check variant), and a third of the RED instructions.
#include <cstdlib>
#include "common.cuh"
#include "kernels.cuh"
#include "render_common.cuh"

namespace rr {

namespace {

constexpr int NV = 9;
constexpr int LIST_WORDS = 8;                   // flag words (of 32 list entries) turned into a visit list per round
constexpr int LIST_CAP = LIST_WORDS * 32;
constexpr int CHUNK = 16;                       // visits staged per buffer (16 x 48 B per warp and buffer keeps 4 CTAs per SM)

__device__ __forceinline__ float2 add2(float2 a, float2 b) {
	float2 r;
	asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
		: "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
	return r;
}

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

// GROUP = visits reduced together through shared memory (0: shuffle butterfly, no transpose rows)
template <bool USE_TMA, int GROUP>
__global__ void __launch_bounds__(RENDER_THREADS) render_bwd_kernel(RenderArgs a)
{
	constexpr bool REDUCE_SHFL = GROUP == 0;
	constexpr int RED_ROWS = GROUP * NV;         // rows of 32 floats per warp

	__shared__ __align__(128) float4 s_rec[RENDER_WARPS][2][CHUNK * 3];   // per warp, double buffered: CHUNK records of 48 B
	__shared__ uint32_t s_gid[RENDER_WARPS][2][CHUNK];                  // Gaussian id of each staged visit
	__shared__ uint16_t s_pos[RENDER_WARPS][LIST_CAP];                  // visit list of the round (positions relative to the round's first entry, descending)
	__shared__ __align__(8) uint64_t s_bar[RENDER_WARPS][2];
	extern __shared__ __align__(16) float s_red_dyn[];                  // transpose rows [warp][visit in group][value][lane] (dynamic: static + this > 48 KB)
	float* const s_red_w = s_red_dyn + (REDUCE_SHFL ? 0 : (threadIdx.x >> 5) * RED_ROWS * 32);

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const int tile = blockIdx.y * a.grid_x + blockIdx.x;
	int wx, wy, lx, ly;
	warp_footprint(warp, lane, wx, wy, lx, ly);
	const int pix_x = blockIdx.x * TILE_X + wx + lx, pix_y = blockIdx.y * TILE_Y + wy + ly;
	const bool inside = pix_x < a.W && pix_y < a.H;
	const float pxf = (float)pix_x, pyf = (float)pix_y;

	const uint2 range = a.ranges[tile];
	const int n = (int)(range.y - range.x);
	const size_t pid = (size_t)a.W * pix_y + pix_x, HW = (size_t)a.H * a.W;
	const uint32_t last_contributor = inside ? a.n_contrib[pid] : 0u;
	const uint32_t wmax = __reduce_max_sync(FULL, last_contributor);
	if (wmax == 0 || n == 0) return;             // warp-uniform: nothing was blended into this warp's pixels

	const float T_final = inside ? a.final_T[pid] : 0.0f;
	float T = T_final;
	float dLp0 = 0, dLp1 = 0, dLp2 = 0;
	if (inside) { dLp0 = a.dL_dpix[pid]; dLp1 = a.dL_dpix[HW + pid]; dLp2 = a.dL_dpix[2 * HW + pid]; }
	float bg_dot_dpixel = 0;
	bg_dot_dpixel += a.bg[0] * dLp0; bg_dot_dpixel += a.bg[1] * dLp1; bg_dot_dpixel += a.bg[2] * dLp2;
	const float neg_Tf_bg = -T_final * bg_dot_dpixel;
	float ar0 = 0, ar1 = 0, ar2 = 0, lc0 = 0, lc1 = 0, lc2 = 0, last_alpha = 0;

	uint64_t* bar = s_bar[warp];
	if (USE_TMA) {
		if (lane == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init(); }
		__syncwarp();
	}
	// which lane adds which reduced value: lanes 0,4,..,28 hold values 0..7 (see warp_reduce9), lane 1 adds the 9th
	const int kslot = (lane & 3) == 0 ? ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1) : (lane == 1 ? 8 : -1);
	const bool adder = kslot >= 0;
	float* __restrict__ acc_lane = a.grad_acc + (adder ? kslot : 0);
	// shared-memory transpose: lane r < 9 * GROUP owns row r = visit * 9 + value of the group
	const int row_visit = lane / NV;
	float* __restrict__ acc_row = a.grad_acc + (lane - row_visit * NV);
	const float4* __restrict__ red_row = reinterpret_cast<const float4*>(s_red_w) + (REDUCE_SHFL ? 0 : lane * 8);

	const int ngroups = (n + 31) >> 5;
	const uint32_t* __restrict__ wf = a.warp_flags + 8ull * ((size_t)(range.x >> 5) + (size_t)tile) + (size_t)warp * ngroups;
	const uint32_t* __restrict__ ids = a.point_list + range.x;
	const float4* __restrict__ grec = reinterpret_cast<const float4*>(a.rec);
	const int nwords = (int)((wmax + 31u) >> 5);
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
			} else if (lane < cc) {
				const float4 r0 = __ldg(grec + (size_t)gid * 3), r1 = __ldg(grec + (size_t)gid * 3 + 1), r2 = __ldg(grec + (size_t)gid * 3 + 2);
				s_rec[warp][buf][lane * 3] = r0; s_rec[warp][buf][lane * 3 + 1] = r1; s_rec[warp][buf][lane * 3 + 2] = r2;
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
			if (USE_TMA) mbar_wait(&bar[buf], (q >> 1) & 1u);
			__syncwarp();
			const float4* __restrict__ rec = s_rec[warp][buf];
			const int cc = min(CHUNK, total - k * CHUNK);
			const uint16_t* __restrict__ cpos = s_pos[warp] + k * CHUNK;
			// per-pixel terms of one visit (reference backward.cu:481-544, moments instead of conic / mean products)
			auto visit = [&](int kk, float (&v)[NV]) {
				const uint32_t pos = ebase + cpos[kk];
				const float4 ra = rec[kk * 3], rb = rec[kk * 3 + 1];
				const float dx = __fsub_rn(ra.x, pxf), dy = __fsub_rn(ra.y, pyf);
				const float qd = __fmaf_rn(dx, __fmul_rn(dx, ra.z), __fmul_rn(dy, __fmul_rn(dy, rb.x)));
				const float power = __fmaf_rn(qd, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, ra.w)));
				const float G = expf(fminf(power, 0.0f));
				const float alpha = fminf(0.99f, __fmul_rn(rb.y, G));
				const bool live = !(power > 0.0f) && !(alpha < 1.0f / 255.0f) && pos < last_contributor;
				float dcdc = 0.0f, u = 0.0f;
				if (live) {
					const float c2 = rec[kk * 3 + 2].x;
					float inv_1ma;      // 1/(1-alpha), 1-alpha in [0.01, 1): the approximate reciprocal (1 ulp) is ample for 1e-4 gradients
					asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_1ma) : "f"(1.f - alpha));
					T = T * inv_1ma;
					dcdc = alpha * T;
					const float one_m_la = 1.f - last_alpha;
					ar0 = last_alpha * lc0 + one_m_la * ar0; lc0 = rb.z;
					ar1 = last_alpha * lc1 + one_m_la * ar1; lc1 = rb.w;
					ar2 = last_alpha * lc2 + one_m_la * ar2; lc2 = c2;
					float dL_dalpha = (rb.z - ar0) * dLp0;
					dL_dalpha += (rb.w - ar1) * dLp1;
					dL_dalpha += (c2 - ar2) * dLp2;
					dL_dalpha = dL_dalpha * T + neg_Tf_bg * inv_1ma;
					last_alpha = alpha;
					u = G * dL_dalpha;
				}
				v[0] = dcdc * dLp0; v[1] = dcdc * dLp1; v[2] = dcdc * dLp2;
				v[3] = u * dx; v[4] = u * dy;
				v[5] = v[3] * dx; v[6] = v[3] * dy; v[7] = v[4] * dy;
				v[8] = u;
			};
			if (REDUCE_SHFL) {
				for (int kk = 0; kk < cc; kk++) {
					float v[NV], y, z;
					visit(kk, v);
					warp_reduce9(v, lane, y, z);
					if (adder) atomicAdd(acc_lane + (size_t)s_gid[warp][buf][kk] * ACC_FLOATS, lane == 1 ? z : y);
				}
			} else {
				float* __restrict__ red = s_red_w;
				for (int g0 = 0; g0 < cc; g0 += GROUP) {
					const int gn = min(GROUP, cc - g0);               // warp-uniform
					#pragma unroll
					for (int j = 0; j < GROUP; j++) {
						if (j < gn) {
							float v[NV];
							visit(g0 + j, v);
							#pragma unroll
							for (int k = 0; k < NV; k++) red[(j * NV + k) * 32 + lane] = v[k];
						}
					}
					__syncwarp();
					if (lane < gn * NV) {
						float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
						#pragma unroll
						for (int i = 0; i < 8; i++) {
							const float4 t = red_row[(i + lane) & 7];
							s0 = add2(s0, make_float2(t.x, t.y)); s1 = add2(s1, make_float2(t.z, t.w));
						}
						s0 = add2(s0, s1);
						atomicAdd(acc_row + (size_t)s_gid[warp][buf][g0 + row_visit] * ACC_FLOATS, s0.x + s0.y);
					}
					__syncwarp();      // rows are free again
				}
			}
			__syncwarp();      // every lane is done with this buffer before the next iteration refills it
		}
	}
}

}  // namespace

int launch_render_bwd2(cudaStream_t stream, const RenderArgs& a);   // render_bwd2.cu: two pixels per lane, packed f32x2 arithmetic

int launch_render_bwd(cudaStream_t stream, const RenderArgs& a)
{
	// RAIN_BWD_KERNEL=1 selects the one-pixel-per-lane kernel of this file (cross-check); default: render_bwd2.cu
	static const bool one_px = []() { const char* e = getenv("RAIN_BWD_KERNEL"); return e && e[0] == '1'; }();
	if (!one_px) return launch_render_bwd2(stream, a);
	const dim3 grid(a.grid_x, a.grid_y);
	// RAIN_BWD_REDUCE=s selects the shuffle-butterfly reduction (cross-check of the shared-memory transpose), 2 / 3 the group size
	static const int group = []() { const char* e = getenv("RAIN_BWD_REDUCE"); return !e ? 3 : e[0] == 's' ? 0 : e[0] == '2' ? 2 : 3; }();
	const bool ldg = a.variant == MODE_PACKED_LDG || a.variant == MODE_GATHER_LDG;
	auto launch = [&](auto kern, int grp) -> int {
		const int red_bytes = RENDER_WARPS * grp * NV * 32 * (int)sizeof(float);
		if (red_bytes && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, red_bytes) != cudaSuccess) {
			set_error("render_bwd: cannot reserve %d bytes of dynamic shared memory", red_bytes); return -1;
		}
		count_launch();
		kern<<<grid, RENDER_THREADS, red_bytes, stream>>>(a);
		return 0;
	};
	int rc;
	if (group == 0) rc = ldg ? launch(render_bwd_kernel<false, 0>, 0) : launch(render_bwd_kernel<true, 0>, 0);
	else if (group == 2) rc = ldg ? launch(render_bwd_kernel<false, 2>, 2) : launch(render_bwd_kernel<true, 2>, 2);
	else rc = ldg ? launch(render_bwd_kernel<false, 3>, 3) : launch(render_bwd_kernel<true, 3>, 3);
	if (rc) return rc;
	return check_launch("render_bwd", stream, false);
}

}  // namespace rr
