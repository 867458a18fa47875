// Tile binning: prefix sum of tiles_touched, (tile, Gaussian) duplicate emission, tile ranges.
// Replaces cub::DeviceScan::InclusiveSum (reference rasterizer_impl.cu:269), duplicateWithKeys (:59-100),
// identifyTileRanges (:105-127) and the cudaMemset of ranges (:302).
//
// Ordering.  The reference sorts N (tile<<32 | depth_bits) keys with 6 radix passes.  Here the P Gaussians are first
// sorted by depth bits (4 passes over P items), duplicates are then emitted in that depth order, and a stable sort
// on the tile id alone (ceil(bits/8) = 2 passes over N at every supported resolution) yields exactly the same
// (tile, depth, index) order: ties in (tile, depth) keep ascending Gaussian index in both schemes, because both
// sorts are stable and the depth sort starts from index order.  The 64-bit key list the reference would hold is
// reconstructed on demand by export_keys_kernel for the parity tests.
#include <cstring>
#include <algorithm>
#include "common.cuh"
#include "kernels.cuh"

namespace rr {

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;        // per thread, consecutive: 4096 per CTA (the look-back walks over 4 x fewer tiles than at 4)
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// Status words of the chained scan: flags in the top two bits, a 30-bit payload.  All sums SATURATE at SC_MASK = 2^30 - 1
// (operands are <= SC_MASK, so a + b cannot wrap 32 bits): a total that does not fit the payload reaches the host as
// exactly SC_MASK, which rr_forward_bin rejects, instead of wrapping silently into a plausible small count.
constexpr uint32_t SC_AGG = 1u << 30, SC_INC = 2u << 30, SC_MASK = (1u << 30) - 1;
__device__ __forceinline__ uint32_t sat_add(uint32_t a, uint32_t b) { return min(a + b, SC_MASK); }

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp /*[8]*/, uint32_t& block_total)
{
	const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
	uint32_t inc = v;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= (unsigned)o) inc = sat_add(inc, t); }
	if (lane == 31) s_warp[warp] = inc;
	__syncthreads();
	uint32_t pre = 0, tot = 0;
	#pragma unroll
	for (int w = 0; w < SCAN_THREADS / 32; w++) { const uint32_t t = s_warp[w]; if (w < (int)warp) pre = sat_add(pre, t); tot = sat_add(tot, t); }
	block_total = tot;
	__syncthreads();
	return sat_add(pre, inc) - v;            // exact below saturation; garbage above it, where the host rejects the total
}

struct ZeroSegs { uint4* p[16]; };
__global__ void __launch_bounds__(256) zero_segments_kernel(ZeroSegs z, size_t n16)
{
	uint4* __restrict__ p = z.p[blockIdx.y];
	for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// Inclusive prefix sum of tiles_touched in DEPTH order (where each Gaussian's duplicates go in the emission list), replacing
// cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:269), in ONE launch: chained scan with decoupled look-back.  Blocks
// take tile ids from an atomic ticket (a block's predecessors are therefore always running or done), publish their tile
// aggregate in a status word, and warp 0 looks back over 32 predecessors per step until it meets an inclusive prefix.
// blockIdx.y = segment: the scans of all views of a step run as one launch (ScanSegs, kernels.cuh).
__global__ void __launch_bounds__(SCAN_THREADS) scan_chained_kernel(int P, int nb, ScanSegs sg)
{
	const uint32_t* __restrict__ order = sg.order[blockIdx.y];
	const uint32_t* __restrict__ tt = sg.tiles_touched[blockIdx.y];
	volatile uint32_t* __restrict__ status = sg.status[blockIdx.y];      // [nb] + ticket at [nb], zero on entry
	uint32_t* __restrict__ offsets = sg.offsets[blockIdx.y];             // inclusive, depth order
	uint32_t* __restrict__ total = sg.total[blockIdx.y];
	uint32_t* __restrict__ total_host = sg.total_host[blockIdx.y];
	__shared__ uint32_t s_warp[SCAN_THREADS / 32];
	__shared__ uint32_t s_tile, s_excl;
	if (threadIdx.x == 0) s_tile = atomicAdd(const_cast<uint32_t*>(status) + nb, 1u);
	__syncthreads();
	const int tile = (int)s_tile;
	uint32_t v[SCAN_ITEMS], sum = 0;
	const int i0 = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
	const bool full = tile * SCAN_TILE + SCAN_TILE <= P;                // warp- and CTA-uniform
	if (full) {
		// the thread's 16 depth-order entries as four 16-byte loads, then 16 independent gathers
		uint32_t g[SCAN_ITEMS];
		#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k += 4) {
			const uint4 o = *reinterpret_cast<const uint4*>(order + i0 + k);
			g[k] = o.x; g[k + 1] = o.y; g[k + 2] = o.z; g[k + 3] = o.w;
		}
		#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k++) v[k] = min(__ldg(tt + g[k]), SC_MASK);
	} else {
		#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k++) v[k] = i0 + k < P ? min(tt[order[i0 + k]], SC_MASK) : 0;
	}
	#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) sum = sat_add(sum, v[k]);
	uint32_t tot;
	const uint32_t ex = block_exclusive_scan(sum, s_warp, tot);
	if (threadIdx.x < 32) {
		const unsigned lane = threadIdx.x;
		if (lane == 0) { status[tile] = tot | (tile == 0 ? SC_INC : SC_AGG); __threadfence(); }
		uint32_t excl = 0;
		int t = tile - 1;
		while (t >= 0) {                                               // warp-uniform
			const int idx = t - (int)lane;
			uint32_t w = 2u << 30; if (idx >= 0) w = status[idx];       // before tile 0: inclusive prefix 0
			const uint32_t f = w >> 30;
			const unsigned not_ready = __ballot_sync(FULL, f == 0), inc = __ballot_sync(FULL, f == 2);
			const int first_nr = not_ready ? __ffs(not_ready) - 1 : 32, first_inc = inc ? __ffs(inc) - 1 : 32;
			const int take = min(first_nr, first_inc + 1);              // lanes [0, take): published, up to and including the first inclusive one
			uint32_t c = (int)lane < take ? (w & SC_MASK) : 0u;
			#pragma unroll
			for (int o = 16; o >= 1; o >>= 1) c = sat_add(c, __shfl_xor_sync(FULL, c, o));
			excl = sat_add(excl, c);
			if (first_inc < first_nr) break;
			t -= take;
		}
		if (lane == 0) {
			const uint32_t incl = sat_add(excl, tot);
			if (tile > 0) { status[tile] = incl | SC_INC; __threadfence(); }
			s_excl = excl;
			if (tile == nb - 1) { *total = incl; if (total_host) { *total_host = incl; __threadfence_system(); } }
		}
	}
	__syncthreads();
	uint32_t run = s_excl + ex;
	if (full) {
		#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k += 4) {
			uint4 o;
			run += v[k]; o.x = run; run += v[k + 1]; o.y = run; run += v[k + 2]; o.z = run; run += v[k + 3]; o.w = run;
			*reinterpret_cast<uint4*>(offsets + i0 + k) = o;
		}
	} else {
		#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k++) {
			run += v[k];
			if (i0 + k < P) offsets[i0 + k] = run;
		}
	}
}

// One warp emits the duplicates of 32 consecutive depth-sorted Gaussians cooperatively: output slot s of the warp's
// contiguous span is owned by the last lane whose exclusive offset is <= s (binary search over 32 shared values),
// so writes are coalesced and a large splat does not serialise on one thread (reference: one thread per rect).
__global__ void __launch_bounds__(256) duplicate_kernel(int P, const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
	const uint2* __restrict__ rect, const unsigned long long* __restrict__ tile_mask, int grid_x, uint32_t* __restrict__ tile_keys,
	uint32_t* __restrict__ gaussian_ids)
{
	__shared__ unsigned long long s_mask[8][32];
	__shared__ uint32_t s_excl[8][32];
	__shared__ uint32_t s_g[8][32];
	__shared__ uint32_t s_xw[8][32];   // x0 | width<<16
	__shared__ uint32_t s_y0[8][32];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int i = blockIdx.x * 256 + threadIdx.x;
	uint32_t g = 0, excl = 0, incl = 0;
	if (i < P) {
		g = order[i];
		incl = offsets[i];
		excl = i == 0 ? 0u : offsets[i - 1];
	}
	const int last_valid = min(31, P - 1 - (blockIdx.x * 256 + warp * 32));   // >= 0 for any warp that has work
	if (last_valid < 0) return;                                                 // warp-uniform
	const uint32_t base = __shfl_sync(FULL, excl, 0);
	const uint32_t end = __shfl_sync(FULL, incl, last_valid);
	const uint32_t total = end - base;
	const uint32_t tt = i < P ? incl - excl : 0u;
	uint32_t x0 = 0, wdt = 1, y0 = 0;
	unsigned long long mask = ~0ull;
	if (tt > 0) {
		const uint2 r = rect[g];
		x0 = r.x & 0xffffu; wdt = (r.x >> 16) - x0; y0 = r.y & 0xffffu;
		if (tile_mask) mask = tile_mask[g];
	}
	s_mask[warp][lane] = mask;
	s_excl[warp][lane] = i < P ? excl - base : total;
	s_g[warp][lane] = g;
	s_xw[warp][lane] = x0 | (wdt << 16);
	s_y0[warp][lane] = y0;
	__syncwarp();
	for (uint32_t s = lane; s < total; s += 32) {
		// last j with s_excl[j] <= s
		int lo = 0;
		#pragma unroll
		for (int step = 16; step >= 1; step >>= 1) if (s_excl[warp][lo + step] <= s) lo += step;
		uint32_t local = s - s_excl[warp][lo];
		const unsigned long long m = s_mask[warp][lo];
		if (m != ~0ull) {
			// culled rect (preprocess.cu): the local-th kept tile
			const uint32_t mlo = (uint32_t)m, mhi = (uint32_t)(m >> 32), c = (uint32_t)__popc(mlo);
			uint32_t word = mlo, n = local, pos = 0;
			if (local >= c) { word = mhi; n = local - c; pos = 32; }
			#pragma unroll
			for (uint32_t wd = 16; wd >= 1; wd >>= 1) {          // position of the n-th set bit of word
				const uint32_t below = (uint32_t)__popc(word & ((1u << wd) - 1u));
				if (n >= below) { n -= below; word >>= wd; pos += wd; }
			}
			local = pos;
		}
		const uint32_t xw = s_xw[warp][lo];
		const uint32_t w = xw >> 16, xx0 = xw & 0xffffu;
		const uint32_t ry = local / w, rx = local - ry * w;
		tile_keys[base + s] = (s_y0[warp][lo] + ry) * (uint32_t)grid_x + xx0 + rx;
		gaussian_ids[base + s] = s_g[warp][lo];        // the sort's value, as in the reference (rasterizer_impl.cu:92)
	}
}

// The pair emission of the 16-bit tile sort (radix_sort.cu::tile_sort16; no tile masks on that path), written for a short
// instruction stream — the kernel is bound by instruction issue, its 6 bytes per pair are far from the HBM rate.  A warp
// takes 32 Gaussians of the depth order, keeps those that touch tiles, and emits their pairs 32 at a time:
//  * which Gaussian owns output slot s is found without a search: every Gaussian whose first pair falls into the current
//    32 slots sets the bit of that slot (one redux.or over the warp); the owner of slot l is then (Gaussians started before
//    the window) + popc(bits up to l) - 1;
//  * tile y / x inside the rect without a division per pair: local / width = umulhi(local, ceil(2^32 / width)), exact for
//    local, width < 2^16 (tile ids are 16-bit on this path); one division per Gaussian.
__global__ void __launch_bounds__(256) duplicate16_kernel(int P, const uint32_t* __restrict__ order, const uint32_t* __restrict__ offsets,
	const uint2* __restrict__ rect, int grid_x, uint16_t* __restrict__ tile_keys, uint32_t* __restrict__ gaussian_ids)
{
	__shared__ uint4 s_a[8][32];       // { first pair (relative to the warp's first), Gaussian, x0 | y0 << 16, ceil(2^32 / width) (0: width 1) }
	__shared__ uint32_t s_w[8][32];    // rect width in tiles
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int i = blockIdx.x * 256 + threadIdx.x;
	uint32_t g = 0, excl = 0, incl = 0;
	if (i < P) {
		g = order[i];
		incl = offsets[i];
		excl = i == 0 ? 0u : offsets[i - 1];
	}
	const int last_valid = min(31, P - 1 - (blockIdx.x * 256 + warp * 32));   // >= 0 for any warp that has work
	if (last_valid < 0) return;                                                 // warp-uniform
	const uint32_t base = __shfl_sync(FULL, excl, 0);
	const uint32_t total = __shfl_sync(FULL, incl, last_valid) - base;
	const bool touches = i < P && incl > excl;
	const unsigned has = __ballot_sync(FULL, touches);
	if (touches) {
		const uint2 r = rect[g];
		const uint32_t x0 = r.x & 0xffffu, w = (r.x >> 16) - x0, y0 = r.y & 0xffffu;
		const int slot = __popc(has & lanemask_lt());
		s_a[warp][slot] = make_uint4(excl - base, g, x0 | (y0 << 16), 0xffffffffu / w + 1u);       // (w == 1 wraps to 0)
		s_w[warp][slot] = w;
	}
	__syncwarp();
	const uint32_t my_start = lane < __popc(has) ? s_a[warp][lane].x : 0xffffffffu;
	const uint32_t le = 0xffffffffu >> (31 - lane);                             // lanes up to and including mine
	uint16_t* __restrict__ kout = tile_keys + base;
	uint32_t* __restrict__ vout = gaussian_ids + base;
	uint32_t started = 0;                                                        // Gaussians whose first pair lies before the window
	for (uint32_t s0 = 0; s0 < total; s0 += 32) {
		const uint32_t rel = my_start - s0;                                      // (wraps to a huge value for a start before the window)
		const uint32_t starts = __reduce_or_sync(FULL, rel < 32u ? 1u << rel : 0u);
		const uint32_t s = s0 + lane;
		const uint32_t lo = started + __popc(starts & le) - 1u;
		started += __popc(starts);
		if (s < total) {
			const uint4 a = s_a[warp][lo];
			const uint32_t w = s_w[warp][lo];
			const uint32_t local = s - a.x;
			const uint32_t ry = a.w ? __umulhi(local, a.w) : local;
			const uint32_t rx = local - ry * w;
			kout[s] = (uint16_t)(((a.z >> 16) + ry) * (uint32_t)grid_x + (a.z & 0xffffu) + rx);
			vout[s] = a.y;                                                       // the sort's value, as in the reference (rasterizer_impl.cu:92)
		}
	}
}

// tile ranges from key changes (identifyTileRanges, rasterizer_impl.cu:105-127); four consecutive keys per thread (one 16-byte load)
__global__ void __launch_bounds__(256) tile_ranges_kernel(uint32_t N, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges)
{
	const uint32_t i4 = (blockIdx.x * 256u + threadIdx.x) * 4u;
	if (i4 >= N) return;
	uint32_t k[4];
	if (i4 + 3u < N) {
		const uint4 v = __ldg(reinterpret_cast<const uint4*>(keys) + (i4 >> 2));
		k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
	} else {
		#pragma unroll
		for (int j = 0; j < 4; j++) k[j] = i4 + j < N ? __ldg(keys + i4 + j) : 0u;
	}
	uint32_t prev = i4 ? __ldg(keys + i4 - 1) : 0u;
	#pragma unroll
	for (int j = 0; j < 4; j++) {
		const uint32_t idx = i4 + j;
		if (idx < N) {
			const uint32_t cur = k[j];
			if (idx == 0) ranges[cur].x = 0;
			else if (cur != prev) { ranges[prev].y = idx; ranges[cur].x = idx; }
			if (idx == N - 1) ranges[cur].y = N;
			prev = cur;
		}
	}
}

// ---- tile plan: per-tile pair counts -> ranges and the digit bases of the tile sort, without touching a pair ----------------
// The pairs of tile t are exactly the visible Gaussians whose rect covers t, so count[t] follows from the rects alone: every
// Gaussian adds +1 / -1 at the four corners of its rect in a (grid_x + 1) x (grid_y + 1) difference grid (four shared-memory
// atomics per Gaussian, whatever the size of its rect), and the 2-D prefix sum of that grid is the count.  From the counts:
// ranges[t] = [start, start + count) (identifyTileRanges, rasterizer_impl.cu:105-127, incl. the zero fill of :302 for empty
// tiles) and the exclusive digit bases of both radix passes — so the tile sort needs no histogram pass over the 8-byte pairs,
// no sorted key array, and no range-detection pass.
constexpr int PLAN_MAX_CELLS = 11264;       // (grid_x + 1) * (grid_y + 1) ints in shared memory (44 KB); 1920x1080 needs 8 349

constexpr int DIFF_THREADS = 512;
// blockIdx.y = view: the plans of all views of a batched geometry pass run as one pair of launches (PlanSegs, kernels.cuh)
__global__ void __launch_bounds__(DIFF_THREADS) tile_diff_kernel(int P, int cells, int gx1, PlanSegs sg)
{
	const uint32_t* __restrict__ tiles_touched = sg.tiles_touched[blockIdx.y];
	const uint2* __restrict__ rect = sg.rect[blockIdx.y];
	int* __restrict__ gdiff = sg.gdiff[blockIdx.y];          // [cells], zero on entry
	__shared__ int s_d[PLAN_MAX_CELLS];
	for (int i = threadIdx.x; i < cells; i += DIFF_THREADS) s_d[i] = 0;
	__syncthreads();
	// one CTA per SM, four Gaussians in flight per thread: few CTAs keep the flush below small (cells atomics per CTA)
	const int stride = gridDim.x * DIFF_THREADS;
	for (int i0 = blockIdx.x * DIFF_THREADS + threadIdx.x; i0 < P; i0 += 4 * stride) {
		uint32_t tt[4];
		uint2 r[4];
		#pragma unroll
		for (int u = 0; u < 4; u++) { const int i = i0 + u * stride; tt[u] = i < P ? __ldg(tiles_touched + i) : 0u; }
		#pragma unroll
		for (int u = 0; u < 4; u++) { const int i = i0 + u * stride; r[u] = tt[u] ? __ldg(rect + i) : make_uint2(0u, 0u); }
		#pragma unroll
		for (int u = 0; u < 4; u++) {
			if (tt[u] == 0) continue;
			const int x0 = r[u].x & 0xffff, x1 = r[u].x >> 16, y0 = r[u].y & 0xffff, y1 = r[u].y >> 16;
			atomicAdd(&s_d[y0 * gx1 + x0], 1); atomicAdd(&s_d[y0 * gx1 + x1], -1);
			atomicAdd(&s_d[y1 * gx1 + x0], -1); atomicAdd(&s_d[y1 * gx1 + x1], 1);
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < cells; i += DIFF_THREADS) { const int v = s_d[i]; if (v) atomicAdd(&gdiff[i], v); }
}

constexpr int PLAN_THREADS = 1024;
__global__ void __launch_bounds__(PLAN_THREADS) tile_plan_kernel(int grid_x, int grid_y, PlanSegs sg, int npass, int shift0, uint32_t mask0, int shift1, uint32_t mask1)
{
	const int* __restrict__ gdiff = sg.gdiff[blockIdx.x];
	uint2* __restrict__ ranges = sg.ranges[blockIdx.x];
	uint32_t* __restrict__ ghist = sg.ghist[blockIdx.x];     // [npass][256] exclusive digit bases out
	__shared__ int s_c[PLAN_MAX_CELLS];
	__shared__ uint32_t s_h[2][256];
	__shared__ uint32_t s_wsum[PLAN_THREADS / 32];
	__shared__ uint32_t s_carry;
	const int tid = threadIdx.x, gx1 = grid_x + 1, gy1 = grid_y + 1, T = grid_x * grid_y;
	for (int i = tid; i < gx1 * gy1; i += PLAN_THREADS) s_c[i] = gdiff[i];
	if (tid < 512) (&s_h[0][0])[tid] = 0;
	if (tid == 0) s_carry = 0;
	__syncthreads();
	for (int y = tid >> 5; y < gy1; y += PLAN_THREADS / 32) {                                                                 // along rows: one warp per row
		int carry = 0;
		for (int xb = 0; xb < gx1; xb += 32) {
			const int x = xb + (tid & 31);
			int v = x < gx1 ? s_c[y * gx1 + x] : 0;
			#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(FULL, v, o); if ((tid & 31) >= o) v += u; }
			if (x < gx1) s_c[y * gx1 + x] = v + carry;
			carry += __shfl_sync(FULL, v, 31);
		}
	}
	__syncthreads();
	if (tid < gx1) { int run = 0; for (int y = 0; y < gy1; y++) { run += s_c[y * gx1 + tid]; s_c[y * gx1 + tid] = run; } }     // along columns
	__syncthreads();
	// exclusive scan of the counts in tile order (t = y * grid_x + x), PLAN_THREADS tiles per round
	for (int base = 0; base < T; base += PLAN_THREADS) {
		const int t = base + tid;
		uint32_t c = 0;
		if (t < T) { const int y = t / grid_x, x = t - y * grid_x; c = (uint32_t)s_c[y * gx1 + x]; }
		uint32_t inc = c;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(FULL, inc, o); if ((tid & 31) >= o) inc += u; }
		if ((tid & 31) == 31) s_wsum[tid >> 5] = inc;
		__syncthreads();
		uint32_t pre = s_carry;
		for (int w = 0; w < (tid >> 5); w++) pre += s_wsum[w];
		const uint32_t start = pre + inc - c;
		if (t < T) {
			ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
			if (c) {
				atomicAdd(&s_h[0][((uint32_t)t >> shift0) & mask0], c);
				if (npass > 1) atomicAdd(&s_h[1][((uint32_t)t >> shift1) & mask1], c);
			}
		}
		__syncthreads();
		if (tid == PLAN_THREADS - 1) s_carry = pre + inc;
		__syncthreads();
	}
	// exclusive digit bases of each pass: warp p scans its 256 bins
	if (tid < 32 * npass) {
		const int p = tid >> 5, lane = tid & 31;
		uint32_t carry = 0;
		for (int b = 0; b < 256; b += 32) {
			const uint32_t c = s_h[p][b + lane];
			uint32_t inc = c;
			#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += u; }
			ghist[p * 256 + b + lane] = carry + inc - c;
			carry += __shfl_sync(FULL, inc, 31);
		}
	}
}

// the 64-bit (tile << 32 | depth bits) key list the reference would hold (point_list_keys, rasterizer_impl.cu:92),
// reconstructed from ranges + point_list for the parity tests; depth bits = float index 9 (c.y) of the 12-float record
__global__ void __launch_bounds__(128) export_keys_kernel(int T, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
	const uint32_t* __restrict__ rec_words, uint64_t* __restrict__ out)
{
	const int t = blockIdx.x;
	if (t >= T) return;
	const uint2 r = ranges[t];
	for (uint32_t j = r.x + threadIdx.x; j < r.y; j += 128)
		out[j] = ((uint64_t)(uint32_t)t << 32) | rec_words[(size_t)point_list[j] * 12 + 9];
}

}  // namespace

size_t scan_ws_elems(int P) { return (size_t)(P + SCAN_TILE - 1) / SCAN_TILE + 1; }

int launch_zero_segments(cudaStream_t stream, int nseg, void* const* ptrs, size_t bytes)
{
	if (nseg <= 0 || bytes == 0) return 0;
	if (nseg == 1) { cudaMemsetAsync(ptrs[0], 0, bytes, stream); return 0; }
	if (nseg > 16) { set_error("zero_segments: %d regions (at most 16)", nseg); return -1; }
	ZeroSegs z;
	for (int g = 0; g < 16; g++) z.p[g] = reinterpret_cast<uint4*>(ptrs[g < nseg ? g : 0]);
	const size_t n16 = (bytes + 15) / 16;
	const unsigned gx = (unsigned)std::min<size_t>((n16 + 255) / 256, 64);
	count_launch(), zero_segments_kernel<<<dim3(gx, nseg), 256, 0, stream>>>(z, n16);
	return check_launch("zero_segments", stream, false);
}

int launch_scan_tiles_segments(cudaStream_t stream, int P, int nseg, const ScanSegs& sg)
{
	if (nseg == 0) return 0;
	if (P == 0) { for (int g = 0; g < nseg; g++) cudaMemsetAsync(sg.total[g], 0, sizeof(uint32_t), stream); return 0; }
	const int nb = (P + SCAN_TILE - 1) / SCAN_TILE;
	{ void* zp[16]; for (int g = 0; g < nseg; g++) zp[g] = sg.status[g]; if (launch_zero_segments(stream, nseg, zp, sizeof(uint32_t) * (size_t)(nb + 1))) return -1; }
	count_launch(), scan_chained_kernel<<<dim3(nb, nseg), SCAN_THREADS, 0, stream>>>(P, nb, sg);
	return check_launch("scan_tiles", stream, false);
}

int launch_scan_tiles(cudaStream_t stream, int P, const uint32_t* order, const uint32_t* tt, uint32_t* offsets, uint32_t* block_sums, uint32_t* num_rendered_dev,
	uint32_t* num_rendered_mapped)
{
	ScanSegs sg; memset(&sg, 0, sizeof(sg));
	sg.order[0] = order; sg.tiles_touched[0] = tt; sg.status[0] = block_sums; sg.offsets[0] = offsets; sg.total[0] = num_rendered_dev; sg.total_host[0] = num_rendered_mapped;
	return launch_scan_tiles_segments(stream, P, 1, sg);
}

int launch_duplicate(cudaStream_t stream, int P, const uint32_t* order, const uint32_t* offsets, const uint2* rect, const unsigned long long* tile_mask,
	int grid_x, uint32_t* tile_keys, uint32_t* gaussian_ids)
{
	if (P == 0) return 0;
	count_launch(), duplicate_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, order, offsets, rect, tile_mask, grid_x, tile_keys, gaussian_ids);
	return check_launch("duplicate", stream, false);
}

int launch_duplicate16(cudaStream_t stream, int P, const uint32_t* order, const uint32_t* offsets, const uint2* rect, int grid_x, uint16_t* tile_keys, uint32_t* gaussian_ids)
{
	if (P == 0) return 0;
	count_launch(), duplicate16_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, order, offsets, rect, grid_x, tile_keys, gaussian_ids);
	return check_launch("duplicate", stream, false);
}

int launch_tile_ranges(cudaStream_t stream, size_t N, const uint32_t* keys, uint2* ranges, int num_tiles)
{
	cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream);
	if (N == 0) return 0;
	count_launch(), tile_ranges_kernel<<<(unsigned)((N + 1023) / 1024), 256, 0, stream>>>((uint32_t)N, keys, ranges);
	return check_launch("tile_ranges", stream, false);
}

bool tile_plan_supported(int grid_x, int grid_y) { return (grid_x + 1) * (grid_y + 1) <= PLAN_MAX_CELLS && grid_x < 65535 && grid_y < 65535; }
size_t tile_plan_ws_bytes(int grid_x, int grid_y) { return sizeof(int) * (size_t)(grid_x + 1) * (grid_y + 1); }

int launch_tile_plan(cudaStream_t stream, int P, int grid_x, int grid_y, int nseg, const PlanSegs& sg, int npass, const int* shift, const uint32_t* mask)
{
	if (nseg == 0) return 0;
	const int cells = (grid_x + 1) * (grid_y + 1);
	{ void* zp[16]; for (int g = 0; g < nseg; g++) zp[g] = sg.gdiff[g]; if (launch_zero_segments(stream, nseg, zp, sizeof(int) * (size_t)cells)) return -1; }
	const int ctas = max(1, min((P + DIFF_THREADS * 4 - 1) / (DIFF_THREADS * 4), 148));
	count_launch(), tile_diff_kernel<<<dim3(ctas, nseg), DIFF_THREADS, 0, stream>>>(P, cells, grid_x + 1, sg);
	count_launch(), tile_plan_kernel<<<nseg, PLAN_THREADS, 0, stream>>>(grid_x, grid_y, sg, npass, shift[0], mask[0], npass > 1 ? shift[1] : 0, npass > 1 ? mask[1] : 0u);
	return check_launch("tile_plan", stream, false);
}

int launch_export_keys(cudaStream_t stream, int T, const uint2* ranges, const uint32_t* point_list, const uint32_t* rec_words, uint64_t* out)
{
	count_launch(), export_keys_kernel<<<T, 128, 0, stream>>>(T, ranges, point_list, rec_words, out);
	return check_launch("export_keys", stream, false);
}

}  // namespace rr
