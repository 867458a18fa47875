// Tile binning WITHOUT a sort over the (tile, Gaussian) pairs.
// Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges of the reference
// (rasterizer_impl.cu:59-127, 295-302) — and this library's own duplicate / onesweep tile sort / ranges passes — by a
// counting placement that exploits what is already known: the Gaussians are in depth order (depth sort over P keys), so
// the list of tile t in (tile, depth, index) order is simply "the Gaussians whose rect covers t, in that order".
//
//   1. tile_count: the depth-ordered Gaussians are cut into C contiguous chunks (C <= one resident wave of the
//      placement kernel).  One CTA per chunk histograms its chunk's (Gaussian, tile) pairs over the tiles in shared
//      memory (warp-cooperative rect enumeration) and writes one row of 16-bit counts: cnt[c][t].
//   2. seg_sum / tile_offsets / chunk_base: column-wise exclusive scan of cnt over the chunks, in 32 segments per
//      column so that 32 x T threads share the work: base[c][t] = start of tile t's list + number of pairs of tile t in
//      chunks before c.  The per-tile totals give ranges[t] directly (identifyTileRanges for free; empty tiles (0,0)).
//   3. tile_place: one WARP per chunk walks its Gaussians in depth order with the chunk's row of base[] in shared
//      memory as running write positions: 32 consecutive pairs per step (same enumeration as step 1); pairs of one step
//      that fall into the same tile are ranked by lane (__match_any_sync — emission order is (Gaussian, tile) order, so a
//      lower lane is an earlier Gaussian); writes point_list[pos] = Gaussian id.  The order inside a tile is the depth
//      order, ties by ascending index: bit-identical to the reference's stable 64-bit sort.
//
// STATUS: opt-in (RAIN_BINNING=count), NOT the default.  It produces the same point_list / ranges bit for bit (the GPU parity
// suite passes with it), but measured on B200 at C3 (8.4M pairs) it is slower than the path it was meant to replace:
//   tile_count 45 us + seg_sum 9 + tile_offsets 23 + chunk_base 12 + tile_place 290-400 us  vs  duplicate 33 + tile sort 197 + ranges 31.
// The placement is the problem: it issues only 52M warp instructions but sits at 29 % issue utilisation — every 32-Gaussian
// batch starts with a chain of three dependent gathers (order -> tiles_touched -> rect), the per-step work is a chain of
// dependent shared-memory accesses (binary search, match, read-modify-write), and the number of independent warps is capped by
// the count matrix (C x T x 6 B).  One warp per chunk: 405 us; eight band-parallel warps per chunk: 290-324 us.
//
// Traffic per render: 6 B x C x T for the count / base matrices (42 MB at 1080p, independent of N) + one 4-byte write
// per pair that lands in L2 (point_list is 34 MB at 8.4M pairs) — against 36 B per pair for the two onesweep passes it
// replaces plus 12 B per pair for duplicate and ranges.
#include "common.cuh"
#include "kernels.cuh"

namespace rr {

namespace {

constexpr int CB_SEGS = 32;             // column-scan segments per tile

// Enumerate the (Gaussian, tile) pairs of 32 consecutive depth-ordered Gaussians cooperatively: pair s of the warp's
// span is owned by the last Gaussian whose exclusive pair offset is <= s (binary search over 32 shared values), so a
// large splat does not serialise on one lane.  f(s, lane-local Gaussian slot j, tile) is called with s in emission order
// per lane stride; all 32 lanes call it the same number of times (tile = 0xffffffff past the end).
struct WarpBatch {
	uint32_t excl[33];      // exclusive pair offsets of the 32 Gaussians, [32] = total
	uint32_t g[32];
	uint32_t xw[32];        // x0 | width << 16
	uint32_t y0[32];
};

__device__ __forceinline__ uint32_t load_batch(WarpBatch& b, int first, int P, int lane, const uint32_t* __restrict__ order,
	const uint32_t* __restrict__ tiles_touched, const uint2* __restrict__ rect)
{
	const int i = first + lane;
	uint32_t g = 0, tt = 0, x0 = 0, wdt = 1, y0 = 0;
	if (i < P) {
		g = __ldg(order + i);
		tt = __ldg(tiles_touched + g);
		if (tt > 0) { const uint2 r = __ldg(rect + g); x0 = r.x & 0xffffu; wdt = (r.x >> 16) - x0; y0 = r.y & 0xffffu; }
	}
	uint32_t inc = tt;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
	b.excl[lane] = inc - tt; b.g[lane] = g; b.xw[lane] = x0 | (wdt << 16); b.y0[lane] = y0;
	const uint32_t total = __shfl_sync(FULL, inc, 31);
	if (lane == 0) b.excl[32] = total;
	__syncwarp();
	return total;
}

__device__ __forceinline__ void pair_of(const WarpBatch& b, uint32_t s, int grid_x, uint32_t& j, uint32_t& tile)
{
	int lo = 0;
	#pragma unroll
	for (int step = 16; step >= 1; step >>= 1) if (b.excl[lo + step] <= s) lo += step;
	const uint32_t local = s - b.excl[lo];
	const uint32_t xw = b.xw[lo], w = xw >> 16, xx0 = xw & 0xffffu;
	const uint32_t ry = local / w, rx = local - ry * w;
	j = (uint32_t)lo;
	tile = (b.y0[lo] + ry) * (uint32_t)grid_x + xx0 + rx;
}

// ---- 1. per-chunk tile histogram ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tile_count_kernel(int P, int T, int grid_x, int chunk, const uint32_t* __restrict__ order,
	const uint32_t* __restrict__ tiles_touched, const uint2* __restrict__ rect, uint16_t* __restrict__ cnt)
{
	extern __shared__ uint32_t s_hist[];                 // [T]
	__shared__ WarpBatch s_b[8];
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	for (int t = tid; t < T; t += 256) s_hist[t] = 0;
	__syncthreads();
	const int c0 = blockIdx.x * chunk, c1 = min(P, c0 + chunk);
	for (int first = c0 + warp * 32; first < c1; first += 256) {
		const uint32_t total = load_batch(s_b[warp], first, c1, lane, order, tiles_touched, rect);
		for (uint32_t s = lane; s < total; s += 32) {
			uint32_t j, tile;
			pair_of(s_b[warp], s, grid_x, j, tile);
			atomicAdd(&s_hist[tile], 1u);
		}
		__syncwarp();
	}
	__syncthreads();
	uint16_t* __restrict__ row = cnt + (size_t)blockIdx.x * T;
	for (int t = tid; t < T; t += 256) row[t] = (uint16_t)s_hist[t];      // <= chunk <= 65535 pairs of one tile per chunk
}

// ---- 2. column scan over the chunks ------------------------------------------------------------------------------
// block (32 tiles, 32 segments): segsum[s][t] = sum of cnt[c][t] over the chunks of segment s
__global__ void __launch_bounds__(1024) seg_sum_kernel(int T, int C, int seg_len, const uint16_t* __restrict__ cnt, uint32_t* __restrict__ segsum)
{
	const int t = blockIdx.x * 32 + threadIdx.x, s = threadIdx.y;
	if (t >= T) return;
	const int c0 = s * seg_len, c1 = min(C, c0 + seg_len);
	uint32_t sum = 0;
	for (int c = c0; c < c1; c++) sum += cnt[(size_t)c * T + t];
	segsum[(size_t)s * T + t] = sum;
}

// one CTA: per-tile totals -> exclusive scan over the tiles -> ranges (their .x is the tile's list start)
__global__ void __launch_bounds__(1024) tile_offsets_kernel(int T, const uint32_t* __restrict__ segsum, uint2* __restrict__ ranges, uint2* __restrict__ tile_span,
	uint32_t* __restrict__ total_out)
{
	__shared__ uint32_t s_warp[32];
	__shared__ uint32_t s_carry;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (tid == 0) s_carry = 0;
	__syncthreads();
	for (int t0 = 0; t0 < T; t0 += 1024) {
		const int t = t0 + tid;
		uint32_t tot = 0;
		if (t < T) {
			#pragma unroll
			for (int s = 0; s < CB_SEGS; s++) tot += __ldg(segsum + (size_t)s * T + t);
		}
		uint32_t inc = tot;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += v; }
		if (lane == 31) s_warp[warp] = inc;
		__syncthreads();
		uint32_t pre = 0;
		for (int w = 0; w < warp; w++) pre += s_warp[w];
		const uint32_t start = s_carry + pre + inc - tot;
		if (t < T) {
			ranges[t] = tot ? make_uint2(start, start + tot) : make_uint2(0u, 0u);     // empty tiles stay (0,0) (rasterizer_impl.cu:302)
			tile_span[t] = make_uint2(start, start + tot);
		}
		__syncthreads();
		if (tid == 1023) s_carry = start + tot;
		__syncthreads();
	}
	if (tid == 0 && total_out) *total_out = s_carry;
}

// block (32 tiles, 32 segments): base[c][t] = start of tile t + pairs of tile t in earlier segments + in the segment's chunks before c
__global__ void __launch_bounds__(1024) chunk_base_kernel(int T, int C, int seg_len, const uint16_t* __restrict__ cnt, const uint32_t* __restrict__ segsum,
	const uint2* __restrict__ tile_span, uint32_t* __restrict__ base)
{
	const int t = blockIdx.x * 32 + threadIdx.x, s = threadIdx.y;
	if (t >= T) return;
	const int c0 = s * seg_len, c1 = min(C, c0 + seg_len);
	uint32_t run = tile_span[t].x;
	#pragma unroll
	for (int sp = 0; sp < CB_SEGS; sp++) { const uint32_t v = __ldg(segsum + (size_t)sp * T + t); if (sp < s) run += v; }
	for (int c = c0; c < c1; c++) { base[(size_t)c * T + t] = run; run += cnt[(size_t)c * T + t]; }
}

// ---- 3. placement: one CTA per chunk, one warp per band of tile rows ---------------------------------------------
// Order only matters inside a tile, so the 8 warps of a chunk's CTA work independently, each on the tile rows
// [band * rows_per_band, ...) : a warp walks the chunk's Gaussians in depth order, clips every rect to its band, and places
// 32 pairs per step with its slice of the chunk's base[] row in shared memory as running write positions.
constexpr int CB_BANDS = 8;

__global__ void __launch_bounds__(32 * CB_BANDS) tile_place_kernel(int P, int T, int grid_x, int rows_per_band, int chunk, const uint32_t* __restrict__ order,
	const uint32_t* __restrict__ tiles_touched, const uint2* __restrict__ rect, const uint32_t* __restrict__ base, uint32_t* __restrict__ point_list)
{
	extern __shared__ uint32_t s_pos_all[];              // [CB_BANDS][rows_per_band * grid_x] running write positions
	__shared__ WarpBatch s_b[CB_BANDS];
	const int lane = threadIdx.x & 31, band = threadIdx.x >> 5;
	const int band_tiles = rows_per_band * grid_x;
	const int t0 = band * band_tiles, t1 = min(T, t0 + band_tiles);
	if (t0 >= T) return;                                  // warp-uniform: no such band in this image
	const uint32_t ya_band = (uint32_t)(band * rows_per_band), yb_band = ya_band + (uint32_t)rows_per_band;
	uint32_t* __restrict__ s_pos = s_pos_all + band * band_tiles;
	const uint32_t* __restrict__ row = base + (size_t)blockIdx.x * T;
	for (int t = t0 + lane; t < t1; t += 32) s_pos[t - t0] = __ldg(row + t);
	__syncwarp();
	WarpBatch& b = s_b[band];
	const int c0 = blockIdx.x * chunk, c1 = min(P, c0 + chunk);
	for (int first = c0; first < c1; first += 32) {
		// this warp's view of the batch: rects clipped to the band's rows
		const int i = first + lane;
		uint32_t g = 0, tt = 0, x0 = 0, wdt = 1, ya = 0;
		if (i < c1) {
			g = __ldg(order + i);
			if (__ldg(tiles_touched + g) > 0) {
				const uint2 r = __ldg(rect + g);
				x0 = r.x & 0xffffu; wdt = (r.x >> 16) - x0;
				ya = max(r.y & 0xffffu, ya_band);
				const uint32_t yb = min(r.y >> 16, yb_band);
				tt = yb > ya ? wdt * (yb - ya) : 0u;
			}
		}
		uint32_t inc = tt;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += v; }
		const uint32_t total = __shfl_sync(FULL, inc, 31);
		if (total == 0) continue;                        // warp-uniform
		b.excl[lane] = inc - tt; b.g[lane] = g; b.xw[lane] = x0 | (wdt << 16); b.y0[lane] = ya;
		if (lane == 0) b.excl[32] = total;
		__syncwarp();
		for (uint32_t s0 = 0; s0 < total; s0 += 32) {
			const uint32_t s = s0 + lane;
			const bool valid = s < total;
			uint32_t j = 0, tile = 0xffffffffu;
			if (valid) { pair_of(b, s, grid_x, j, tile); tile -= (uint32_t)t0; }
			// pairs of this step in the same tile: earlier lane = earlier Gaussian (a Gaussian covers a tile at most once)
			const unsigned peers = __match_any_sync(FULL, tile);
			const uint32_t old = valid ? s_pos[tile] : 0u;
			__syncwarp();
			if (valid) {
				if (lane == __ffs(peers) - 1) s_pos[tile] = old + __popc(peers);
				point_list[old + __popc(peers & lanemask_lt())] = b.g[j];
			}
			__syncwarp();
		}
	}
}

// the 64-bit (tile << 32 | depth bits) key list the reference would hold, reconstructed from ranges + point_list (parity tests)
__global__ void __launch_bounds__(128) export_keys_ranges_kernel(int T, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
	const uint32_t* __restrict__ rec_words, uint64_t* __restrict__ out)
{
	const int t = blockIdx.x;
	if (t >= T) return;
	const uint2 r = ranges[t];
	for (uint32_t j = r.x + threadIdx.x; j < r.y; j += 128)
		out[j] = ((uint64_t)(uint32_t)t << 32) | rec_words[(size_t)point_list[j] * 12 + 9];
}

}  // namespace

int count_binning_chunks(int P, int* chunk_out)
{
	// one resident wave of single-warp placement CTAs: 6 per SM (32 KB of shared memory each at 8 160 tiles)
	int dev = 0, sms = 148;
	if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	const int cmax = sms * 6 < CB_MAX_CHUNKS ? sms * 6 : CB_MAX_CHUNKS;
	int chunk = (P + cmax - 1) / cmax;
	chunk = (chunk + 31) / 32 * 32;
	if (chunk < 32) chunk = 32;
	if (chunk > 65504) chunk = 65504;                    // 16-bit counts; more Gaussians than 58M would need more chunks than CB_MAX_CHUNKS
	*chunk_out = chunk;
	return (P + chunk - 1) / chunk;
}

size_t count_binning_ws_bytes(int T)
{
	return align_up((size_t)CB_MAX_CHUNKS * T * sizeof(uint16_t), 256) + align_up((size_t)CB_MAX_CHUNKS * T * sizeof(uint32_t), 256) +
		align_up((size_t)CB_SEGS * T * sizeof(uint32_t), 256) + align_up((size_t)T * sizeof(uint2), 256) + 256;
}

int launch_count_binning(cudaStream_t stream, int P, int T, int grid_x, const uint32_t* order, const uint32_t* tiles_touched, const uint2* rect,
	void* ws, uint2* ranges, uint32_t* point_list)
{
	int chunk;
	const int C = count_binning_chunks(P, &chunk);
	if (C > CB_MAX_CHUNKS || T > CB_MAX_TILES) { set_error("count binning: %d chunks / %d tiles exceed the limits", C, T); return -1; }
	char* w = reinterpret_cast<char*>(ws);
	uint16_t* cnt = reinterpret_cast<uint16_t*>(w); w += align_up((size_t)CB_MAX_CHUNKS * T * sizeof(uint16_t), 256);
	uint32_t* base = reinterpret_cast<uint32_t*>(w); w += align_up((size_t)CB_MAX_CHUNKS * T * sizeof(uint32_t), 256);
	uint32_t* segsum = reinterpret_cast<uint32_t*>(w); w += align_up((size_t)CB_SEGS * T * sizeof(uint32_t), 256);
	uint2* tile_span = reinterpret_cast<uint2*>(w);
	const int smem = T * (int)sizeof(uint32_t);
	static const bool attr_ok = []() {
		// the carveout matters: with the default preference only one or two of these small CTAs are resident per SM
		return cudaFuncSetAttribute(tile_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CB_MAX_TILES * 4) == cudaSuccess &&
			cudaFuncSetAttribute(tile_place_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CB_MAX_TILES * 4 + 8192) == cudaSuccess &&
			cudaFuncSetAttribute(tile_count_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) == cudaSuccess &&
			cudaFuncSetAttribute(tile_place_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) == cudaSuccess;
	}();
	if (!attr_ok) { set_error("count binning: cannot reserve dynamic shared memory"); return -1; }
	const int seg_len = (C + CB_SEGS - 1) / CB_SEGS;
	const dim3 cb(32, CB_SEGS);
	count_launch(), tile_count_kernel<<<C, 256, smem, stream>>>(P, T, grid_x, chunk, order, tiles_touched, rect, cnt);
	count_launch(), seg_sum_kernel<<<(T + 31) / 32, cb, 0, stream>>>(T, C, seg_len, cnt, segsum);
	count_launch(), tile_offsets_kernel<<<1, 1024, 0, stream>>>(T, segsum, ranges, tile_span, nullptr);
	count_launch(), chunk_base_kernel<<<(T + 31) / 32, cb, 0, stream>>>(T, C, seg_len, cnt, segsum, tile_span, base);
	const int grid_y = (T + grid_x - 1) / grid_x, rows_per_band = (grid_y + CB_BANDS - 1) / CB_BANDS;
	count_launch(), tile_place_kernel<<<C, 32 * CB_BANDS, CB_BANDS * rows_per_band * grid_x * (int)sizeof(uint32_t), stream>>>(P, T, grid_x, rows_per_band, chunk, order,
		tiles_touched, rect, base, point_list);
	return check_launch("count_binning", stream, false);
}

int launch_export_keys_ranges(cudaStream_t stream, int T, const uint2* ranges, const uint32_t* point_list, const uint32_t* rec_words, uint64_t* out)
{
	count_launch(), export_keys_ranges_kernel<<<T, 128, 0, stream>>>(T, ranges, point_list, rec_words, out);
	return check_launch("export_keys", stream, false);
}

}  // namespace rr
