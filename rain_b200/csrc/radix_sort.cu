// Hand-written onesweep LSD radix sort (key/value pairs, stable, ascending) for sm_100a.
//
// Replaces the reference's cub::DeviceRadixSort::SortPairs calls (rasterizer_impl.cu:295-300 — u64 tile|depth keys,
// u32 Gaussian index values, bits [0, 32+bit); simple_knn.cu:199-202 — u32 Morton keys).  Structure:
//   1. one histogram kernel builds the 256-bin digit histograms of ALL passes in a single read of the keys;
//   2. a tiny kernel turns them into exclusive digit bases;
//   3. one kernel per 8-bit pass ("onesweep"): every CTA takes a tile of keys (tile ids are handed out by an atomic
//      ticket so a CTA's predecessors are always already running), ranks its keys per warp (lanes with the same digit
//      meet in a shared-memory word through atomicOr; stable: items are visited in memory order), combines the warp
//      histograms, reorders the tile through shared memory so the global scatter is written in runs, resolves its
//      global digit offsets by DECOUPLED LOOK-BACK over per-tile status words (aggregate / inclusive flags in the top
//      two bits), and writes keys and values to the other buffer.
// Memory traffic per pass: read 1x + write 1x of (key,value); histogram: read keys once.
#include <cstring>
#include "common.cuh"
#include "kernels.cuh"
#include "radix_sort.cuh"

namespace rr {

namespace {

#ifndef SORT_BALANCED_DIGITS
#define SORT_BALANCED_DIGITS 1
#endif
#ifndef SORT_LOOKBACK
#define SORT_LOOKBACK 8         // status words a look-back round loads at once
#endif
#ifndef SORT_MIN_CTAS
#define SORT_MIN_CTAS 4         // resident CTAs per SM the onesweep pass is compiled for (registers: 64 at 4)
#endif
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int MAX_PASSES = 8;
constexpr uint32_t FLAG_SHIFT = 30;
constexpr uint32_t FLAG_AGG = 1u << FLAG_SHIFT;      // tile aggregate available
constexpr uint32_t FLAG_INC = 2u << FLAG_SHIFT;      // inclusive prefix available
constexpr uint32_t VALUE_MASK = (1u << FLAG_SHIFT) - 1;

template <class K> struct Cfg;
template <> struct Cfg<uint32_t> { static constexpr int IPT = 16; };
template <> struct Cfg<uint64_t> { static constexpr int IPT = 12; };
template <> struct Cfg<uint16_t> { static constexpr int IPT = 16; };      // tile ids of the binning stage's sort

template <class K> __device__ __forceinline__ uint32_t digit_of(K key, int shift, uint32_t mask) {
	return (uint32_t)(key >> shift) & mask;
}

struct PassDesc { int shift[MAX_PASSES]; uint32_t mask[MAX_PASSES]; int npass; };

// Up to SORT_MAX_SEGS independent sorts of the same length and bit range run as ONE set of launches (blockIdx.y = segment):
// the depth sorts of all views of a step (1M keys each are latency bound at 245 CTAs per pass; eight of them together fill
// the GPU).  A plain sort is one segment.
template <class K> struct SegPtrs {
	K* keys_a[SORT_MAX_SEGS]; K* keys_b[SORT_MAX_SEGS];
	uint32_t* vals_a[SORT_MAX_SEGS]; uint32_t* vals_b[SORT_MAX_SEGS];
	uint32_t* ghist[SORT_MAX_SEGS]; uint32_t* tickets[SORT_MAX_SEGS]; uint32_t* status[SORT_MAX_SEGS];
};

// ---- 1. histograms of all passes ------------------------------------------------------------------------------
template <class K>
__global__ void __launch_bounds__(SORT_THREADS) histogram_kernel(SegPtrs<K> sp, uint32_t n, PassDesc pd)
{
	const K* __restrict__ keys = sp.keys_a[blockIdx.y];
	uint32_t* __restrict__ ghist = sp.ghist[blockIdx.y];
	constexpr int HV = 4;                       // keys in flight per thread and iteration (independent loads)
	__shared__ uint32_t sh[MAX_PASSES * RADIX];
	for (int i = threadIdx.x; i < pd.npass * RADIX; i += SORT_THREADS) sh[i] = 0;
	__syncthreads();
	const uint32_t stride = gridDim.x * SORT_THREADS * HV;
	// uniform trip count so the warp-wide votes below always see all 32 lanes
	for (uint32_t base = blockIdx.x * SORT_THREADS * HV; base < n; base += stride) {
		K key[HV];
		bool valid[HV];
		#pragma unroll
		for (int j = 0; j < HV; j++) {
			const uint32_t i = base + j * SORT_THREADS + threadIdx.x;
			valid[j] = i < n;
			key[j] = valid[j] ? keys[i] : (K)0;
		}
		#pragma unroll
		for (int p = 0; p < MAX_PASSES; p++) {
			if (p >= pd.npass) break;
			#pragma unroll
			for (int j = 0; j < HV; j++) {
				// warp-uniform digit (the usual case for the high digits of tile ids / depth exponents): one add of 32;
				// otherwise plain shared-memory atomics, which are conflict-light exactly when the digits differ
				const uint32_t d = valid[j] ? digit_of(key[j], pd.shift[p], pd.mask[p]) : RADIX;
				const uint32_t d0 = __shfl_sync(FULL, d, 0);
				if (__all_sync(FULL, d == d0)) { if (lane_id() == 0 && valid[j]) atomicAdd(&sh[p * RADIX + d0], 32u); }
				else if (valid[j]) atomicAdd(&sh[p * RADIX + d], 1u);
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < pd.npass * RADIX; i += SORT_THREADS) {
		const uint32_t v = sh[i];
		if (v) atomicAdd(&ghist[i], v);
	}
}

// ---- 2. exclusive scan of each 256-bin histogram (one CTA per pass) ---------------------------------------------
template <class K>
__global__ void __launch_bounds__(RADIX) digit_base_kernel(SegPtrs<K> sp)
{
	__shared__ uint32_t warp_tot[RADIX / 32];
	uint32_t* h = sp.ghist[blockIdx.y] + blockIdx.x * RADIX;
	const uint32_t v = h[threadIdx.x];
	uint32_t inc = v;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane_id() >= (unsigned)o) inc += t; }
	if (lane_id() == 31) warp_tot[threadIdx.x >> 5] = inc;
	__syncthreads();
	uint32_t pre = 0;
	for (int w = 0; w < (int)(threadIdx.x >> 5); w++) pre += warp_tot[w];
	h[threadIdx.x] = pre + inc - v;
}

// ---- 3. one onesweep pass ---------------------------------------------------------------------------------------
// The kernel is bound by instruction issue, not by HBM (ncu: 65 % issue-active at 11 % of the DRAM rate), so it is written
// for a short instruction stream: no per-item validity tests (a partial tile — only the last one — is padded with all-ones
// keys, which rank behind every real key of the highest digit and are dropped at the scatter), digit and rank packed into one
// register per item, one shared-memory table lookup per item at the reorder.
template <class K, int IPT>
__global__ void __launch_bounds__(SORT_THREADS, SORT_MIN_CTAS) onesweep_kernel(SegPtrs<K> sp, int pass, bool iota_vals, bool write_keys, uint32_t n, uint32_t ntiles, int shift, uint32_t mask)
{
	// this segment's buffers: even passes read the *_a buffers and write *_b, odd passes the other way round
	const int seg = blockIdx.y;
	const K* __restrict__ keys_in = (pass & 1) ? sp.keys_b[seg] : sp.keys_a[seg];
	K* __restrict__ keys_out = (pass & 1) ? sp.keys_a[seg] : sp.keys_b[seg];
	const uint32_t* __restrict__ vals_in = (pass == 0 && iota_vals) ? nullptr : ((pass & 1) ? sp.vals_b[seg] : sp.vals_a[seg]);
	uint32_t* __restrict__ vals_out = (pass & 1) ? sp.vals_a[seg] : sp.vals_b[seg];
	const uint32_t* __restrict__ digit_base = sp.ghist[seg] + pass * RADIX;                       // [RADIX] exclusive global base of each digit
	uint32_t* __restrict__ ticket = sp.tickets[seg] + pass;                                       // dynamic tile id
	volatile uint32_t* __restrict__ status = sp.status[seg] + (size_t)pass * ntiles * RADIX;      // [ntiles][RADIX]
	constexpr int TILE = SORT_THREADS * IPT;
	constexpr int WARP_ITEMS = 32 * IPT;
	constexpr int MATCH_WORDS = 2 * SORT_WARPS * RADIX;      // two digit -> lane-mask tables per warp (even / odd items)
	__shared__ uint32_t s_tile;
	__shared__ __align__(16) uint32_t s_whist[SORT_WARPS * RADIX];   // ranking: per-warp digit counts; afterwards: tile-local start of (warp, digit)
	__shared__ uint32_t s_goff[RADIX];
	__shared__ uint32_t s_wtot[SORT_WARPS];
	// the tile's keys, then its values, in tile-local sorted order; during the ranking the match tables live here
	__shared__ __align__(16) unsigned char s_buf[TILE * (sizeof(K) + sizeof(uint32_t))];
	static_assert(sizeof(s_buf) >= sizeof(uint32_t) * MATCH_WORDS, "the match tables do not fit into the reorder buffer");
	K* const s_keys = reinterpret_cast<K*>(s_buf);
	uint32_t* const s_vals = reinterpret_cast<uint32_t*>(s_buf + TILE * sizeof(K));

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	if (tid == 0) s_tile = atomicAdd(ticket, 1u);
	{
		const uint4 z = make_uint4(0u, 0u, 0u, 0u);
		#pragma unroll
		for (int i = tid; i < SORT_WARPS * RADIX / 4; i += SORT_THREADS) reinterpret_cast<uint4*>(s_whist)[i] = z;
		#pragma unroll
		for (int i = tid; i < MATCH_WORDS / 4; i += SORT_THREADS) reinterpret_cast<uint4*>(s_buf)[i] = z;
	}
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t tile_base = tile * (uint32_t)TILE;
	const uint32_t tile_count = min((uint32_t)TILE, n - tile_base);
	const bool full = tile_count == (uint32_t)TILE;
	const uint32_t my0 = (uint32_t)(warp * WARP_ITEMS + lane);       // tile-local index of this lane's item 0; item k is 32 k further

	// the values are only fetched after the ranking (registers); pull their lines into L2 now (one 128-byte line per thread)
	if (vals_in) {
		const uint32_t line = (uint32_t)tid * 32u;
		if (line < tile_count) asm volatile("prefetch.global.L2 [%0];" ::"l"(vals_in + tile_base + line));
	}
	// load (warp-striped: item k of lane l of warp w sits at w*WARP_ITEMS + k*32 + l, i.e. memory order within the warp)
	K key[IPT];
	{
		const K* __restrict__ kin = keys_in + tile_base + my0;
		if (full) {
			#pragma unroll
			for (int k = 0; k < IPT; k++) key[k] = kin[k * 32];
		} else {
			#pragma unroll
			for (int k = 0; k < IPT; k++) key[k] = (my0 + k * 32 < tile_count) ? kin[k * 32] : (K) ~(K)0;
		}
	}
	// per-warp stable ranking: pd[k] = (digit << 16) | rank of the item among the warp's items of the same digit
	uint32_t pd[IPT];
	{
		uint32_t* const wh = s_whist + warp * RADIX;
		const uint32_t lt = lanemask_lt();
		// lanes holding the same digit meet in one shared word (atomicOr of the lane bit; match.any is a slow MIO-queue
		// operation on sm_100a and a ballot per digit bit costs more issue slots and latency than it saves shared-memory
		// traffic — both measured).  Items alternate between two tables, so the word an item's first lane clears is not one
		// the next item's lanes are setting bits in.
		uint32_t* const t0 = reinterpret_cast<uint32_t*>(s_buf) + warp * RADIX;
		const uint32_t lanebit = 1u << lane;
		#pragma unroll
		for (int k = 0; k < IPT; k++) {
			const uint32_t d = digit_of(key[k], shift, mask);
			uint32_t* const t = t0 + (k & 1) * (SORT_WARPS * RADIX);
			atomicOr(&t[d], lanebit);
			__syncwarp();
			const uint32_t peers = t[d], old = wh[d];
			__syncwarp();
			const uint32_t below = peers & lt;
			if (below == 0) { t[d] = 0; wh[d] = old + __popc(peers); }       // the first lane of the group
			pd[k] = (old + __popc(below)) | (d << 16);
		}
	}
	__syncthreads();

	// thread d owns digit d: the tile's digit count is published at once (AGG), the tile-local start of every (warp, digit)
	// group follows from a block scan.  The look-back itself comes only after the reorder below, which does not need it: by
	// then the tiles just ahead of this one have published their counts too, so it polls less.
	const int d = tid;
	uint32_t sum = 0, dstart;
	{
		uint32_t cnt[SORT_WARPS];
		#pragma unroll
		for (int w = 0; w < SORT_WARPS; w++) { cnt[w] = s_whist[w * RADIX + d]; sum += cnt[w]; }
		status[(size_t)tile * RADIX + d] = sum | (tile == 0 ? FLAG_INC : FLAG_AGG);
		// exclusive scan of the tile's digit counts -> start of each digit in tile-local sorted order
		uint32_t inc = sum;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
		if (lane == 31) s_wtot[warp] = inc;
		__syncthreads();
		uint32_t pre = 0;
		for (int w = 0; w < warp; w++) pre += s_wtot[w];
		dstart = pre + inc - sum;
		uint32_t run = dstart;
		#pragma unroll
		for (int w = 0; w < SORT_WARPS; w++) { s_whist[w * RADIX + d] = run; run += cnt[w]; }
	}
	__syncthreads();

	// reorder through shared memory into tile-local sorted order: keys first (their registers are free afterwards), then the
	// values, which are fetched only now
	{
		const uint32_t* const wh = s_whist + warp * RADIX;
		#pragma unroll
		for (int k = 0; k < IPT; k++) {
			const uint32_t p = wh[pd[k] >> 16] + (pd[k] & 0xffffu);
			pd[k] = p;
			s_keys[p] = key[k];
		}
		uint32_t val[IPT];
		if (vals_in) {
			const uint32_t* __restrict__ vin = vals_in + tile_base + my0;
			if (full) {
				#pragma unroll
				for (int k = 0; k < IPT; k++) val[k] = vin[k * 32];
			} else {
				#pragma unroll
				for (int k = 0; k < IPT; k++) val[k] = (my0 + k * 32 < tile_count) ? vin[k * 32] : 0u;
			}
		} else {
			#pragma unroll
			for (int k = 0; k < IPT; k++) val[k] = tile_base + my0 + k * 32;
		}
		#pragma unroll
		for (int k = 0; k < IPT; k++) s_vals[pd[k]] = val[k];
	}
	// decoupled look-back, LB predecessors per round: the status words of a round are loaded with independent loads (one L2
	// round trip per round instead of one per predecessor), then consumed in order; at a word that is not published yet the
	// thread polls that one word (with a pause, so that it does not take issue slots from the warps it waits for) and
	// resumes from there.
	{
		uint32_t excl = 0;
		if (tile > 0 && (uint32_t)d <= mask) {                 // digits above the mask do not occur
			constexpr int LB = SORT_LOOKBACK;
			int64_t t = (int64_t)tile - 1;
			bool found = false;
			while (!found) {
				uint32_t v[LB];
				#pragma unroll
				for (int j = 0; j < LB; j++) { v[j] = 2u << FLAG_SHIFT; if (t - j >= 0) v[j] = status[(size_t)(t - j) * RADIX + d]; }   // before tile 0: inclusive 0
				int used = LB;
				#pragma unroll
				for (int j = 0; j < LB; j++) {
					const uint32_t f = v[j] >> FLAG_SHIFT;
					if (f == 0) { used = j; break; }
					excl += v[j] & VALUE_MASK;
					if (f == 2) { found = true; break; }
				}
				t -= used;
				if (!found && used < LB) {
					while ((status[(size_t)t * RADIX + d] >> FLAG_SHIFT) == 0) __nanosleep(100);
				}
			}
			status[(size_t)tile * RADIX + d] = (excl + sum) | FLAG_INC;
		}
		s_goff[d] = digit_base[d] + excl - dstart;
	}
	__syncthreads();
	// scatter: consecutive threads hold consecutive items of the sorted tile, so every digit's run is written contiguously
	if (full) {
		#pragma unroll
		for (int k = 0; k < IPT; k++) {
			const uint32_t p = k * SORT_THREADS + tid;
			const K kk = s_keys[p];
			const uint32_t g = s_goff[digit_of(kk, shift, mask)] + p;
			if (write_keys) keys_out[g] = kk;
			vals_out[g] = s_vals[p];
		}
	} else {
		#pragma unroll
		for (int k = 0; k < IPT; k++) {
			const uint32_t p = k * SORT_THREADS + tid;
			if (p < tile_count) {
				const K kk = s_keys[p];
				const uint32_t g = s_goff[digit_of(kk, shift, mask)] + p;
				if (write_keys) keys_out[g] = kk;
				vals_out[g] = s_vals[p];
			}
		}
	}
}

struct WsLayout { uint32_t* ghist; uint32_t* tickets; uint32_t* status; size_t zero_bytes; size_t total; uint32_t ntiles; };

template <class K>
WsLayout ws_layout(void* ws, size_t n, int npass) {
	constexpr int TILE = SORT_THREADS * Cfg<K>::IPT;
	WsLayout L;
	L.ntiles = (uint32_t)((n + TILE - 1) / TILE);
	Carver c(ws);
	L.ghist = c.take<uint32_t>((size_t)MAX_PASSES * RADIX);
	L.tickets = c.take<uint32_t>(MAX_PASSES);
	L.status = c.take<uint32_t>((size_t)npass * L.ntiles * RADIX);
	L.zero_bytes = c.off;
	L.total = c.used();
	return L;
}

PassDesc make_passes(int begin_bit, int end_bit) {
	PassDesc pd;
	pd.npass = 0;
#if SORT_BALANCED_DIGITS
	// the bits are split evenly over the passes (13 tile bits: 7 + 6 instead of 8 + 5): a pass with fewer bins scatters in
	// longer runs, and the wide pass is the slow one
	const int total = (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS;
	for (int b = begin_bit; b < end_bit && pd.npass < MAX_PASSES;) {
		const int left = total - pd.npass;
		const int nb = (end_bit - b + left - 1) / left;
		pd.shift[pd.npass] = b;
		pd.mask[pd.npass] = (1u << nb) - 1;
		pd.npass++;
		b += nb;
	}
#else
	for (int b = begin_bit; b < end_bit && pd.npass < MAX_PASSES; b += RADIX_BITS) {
		const int nb = (end_bit - b) < RADIX_BITS ? (end_bit - b) : RADIX_BITS;
		pd.shift[pd.npass] = b;
		pd.mask[pd.npass] = (1u << nb) - 1;
		pd.npass++;
	}
#endif
	return pd;
}

}  // namespace

template <class K>
size_t sort_ws_bytes(size_t n, int begin_bit, int end_bit) {
	const PassDesc pd = make_passes(begin_bit, end_bit);
	return ws_layout<K>(nullptr, n, pd.npass).total + 256;
}

template <class K>
int sort_pairs_segments(cudaStream_t stream, int nseg, K* const* keys_a, K* const* keys_b, uint32_t* const* vals_a, uint32_t* const* vals_b, size_t n,
	int begin_bit, int end_bit, void* const* ws, bool iota_vals)
{
	if (n == 0 || nseg == 0) return 0;
	if (nseg < 0 || nseg > SORT_MAX_SEGS) { set_error("sort_pairs: %d segments (at most %d)", nseg, SORT_MAX_SEGS); return -1; }
	if (n >= (size_t)VALUE_MASK) { set_error("sort_pairs: n=%zu exceeds the 2^30-1 item limit of the look-back status word", n); return -1; }
	if (end_bit <= begin_bit || end_bit > (int)sizeof(K) * 8 || (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS > MAX_PASSES) {
		set_error("sort_pairs: bad bit range [%d,%d)", begin_bit, end_bit); return -1;
	}
	constexpr int IPT = Cfg<K>::IPT;
	const PassDesc pd = make_passes(begin_bit, end_bit);
	SegPtrs<K> sp;
	memset(&sp, 0, sizeof(sp));
	uint32_t ntiles = 0;
	size_t zero_bytes = 0;
	for (int g = 0; g < nseg; g++) {
		const WsLayout L = ws_layout<K>(ws[g], n, pd.npass);
		zero_bytes = L.zero_bytes;
		sp.keys_a[g] = keys_a[g]; sp.keys_b[g] = keys_b[g]; sp.vals_a[g] = vals_a[g]; sp.vals_b[g] = vals_b[g];
		sp.ghist[g] = L.ghist; sp.tickets[g] = L.tickets; sp.status[g] = L.status;
		ntiles = L.ntiles;
	}
	if (launch_zero_segments(stream, nseg, ws, zero_bytes)) return -1;
	const uint32_t hgrid = (uint32_t)((n + SORT_THREADS * 16 - 1) / (SORT_THREADS * 16));
	const uint32_t hcap = nseg > 1 ? (1184u + nseg - 1) / nseg : 1184u;
	count_launch(), histogram_kernel<K><<<dim3(hgrid < hcap ? hgrid : hcap, nseg), SORT_THREADS, 0, stream>>>(sp, (uint32_t)n, pd);
	count_launch(), digit_base_kernel<K><<<dim3(pd.npass, nseg), RADIX, 0, stream>>>(sp);
	for (int p = 0; p < pd.npass; p++) {
		count_launch(), onesweep_kernel<K, IPT><<<dim3(ntiles, nseg), SORT_THREADS, 0, stream>>>(sp, p, iota_vals, true, (uint32_t)n, ntiles, pd.shift[p], pd.mask[p]);
	}
	if (check_launch("sort_pairs", stream, false)) return -1;
	return pd.npass & 1;   // odd number of passes -> result sits in the *_b buffers
}

// ---- the tile sort of the binning stage: 16-bit tile ids, digit bases and ranges from the tile plan, value-only last pass ----
size_t tile_sort_ws_bytes(size_t n, int bits) { return sort_ws_bytes<uint16_t>(n, 0, bits); }

int tile_sort_passes(int bits, int* shift, uint32_t* mask)
{
	const PassDesc pd = make_passes(0, bits);
	for (int p = 0; p < pd.npass && p < 2; p++) { shift[p] = pd.shift[p]; mask[p] = pd.mask[p]; }
	return pd.npass;
}

int tile_sort16(cudaStream_t stream, int P, size_t N, int bits, int grid_x, const uint32_t* order, const uint32_t* offsets, const uint2* rect,
	const uint32_t* digit_bases, uint16_t* keys_a, uint16_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, void* ws)
{
	if (bits < 1 || bits > 16) { set_error("tile_sort16: %d tile-id bits (1..16)", bits); return -1; }
	if (N >= (size_t)VALUE_MASK) { set_error("tile_sort16: N=%zu exceeds the 2^30-1 item limit", N); return -1; }
	const PassDesc pd = make_passes(0, bits);                       // 1 or 2 passes
	const WsLayout L = ws_layout<uint16_t>(ws, N, pd.npass);
	cudaMemsetAsync(ws, 0, L.zero_bytes, stream);
	// the exclusive digit bases of both passes (and the ranges) come from the tile plan of the geometry half (binning.cu): no
	// histogram pass over the pairs, no sorted key array, no range-detection pass
	if (N == 0) return pd.npass & 1;
	// duplicateWithKeys (rasterizer_impl.cu:59-100) in depth order with 2-byte tile ids: 6 bytes per pair instead of 12
	if (launch_duplicate16(stream, P, order, offsets, rect, grid_x, keys_a, vals_a)) return -1;
	SegPtrs<uint16_t> sp;
	memset(&sp, 0, sizeof(sp));
	sp.keys_a[0] = keys_a; sp.keys_b[0] = keys_b; sp.vals_a[0] = vals_a; sp.vals_b[0] = vals_b;
	sp.ghist[0] = const_cast<uint32_t*>(digit_bases); sp.tickets[0] = L.tickets; sp.status[0] = L.status;
	constexpr int IPT = Cfg<uint16_t>::IPT;
	for (int p = 0; p < pd.npass; p++) {
		const bool wk = p + 1 < pd.npass;                             // the last pass writes the Gaussian ids (point_list) only
		count_launch(), onesweep_kernel<uint16_t, IPT><<<dim3(L.ntiles, 1), SORT_THREADS, 0, stream>>>(sp, p, false, wk, (uint32_t)N, L.ntiles, pd.shift[p], pd.mask[p]);
	}
	if (check_launch("tile_sort16", stream, false)) return -1;
	return pd.npass & 1;   // 1: the sorted ids are in vals_b, 0: in vals_a
}

template <class K>
int sort_pairs(cudaStream_t stream, K* keys_a, K* keys_b, uint32_t* vals_a, uint32_t* vals_b, size_t n,
	int begin_bit, int end_bit, void* ws, bool iota_vals)
{
	return sort_pairs_segments<K>(stream, 1, &keys_a, &keys_b, &vals_a, &vals_b, n, begin_bit, end_bit, &ws, iota_vals);
}

template size_t sort_ws_bytes<uint32_t>(size_t, int, int);
template size_t sort_ws_bytes<uint64_t>(size_t, int, int);
template int sort_pairs<uint32_t>(cudaStream_t, uint32_t*, uint32_t*, uint32_t*, uint32_t*, size_t, int, int, void*, bool);
template int sort_pairs<uint64_t>(cudaStream_t, uint64_t*, uint64_t*, uint32_t*, uint32_t*, size_t, int, int, void*, bool);
template int sort_pairs_segments<uint32_t>(cudaStream_t, int, uint32_t* const*, uint32_t* const*, uint32_t* const*, uint32_t* const*, size_t, int, int, void* const*, bool);

}  // namespace rr

// ---- rr_sort_pairs_u32_noalloc: the u32 pair sort with its workspace in a library-owned device array ----------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the same kernels as
// rr_sort_pairs_u32, handed a file-scope device array instead of a caller-supplied workspace.  The array's name is only a
// host-side handle here: its device address is looked up once per device with cudaGetSymbolAddress (the device current at the
// call) and that address is what the kernels get.  No cudaMalloc / cudaFree / allocating library call on this path.
namespace rr {
namespace {

constexpr size_t NOALLOC_MAX_PAIRS = (size_t)1 << 22;                      // 4 194 304 pairs
// histograms + tickets (MAX_PASSES x 256 words, padded) + one 256-word status row per 4096-key tile and pass (4 passes of a
// 32-bit key): 4 MiB + 16 KiB, rounded up; sort_noalloc checks the real figure against sizeof at every call
constexpr size_t NOALLOC_WS_BYTES = 4 * (NOALLOC_MAX_PAIRS / 4096) * RADIX * sizeof(uint32_t) + (64u << 10);
static __device__ __align__(256) unsigned char g_noalloc_ws[NOALLOC_WS_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES];       // device address of g_noalloc_ws on device i (null: not looked up yet)

void* noalloc_ws_of_current_device()
{
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) { set_error("sort_noalloc: cudaGetDevice: %s", cudaGetErrorString(e)); return nullptr; }
	if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { set_error("sort_noalloc: device %d (at most %d devices)", dev, NOALLOC_MAX_DEVICES); return nullptr; }
	void* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
	if (p) return p;
	e = cudaGetSymbolAddress(&p, g_noalloc_ws);           // this device's instance of the array; same answer from every thread
	if (e != cudaSuccess || !p) { set_error("sort_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return nullptr; }
	__atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
	return p;
}

}  // namespace
}  // namespace rr

extern "C" size_t rr_sort_noalloc_max_pairs(void) { return rr::NOALLOC_MAX_PAIRS; }

extern "C" int rr_sort_pairs_u32_noalloc(void* stream_, int64_t n, uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_in, uint32_t* vals_out,
	int begin_bit, int end_bit)
{
	using namespace rr;
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (n < 0) { set_error("sort_noalloc: bad pair count %lld", (long long)n); return -1; }
	if (n == 0) return 0;
	if ((size_t)n > NOALLOC_MAX_PAIRS) {
		set_error("sort_noalloc: %lld pairs exceed the %zu-pair limit of the built-in workspace (use rr_sort_pairs_u32 with your own)", (long long)n, NOALLOC_MAX_PAIRS);
		return -1;
	}
	if (!keys_in || !keys_out || !vals_in || !vals_out) { set_error("sort_noalloc: keys_in, keys_out, vals_in and vals_out must not be null"); return -1; }
	if (end_bit <= begin_bit || begin_bit < 0 || end_bit > 32) { set_error("sort_noalloc: bad bit range [%d,%d)", begin_bit, end_bit); return -1; }
	if (sort_ws_bytes<uint32_t>((size_t)n, begin_bit, end_bit) > NOALLOC_WS_BYTES) {
		set_error("sort_noalloc: workspace of %zu bytes exceeds the built-in %zu", sort_ws_bytes<uint32_t>((size_t)n, begin_bit, end_bit), NOALLOC_WS_BYTES);
		return -1;
	}
	void* ws = noalloc_ws_of_current_device();
	if (!ws) return -1;
	// sort_pairs clears the histogram / ticket / status words of this call on the stream before its first kernel, so what the
	// previous call left in the array does not matter
	const int which = sort_pairs<uint32_t>(stream, keys_in, keys_out, vals_in, vals_out, (size_t)n, begin_bit, end_bit, ws, false);
	if (which < 0) return -1;
	if (which == 0) {
		cudaMemcpyAsync(keys_out, keys_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
		cudaMemcpyAsync(vals_out, vals_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
	}
	return check_launch("sort_pairs_noalloc", stream, false);
}
