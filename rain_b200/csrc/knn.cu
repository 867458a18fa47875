// distCUDA2: mean squared distance of every point to its 3 nearest other points.
// Replaces SimpleKNN::knn (reference simple-knn/simple_knn.cu:174-210: CustomMin/Max reduce :18-32,182-189,
// coord2Morton :34-59, boxMinMax :67-106, boxMeanDist :136-172).
//
// Same plan as the reference — Morton order, boxes of consecutive points, exact pruned 3-NN — redesigned:
//  * no host round trips (the reference copies min and max to the host twice and cudaMallocs per call): the bounding
//    box stays in device memory and the caller owns the workspace;
//  * the Morton sort is the hand-written onesweep of radix_sort.cu;
//  * points are gathered into Morton order once (float4: xyz + original index), so box contents are contiguous;
//  * boxes hold 128 points (reference: 1024) and carry a 3-level AABB hierarchy (box / 32 boxes / 1024 boxes), so a
//    query prunes with O(log) AABB tests instead of testing every box (reference: ceil(P/1024) tests per point);
//  * each CTA is one box: it scans its own box from shared memory first, which gives a tight initial bound.
// The result is the exact 3-NN value: squared distances use the reference build's rounding,
// fma(dz,dz, fma(dx,dx, dy*dy)) with d = other - self, and the mean is ((b0 + b1) + b2) / 3 with b ascending.
#include <cfloat>
#include "common.cuh"
#include "kernels.cuh"
#include "radix_sort.cuh"

namespace rr {

namespace {

constexpr int BOX = 128;
constexpr int FAN = 32;

struct Aabb { float lo[3], hi[3]; };

__device__ __forceinline__ void insert3(float d, float& b0, float& b1, float& b2) {
	// updateKBest<3>: keep the three smallest, ascending
	if (b0 > d) { const float t = b0; b0 = d; d = t; }
	if (b1 > d) { const float t = b1; b1 = d; d = t; }
	if (b2 > d) { b2 = d; }
}

__device__ __forceinline__ float sqdist(float ox, float oy, float oz, float mx, float my, float mz) {
	const float dx = __fsub_rn(ox, mx), dy = __fsub_rn(oy, my), dz = __fsub_rn(oz, mz);
	return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// squared distance from p to an AABB (0 inside); scaled down so rounding can never prune a true neighbour
__device__ __forceinline__ float box_dist(const Aabb& b, float x, float y, float z) {
	const float dx = fmaxf(fmaxf(b.lo[0] - x, x - b.hi[0]), 0.0f);
	const float dy = fmaxf(fmaxf(b.lo[1] - y, y - b.hi[1]), 0.0f);
	const float dz = fmaxf(fmaxf(b.lo[2] - z, z - b.hi[2]), 0.0f);
	return (dx * dx + dy * dy + dz * dz) * (1.0f - 1e-6f);
}

__global__ void __launch_bounds__(256) minmax_partial_kernel(int P, const float* __restrict__ pts, float* __restrict__ partial /*[grid][6]*/)
{
	float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
		#pragma unroll
		for (int a = 0; a < 3; a++) { const float v = pts[3 * (size_t)i + a]; lo[a] = fminf(lo[a], v); hi[a] = fmaxf(hi[a], v); }
	}
	__shared__ float s[8][6];
	#pragma unroll
	for (int a = 0; a < 3; a++) {
		#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) { lo[a] = fminf(lo[a], __shfl_xor_sync(FULL, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(FULL, hi[a], o)); }
	}
	if (lane_id() == 0) { for (int a = 0; a < 3; a++) { s[threadIdx.x >> 5][a] = lo[a]; s[threadIdx.x >> 5][3 + a] = hi[a]; } }
	__syncthreads();
	if (threadIdx.x < 6) {
		float v = s[0][threadIdx.x];
		for (int w = 1; w < 8; w++) v = threadIdx.x < 3 ? fminf(v, s[w][threadIdx.x]) : fmaxf(v, s[w][threadIdx.x]);
		partial[blockIdx.x * 6 + threadIdx.x] = v;
	}
}

// final reduce; the reference seeds both reductions with {0,0,0} (simple_knn.cu:180-188), i.e. the Morton cube always
// contains the origin — reproduced so the Morton order is the reference's
__global__ void minmax_final_kernel(int nblocks, const float* __restrict__ partial, float* __restrict__ out /*[6]*/)
{
	if (threadIdx.x < 6) {
		float v = 0.0f;
		for (int b = 0; b < nblocks; b++) v = threadIdx.x < 3 ? fminf(v, partial[b * 6 + threadIdx.x]) : fmaxf(v, partial[b * 6 + threadIdx.x]);
		out[threadIdx.x] = v;
	}
}

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {
	x = (x | (x << 16)) & 0x030000FFu;
	x = (x | (x << 8)) & 0x0300F00Fu;
	x = (x | (x << 4)) & 0x030C30C3u;
	x = (x | (x << 2)) & 0x09249249u;
	return x;
}

__global__ void __launch_bounds__(256) morton_kernel(int P, const float* __restrict__ pts, const float* __restrict__ mm, uint32_t* __restrict__ codes)
{
	const int i = blockIdx.x * 256 + threadIdx.x;
	if (i >= P) return;
	uint32_t c[3];
	#pragma unroll
	for (int a = 0; a < 3; a++) {
		const float v = __fmul_rn(__fdiv_rn(__fsub_rn(pts[3 * (size_t)i + a], mm[a]), __fsub_rn(mm[3 + a], mm[a])), 1023.0f);
		c[a] = prep_morton((uint32_t)v);
	}
	codes[i] = c[0] | (c[1] << 1) | (c[2] << 2);
}

__global__ void __launch_bounds__(256) gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order, float4* __restrict__ sorted)
{
	const int i = blockIdx.x * 256 + threadIdx.x;
	if (i >= P) return;
	const uint32_t g = order[i];
	sorted[i] = make_float4(pts[3 * (size_t)g], pts[3 * (size_t)g + 1], pts[3 * (size_t)g + 2], __uint_as_float(g));
}

// one warp per box of BOX consecutive sorted points
__global__ void __launch_bounds__(256) box_aabb_kernel(int P, int nboxes, const float4* __restrict__ sorted, Aabb* __restrict__ boxes)
{
	const int box = blockIdx.x * 8 + (threadIdx.x >> 5);
	if (box >= nboxes) return;
	float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	for (int k = lane_id(); k < BOX; k += 32) {
		const int i = box * BOX + k;
		if (i < P) { const float4 p = sorted[i]; lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x); lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y); lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z); }
	}
	#pragma unroll
	for (int a = 0; a < 3; a++) {
		#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) { lo[a] = fminf(lo[a], __shfl_xor_sync(FULL, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(FULL, hi[a], o)); }
	}
	if (lane_id() == 0) { Aabb b; for (int a = 0; a < 3; a++) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; } boxes[box] = b; }
}

// one warp per parent: union of up to FAN children
__global__ void __launch_bounds__(256) merge_aabb_kernel(int nchild, int nparent, const Aabb* __restrict__ child, Aabb* __restrict__ parent)
{
	const int p = blockIdx.x * 8 + (threadIdx.x >> 5);
	if (p >= nparent) return;
	float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	const int c = p * FAN + lane_id();
	if (c < nchild) { const Aabb b = child[c]; for (int a = 0; a < 3; a++) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; } }
	#pragma unroll
	for (int a = 0; a < 3; a++) {
		#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) { lo[a] = fminf(lo[a], __shfl_xor_sync(FULL, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(FULL, hi[a], o)); }
	}
	if (lane_id() == 0) { Aabb b; for (int a = 0; a < 3; a++) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; } parent[p] = b; }
}

__global__ void __launch_bounds__(BOX) knn_search_kernel(int P, int nboxes, int ngroups, int nsuper,
	const float4* __restrict__ sorted, const Aabb* __restrict__ boxes, const Aabb* __restrict__ groups, const Aabb* __restrict__ supers,
	float* __restrict__ out)
{
	__shared__ float4 s_own[BOX];
	const int mybox = blockIdx.x;
	const int i = mybox * BOX + threadIdx.x;
	const bool valid = i < P;
	const float4 me = valid ? sorted[i] : make_float4(0, 0, 0, 0);
	s_own[threadIdx.x] = me;
	__syncthreads();
	if (!valid) return;
	const int own_cnt = min(BOX, P - mybox * BOX);
	float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
	for (int k = 0; k < own_cnt; k++) {
		if (k == (int)threadIdx.x) continue;
		const float4 o = s_own[k];
		insert3(sqdist(o.x, o.y, o.z, me.x, me.y, me.z), b0, b1, b2);
	}
	for (int s = 0; s < nsuper; s++) {
		if (box_dist(supers[s], me.x, me.y, me.z) > b2) continue;
		const int g_end = min(ngroups, (s + 1) * FAN);
		for (int g = s * FAN; g < g_end; g++) {
			if (box_dist(groups[g], me.x, me.y, me.z) > b2) continue;
			const int b_end = min(nboxes, (g + 1) * FAN);
			for (int b = g * FAN; b < b_end; b++) {
				if (b == mybox) continue;
				if (box_dist(boxes[b], me.x, me.y, me.z) > b2) continue;
				const int cnt = min(BOX, P - b * BOX);
				const float4* __restrict__ pts = sorted + (size_t)b * BOX;
				for (int k = 0; k < cnt; k++) {
					const float4 o = __ldg(pts + k);
					insert3(sqdist(o.x, o.y, o.z, me.x, me.y, me.z), b0, b1, b2);
				}
			}
		}
	}
	out[__float_as_uint(me.w)] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

struct KnnWs {
	uint32_t *keys_a, *keys_b, *vals_a, *vals_b;
	float4* sorted; Aabb *boxes, *groups, *supers; float *mm_partial, *mm; void* sort_ws;
	int nboxes, ngroups, nsuper; size_t total;
};

KnnWs knn_layout(void* ws, int P) {
	KnnWs L;
	Carver c(ws);
	L.nboxes = (P + BOX - 1) / BOX; L.ngroups = (L.nboxes + FAN - 1) / FAN; L.nsuper = (L.ngroups + FAN - 1) / FAN;
	L.keys_a = c.take<uint32_t>(P); L.keys_b = c.take<uint32_t>(P); L.vals_a = c.take<uint32_t>(P); L.vals_b = c.take<uint32_t>(P);
	L.sorted = c.take<float4>(P);
	L.boxes = c.take<Aabb>(L.nboxes); L.groups = c.take<Aabb>(L.ngroups); L.supers = c.take<Aabb>(L.nsuper);
	L.mm_partial = c.take<float>(1024 * 6); L.mm = c.take<float>(8);
	L.sort_ws = c.take<char>(sort_ws_bytes<uint32_t>((size_t)P, 0, 30));
	L.total = c.used();
	return L;
}

}  // namespace

size_t knn_ws_bytes(int P) { return knn_layout(nullptr, P < 1 ? 1 : P).total + 256; }

int launch_dist2(cudaStream_t stream, int P, const float* points, float* out, void* ws, size_t ws_bytes)
{
	if (P == 0) return 0;
	const KnnWs L = knn_layout(ws, P);
	if (L.total > ws_bytes) { set_error("dist2: workspace too small (%zu < %zu)", ws_bytes, L.total); return -1; }
	const int rb = min(1024, (P + 255) / 256);
	count_launch(), minmax_partial_kernel<<<rb, 256, 0, stream>>>(P, points, L.mm_partial);
	count_launch(), minmax_final_kernel<<<1, 32, 0, stream>>>(rb, L.mm_partial, L.mm);
	count_launch(), morton_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, points, L.mm, L.keys_a);
	const int which = sort_pairs<uint32_t>(stream, L.keys_a, L.keys_b, L.vals_a, L.vals_b, (size_t)P, 0, 30, L.sort_ws, true);
	if (which < 0) return -1;
	const uint32_t* order = which ? L.vals_b : L.vals_a;
	count_launch(), gather_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, points, order, L.sorted);
	count_launch(), box_aabb_kernel<<<(L.nboxes + 7) / 8, 256, 0, stream>>>(P, L.nboxes, L.sorted, L.boxes);
	count_launch(), merge_aabb_kernel<<<(L.ngroups + 7) / 8, 256, 0, stream>>>(L.nboxes, L.ngroups, L.boxes, L.groups);
	count_launch(), merge_aabb_kernel<<<(L.nsuper + 7) / 8, 256, 0, stream>>>(L.ngroups, L.nsuper, L.groups, L.supers);
	count_launch(), knn_search_kernel<<<L.nboxes, BOX, 0, stream>>>(P, L.nboxes, L.ngroups, L.nsuper, L.sorted, L.boxes, L.groups, L.supers, out);
	return check_launch("dist2", stream, false);
}

}  // namespace rr
