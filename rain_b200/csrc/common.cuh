// Shared definitions for the sm_100a splat rasterizer kernels.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

namespace rr {

constexpr int TILE_X = 16;   // reference BLOCK_X / BLOCK_Y (config.h:5-6); fixed by the key/range parity contract
constexpr int TILE_Y = 16;
constexpr unsigned FULL = 0xffffffffu;

// One projected Gaussian as the render kernels consume it: 48 bytes, 16-byte aligned so a tile's list is a
// contiguous span a 1-D bulk copy (cp.async.bulk) can stage into shared memory.
//   a = { pix.x, pix.y, conic.x, conic.y }   b = { conic.z, opacity, r, g }   c = { b, depth, ext.x, ext.y }
// ext = conservative half-extent (in pixels) of the region where alpha can reach 1/255 (used only for culling;
// +inf = never cull, -inf = can never contribute).
struct __align__(16) SplatRec {
	float4 a, b, c;
};
static_assert(sizeof(SplatRec) == 48, "SplatRec must be 48 bytes");

// Per (tile, Gaussian) gradient partial written by the backward render, 10 floats = 40 bytes (8-byte aligned):
//   0..2 dL_dcolor, 3..4 dL_dmean2D.xy, 5..7 dL_dconic.{x,y,w}, 8 dL_dopacity, 9 unused
constexpr int PARTIAL_FLOATS = 10;

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bump allocator over a caller-owned blob (the replacement for the reference's obtain()/fromChunk carving,
// rasterizer_impl.h:10-17): 256-byte aligned slots.
struct Carver {
	char* base; size_t off;
	__host__ explicit Carver(void* p) : base(reinterpret_cast<char*>(p)), off(0) {}
	template <class T> __host__ T* take(size_t count) {
		off = align_up(off, 256);
		T* p = reinterpret_cast<T*>(base + off);
		off += count * sizeof(T);
		return p;
	}
	__host__ size_t used() const { return align_up(off, 256); }
};

void set_error(const char* fmt, ...);
int check_launch(const char* what, cudaStream_t stream, bool sync);
// zero `bytes` (rounded up to 16; the regions are 256-byte-aligned carvings with room to the next one) at each of nseg device
// pointers (host array): one memset for a single region, ONE kernel for several (instead of a memset launch per region)
int launch_zero_segments(cudaStream_t stream, int nseg, void* const* ptrs, size_t bytes);
void count_launch();   // every kernel launch of the library is counted (rr_launch_count), bench.py reports it

// element idx of a [n,4] float array as a float4: one 16-byte load when the base is 16-byte aligned (a tensor that is a view
// at an odd offset of a larger buffer need not be), four scalar loads otherwise; the branch is uniform
__device__ __forceinline__ float4 load_f4(const float* __restrict__ base, size_t idx) {
	const float* p = base + 4 * idx;
	if ((reinterpret_cast<uintptr_t>(base) & 15) == 0) return *reinterpret_cast<const float4*>(p);
	return make_float4(p[0], p[1], p[2], p[3]);
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

}  // namespace rr
