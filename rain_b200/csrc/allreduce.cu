// One-kernel all-reduce (sum, f32, in place) of the per-rank gradient bucket over NVLink 5 / NVSwitch peer memory.
// No reference counterpart (sharonal10/rain is single-GPU); replaces the ncclAllReduce of the view-sharded step
//  when the bucket lives in symmetric memory (every rank's buffer mapped into every rank's address space,
// plus — where the fabric offers it — one MULTICAST address that aliases all of them).
//
// Two-shot: rank r owns the 1/world slice r of the buffer.
//   MULTIMEM (NVLink SHARP, in-switch reduction):  v = multimem.ld_reduce.add.v4.f32 [mc + i]   — the switch pulls element i from
//       every rank's buffer and returns the sum;  multimem.st.v4.f32 [mc + i], v — the switch writes it into every rank's buffer.
//       A GPU sends and receives ~one bucket in total instead of 2 (world - 1) / world buckets per direction of a ring.
//   PEER (no multicast): the owner loads element i from every peer's buffer (ld.global over NVLink), sums in rank order
//       (so every replica of the job computes the same bits), and stores the sum into every peer's buffer.
// Synchronisation is per CTA, across ranks: CTA b of every rank handles the same sub-range of every slice and meets CTA b of
// the other ranks at a signal-flag barrier before (all buckets final) and after (all slices broadcast) — epoch-numbered flags
// in a small symmetric buffer, st.release.sys / ld.acquire.sys, no reset between calls.  Each rank launches the same grid,
// and CTA b of a rank only ever waits for CTA b of its peers, which run on OTHER GPUs: no co-residency requirement inside
// one GPU.  A spin that does not complete within ~2 s sets *error_flag instead of hanging the device.
#include "common.cuh"
#include "kernels.cuh"

namespace rr {

namespace {

constexpr int AR_THREADS = 512;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

// CTA b of this rank meets CTA b of every other rank: flag word [b * world + src] of rank dst's signal buffer = latest epoch src's CTA b reached
__device__ __forceinline__ void cross_rank_barrier(const AllReduceArgs& a, uint32_t epoch)
{
	__syncthreads();
	if (threadIdx.x < (unsigned)a.world) {
		const int peer = threadIdx.x;
		__threadfence_system();
		st_release_sys(a.signal[peer] + (size_t)blockIdx.x * a.world + a.rank, epoch);
		const uint32_t* mine = a.signal[a.rank] + (size_t)blockIdx.x * a.world + peer;
		const long long t0 = clock64();
		while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
			if (clock64() - t0 > 4000000000ll) { atomicExch(a.error_flag, 1u); break; }     // ~2 s at 2 GHz: report, do not hang
			__nanosleep(64);
		}
	}
	__syncthreads();
}

// peer-memory accesses at system scope (never served from a stale L1 line)
__device__ __forceinline__ float4 ld_peer4(const float* p) {
	float4 v;
	asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_peer4(float* p, float4 v) {
	asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float ld_peer1(const float* p) { float v; asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_peer1(float* p, float v) { asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

__device__ __forceinline__ float4 mm_ld_reduce(const float* mc) {
	float4 v;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
	return v;
}
__device__ __forceinline__ void mm_st(float* mc, float4 v) {
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <bool MULTIMEM>
__global__ void __launch_bounds__(AR_THREADS) allreduce_kernel(AllReduceArgs a)
{
	cross_rank_barrier(a, a.epoch);                         // every rank's bucket (these ranges of it) is final
	constexpr int U = 4;                                    // independent vectors in flight per thread
	const int nthreads = blockDim.x;
	for (int rg = 0; rg < a.nranges; rg++) {
		// float4 vectors [lo, hi) of this rank's slice of the range; CTA b takes a contiguous share, its threads stride through it
		const int64_t v0 = a.off[rg] / 4, n4 = a.len[rg] / 4;
		const int64_t lo = v0 + n4 * a.rank / a.world, hi = v0 + n4 * (a.rank + 1) / a.world;
		const int64_t per = (hi - lo + gridDim.x - 1) / gridDim.x;
		const int64_t b0 = lo + per * blockIdx.x, b1 = min(hi, b0 + per);
		for (int64_t i = b0 + threadIdx.x; i < b1; i += (int64_t)nthreads * U) {
			float4 v[U];
			#pragma unroll
			for (int u = 0; u < U; u++) {
				const int64_t j = i + (int64_t)u * nthreads;
				if (j < b1) {
					if (MULTIMEM) v[u] = mm_ld_reduce(a.multicast + 4 * j);
					else {
						float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
						for (int r = 0; r < a.world; r++) {           // rank order: the same sum on every replica of the job
							const float4 t = ld_peer4(a.buffers[r] + 4 * j);
							s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
						}
						v[u] = s;
					}
				}
			}
			#pragma unroll
			for (int u = 0; u < U; u++) {
				const int64_t j = i + (int64_t)u * nthreads;
				if (j < b1) {
					if (MULTIMEM) mm_st(a.multicast + 4 * j, v[u]);
					else for (int r = 0; r < a.world; r++) st_peer4(a.buffers[r] + 4 * j, v[u]);
				}
			}
		}
		// the len % 4 tail floats of the range (single-range calls only): rank 0, CTA 0, through the peer pointers
		if (a.rank == 0 && blockIdx.x == 0 && threadIdx.x < (unsigned)(a.len[rg] & 3)) {
			const int64_t j = a.off[rg] + (a.len[rg] & ~(int64_t)3) + threadIdx.x;
			float s = 0.f;
			for (int r = 0; r < a.world; r++) s += ld_peer1(a.buffers[r] + j);
			for (int r = 0; r < a.world; r++) st_peer1(a.buffers[r] + j, s);
		}
	}
	cross_rank_barrier(a, a.epoch + 1);                     // every slice has been broadcast: the bucket may be read
}

}  // namespace

int launch_allreduce(cudaStream_t stream, const AllReduceArgs& a, int ctas, int threads)
{
	if (a.world <= 1) return 0;
	if (threads <= 0 || threads > AR_THREADS) threads = AR_THREADS;
	threads = (threads + 31) / 32 * 32;
	count_launch();
	if (a.multicast) allreduce_kernel<true><<<ctas, threads, 0, stream>>>(a);
	else allreduce_kernel<false><<<ctas, threads, 0, stream>>>(a);
	return check_launch("allreduce", stream, false);
}

}  // namespace rr
