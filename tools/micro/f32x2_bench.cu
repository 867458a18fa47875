// Microbenchmark: issue throughput of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_bench f32x2_bench.cu && ./f32x2_bench
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void fma2(float& x, float& y, float a, float b, float c, float d) {
	asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%0,%1}; mov.b64 rb, {%2,%3}; mov.b64 rc, {%4,%5}; fma.rn.f32x2 ra, ra, rb, rc; mov.b64 {%0,%1}, ra;}"
		: "+f"(x), "+f"(y) : "f"(a), "f"(b), "f"(c), "f"(d));
}
template <int MODE> __global__ void k(float* out, float a, float b, int iters) {
	float v[16];
	#pragma unroll
	for (int i = 0; i < 16; i++) v[i] = threadIdx.x * 0.001f + i;
	for (int it = 0; it < iters; it++) {
		if (MODE == 0) {
			#pragma unroll
			for (int i = 0; i < 16; i++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(a), "f"(b));
		} else {
			#pragma unroll
			for (int i = 0; i < 16; i += 2) fma2(v[i], v[i + 1], a, a, b, b);
		}
	}
	float s = 0;
	#pragma unroll
	for (int i = 0; i < 16; i++) s += v[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
	float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	const int iters = 20000;
	for (int mode = 0; mode < 2; mode++) {
		for (int rep = 0; rep < 2; rep++) {
			cudaEventRecord(e0);
			if (mode == 0) k<0><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters); else k<1><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
			cudaEventRecord(e1); cudaEventSynchronize(e1);
			float ms; cudaEventElapsedTime(&ms, e0, e1);
			const double fma = 148.0 * 8 * 256 * 16.0 * iters;
			if (rep) printf("%s: %.3f ms, %.1f TFLOP/s fp32 (%.1f Gfma/s)\n", mode ? "FFMA2" : "FFMA ", ms, 2 * fma / ms * 1e-9, fma / ms * 1e-6);
		}
	}
	return 0;
}
