// Reference point for the hand-written onesweep (csrc/radix_sort.cu): cub::DeviceRadixSort::SortPairs from the CUDA
// toolkit (CCCL, sm_100-tuned onesweep) on the same shapes, same B200.
//   a) what the reference rasterizer does per render (rasterizer_impl.cu:295-300): N pairs, 64-bit tile|depth keys, bits [0, 45)
//   b) 32-bit keys, 13 bits, N pairs (this library's tile sort)      c) 32-bit keys, 32 bits, P pairs (this library's depth sort)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cub_sort_bench cub_sort_bench.cu && ./cub_sort_bench [N] [P]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cub/cub.cuh>

template <class K>
float run(size_t n, int begin_bit, int end_bit, uint64_t key_mask, int reps)
{
	std::vector<K> h(n);
	uint64_t s = 88172645463325252ull;
	for (size_t i = 0; i < n; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (K)(s & key_mask); }
	K *k_in, *k_out; uint32_t *v_in, *v_out;
	cudaMalloc(&k_in, n * sizeof(K)); cudaMalloc(&k_out, n * sizeof(K)); cudaMalloc(&v_in, n * 4); cudaMalloc(&v_out, n * 4);
	cudaMemcpy(k_in, h.data(), n * sizeof(K), cudaMemcpyHostToDevice); cudaMemset(v_in, 0, n * 4);
	size_t tmp_bytes = 0; void* tmp = nullptr;
	cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit);
	cudaMalloc(&tmp, tmp_bytes);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	for (int i = 0; i < 3; i++) cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit);
	cudaEventRecord(e0);
	for (int i = 0; i < reps; i++) cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit);
	cudaEventRecord(e1); cudaEventSynchronize(e1);
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	cudaFree(k_in); cudaFree(k_out); cudaFree(v_in); cudaFree(v_out); cudaFree(tmp);
	return ms / reps;
}

int main(int argc, char** argv)
{
	const size_t N = argc > 1 ? strtoull(argv[1], 0, 10) : 8435457, P = argc > 2 ? strtoull(argv[2], 0, 10) : 1000000;
	const float a = run<uint64_t>(N, 0, 45, (1ull << 45) - 1, 20);
	const float b = run<uint32_t>(N, 0, 13, (1u << 13) - 1, 20);
	const float c = run<uint32_t>(P, 0, 32, 0xffffffffu, 20);
	printf("{\"cub_u64_45bit_N_ms\": %.4f, \"cub_u32_13bit_N_ms\": %.4f, \"cub_u32_32bit_P_ms\": %.4f, \"N\": %zu, \"P\": %zu, "
		"\"cub_u64_gbs\": %.1f, \"cub_u32_13bit_gbs\": %.1f}\n", a, b, c, N, P, 152.0 * N / a / 1e6, 36.0 * N / b / 1e6);
	return 0;
}
