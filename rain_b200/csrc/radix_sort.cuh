#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace rr {

// Workspace bytes for sorting n pairs over key bits [begin_bit, end_bit).
template <class K> size_t sort_ws_bytes(size_t n, int begin_bit, int end_bit);

// Stable ascending LSD sort of (key, value) pairs on key bits [begin_bit, end_bit).  Ping-pongs between the *_a and
// *_b buffers; returns 0 if the result is in the *_a buffers, 1 if it is in the *_b buffers, -1 on error.
// iota_vals: treat the input values as 0..n-1 without reading vals_a (vals_a is still used as a ping-pong buffer).
template <class K>
int sort_pairs(cudaStream_t stream, K* keys_a, K* keys_b, uint32_t* vals_a, uint32_t* vals_b, size_t n,
	int begin_bit, int end_bit, void* ws, bool iota_vals);

// nseg independent sorts of the same length and bit range as ONE set of launches (blockIdx.y = segment); arrays of nseg
// pointers (host memory); same return value as sort_pairs (all segments land in the same buffers).
constexpr int SORT_MAX_SEGS = 16;
template <class K>
int sort_pairs_segments(cudaStream_t stream, int nseg, K* const* keys_a, K* const* keys_b, uint32_t* const* vals_a, uint32_t* const* vals_b, size_t n,
	int begin_bit, int end_bit, void* const* ws, bool iota_vals);

}  // namespace rr
