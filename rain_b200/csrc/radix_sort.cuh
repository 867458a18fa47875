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

// The tile binning after the prefix sum (replaces duplicateWithKeys + SortPairs, rasterizer_impl.cu:275-300): pair emission with
// 16-bit tile ids and two 16-bit-key radix passes of which the last writes the Gaussian ids only.  digit_bases [npass][256] are
// the exclusive digit bases the tile plan of the geometry half computed from the rects (binning.cu; it also wrote the ranges,
// so there is no identifyTileRanges pass).  keys_a / keys_b: N uint16 each; vals_a / vals_b: N uint32 each; returns 1 if the
// sorted ids (point_list) are in vals_b, 0 if in vals_a, -1 on error.  bits <= 16.
size_t tile_sort_ws_bytes(size_t n, int bits);
int tile_sort_passes(int bits, int* shift /*[2]*/, uint32_t* mask /*[2]*/);      // the sort's digit split: returns the number of passes (1 or 2)
int tile_sort16(cudaStream_t stream, int P, size_t N, int bits, int grid_x, const uint32_t* order, const uint32_t* offsets, const uint2* rect,
	const uint32_t* digit_bases, uint16_t* keys_a, uint16_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, void* ws);

}  // namespace rr
