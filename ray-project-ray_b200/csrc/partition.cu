// Key encoding and range-partition boundary search on a sorted block.
//
// Replaces find_partition_index / find_partitions
// (python/ray/data/_internal/util.py:719-758, :761-779), which run
// numpy.searchsorted per boundary on the sorted key column, and the key
// normalisation that pyarrow's sort does internally.
//
// Because the sorted block is held as order-preserving 64-bit codes that are
// ascending for BOTH sort directions (descending complements the code), the two
// reference cases collapse to plain lower/upper bounds on the codes:
//   ascending : index = #rows <  b  = lower_bound(codes, code(b))   (rows == b go right)
//   descending: index = #rows >= b  = upper_bound(codes, code(b))   (rows == b go left)
#include "common.cuh"
#include "../../include/b200x.h"

namespace b200x {

__global__ void __launch_bounds__(256)
encode_kernel(const void* __restrict__ raw, int kt, int desc, uint64_t n, uint64_t* __restrict__ codes) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    codes[i] = load_encode(raw, i, kt, desc != 0);
}

__global__ void __launch_bounds__(128)
search_kernel(const uint64_t* __restrict__ codes, uint64_t n, const uint64_t* __restrict__ bounds, int nb, int upper,
              uint64_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const uint64_t x = bounds[b];
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    const uint64_t v = codes[mid];
    const bool go_right = upper ? (v <= x) : (v < x);
    if (go_right) lo = mid + 1; else hi = mid;
  }
  out[b] = lo;
}

// Same search on a SORTED RAW key column: the code of keys[mid] is computed on the fly, so a
// sorted block needs no materialised code array to be cut at boundaries.
__global__ void __launch_bounds__(128)
search_keys_kernel(const void* __restrict__ keys, int kt, int desc, uint64_t n, const uint64_t* __restrict__ bounds,
                   int nb, int upper, uint64_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const uint64_t x = bounds[b];
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    const uint64_t v = load_encode(keys, mid, kt, desc != 0);
    const bool go_right = upper ? (v <= x) : (v < x);
    if (go_right) lo = mid + 1; else hi = mid;
  }
  out[b] = lo;
}

// Cut positions that split a sorted key column into `parts` near-equal blocks without splitting a
// run of equal keys: cut i = first row whose key equals the key at row floor(i*n/parts).
__global__ void __launch_bounds__(128)
even_cuts_kernel(const void* __restrict__ keys, int kt, int desc, uint64_t n, int parts, uint64_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= parts - 1) return;
  uint64_t pos = (uint64_t)(b + 1) * n / (uint64_t)parts;
  if (pos >= n) pos = n - 1;
  const uint64_t x = load_encode(keys, pos, kt, desc != 0);
  uint64_t lo = 0, hi = pos;  // the answer is <= pos
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    if (load_encode(keys, mid, kt, desc != 0) < x) lo = mid + 1; else hi = mid;
  }
  out[b] = lo;
}

}  // namespace b200x

using namespace b200x;

extern "C" int b200x_even_cuts(const void* keys_sorted, int key_type, int descending, size_t n, int parts,
                               uint64_t* out_index, void* stream_) {
  if (parts <= 1) return B200X_OK;
  if (n == 0 || keys_sorted == nullptr || out_index == nullptr || key_type < 0 || key_type >= KEY_TYPE_COUNT) {
    set_error("b200x_even_cuts: bad arguments");
    return B200X_ERR_ARG;
  }
  even_cuts_kernel<<<(parts - 1 + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(keys_sorted, key_type, descending, n, parts,
                                                                              out_index);
  B200X_LAUNCH_CHECK("even_cuts_kernel");
  return B200X_OK;
}

extern "C" int b200x_search_sorted_keys(const void* keys_sorted, int key_type, int descending, size_t n,
                                        const uint64_t* bound_codes, int num_bounds, int upper, uint64_t* out_index,
                                        void* stream_) {
  if (num_bounds == 0) return B200X_OK;
  if (bound_codes == nullptr || out_index == nullptr || num_bounds < 0 || (n > 0 && keys_sorted == nullptr) ||
      key_type < 0 || key_type >= KEY_TYPE_COUNT) {
    set_error("b200x_search_sorted_keys: bad arguments");
    return B200X_ERR_ARG;
  }
  search_keys_kernel<<<(num_bounds + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(keys_sorted, key_type, descending, n,
                                                                               bound_codes, num_bounds, upper, out_index);
  B200X_LAUNCH_CHECK("search_keys_kernel");
  return B200X_OK;
}

extern "C" int b200x_encode_keys(const void* keys, int key_type, int descending, size_t n, uint64_t* codes_out,
                                 void* stream_) {
  if (n == 0) return B200X_OK;
  if (keys == nullptr || codes_out == nullptr || key_type < 0 || key_type >= KEY_TYPE_COUNT) {
    set_error("b200x_encode_keys: bad arguments");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  encode_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(keys, key_type, descending, n, codes_out);
  B200X_LAUNCH_CHECK("encode_kernel");
  return B200X_OK;
}

// Fixed-width byte-string keys: bytes [offset, offset + width) of every row as one big-endian unsigned
// integer, left-aligned in 64 bits (zero-padded on the right), so that comparing the integers compares
// the byte strings lexicographically (the order pyarrow gives binary / fixed_size_binary columns).
__global__ void __launch_bounds__(256)
pack_key_bytes_kernel(const unsigned char* __restrict__ col, uint64_t row_bytes, uint32_t offset, uint32_t width,
                      uint64_t n, uint64_t* __restrict__ out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const unsigned char* p = col + i * row_bytes + offset;
    uint64_t v = 0;
#pragma unroll
    for (uint32_t b = 0; b < 8; ++b) v = (v << 8) | (b < width ? (uint64_t)p[b] : 0ull);
    out[i] = v;
  }
}

extern "C" int b200x_pack_key_bytes(const void* col, size_t row_bytes, size_t offset, int width, size_t n,
                                    uint64_t* out, void* stream_) {
  if (n == 0) return B200X_OK;
  if (col == nullptr || out == nullptr || width < 1 || width > 8 || offset + (size_t)width > row_bytes) {
    set_error("b200x_pack_key_bytes: bad arguments (row_bytes=%zu offset=%zu width=%d)", row_bytes, offset, width);
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  pack_key_bytes_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>((const unsigned char*)col, row_bytes,
                                                                            (uint32_t)offset, (uint32_t)width, n, out);
  B200X_LAUNCH_CHECK("pack_key_bytes_kernel");
  return B200X_OK;
}

extern "C" int b200x_search_sorted_codes(const uint64_t* codes, size_t n, const uint64_t* bound_codes, int num_bounds,
                                         int upper, uint64_t* out_index, void* stream_) {
  if (num_bounds == 0) return B200X_OK;
  if (bound_codes == nullptr || out_index == nullptr || num_bounds < 0 || (n > 0 && codes == nullptr)) {
    set_error("b200x_search_sorted_codes: bad arguments");
    return B200X_ERR_ARG;
  }
  search_kernel<<<(num_bounds + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(codes, n, bound_codes, num_bounds, upper,
                                                                          out_index);
  B200X_LAUNCH_CHECK("search_kernel");
  return B200X_OK;
}
