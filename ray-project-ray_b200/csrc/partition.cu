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
  B200X_ON_DEVICE_OF(out_index);
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
  B200X_ON_DEVICE_OF(bound_codes);
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
  B200X_ON_DEVICE_OF(keys);
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

// Number of positions i >= 1 with keys[i] == keys[i - 1] (as codes: -0.0 == +0.0, NaN == NaN) in a sorted column:
// zero means every key is distinct, i.e. lower-priority key columns cannot influence the order.
__global__ void __launch_bounds__(256)
count_adjacent_equal_kernel(const void* __restrict__ keys, int kt, uint64_t n, unsigned long long* __restrict__ out) {
  unsigned long long c = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    c += load_encode(keys, i, kt, false) == load_encode(keys, i - 1, kt, false) ? 1ull : 0ull;
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

extern "C" int b200x_count_adjacent_equal(const void* keys, int key_type, size_t n, uint64_t* count_out, void* stream_) {
  B200X_ON_DEVICE_OF(keys);
  if (count_out == nullptr || key_type < 0 || key_type >= KEY_TYPE_COUNT || (n > 0 && keys == nullptr)) {
    set_error("b200x_count_adjacent_equal: bad arguments");
    return B200X_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream_;
  B200X_CUDA(cudaMemsetAsync(count_out, 0, sizeof(uint64_t), st));
  if (n < 2) return B200X_OK;
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  count_adjacent_equal_kernel<<<(uint32_t)blocks, 256, 0, st>>>(keys, key_type, n, (unsigned long long*)count_out);
  B200X_LAUNCH_CHECK("count_adjacent_equal_kernel");
  return B200X_OK;
}

// Multi-column sort, first column almost distinct: after a stable sort by the first key column alone the
// order is already final everywhere except inside runs of equal first keys. Every such run (found by its
// first position) is put in order by the remaining key columns with a stable insertion sort on the
// permutation entries - the remaining columns are read through the permutation, so nothing but `perm`
// changes. Runs longer than FIX_MAX_RUN are left alone and reported: the caller then sorts the usual way.
constexpr int FIX_MAX_RUN = 64;
constexpr int FIX_MAX_COLS = 3;
struct TieCols {
  const void* col[FIX_MAX_COLS];
  int kt[FIX_MAX_COLS];
  unsigned long long stride[FIX_MAX_COLS];  // bytes from one row's value to the next
  int be[FIX_MAX_COLS];                     // > 0: the value is `be` big-endian bytes (a piece of a byte-string key)
  int n;
};

__device__ __forceinline__ uint64_t tie_code(const TieCols& tc, int c, uint32_t row, bool desc) {
  const unsigned char* p = (const unsigned char*)tc.col[c] + (uint64_t)row * tc.stride[c];
  if (tc.be[c] > 0) {
    uint64_t v = 0;
    for (int b = 0; b < tc.be[c]; ++b) v = (v << 8) | p[b];
    return desc ? ~v : v;
  }
  return load_encode(p, 0, tc.kt[c], desc);
}

__device__ __forceinline__ bool tie_less(const TieCols& tc, bool desc, uint32_t a, uint32_t b) {  // row a sorts before row b
  for (int c = 0; c < tc.n; ++c) {
    const uint64_t ca = tie_code(tc, c, a, desc), cb = tie_code(tc, c, b, desc);
    if (ca != cb) return ca < cb;
  }
  return false;
}

__global__ void __launch_bounds__(256)
fix_ties_kernel(const void* __restrict__ keys, int kt, uint64_t n, uint32_t* __restrict__ perm, const TieCols tc, int desc,
                unsigned long long* __restrict__ stats /* [0] rows in tie runs, [1] longest run */) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t k = load_encode(keys, i, kt, false);
    if (load_encode(keys, i + 1, kt, false) != k) continue;
    if (i > 0 && load_encode(keys, i - 1, kt, false) == k) continue;  // not the first row of its run
    uint64_t len = 2;
    while (i + len < n && len <= FIX_MAX_RUN && load_encode(keys, i + len, kt, false) == k) ++len;
    atomicAdd(&stats[0], (unsigned long long)len);
    atomicMax(&stats[1], (unsigned long long)len);
    if (len > FIX_MAX_RUN) continue;
    uint32_t* p = perm + i;
    for (uint32_t j = 1; j < (uint32_t)len; ++j) {
      const uint32_t pj = p[j];
      uint32_t q = j;
      while (q > 0) {
        const uint32_t pk = p[q - 1];
        if (!tie_less(tc, desc != 0, pj, pk)) break;   // stable: equal rows keep their order
        p[q] = pk;
        --q;
      }
      p[q] = pj;
    }
  }
}

extern "C" int b200x_fix_ties(const void* keys_sorted, int key_type, size_t n, uint32_t* perm, const void* const* cols_host,
                              const int* types_host, const unsigned long long* strides_host, const int* be_bytes_host,
                              int num_cols, int descending, unsigned long long* stats_out, void* stream_) {
  B200X_ON_DEVICE_OF(keys_sorted);
  if (stats_out == nullptr || key_type < 0 || key_type >= KEY_TYPE_COUNT || num_cols < 1 || num_cols > FIX_MAX_COLS ||
      cols_host == nullptr || types_host == nullptr || (n > 0 && (keys_sorted == nullptr || perm == nullptr))) {
    set_error("b200x_fix_ties: bad arguments (1..%d further key columns)", FIX_MAX_COLS);
    return B200X_ERR_ARG;
  }
  TieCols tc;
  tc.n = num_cols;
  for (int c = 0; c < FIX_MAX_COLS; ++c) {
    tc.col[c] = c < num_cols ? cols_host[c] : nullptr;
    tc.kt[c] = c < num_cols ? types_host[c] : 0;
    tc.be[c] = (c < num_cols && be_bytes_host != nullptr) ? be_bytes_host[c] : 0;
    if (c < num_cols && (tc.col[c] == nullptr || tc.kt[c] < 0 || tc.kt[c] >= KEY_TYPE_COUNT || tc.be[c] < 0 || tc.be[c] > 8)) {
      set_error("b200x_fix_ties: key column %d is NULL or has a bad type", c);
      return B200X_ERR_ARG;
    }
    const unsigned long long width = tc.be[c] > 0 ? (unsigned long long)tc.be[c] : (unsigned long long)key_type_bytes(tc.kt[c]);
    tc.stride[c] = (c < num_cols && strides_host != nullptr && strides_host[c] != 0) ? strides_host[c] : width;
    if (c < num_cols && tc.be[c] == 0 && (((uintptr_t)tc.col[c] | tc.stride[c]) & (width - 1)) != 0) {
      set_error("b200x_fix_ties: key column %d: address and stride must be multiples of the element size", c);
      return B200X_ERR_ALIGN;
    }
  }
  cudaStream_t st = (cudaStream_t)stream_;
  B200X_CUDA(cudaMemsetAsync(stats_out, 0, 2 * sizeof(unsigned long long), st));
  if (n < 2) return B200X_OK;
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  fix_ties_kernel<<<(uint32_t)blocks, 256, 0, st>>>(keys_sorted, key_type, n, perm, tc, descending, stats_out);
  B200X_LAUNCH_CHECK("fix_ties_kernel");
  return B200X_OK;
}

// Fixed-width byte-string keys: bytes [offset, offset + width) of every row as one big-endian unsigned
// integer, left-aligned in OUT (zero-padded on the right), so that comparing the integers compares
// the byte strings lexicographically (the order pyarrow gives binary / fixed_size_binary columns).
template <typename OUT>
__global__ void __launch_bounds__(256)
pack_key_bytes_kernel(const unsigned char* __restrict__ col, uint64_t row_bytes, uint32_t offset, uint32_t width,
                      uint64_t n, OUT* __restrict__ out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const unsigned char* p = col + i * row_bytes + offset;
    uint64_t v = 0;
#pragma unroll
    for (uint32_t b = 0; b < sizeof(OUT); ++b) v = (v << 8) | (b < width ? (uint64_t)p[b] : 0ull);
    out[i] = (OUT)v;
  }
}

// The inverse: the first `width` big-endian bytes of every chunk back into bytes [offset, offset + width) of the rows.
template <typename IN>
__global__ void __launch_bounds__(256)
unpack_key_bytes_kernel(const IN* __restrict__ chunk, uint32_t width, unsigned char* __restrict__ col,
                        uint64_t row_bytes, uint32_t offset, uint64_t n) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t v = (uint64_t)chunk[i];
    unsigned char* p = col + i * row_bytes + offset;
#pragma unroll
    for (uint32_t b = 0; b < sizeof(IN); ++b)
      if (b < width) p[b] = (unsigned char)(v >> (8 * (sizeof(IN) - 1 - b)));
  }
}

// All 8-byte chunks of a byte-string key in ONE pass, for key columns whose rows lie at a pitch that is a
// multiple of 16 (record members): a row's key is fetched with the one to three aligned 16-byte loads that
// cover it - one DRAM sector per row for a gensort key - instead of one byte load per key byte and chunk
// (125 M records: 3.5 ms + 2.4 ms for the two chunks of a 10-byte key with pack_key_bytes_kernel).
// A = (address of the key) mod 16, the same for every row; a template parameter so that every byte position is a
// compile-time register index.
constexpr int PK_MAX_CHUNKS = 4;
struct PackOut {
  void* out[PK_MAX_CHUNKS];
  int out_bytes[PK_MAX_CHUNKS];
};

template <int A>
__global__ void __launch_bounds__(256)
pack_key_chunks_kernel(const uint4* __restrict__ rec /* 16-byte aligned: key address - A */, uint64_t pitch_vec,
                       uint32_t width, uint64_t n, const PackOut po) {
  const uint32_t nvec = (A + width + 15u) / 16u;
  const uint32_t nchunks = (width + 7u) / 8u;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t w[12];
#pragma unroll
    for (int v = 0; v < 3; ++v) {
      uint4 x = make_uint4(0u, 0u, 0u, 0u);
      if ((uint32_t)v < nvec) x = rec[i * pitch_vec + v];
      w[4 * v] = x.x;
      w[4 * v + 1] = x.y;
      w[4 * v + 2] = x.z;
      w[4 * v + 3] = x.w;
    }
#pragma unroll
    for (int k = 0; k < PK_MAX_CHUNKS; ++k) {
      if ((uint32_t)k < nchunks) {
        uint64_t val = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int pos = A + 8 * k + j;          // compile-time byte position inside the loaded vectors
          const uint32_t byte = (w[pos >> 2] >> (8 * (pos & 3))) & 0xFFu;
          val = (val << 8) | ((uint32_t)(8 * k + j) < width ? (uint64_t)byte : 0ull);
        }
        const int ob = po.out_bytes[k];
        if (ob == 8) ((uint64_t*)po.out[k])[i] = val;
        else if (ob == 4) ((uint32_t*)po.out[k])[i] = (uint32_t)(val >> 32);
        else ((uint16_t*)po.out[k])[i] = (uint16_t)(val >> 48);
      }
    }
  }
}

extern "C" int b200x_pack_key_chunks(const void* col, size_t row_pitch, int width, size_t n, void* const* outs_host,
                                     const int* out_bytes_host, void* stream_) {
  B200X_ON_DEVICE_OF(col);
  if (n == 0) return B200X_OK;
  const int nchunks = (width + 7) / 8;
  if (col == nullptr || outs_host == nullptr || out_bytes_host == nullptr || width < 1 || nchunks > PK_MAX_CHUNKS ||
      (size_t)width > row_pitch || (row_pitch & 15) != 0) {
    set_error("b200x_pack_key_chunks: bad arguments (width=%d, 1..%d; row_pitch=%zu must be a multiple of 16 and >= width)",
              width, 8 * PK_MAX_CHUNKS, row_pitch);
    return B200X_ERR_ARG;
  }
  PackOut po;
  for (int k = 0; k < PK_MAX_CHUNKS; ++k) {
    po.out[k] = k < nchunks ? outs_host[k] : nullptr;
    po.out_bytes[k] = k < nchunks ? out_bytes_host[k] : 8;
    const int last = width - 8 * k;   // bytes of this chunk
    if (k < nchunks && (po.out[k] == nullptr || (po.out_bytes[k] != 2 && po.out_bytes[k] != 4 && po.out_bytes[k] != 8) ||
                        (last < 8 && last > po.out_bytes[k]) || (last >= 8 && po.out_bytes[k] != 8))) {
      set_error("b200x_pack_key_chunks: chunk %d has a NULL or too narrow output", k);
      return B200X_ERR_ARG;
    }
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const uint32_t g = (uint32_t)blocks;
  cudaStream_t st = (cudaStream_t)stream_;
  const int a = (int)((uintptr_t)col & 15);
  const uint4* rec = (const uint4*)((const unsigned char*)col - a);
  const uint64_t pv = row_pitch / 16;
  const uint32_t w = (uint32_t)width;
#define PK_CASE(A) case A: pack_key_chunks_kernel<A><<<g, 256, 0, st>>>(rec, pv, w, n, po); break;
  switch (a) {
    PK_CASE(0) PK_CASE(1) PK_CASE(2) PK_CASE(3) PK_CASE(4) PK_CASE(5) PK_CASE(6) PK_CASE(7)
    PK_CASE(8) PK_CASE(9) PK_CASE(10) PK_CASE(11) PK_CASE(12) PK_CASE(13) PK_CASE(14) PK_CASE(15)
  }
#undef PK_CASE
  B200X_LAUNCH_CHECK("pack_key_chunks_kernel");
  return B200X_OK;
}

static bool chunk_args_ok(size_t row_bytes, size_t offset, int width, int chunk_bytes) {
  return width >= 1 && (chunk_bytes == 2 || chunk_bytes == 4 || chunk_bytes == 8) && width <= chunk_bytes &&
         offset + (size_t)width <= row_bytes;
}

extern "C" int b200x_pack_key_bytes(const void* col, size_t row_bytes, size_t offset, int width, size_t n, void* out,
                                    int out_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(col);
  if (n == 0) return B200X_OK;
  if (col == nullptr || out == nullptr || !chunk_args_ok(row_bytes, offset, width, out_bytes)) {
    set_error("b200x_pack_key_bytes: bad arguments (row_bytes=%zu offset=%zu width=%d out_bytes=%d)", row_bytes, offset,
              width, out_bytes);
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const unsigned char* c = (const unsigned char*)col;
  cudaStream_t st = (cudaStream_t)stream_;
  const uint32_t g = (uint32_t)blocks, off = (uint32_t)offset, w = (uint32_t)width;
  if (out_bytes == 8) pack_key_bytes_kernel<uint64_t><<<g, 256, 0, st>>>(c, row_bytes, off, w, n, (uint64_t*)out);
  else if (out_bytes == 4) pack_key_bytes_kernel<uint32_t><<<g, 256, 0, st>>>(c, row_bytes, off, w, n, (uint32_t*)out);
  else pack_key_bytes_kernel<uint16_t><<<g, 256, 0, st>>>(c, row_bytes, off, w, n, (uint16_t*)out);
  B200X_LAUNCH_CHECK("pack_key_bytes_kernel");
  return B200X_OK;
}

extern "C" int b200x_unpack_key_bytes(const void* chunk, int chunk_bytes, int width, void* col, size_t row_bytes,
                                      size_t offset, size_t n, void* stream_) {
  B200X_ON_DEVICE_OF(chunk);
  if (n == 0) return B200X_OK;
  if (col == nullptr || chunk == nullptr || !chunk_args_ok(row_bytes, offset, width, chunk_bytes)) {
    set_error("b200x_unpack_key_bytes: bad arguments (row_bytes=%zu offset=%zu width=%d chunk_bytes=%d)", row_bytes,
              offset, width, chunk_bytes);
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  unsigned char* c = (unsigned char*)col;
  cudaStream_t st = (cudaStream_t)stream_;
  const uint32_t g = (uint32_t)blocks, off = (uint32_t)offset, w = (uint32_t)width;
  if (chunk_bytes == 8) unpack_key_bytes_kernel<uint64_t><<<g, 256, 0, st>>>((const uint64_t*)chunk, w, c, row_bytes, off, n);
  else if (chunk_bytes == 4) unpack_key_bytes_kernel<uint32_t><<<g, 256, 0, st>>>((const uint32_t*)chunk, w, c, row_bytes, off, n);
  else unpack_key_bytes_kernel<uint16_t><<<g, 256, 0, st>>>((const uint16_t*)chunk, w, c, row_bytes, off, n);
  B200X_LAUNCH_CHECK("unpack_key_bytes_kernel");
  return B200X_OK;
}

extern "C" int b200x_search_sorted_codes(const uint64_t* codes, size_t n, const uint64_t* bound_codes, int num_bounds,
                                         int upper, uint64_t* out_index, void* stream_) {
  B200X_ON_DEVICE_OF(bound_codes);
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
