// Row / column gather: out[i] = in[idx[i]].
//
// Replaces pyarrow.Table.take(indices) as used by the reference after every
// sort_indices / permutation:
//   transform_pyarrow.take_table     python/ray/data/_internal/arrow_ops/transform_pyarrow.py:24-50
//   ArrowBlockAccessor.take          python/ray/data/_internal/arrow_block.py:346-355
//   ArrowBlockAccessor.random_shuffle  (permutation + take)            arrow_block.py:249-255
//
// Layout: a fixed-width column is a dense [n, row_bytes] byte matrix (row_bytes
// = 8 for an int64/float64 column, 64/90/100 for a packed payload). Each row is
// moved by row_bytes / V consecutive lanes using V-byte accesses, V the widest
// of 16/8/4/2/1 that divides row_bytes and the base alignment, so that stores
// are fully coalesced and loads are coalesced per source row.
#include "common.cuh"
#include "../../include/b200x.h"

namespace b200x {

template <typename V>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const V* __restrict__ src, const uint32_t* __restrict__ idx, V* __restrict__ dst, uint64_t n,
                   uint32_t vec_per_row) {
  // One thread per V-sized chunk, grid-stride in units of chunks, U independent (index -> load ->
  // store) chains in flight per thread: the narrower the chunk, the more chains, so that rows whose
  // width only allows 2- or 4-byte chunks (90-byte gensort payloads) still keep enough bytes in
  // flight. (row, chunk) of a thread's next element follow from the previous one by additions - a
  // 64-bit division per chunk made 90-byte rows 1.7x slower than 96-byte ones.
  constexpr int U = sizeof(V) >= 8 ? 4 : (sizeof(V) == 4 ? 8 : 16);
  const uint64_t total = n * vec_per_row;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t q = stride / vec_per_row;                 // rows per stride
  const uint32_t m = (uint32_t)(stride - q * vec_per_row);  // chunks per stride beyond whole rows
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = t / vec_per_row;
  uint32_t c = (uint32_t)(t - r * vec_per_row);
  for (; t + (U - 1) * stride < total; t += U * stride) {
    uint64_t s[U];
    uint32_t cc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      cc[u] = c;
      s[u] = idx[r];
      c += m;
      r += q;
      if (c >= vec_per_row) {
        c -= vec_per_row;
        ++r;
      }
    }
    V v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = src[s[u] * vec_per_row + cc[u]];
#pragma unroll
    for (int u = 0; u < U; ++u) dst[t + u * stride] = v[u];
  }
  for (; t < total; t += stride) {
    dst[t] = src[(uint64_t)idx[r] * vec_per_row + c];
    c += m;
    r += q;
    if (c >= vec_per_row) {
      c -= vec_per_row;
      ++r;
    }
  }
}

// Specialisation for rows of exactly one vector (8- or 16-byte columns): four
// independent gathers in flight per thread.
template <typename V>
__global__ void __launch_bounds__(256)
gather_elems_kernel(const V* __restrict__ src, const uint32_t* __restrict__ idx, V* __restrict__ dst, uint64_t n) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    const uint32_t s0 = idx[i], s1 = idx[i + stride], s2 = idx[i + 2 * stride], s3 = idx[i + 3 * stride];
    const V v0 = src[s0], v1 = src[s1], v2 = src[s2], v3 = src[s3];
    dst[i] = v0;
    dst[i + stride] = v1;
    dst[i + 2 * stride] = v2;
    dst[i + 3 * stride] = v3;
  }
  for (; i < n; i += stride) dst[i] = src[idx[i]];
}

constexpr int MAX_GATHER_COLS = 16;
struct ColPtrs {
  const uint64_t* src[MAX_GATHER_COLS];
  uint64_t* dst[MAX_GATHER_COLS];
};

// Several 8-byte columns that share one index vector: the index is read once
// and ncols independent loads are in flight per thread.
__global__ void __launch_bounds__(256)
gather_cols8_kernel(const ColPtrs p, int ncols, const uint32_t* __restrict__ idx, uint64_t n) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t s = idx[i];
    uint64_t v[MAX_GATHER_COLS];
#pragma unroll
    for (int c = 0; c < MAX_GATHER_COLS; ++c)
      if (c < ncols) v[c] = p.src[c][s];
#pragma unroll
    for (int c = 0; c < MAX_GATHER_COLS; ++c)
      if (c < ncols) p.dst[c][i] = v[c];
  }
}

// Rows kept at a pitch: record r of a buffer starts at byte r * pitch (pitch a multiple of 16, base 16-byte
// aligned) and only its first vec_per_row 16-byte vectors are moved. This is how a block keeps several
// fixed-width columns (a gensort record = 10-byte key + 90-byte payload at a 128-byte pitch) as ONE record
// per row: every access is an aligned 16-byte vector whatever the column widths, a 128-byte record is
// exactly one DRAM line, and the whole row moves in one pass instead of one pass per column.
__global__ void __launch_bounds__(256)
gather_rows_pitched_kernel(const uint4* __restrict__ src, const uint32_t* __restrict__ idx, uint4* __restrict__ dst,
                           uint64_t n, uint32_t vec_per_row, uint32_t src_pitch_vec, uint32_t dst_pitch_vec) {
  constexpr int U = 4;
  const uint64_t total = n * vec_per_row;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t q = stride / vec_per_row;
  const uint32_t m = (uint32_t)(stride - q * vec_per_row);
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = t / vec_per_row;
  uint32_t c = (uint32_t)(t - r * vec_per_row);
  for (; t + (U - 1) * stride < total; t += U * stride) {
    uint64_t s[U], d[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      s[u] = (uint64_t)idx[r] * src_pitch_vec + c;
      d[u] = r * dst_pitch_vec + c;
      c += m;
      r += q;
      if (c >= vec_per_row) {
        c -= vec_per_row;
        ++r;
      }
    }
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = src[s[u]];
#pragma unroll
    for (int u = 0; u < U; ++u) dst[d[u]] = v[u];
  }
  for (; t < total; t += stride) {
    dst[r * dst_pitch_vec + c] = src[(uint64_t)idx[r] * src_pitch_vec + c];
    c += m;
    r += q;
    if (c >= vec_per_row) {
      c -= vec_per_row;
      ++r;
    }
  }
}

// dst[r, 0:row_bytes) = src[r, 0:row_bytes) between two row pitches (column <-> record conversion at
// ingest / egress). V = widest access dividing both pitches, both bases and row_bytes.
template <typename V>
__global__ void __launch_bounds__(256)
copy_rows_2d_kernel(const unsigned char* __restrict__ src, uint64_t src_pitch, unsigned char* __restrict__ dst,
                    uint64_t dst_pitch, uint32_t vec_per_row, uint64_t n) {
  const uint64_t total = n * vec_per_row;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t q = stride / vec_per_row;
  const uint32_t m = (uint32_t)(stride - q * vec_per_row);
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = t / vec_per_row;
  uint32_t c = (uint32_t)(t - r * vec_per_row);
  for (; t < total; t += stride) {
    *reinterpret_cast<V*>(dst + r * dst_pitch + (uint64_t)c * sizeof(V)) =
        *reinterpret_cast<const V*>(src + r * src_pitch + (uint64_t)c * sizeof(V));
    c += m;
    r += q;
    if (c >= vec_per_row) {
      c -= vec_per_row;
      ++r;
    }
  }
}

__global__ void __launch_bounds__(256) iota_kernel(uint32_t* out, uint64_t n, uint32_t start) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = start + (uint32_t)i;
}

static inline uint32_t grid_for(uint64_t work_items, int threads, int per_sm) {
  uint64_t blocks = (work_items + threads - 1) / threads;
  const uint64_t cap = (uint64_t)sm_count() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return (uint32_t)blocks;
}

}  // namespace b200x

using namespace b200x;

extern "C" int b200x_gather_rows(const void* src, size_t row_bytes, const uint32_t* idx, size_t n, void* dst,
                                 void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || idx == nullptr || dst == nullptr) {
    set_error("b200x_gather_rows: NULL pointer");
    return B200X_ERR_ARG;
  }
  if (row_bytes > (1u << 20)) {
    set_error("b200x_gather_rows: row_bytes=%zu too large", row_bytes);
    return B200X_ERR_TOO_LARGE;
  }
  const uintptr_t al = (uintptr_t)src | (uintptr_t)dst | (uintptr_t)row_bytes;
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  const uint64_t total = (uint64_t)n * vpr;
  if (vpr == 1 && v >= 8) {
    const uint32_t g = grid_for(n / 4 + 1, 256, 16);
    if (v == 16)
      gather_elems_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, idx, (uint4*)dst, n);
    else
      gather_elems_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, idx, (uint64_t*)dst, n);
    B200X_LAUNCH_CHECK("gather_elems_kernel");
    return B200X_OK;
  }
  const uint32_t g = grid_for(total, 256, 16);
  prof_begin(stream, PROF_GATHER);
  switch (v) {
    case 16: gather_rows_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, idx, (uint4*)dst, n, vpr); break;
    case 8: gather_rows_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, idx, (uint64_t*)dst, n, vpr); break;
    case 4: gather_rows_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t*)src, idx, (uint32_t*)dst, n, vpr); break;
    case 2: gather_rows_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t*)src, idx, (uint16_t*)dst, n, vpr); break;
    default: gather_rows_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t*)src, idx, (uint8_t*)dst, n, vpr); break;
  }
  prof_end(stream);
  B200X_LAUNCH_CHECK("gather_rows_kernel");
  return B200X_OK;
}

extern "C" int b200x_gather_rows_pitched(const void* src, size_t src_pitch, const uint32_t* idx, size_t n, void* dst,
                                         size_t dst_pitch, size_t row_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || idx == nullptr || dst == nullptr) {
    set_error("b200x_gather_rows_pitched: NULL pointer");
    return B200X_ERR_ARG;
  }
  if (row_bytes > src_pitch || row_bytes > dst_pitch || row_bytes > (1u << 20)) {
    set_error("b200x_gather_rows_pitched: row_bytes=%zu exceeds a pitch (%zu, %zu) or 1 MiB", row_bytes, src_pitch, dst_pitch);
    return B200X_ERR_ARG;
  }
  if ((((uintptr_t)src | (uintptr_t)dst | src_pitch | dst_pitch | row_bytes) & 15) != 0) {
    set_error("b200x_gather_rows_pitched: bases, pitches and row_bytes must be multiples of 16");
    return B200X_ERR_ALIGN;
  }
  if (src_pitch > (1u << 24) || dst_pitch > (1u << 24)) {
    set_error("b200x_gather_rows_pitched: pitch (%zu, %zu) above 16 MiB", src_pitch, dst_pitch);
    return B200X_ERR_TOO_LARGE;
  }
  const uint32_t vpr = (uint32_t)(row_bytes / 16);
  const uint32_t g = grid_for((uint64_t)n * vpr, 256, 16);
  prof_begin(stream, PROF_GATHER);
  gather_rows_pitched_kernel<<<g, 256, 0, stream>>>((const uint4*)src, idx, (uint4*)dst, n, vpr, (uint32_t)(src_pitch / 16),
                                                    (uint32_t)(dst_pitch / 16));
  prof_end(stream);
  B200X_LAUNCH_CHECK("gather_rows_pitched_kernel");
  return B200X_OK;
}

extern "C" int b200x_copy_rows_2d(const void* src, size_t src_pitch, void* dst, size_t dst_pitch, size_t row_bytes,
                                  size_t n, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || dst == nullptr || row_bytes > src_pitch || row_bytes > dst_pitch || row_bytes > (1u << 20)) {
    set_error("b200x_copy_rows_2d: NULL pointer or row_bytes=%zu exceeds a pitch (%zu, %zu) or 1 MiB", row_bytes, src_pitch,
              dst_pitch);
    return B200X_ERR_ARG;
  }
  const uintptr_t al = (uintptr_t)src | (uintptr_t)dst | src_pitch | dst_pitch | row_bytes;
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  const uint32_t g = grid_for((uint64_t)n * vpr, 256, 16);
  const unsigned char* s = (const unsigned char*)src;
  unsigned char* d = (unsigned char*)dst;
  switch (v) {
    case 16: copy_rows_2d_kernel<uint4><<<g, 256, 0, stream>>>(s, src_pitch, d, dst_pitch, vpr, n); break;
    case 8: copy_rows_2d_kernel<uint64_t><<<g, 256, 0, stream>>>(s, src_pitch, d, dst_pitch, vpr, n); break;
    case 4: copy_rows_2d_kernel<uint32_t><<<g, 256, 0, stream>>>(s, src_pitch, d, dst_pitch, vpr, n); break;
    case 2: copy_rows_2d_kernel<uint16_t><<<g, 256, 0, stream>>>(s, src_pitch, d, dst_pitch, vpr, n); break;
    default: copy_rows_2d_kernel<uint8_t><<<g, 256, 0, stream>>>(s, src_pitch, d, dst_pitch, vpr, n); break;
  }
  B200X_LAUNCH_CHECK("copy_rows_2d_kernel");
  return B200X_OK;
}

extern "C" int b200x_gather_columns8(const void* const* srcs, void* const* dsts, int ncols, const uint32_t* idx,
                                     size_t n, void* stream_) {
  B200X_ON_DEVICE_OF(idx);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n == 0 || ncols == 0) return B200X_OK;
  if (srcs == nullptr || dsts == nullptr || idx == nullptr || ncols < 0) {
    set_error("b200x_gather_columns8: bad arguments");
    return B200X_ERR_ARG;
  }
  for (int c0 = 0; c0 < ncols; c0 += MAX_GATHER_COLS) {
    const int m = (ncols - c0 < MAX_GATHER_COLS) ? ncols - c0 : MAX_GATHER_COLS;
    ColPtrs p;
    for (int c = 0; c < MAX_GATHER_COLS; ++c) {
      p.src[c] = (c < m) ? (const uint64_t*)srcs[c0 + c] : nullptr;
      p.dst[c] = (c < m) ? (uint64_t*)dsts[c0 + c] : nullptr;
      if (c < m && (p.src[c] == nullptr || p.dst[c] == nullptr || (((uintptr_t)p.src[c] | (uintptr_t)p.dst[c]) & 7))) {
        set_error("b200x_gather_columns8: column %d is NULL or not 8-byte aligned", c0 + c);
        return B200X_ERR_ALIGN;
      }
    }
    gather_cols8_kernel<<<grid_for(n, 256, 16), 256, 0, stream>>>(p, m, idx, n);
    B200X_LAUNCH_CHECK("gather_cols8_kernel");
  }
  return B200X_OK;
}

extern "C" int b200x_iota_u32(uint32_t* out, size_t n, uint32_t start, void* stream_) {
  B200X_ON_DEVICE_OF(out);
  if (n == 0) return B200X_OK;
  if (out == nullptr) {
    set_error("b200x_iota_u32: NULL pointer");
    return B200X_ERR_ARG;
  }
  iota_kernel<<<grid_for(n, 256, 16), 256, 0, (cudaStream_t)stream_>>>(out, n, start);
  B200X_LAUNCH_CHECK("iota_kernel");
  return B200X_OK;
}
