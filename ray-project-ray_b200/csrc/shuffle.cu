// Seeded random permutation without an index array: a Feistel bijection on [0, n)
// evaluated per output row, fused with the row gather.
//
// Replaces ArrowBlockAccessor.random_shuffle
// (python/ray/data/_internal/arrow_block.py:249-255:
//  numpy.random.RandomState(seed).permutation(n) followed by take), as called by
// ShuffleTaskSpec.map / reduce
// (python/ray/data/_internal/planner/exchange/shuffle_task_spec.py:81-84, :119-123).
//
// The reference materialises an int64 index vector with a sequential
// Fisher-Yates walk over the MT19937 stream; neither is reproducible in parallel
// nor part of the parity contract (pull- and push-based shuffles already give
// different permutations for one seed). What is kept: an exact permutation,
// fully determined by (seed, n).
//
// pi(x): x is split into a low part R (ceil(bits/2) bits) and a high part L (floor(bits/2) bits),
// bits = ceil(log2 n), and run through ROUNDS alternating Feistel rounds (even rounds:
// L ^= F(R), odd rounds: R ^= F(L)) with a murmur3-finalizer round function keyed from
// splitmix64(seed); the result is "cycle walked" (re-encrypted) while it is >= n. The domain
// 2^bits is < 2n, so the expected walk is < 2 steps. Every round is its own inverse, so
// pi^-1 runs the rounds backwards.
#include "common.cuh"
#include "../../include/b200x.h"

namespace b200x {

constexpr int FEISTEL_ROUNDS = 6;

struct PermSpec {
  uint64_t n;
  uint32_t lo_bits;   // width of R
  uint32_t lo_mask;
  uint32_t hi_mask;   // mask of L
  uint32_t key[FEISTEL_ROUNDS];
};

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

__host__ __device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return h;
}

static PermSpec make_spec(uint64_t n, uint64_t seed) {
  PermSpec s;
  s.n = n;
  int bits = 0;
  while (bits < 32 && (n > 1) && (((uint64_t)1 << bits) < n)) ++bits;  // bits = ceil(log2(n))
  if (bits < 2) bits = 2;
  const uint32_t lo = (uint32_t)((bits + 1) / 2), hi = (uint32_t)bits - lo;
  s.lo_bits = lo;
  s.lo_mask = (1u << lo) - 1u;
  s.hi_mask = (1u << hi) - 1u;
  uint64_t st = seed;
  for (int r = 0; r < FEISTEL_ROUNDS; ++r) {
    st = splitmix64(st + (uint64_t)r);
    s.key[r] = (uint32_t)(st >> 32);
  }
  return s;
}

__host__ __device__ __forceinline__ uint32_t feistel_once(const PermSpec& s, uint32_t x, bool inverse) {
  uint32_t L = (x >> s.lo_bits) & s.hi_mask;
  uint32_t R = x & s.lo_mask;
#pragma unroll
  for (int i = 0; i < FEISTEL_ROUNDS; ++i) {
    const int r = inverse ? (FEISTEL_ROUNDS - 1 - i) : i;
    if ((r & 1) == 0) L ^= fmix32(R * 0x9E3779B1u + s.key[r]) & s.hi_mask;
    else R ^= fmix32(L * 0x9E3779B1u + s.key[r]) & s.lo_mask;
  }
  return (L << s.lo_bits) | R;
}

__host__ __device__ __forceinline__ uint64_t perm_eval(const PermSpec& s, uint64_t x, bool inverse) {
  if (s.n <= 1) return x;
  uint32_t y = (uint32_t)x;
  do {
    y = feistel_once(s, y, inverse);
  } while (y >= s.n);
  return y;
}

__global__ void __launch_bounds__(256)
perm_indices_kernel(const PermSpec s, int inverse, uint64_t start, uint64_t count, uint32_t* __restrict__ out) {
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < count; t += (uint64_t)gridDim.x * blockDim.x)
    out[t] = (uint32_t)perm_eval(s, start + t, inverse != 0);
}

// dst[j] = src[pi(j)], rows of vec_per_row V-sized chunks. A warp takes 32 consecutive output
// rows: each lane evaluates pi for one of them (so the bijection costs one evaluation per row,
// not one per 16-byte chunk), then groups of lanes_per_row lanes copy the rows, the source row
// number being broadcast with a shuffle.
template <typename V>
__global__ void __launch_bounds__(256)
permute_rows_kernel(const V* __restrict__ src, V* __restrict__ dst, const PermSpec s, int inverse,
                    uint32_t vec_per_row, uint32_t lanes_per_row, uint32_t src_pitch_vec, uint32_t dst_pitch_vec) {
  const uint64_t n = s.n;
  const unsigned lane = threadIdx.x & 31;
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t groups = 32 / lanes_per_row;      // rows copied concurrently by one warp
  const uint32_t g = lane / lanes_per_row, c0 = lane % lanes_per_row;
  for (uint64_t r0 = warp * 32; r0 < n; r0 += nwarps * 32) {
    const uint64_t mine = r0 + lane;
    const uint32_t sr = mine < n ? (uint32_t)perm_eval(s, mine, inverse != 0) : 0u;
    for (uint32_t i = 0; i < 32; i += groups) {
      const uint32_t ri = i + g;                   // row of this lane group within the warp's 32
      const uint32_t srow = __shfl_sync(0xffffffffu, sr, ri);
      const uint64_t r = r0 + ri;
      if (r < n) {
        const V* sp = src + (uint64_t)srow * src_pitch_vec;
        V* dp = dst + r * dst_pitch_vec;
        for (uint32_t c = c0; c < vec_per_row; c += lanes_per_row) dp[c] = sp[c];
      }
    }
  }
}

}  // namespace b200x

using namespace b200x;

static int check_perm_n(const char* fn, size_t n) {
  if (n > 0xFFFFFFFFull) {
    set_error("%s: n=%zu exceeds 2^32-1 rows", fn, n);
    return B200X_ERR_TOO_LARGE;
  }
  return B200X_OK;
}

extern "C" int b200x_perm_indices(size_t n, uint64_t seed, int inverse, size_t start, size_t count, uint32_t* out,
                                  void* stream_) {
  B200X_ON_DEVICE_OF(out);
  if (int rc = check_perm_n("b200x_perm_indices", n)) return rc;
  if (count == 0) return B200X_OK;
  if (out == nullptr || start + count > n) {
    set_error("b200x_perm_indices: bad range [%zu, %zu) of %zu or NULL out", start, start + count, n);
    return B200X_ERR_ARG;
  }
  const PermSpec s = make_spec(n, seed);
  uint64_t blocks = (count + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  perm_indices_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(s, inverse, start, count, out);
  B200X_LAUNCH_CHECK("perm_indices_kernel");
  return B200X_OK;
}

extern "C" int b200x_perm_indices_host(size_t n, uint64_t seed, int inverse, size_t start, size_t count,
                                       uint32_t* out_host) {
  if (int rc = check_perm_n("b200x_perm_indices_host", n)) return rc;
  if (count == 0) return B200X_OK;
  if (out_host == nullptr || start + count > n) {
    set_error("b200x_perm_indices_host: bad range or NULL out");
    return B200X_ERR_ARG;
  }
  const PermSpec s = make_spec(n, seed);
  for (size_t t = 0; t < count; ++t) out_host[t] = (uint32_t)perm_eval(s, start + t, inverse != 0);
  return B200X_OK;
}

extern "C" int b200x_permute_rows(const void* src, size_t row_bytes, size_t n, uint64_t seed, int inverse, void* dst,
                                  void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_perm_n("b200x_permute_rows", n)) return rc;
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || dst == nullptr) {
    set_error("b200x_permute_rows: NULL pointer");
    return B200X_ERR_ARG;
  }
  if (row_bytes > (1u << 20)) {
    set_error("b200x_permute_rows: row_bytes=%zu too large", row_bytes);
    return B200X_ERR_TOO_LARGE;
  }
  const PermSpec s = make_spec(n, seed);
  const uintptr_t al = (uintptr_t)src | (uintptr_t)dst | (uintptr_t)row_bytes;
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  uint32_t lpr = 1;
  while (lpr < vpr && lpr < 32) lpr <<= 1;
  const uint64_t threads_needed = (uint64_t)n * lpr;
  uint64_t blocks = (threads_needed + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const uint32_t g = (uint32_t)blocks;
  switch (v) {
    case 16: permute_rows_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, (uint4*)dst, s, inverse, vpr, lpr, vpr, vpr); break;
    case 8: permute_rows_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, (uint64_t*)dst, s, inverse, vpr, lpr, vpr, vpr); break;
    case 4: permute_rows_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t*)src, (uint32_t*)dst, s, inverse, vpr, lpr, vpr, vpr); break;
    case 2: permute_rows_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t*)src, (uint16_t*)dst, s, inverse, vpr, lpr, vpr, vpr); break;
    default: permute_rows_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t*)src, (uint8_t*)dst, s, inverse, vpr, lpr, vpr, vpr); break;
  }
  B200X_LAUNCH_CHECK("permute_rows_kernel");
  return B200X_OK;
}

// The same permutation for rows kept at a pitch (records, see b200x_gather_rows_pitched): only the first
// row_bytes bytes of every record move; bases, pitches and row_bytes are multiples of 16.
extern "C" int b200x_permute_rows_pitched(const void* src, size_t src_pitch, size_t n, uint64_t seed, int inverse,
                                          void* dst, size_t dst_pitch, size_t row_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_perm_n("b200x_permute_rows_pitched", n)) return rc;
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || dst == nullptr || row_bytes > src_pitch || row_bytes > dst_pitch || row_bytes > (1u << 20) ||
      src_pitch > (1u << 24) || dst_pitch > (1u << 24)) {
    set_error("b200x_permute_rows_pitched: NULL pointer, or row_bytes=%zu / pitches (%zu, %zu) out of range", row_bytes,
              src_pitch, dst_pitch);
    return B200X_ERR_ARG;
  }
  if ((((uintptr_t)src | (uintptr_t)dst | src_pitch | dst_pitch | row_bytes) & 15) != 0) {
    set_error("b200x_permute_rows_pitched: bases, pitches and row_bytes must be multiples of 16");
    return B200X_ERR_ALIGN;
  }
  const PermSpec s = make_spec(n, seed);
  const uint32_t vpr = (uint32_t)(row_bytes / 16);
  uint32_t lpr = 1;
  while (lpr < vpr && lpr < 32) lpr <<= 1;
  uint64_t blocks = ((uint64_t)n * lpr + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  prof_begin(stream, PROF_PERMUTE);
  permute_rows_kernel<uint4><<<(uint32_t)blocks, 256, 0, stream>>>((const uint4*)src, (uint4*)dst, s, inverse, vpr, lpr,
                                                                  (uint32_t)(src_pitch / 16), (uint32_t)(dst_pitch / 16));
  prof_end(stream);
  B200X_LAUNCH_CHECK("permute_rows_kernel");
  return B200X_OK;
}

// Shuffle across GPUs as ONE scatter: local row i is row first_global + i of the whole dataset (total rows over all
// ranks); its place in the shuffled dataset is pi(first_global + i), which lies on rank pi / rows_per_rank at row
// pi % rows_per_rank - the row is stored straight there through the peer mapping. Every rank evaluates the same
// bijection, so no counts, send buffers or second local shuffle are needed: the receive buffer IS the result.
namespace b200x {
constexpr int PS_MAX_PEERS = 16;
struct PeerBases {
  unsigned char* base[PS_MAX_PEERS];
};

template <typename V>
__global__ void __launch_bounds__(256)
permute_scatter_kernel(const V* __restrict__ src, uint32_t src_pitch_vec, uint32_t vec_per_row, uint32_t lanes_per_row,
                       uint64_t n_local, uint64_t first_global, const PermSpec s, uint64_t rows_per_rank, const PeerBases pb,
                       uint64_t dst_pitch_vec) {
  const unsigned lane = threadIdx.x & 31;
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t groups = 32 / lanes_per_row;
  const uint32_t g = lane / lanes_per_row, c0 = lane % lanes_per_row;
  for (uint64_t r0 = warp * 32; r0 < n_local; r0 += nwarps * 32) {
    const uint64_t mine = r0 + lane;
    const uint64_t p = mine < n_local ? perm_eval(s, first_global + mine, true) : 0ull;   // where my row goes
    for (uint32_t i = 0; i < 32; i += groups) {
      const uint32_t ri = i + g;
      const uint64_t pr = __shfl_sync(0xffffffffu, p, ri);
      const uint64_t r = r0 + ri;
      if (r < n_local) {
        const uint64_t d = pr / rows_per_rank, slot = pr - d * rows_per_rank;
        const V* sp = src + r * src_pitch_vec;
        V* dp = reinterpret_cast<V*>(pb.base[d]) + slot * dst_pitch_vec;
        for (uint32_t c = c0; c < vec_per_row; c += lanes_per_row) dp[c] = sp[c];
      }
    }
  }
}
}  // namespace b200x

extern "C" int b200x_permute_scatter_to_peers(const void* src, size_t src_pitch, size_t row_bytes, size_t n_local,
                                              size_t first_global, size_t total_rows, uint64_t seed, size_t rows_per_rank,
                                              void* const* peer_base_host, size_t dst_pitch, int world, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_perm_n("b200x_permute_scatter_to_peers", total_rows)) return rc;
  if (n_local == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || peer_base_host == nullptr || world < 1 || world > PS_MAX_PEERS || rows_per_rank == 0 ||
      first_global + n_local > total_rows || rows_per_rank * (size_t)world < total_rows || row_bytes > src_pitch ||
      row_bytes > dst_pitch || row_bytes > (1u << 20)) {
    set_error("b200x_permute_scatter_to_peers: bad arguments (rows [%zu, %zu) of %zu, %zu rows per rank x %d ranks, row_bytes %zu)",
              first_global, first_global + n_local, total_rows, rows_per_rank, world, row_bytes);
    return B200X_ERR_ARG;
  }
  PeerBases pb;
  uintptr_t al = (uintptr_t)src | src_pitch | dst_pitch | row_bytes;
  for (int d = 0; d < PS_MAX_PEERS; ++d) {
    pb.base[d] = d < world ? (unsigned char*)peer_base_host[d] : nullptr;
    if (d < world) al |= (uintptr_t)pb.base[d];
    if (d < world && pb.base[d] == nullptr) {
      set_error("b200x_permute_scatter_to_peers: receive buffer %d is NULL", d);
      return B200X_ERR_ARG;
    }
  }
  // scatter form of out[q] = in[pi(q)] (b200x_permute_rows): row G goes to the position q with pi(q) = G, q = pi^-1(G)
  const PermSpec s = make_spec(total_rows, seed);
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  uint32_t lpr = 1;
  while (lpr < vpr && lpr < 32) lpr <<= 1;
  uint64_t blocks = ((uint64_t)n_local * lpr + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const uint32_t g = (uint32_t)blocks, spv = (uint32_t)(src_pitch / v);
  const uint64_t dpv = dst_pitch / v;
  prof_begin(stream, PROF_PERMUTE);
  switch (v) {
    case 16: permute_scatter_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, spv, vpr, lpr, n_local, first_global, s, rows_per_rank, pb, dpv); break;
    case 8: permute_scatter_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, spv, vpr, lpr, n_local, first_global, s, rows_per_rank, pb, dpv); break;
    case 4: permute_scatter_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t*)src, spv, vpr, lpr, n_local, first_global, s, rows_per_rank, pb, dpv); break;
    case 2: permute_scatter_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t*)src, spv, vpr, lpr, n_local, first_global, s, rows_per_rank, pb, dpv); break;
    default: permute_scatter_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t*)src, spv, vpr, lpr, n_local, first_global, s, rows_per_rank, pb, dpv); break;
  }
  prof_end(stream);
  B200X_LAUNCH_CHECK("permute_scatter_kernel");
  return B200X_OK;
}
