// Sort-based groupby: group boundaries + segmented reduction over a sorted block.
//
// Replaces the two interpreter-bound loops of the reference's aggregate path:
//   ArrowBlockAccessor.combine                  python/ray/data/_internal/arrow_block.py:466-548
//     (walk sorted rows, cut groups on key change, pac.sum / pac.count / pac.min_max per slice)
//   ArrowBlockAccessor.aggregate_combined_blocks  arrow_block.py:569-680
//     (heapq.merge of partial rows, AggregateFn.merge, AggregateFn.finalize)
// and the accumulator arithmetic of Sum / Min / Max / Mean / Std / Count
//   python/ray/data/aggregate/_aggregate.py:90-319.
//
// Layout: the block has been argsorted by key (b200x_radix_argsort): `codes` are
// the ascending key codes, `perm` the row order. Groups are maximal runs of equal
// codes. Each 2048-row tile reduces the runs that lie wholly inside it and
// records the partial sums of its first and last (possibly open) run; a
// single-block pass then stitches runs that span tiles in a fixed order, so float
// results do not depend on scheduling.
#include "common.cuh"
#include "../../include/b200x.h"
#include <math_constants.h>

namespace b200x {

constexpr int G_THREADS = 256;
constexpr int G_IPT = 8;
constexpr int G_TILE = G_THREADS * G_IPT;
constexpr int SCAN_THREADS = 1024;

// ------------------------------- boundaries --------------------------------

constexpr int MAX_GROUP_COLS = 4;
struct CodeCols {
  const uint64_t* c[MAX_GROUP_COLS];
  int n;
};

// a group starts where ANY key column changes
__device__ __forceinline__ bool is_head(const CodeCols& k, uint64_t i) {
  if (i == 0) return true;
  bool h = false;
#pragma unroll
  for (int j = 0; j < MAX_GROUP_COLS; ++j)
    if (j < k.n) h |= k.c[j][i] != k.c[j][i - 1];
  return h;
}

__global__ void __launch_bounds__(G_THREADS)
heads_count_kernel(const CodeCols codes, uint64_t n, uint32_t* __restrict__ tile_heads) {
  __shared__ uint32_t s_warp[G_THREADS / 32];
  const uint64_t base = (uint64_t)blockIdx.x * G_TILE;
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < G_IPT; ++j) {
    const uint64_t i = base + (uint64_t)j * G_THREADS + threadIdx.x;
    if (i < n && is_head(codes, i)) ++c;
  }
  c = __reduce_add_sync(0xffffffffu, c);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < G_THREADS / 32; ++w) t += s_warp[w];
    tile_heads[blockIdx.x] = t;
  }
}

// exclusive scan of tile_heads[0..T) in place (-> first group index of each tile); total -> *num_groups
__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(uint32_t* __restrict__ tile_heads, uint32_t T, uint32_t* __restrict__ num_groups) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  const uint32_t per = (T + SCAN_THREADS - 1) / SCAN_THREADS;
  const uint32_t lo = min(T, threadIdx.x * per), hi = min(T, lo + per);
  uint32_t sum = 0;
  for (uint32_t t = lo; t < hi; ++t) sum += tile_heads[t];
  uint32_t x = sum;
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  uint32_t off = 0;
  for (unsigned k = 0; k < w; ++k) off += s_warp[k];
  uint32_t run = off + x - sum;
  for (uint32_t t = lo; t < hi; ++t) {
    const uint32_t c = tile_heads[t];
    tile_heads[t] = run;
    run += c;
  }
  if (threadIdx.x == SCAN_THREADS - 1) *num_groups = off + x;
}

__global__ void __launch_bounds__(G_THREADS)
group_starts_kernel(const CodeCols codes, uint64_t n, const uint32_t* __restrict__ tile_base,
                    uint32_t* __restrict__ group_starts) {
  __shared__ uint32_t s_warp[G_THREADS / 32];
  const uint64_t base = (uint64_t)blockIdx.x * G_TILE;
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  // blocked arrangement so that ranks follow row order
  bool h[G_IPT];
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < G_IPT; ++j) {
    const uint64_t i = base + (uint64_t)threadIdx.x * G_IPT + j;
    h[j] = i < n && is_head(codes, i);
    c += h[j];
  }
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  uint32_t off = tile_base[blockIdx.x];
  for (unsigned k = 0; k < w; ++k) off += s_warp[k];
  uint32_t r = off + x - c;
#pragma unroll
  for (int j = 0; j < G_IPT; ++j)
    if (h[j]) group_starts[r++] = (uint32_t)(base + (uint64_t)threadIdx.x * G_IPT + j);
}

// ---------------------------- segmented reduce -----------------------------

template <typename T> struct Acc { T sum, mn, mx; };

template <typename T> __device__ __forceinline__ T type_max();
template <typename T> __device__ __forceinline__ T type_lowest();
template <> __device__ __forceinline__ int64_t type_max<int64_t>() { return INT64_MAX; }
template <> __device__ __forceinline__ int64_t type_lowest<int64_t>() { return INT64_MIN; }
template <> __device__ __forceinline__ double type_max<double>() { return CUDART_INF; }
template <> __device__ __forceinline__ double type_lowest<double>() { return -CUDART_INF; }

template <typename T> __device__ __forceinline__ Acc<T> acc_identity() {
  Acc<T> a;
  a.sum = (T)0;
  a.mn = type_max<T>();
  a.mx = type_lowest<T>();
  return a;
}
__device__ __forceinline__ int64_t add_t(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }
__device__ __forceinline__ double add_t(double a, double b) { return a + b; }
__device__ __forceinline__ int64_t min_t(int64_t a, int64_t b) { return a < b ? a : b; }
__device__ __forceinline__ int64_t max_t(int64_t a, int64_t b) { return a > b ? a : b; }
__device__ __forceinline__ double min_t(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ double max_t(double a, double b) { return fmax(a, b); }

template <typename T> __device__ __forceinline__ Acc<T> acc_merge(const Acc<T>& a, const Acc<T>& b) {
  Acc<T> r;
  r.sum = add_t(a.sum, b.sum);
  r.mn = min_t(a.mn, b.mn);
  r.mx = max_t(a.mx, b.mx);
  return r;
}
template <typename T> __device__ __forceinline__ Acc<T> acc_add(const Acc<T>& a, T v) {
  Acc<T> r;
  r.sum = add_t(a.sum, v);
  r.mn = min_t(a.mn, v);
  r.mx = max_t(a.mx, v);
  return r;
}

// segmented-scan element: (flag = "a group starts in here", value = reduction of the open tail)
template <typename T> struct Seg { uint32_t f; Acc<T> a; };
template <typename T> __device__ __forceinline__ Seg<T> seg_op(const Seg<T>& l, const Seg<T>& r) {
  Seg<T> o;
  o.f = l.f | r.f;
  o.a = r.f ? r.a : acc_merge(l.a, r.a);
  return o;
}
template <typename T> __device__ __forceinline__ Seg<T> seg_shfl_up(const Seg<T>& s, int o) {
  Seg<T> r;
  r.f = __shfl_up_sync(0xffffffffu, s.f, o);
  r.a.sum = __shfl_up_sync(0xffffffffu, s.a.sum, o);
  r.a.mn = __shfl_up_sync(0xffffffffu, s.a.mn, o);
  r.a.mx = __shfl_up_sync(0xffffffffu, s.a.mx, o);
  return r;
}

template <typename T> struct TileSummary {
  Acc<T> first;   // rows before the first group start of the tile (whole tile if none)
  Acc<T> last;    // rows from the last group start to the end of the tile (whole tile if none)
  uint32_t has_head;
  int32_t open_gid;  // group that is open when the tile begins (lo - 1)
};

template <typename T> struct OutPtrs { T* sum; T* mn; T* mx; };
template <typename T> __device__ __forceinline__ void emit(const OutPtrs<T>& o, int64_t gid, const Acc<T>& a) {
  if (gid < 0) return;
  if (o.sum) o.sum[gid] = a.sum;
  if (o.mn) o.mn[gid] = a.mn;
  if (o.mx) o.mx[gid] = a.mx;
}

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* __restrict__ a, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if ((uint64_t)a[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// VIn: element type of `values`; T: accumulator type (VIn, or double when centering)
template <typename VIn, typename T, bool CENTER>
__global__ void __launch_bounds__(G_THREADS)
seg_reduce_kernel(const uint32_t* __restrict__ group_starts, const uint32_t* __restrict__ num_groups_p,
                  const uint32_t* __restrict__ perm, const VIn* __restrict__ values, uint64_t n,
                  const double* __restrict__ center, const int64_t* __restrict__ weights,
                  const double* __restrict__ addend, OutPtrs<T> out, TileSummary<T>* __restrict__ summ) {
  __shared__ uint8_t s_flag[G_TILE];
  __shared__ uint32_t s_lohi[2];
  __shared__ uint32_t s_wcnt[G_THREADS / 32];
  __shared__ Seg<T> s_wseg[G_THREADS / 32];

  const uint32_t G = *num_groups_p;
  const uint64_t tbase = (uint64_t)blockIdx.x * G_TILE;
  const uint64_t tend = min(n, tbase + G_TILE);
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;

  for (int i = threadIdx.x; i < G_TILE / 4; i += G_THREADS) ((uint32_t*)s_flag)[i] = 0;
  if (threadIdx.x < 2) s_lohi[threadIdx.x] = lower_bound_u32(group_starts, G, threadIdx.x == 0 ? tbase : tend);
  __syncthreads();
  const uint32_t lo = s_lohi[0], hi = s_lohi[1];
  for (uint32_t g = lo + threadIdx.x; g < hi; g += G_THREADS) s_flag[group_starts[g] - tbase] = 1;

  // values of this thread's 8 consecutive rows (issued before the flag barrier)
  VIn v[G_IPT];
  double wgt[G_IPT], add[G_IPT];
  const uint64_t r0 = tbase + (uint64_t)threadIdx.x * G_IPT;
#pragma unroll
  for (int j = 0; j < G_IPT; ++j) {
    const uint64_t i = r0 + j;
    v[j] = (VIn)0;
    wgt[j] = 1.0;
    add[j] = 0.0;
    if (i < tend) {
      const uint64_t src = perm ? (uint64_t)perm[i] : i;
      v[j] = values[src];
      if (CENTER && weights) wgt[j] = (double)weights[src];
      if (CENTER && addend) add[j] = addend[src];
    }
  }
  __syncthreads();

  uint32_t hcnt = 0;
  uint8_t fl[G_IPT];
#pragma unroll
  for (int j = 0; j < G_IPT; ++j) {
    fl[j] = s_flag[threadIdx.x * G_IPT + j];
    hcnt += fl[j];
  }
  // exclusive scan of head counts over threads -> index of the run open at this thread's first row
  uint32_t x = hcnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_wcnt[w] = x;
  __syncthreads();
  uint32_t heads_before = x - hcnt;
  for (unsigned k = 0; k < w; ++k) heads_before += s_wcnt[k];
  const int64_t run0 = (int64_t)lo - 1 + heads_before;

  // sequential pass over the thread's rows
  Acc<T> acc = acc_identity<T>(), first_part = acc_identity<T>();
  uint32_t seen = 0;
#pragma unroll
  for (int j = 0; j < G_IPT; ++j) {
    if (fl[j]) {
      if (seen == 0) first_part = acc;
      else emit(out, run0 + seen, acc);  // run opened and closed inside this thread
      ++seen;
      acc = acc_identity<T>();
    }
    if (r0 + j < tend) {
      T val;
      if (CENTER) {
        const double d = (double)v[j] - center[run0 + seen];
        val = (T)(d * d * wgt[j] + add[j]);
      } else {
        val = (T)v[j];
      }
      acc = acc_add(acc, val);
    }
  }

  // block-wide segmented inclusive scan of (seen != 0, tail)
  Seg<T> me;
  me.f = seen ? 1u : 0u;
  me.a = acc;
  Seg<T> inc = me;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Seg<T> up = seg_shfl_up(inc, o);
    if (lane >= (unsigned)o) inc = seg_op(up, inc);
  }
  if (lane == 31) s_wseg[w] = inc;
  __syncthreads();
  Seg<T> wprefix;
  wprefix.f = 0;
  wprefix.a = acc_identity<T>();
  for (unsigned k = 0; k < w; ++k) wprefix = seg_op(wprefix, s_wseg[k]);
  Seg<T> prev = seg_shfl_up(inc, 1);
  if (lane == 0) {
    prev.f = 0;
    prev.a = acc_identity<T>();
  }
  const Seg<T> excl = seg_op(wprefix, prev);  // everything before this thread inside the tile

  if (seen) {
    const Acc<T> closed = acc_merge(excl.a, first_part);
    if (heads_before == 0) {
      summ[blockIdx.x].first = closed;  // first run of the tile: finished by the stitch pass
    } else {
      emit(out, run0, closed);
    }
  }
  if (threadIdx.x == G_THREADS - 1) {
    const Seg<T> total = seg_op(excl, me);
    summ[blockIdx.x].last = total.a;
    summ[blockIdx.x].has_head = total.f;
    summ[blockIdx.x].open_gid = (int32_t)lo - 1;
    if (!total.f) summ[blockIdx.x].first = total.a;
  }
}

// One block: segmented exclusive scan over the tile summaries in tile order, then close the
// runs that end at a tile's first group start, and the very last run.
template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
stitch_kernel(const TileSummary<T>* __restrict__ summ, uint32_t Tn, const uint32_t* __restrict__ num_groups_p,
              OutPtrs<T> out) {
  __shared__ Seg<T> s_wseg[SCAN_THREADS / 32];
  const uint32_t per = (Tn + SCAN_THREADS - 1) / SCAN_THREADS;
  const uint32_t lo = min(Tn, threadIdx.x * per), hi = min(Tn, lo + per);
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;

  Seg<T> me;
  me.f = 0;
  me.a = acc_identity<T>();
  for (uint32_t t = lo; t < hi; ++t) {
    Seg<T> e;
    e.f = summ[t].has_head;
    e.a = summ[t].last;  // == whole tile when it has no head
    me = seg_op(me, e);
  }
  Seg<T> inc = me;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Seg<T> up = seg_shfl_up(inc, o);
    if (lane >= (unsigned)o) inc = seg_op(up, inc);
  }
  if (lane == 31) s_wseg[w] = inc;
  __syncthreads();
  Seg<T> carry;
  carry.f = 0;
  carry.a = acc_identity<T>();
  for (unsigned k = 0; k < w; ++k) carry = seg_op(carry, s_wseg[k]);
  Seg<T> prev = seg_shfl_up(inc, 1);
  if (lane == 0) {
    prev.f = 0;
    prev.a = acc_identity<T>();
  }
  carry = seg_op(carry, prev);

  const uint32_t G = *num_groups_p;
  for (uint32_t t = lo; t < hi; ++t) {
    const TileSummary<T> s = summ[t];
    if (s.has_head) emit(out, (int64_t)s.open_gid, acc_merge(carry.a, s.first));
    if (t == Tn - 1) emit(out, (int64_t)G - 1, s.has_head ? s.last : acc_merge(carry.a, s.first));
    Seg<T> e;
    e.f = s.has_head;
    e.a = s.last;
    carry = seg_op(carry, e);
  }
}

// ------------------------------- finalizers --------------------------------

__global__ void __launch_bounds__(256)
group_counts_kernel(const uint32_t* __restrict__ group_starts, const uint32_t* __restrict__ num_groups_p, uint64_t n,
                    int64_t* __restrict__ counts) {
  const uint32_t G = *num_groups_p;
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < G; g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t e = (g + 1 < G) ? group_starts[g + 1] : n;
    counts[g] = (int64_t)(e - group_starts[g]);
  }
}

// mode 0: mean = sum / count; mode 1: std = count < 2 ? 0 : sqrt(m2 / (count - ddof))
__global__ void __launch_bounds__(256)
finalize_kernel(const void* __restrict__ sum, int vt, const int64_t* __restrict__ counts, uint64_t G, int mode, int ddof,
                double* __restrict__ out) {
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < G; g += (uint64_t)gridDim.x * blockDim.x) {
    const double s = (vt == B200X_VAL_I64) ? (double)((const int64_t*)sum)[g] : ((const double*)sum)[g];
    const double c = (double)counts[g];
    if (mode == 0) out[g] = s / c;
    else out[g] = (counts[g] < 2) ? 0.0 : sqrt(s / (c - (double)ddof));
  }
}

__global__ void __launch_bounds__(256)
absmax_kernel(const void* __restrict__ mn, const void* __restrict__ mx, int vt, uint64_t G, void* __restrict__ out) {
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < G; g += (uint64_t)gridDim.x * blockDim.x) {
    if (vt == B200X_VAL_I64) {
      const int64_t a = ((const int64_t*)mn)[g], b = ((const int64_t*)mx)[g];
      const int64_t aa = a < 0 ? -a : a, bb = b < 0 ? -b : b;
      ((int64_t*)out)[g] = aa > bb ? aa : bb;
    } else {
      ((double*)out)[g] = fmax(fabs(((const double*)mn)[g]), fabs(((const double*)mx)[g]));
    }
  }
}

// Quantile of each group by linear interpolation between the two nearest ranks. `perm` visits the
// rows ordered by (group key, value), so a group's values are ascending inside [start, end).
// The arithmetic is the reference's, step by step, without fused multiply-adds:
// k = (m - 1) * q; f = floor(k); c = ceil(k); f == c ? v[f] : round(v[f] * (c - k) + v[c] * (k - f), 5)
__global__ void __launch_bounds__(256)
group_quantile_kernel(const uint32_t* __restrict__ group_starts, const uint32_t* __restrict__ num_groups_p,
                      const uint32_t* __restrict__ perm, const void* __restrict__ values, int vt, uint64_t n, double q,
                      double* __restrict__ out) {
  const uint32_t G = *num_groups_p;
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < G; g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t s = group_starts[g];
    const uint64_t e = (g + 1 < G) ? group_starts[g + 1] : n;
    const uint64_t m = e - s;
    const double k = __dmul_rn((double)(m - 1), q);
    const double f = floor(k), c = ceil(k);
    auto value_at = [&](uint64_t r) -> double {
      const uint64_t row = perm ? (uint64_t)perm[s + r] : s + r;
      return (vt == B200X_VAL_I64) ? (double)((const int64_t*)values)[row] : ((const double*)values)[row];
    };
    double r;
    if (f == c) {
      r = value_at((uint64_t)k);
    } else {
      const double d0 = __dmul_rn(value_at((uint64_t)f), __dsub_rn(c, k));
      const double d1 = __dmul_rn(value_at((uint64_t)c), __dsub_rn(k, f));
      r = __dadd_rn(d0, d1);
      // round(x, 5): beyond 2^52 / 1e5 the scaled value is an integer already
      if (fabs(r) < 4.0e10) r = __ddiv_rn(rint(__dmul_rn(r, 1.0e5)), 1.0e5);
    }
    out[g] = r;
  }
}

static inline uint32_t tiles_of(size_t n) { return (uint32_t)((n + G_TILE - 1) / G_TILE); }

template <typename VIn, typename T, bool CENTER>
static int launch_seg_reduce(const uint32_t* group_starts, const uint32_t* num_groups, const uint32_t* perm,
                             const void* values, size_t n, const double* center, const int64_t* weights,
                             const double* addend, void* sum_out, void* min_out, void* max_out, void* temp,
                             cudaStream_t stream) {
  const uint32_t Tn = tiles_of(n);
  OutPtrs<T> out{(T*)sum_out, (T*)min_out, (T*)max_out};
  TileSummary<T>* summ = (TileSummary<T>*)temp;
  seg_reduce_kernel<VIn, T, CENTER><<<Tn, G_THREADS, 0, stream>>>(group_starts, num_groups, perm, (const VIn*)values, n,
                                                                  center, weights, addend, out, summ);
  B200X_LAUNCH_CHECK("seg_reduce_kernel");
  stitch_kernel<T><<<1, SCAN_THREADS, 0, stream>>>(summ, Tn, num_groups, out);
  B200X_LAUNCH_CHECK("stitch_kernel");
  return B200X_OK;
}

}  // namespace b200x

using namespace b200x;

extern "C" int b200x_group_boundaries_multi(const uint64_t* const* codes_host, int num_cols, size_t n,
                                            uint32_t* group_starts, uint32_t* num_groups_out, void* temp,
                                            size_t* temp_bytes, void* stream_);

extern "C" int b200x_group_boundaries(const uint64_t* codes, size_t n, uint32_t* group_starts, uint32_t* num_groups_out,
                                      void* temp, size_t* temp_bytes, void* stream_) {
  const uint64_t* one[1] = {codes};
  return b200x_group_boundaries_multi(one, 1, n, group_starts, num_groups_out, temp, temp_bytes, stream_);
}

extern "C" int b200x_group_boundaries_multi(const uint64_t* const* codes_host, int num_cols, size_t n,
                                            uint32_t* group_starts, uint32_t* num_groups_out, void* temp,
                                            size_t* temp_bytes, void* stream_) {
  if (codes_host == nullptr || num_cols < 1 || num_cols > MAX_GROUP_COLS) {
    set_error("b200x_group_boundaries_multi: 1..%d key columns", MAX_GROUP_COLS);
    return B200X_ERR_ARG;
  }
  CodeCols cc;
  cc.n = num_cols;
  for (int j = 0; j < MAX_GROUP_COLS; ++j) cc.c[j] = j < num_cols ? codes_host[j] : nullptr;
  const uint64_t* codes = cc.c[0];
  cudaStream_t stream = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_group_boundaries: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  if (n > 0xFFFFFFFFull) {
    set_error("b200x_group_boundaries: n=%zu exceeds 2^32-1", n);
    return B200X_ERR_TOO_LARGE;
  }
  const uint32_t Tn = tiles_of(n);
  const size_t need = align_up((size_t)(Tn + 1) * sizeof(uint32_t), 256);
  if (temp == nullptr) {
    *temp_bytes = need;
    return B200X_OK;
  }
  if (*temp_bytes < need) {
    set_error("b200x_group_boundaries: temp buffer %zu < required %zu bytes", *temp_bytes, need);
    return B200X_ERR_TEMP;
  }
  if (num_groups_out == nullptr) {
    set_error("b200x_group_boundaries: num_groups_out is NULL");
    return B200X_ERR_ARG;
  }
  if (n == 0) {
    B200X_CUDA(cudaMemsetAsync(num_groups_out, 0, sizeof(uint32_t), stream));
    return B200X_OK;
  }
  for (int j = 1; j < num_cols; ++j)
    if (cc.c[j] == nullptr) codes = nullptr;
  if (codes == nullptr || group_starts == nullptr) {
    set_error("b200x_group_boundaries: NULL pointer");
    return B200X_ERR_ARG;
  }
  uint32_t* tile_heads = (uint32_t*)temp;
  heads_count_kernel<<<Tn, G_THREADS, 0, stream>>>(cc, n, tile_heads);
  B200X_LAUNCH_CHECK("heads_count_kernel");
  scan_tiles_kernel<<<1, SCAN_THREADS, 0, stream>>>(tile_heads, Tn, num_groups_out);
  B200X_LAUNCH_CHECK("scan_tiles_kernel");
  group_starts_kernel<<<Tn, G_THREADS, 0, stream>>>(cc, n, tile_heads, group_starts);
  B200X_LAUNCH_CHECK("group_starts_kernel");
  return B200X_OK;
}

extern "C" int b200x_segmented_reduce(const uint32_t* group_starts, const uint32_t* num_groups, const uint32_t* perm,
                                      const void* values, int value_type, size_t n, const double* center,
                                      const int64_t* weights, const double* addend, void* sum_out, void* min_out,
                                      void* max_out, void* temp, size_t* temp_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_segmented_reduce: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  if (value_type != B200X_VAL_I64 && value_type != B200X_VAL_F64) {
    set_error("b200x_segmented_reduce: bad value_type %d", value_type);
    return B200X_ERR_ARG;
  }
  if (n > 0xFFFFFFFFull) {
    set_error("b200x_segmented_reduce: n=%zu exceeds 2^32-1", n);
    return B200X_ERR_TOO_LARGE;
  }
  const size_t need = align_up((size_t)(tiles_of(n) + 1) * sizeof(TileSummary<double>), 256);
  if (temp == nullptr) {
    *temp_bytes = need;
    return B200X_OK;
  }
  if (*temp_bytes < need) {
    set_error("b200x_segmented_reduce: temp buffer %zu < required %zu bytes", *temp_bytes, need);
    return B200X_ERR_TEMP;
  }
  if (n == 0) return B200X_OK;
  if (group_starts == nullptr || num_groups == nullptr || values == nullptr) {
    set_error("b200x_segmented_reduce: NULL pointer");
    return B200X_ERR_ARG;
  }
  if (center != nullptr) {
    if (value_type == B200X_VAL_I64)
      return launch_seg_reduce<int64_t, double, true>(group_starts, num_groups, perm, values, n, center, weights, addend, sum_out, min_out, max_out, temp, stream);
    return launch_seg_reduce<double, double, true>(group_starts, num_groups, perm, values, n, center, weights, addend, sum_out, min_out, max_out, temp, stream);
  }
  if (value_type == B200X_VAL_I64)
    return launch_seg_reduce<int64_t, int64_t, false>(group_starts, num_groups, perm, values, n, nullptr, nullptr, nullptr, sum_out, min_out, max_out, temp, stream);
  return launch_seg_reduce<double, double, false>(group_starts, num_groups, perm, values, n, nullptr, nullptr, nullptr, sum_out, min_out, max_out, temp, stream);
}

extern "C" int b200x_group_counts(const uint32_t* group_starts, const uint32_t* num_groups, size_t max_groups, size_t n,
                                  int64_t* counts_out, void* stream_) {
  if (max_groups == 0) return B200X_OK;
  if (group_starts == nullptr || num_groups == nullptr || counts_out == nullptr) {
    set_error("b200x_group_counts: NULL pointer");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (max_groups + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  group_counts_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(group_starts, num_groups, n, counts_out);
  B200X_LAUNCH_CHECK("group_counts_kernel");
  return B200X_OK;
}

extern "C" int b200x_finalize_groups(const void* sum, int value_type, const int64_t* counts, size_t num_groups, int mode,
                                     int ddof, double* out, void* stream_) {
  if (num_groups == 0) return B200X_OK;
  if (sum == nullptr || counts == nullptr || out == nullptr || (mode != 0 && mode != 1) ||
      (value_type != B200X_VAL_I64 && value_type != B200X_VAL_F64)) {
    set_error("b200x_finalize_groups: bad arguments");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (num_groups + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  finalize_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(sum, value_type, counts, num_groups, mode, ddof, out);
  B200X_LAUNCH_CHECK("finalize_kernel");
  return B200X_OK;
}

extern "C" int b200x_absmax_groups(const void* min_in, const void* max_in, int value_type, size_t num_groups, void* out,
                                   void* stream_) {
  if (num_groups == 0) return B200X_OK;
  if (min_in == nullptr || max_in == nullptr || out == nullptr ||
      (value_type != B200X_VAL_I64 && value_type != B200X_VAL_F64)) {
    set_error("b200x_absmax_groups: bad arguments");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (num_groups + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  absmax_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(min_in, max_in, value_type, num_groups, out);
  B200X_LAUNCH_CHECK("absmax_kernel");
  return B200X_OK;
}

extern "C" int b200x_group_quantile(const uint32_t* group_starts, const uint32_t* num_groups, size_t max_groups,
                                    const uint32_t* perm, const void* values, int value_type, size_t n, double q,
                                    double* out, void* stream_) {
  if (max_groups == 0 || n == 0) return B200X_OK;
  if (group_starts == nullptr || num_groups == nullptr || values == nullptr || out == nullptr ||
      (value_type != B200X_VAL_I64 && value_type != B200X_VAL_F64)) {
    set_error("b200x_group_quantile: bad arguments");
    return B200X_ERR_ARG;
  }
  if (!(q >= 0.0 && q <= 1.0)) {
    set_error("b200x_group_quantile: q=%g outside [0, 1]", q);
    return B200X_ERR_ARG;
  }
  if (n > 0xFFFFFFFFull) {
    set_error("b200x_group_quantile: n=%zu exceeds 2^32-1", n);
    return B200X_ERR_TOO_LARGE;
  }
  uint64_t blocks = (max_groups + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  group_quantile_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>(group_starts, num_groups, perm, values,
                                                                             value_type, n, q, out);
  B200X_LAUNCH_CHECK("group_quantile_kernel");
  return B200X_OK;
}
