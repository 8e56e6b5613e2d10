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
    if (mode == 0) out[g] = c == 0 ? 0.0 : s / c;   // (no non-null value: the column's validity says so; keeps merges NaN-free)
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

// ascending code -> value of the integer key column it came from
__device__ __forceinline__ void store_decoded_key(void* out, size_t i, uint64_t code, int kt) {
  switch (kt) {
    case KEY_I64: ((uint64_t*)out)[i] = code ^ 0x8000000000000000ull; break;
    case KEY_U64: ((uint64_t*)out)[i] = code; break;
    case KEY_I32: ((uint32_t*)out)[i] = (uint32_t)code ^ 0x80000000u; break;
    case KEY_U32: ((uint32_t*)out)[i] = (uint32_t)code; break;
    case KEY_I16: ((uint16_t*)out)[i] = (uint16_t)((uint16_t)code ^ 0x8000u); break;
    case KEY_U16: ((uint16_t*)out)[i] = (uint16_t)code; break;
    case KEY_I8: ((uint8_t*)out)[i] = (uint8_t)((uint8_t)code ^ 0x80u); break;
    default: ((uint8_t*)out)[i] = (uint8_t)code; break;
  }
}

// ---------------------------------------------------------------------------
// Dense-range aggregation: integer keys whose codes span a small range
// ---------------------------------------------------------------------------
// When max(code) - min(code) is small (ids, categories, dates: BASELINE configs[2] groups 200 M rows by 1 M
// distinct int64 keys) the group of a row is simply code - min: no sort is needed to bring equal keys together.
// Every row adds its value to accumulator tables indexed by that number (count, and per value column sum / min /
// max as int64); the non-empty entries, read in index order, ARE the groups in ascending key order. Integer
// sums, counts, minima and maxima do not depend on the order of the additions, so the result is bit-identical
// to the sorted path; float columns keep the sorted path (a fixed reduction order is what makes them
// reproducible). Heavy keys (Zipf) would serialise on one table entry: a block first combines the rows of one
// tile (DG_TILE rows) in a shared-memory hash table that is emptied after every tile, so that a key costs one
// global atomic per tile it occurs in, however often it occurs there.
constexpr int DG_THREADS = 256;
constexpr int DG_IPT = 4;
constexpr int DG_TILE = DG_THREADS * DG_IPT;
constexpr int DG_SLOTS = 2048;  // power of two, 2x the tile

__global__ void __launch_bounds__(256)
key_code_range_kernel(const void* __restrict__ keys, int kt, uint64_t n, unsigned long long* __restrict__ out) {
  unsigned long long mn = ~0ull, mx = 0ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const unsigned long long c = load_encode(keys, i, kt, false);
    mn = c < mn ? c : mn;
    mx = c > mx ? c : mx;
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, mn, d), b = __shfl_xor_sync(0xffffffffu, mx, d);
    mn = a < mn ? a : mn;
    mx = b > mx ? b : mx;
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(&out[0], mn);
    atomicMax(&out[1], mx);
  }
}

__global__ void __launch_bounds__(256)
dense_init_kernel(uint64_t range, long long* cnt, long long* sum, long long* mn, long long* mx) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < range; i += (uint64_t)gridDim.x * blockDim.x) {
    if (cnt) cnt[i] = 0;
    if (sum) sum[i] = 0;
    if (mn) mn[i] = INT64_MAX;
    if (mx) mx[i] = INT64_MIN;
  }
}

__global__ void __launch_bounds__(DG_THREADS)
dense_accumulate_kernel(const void* __restrict__ keys, int kt, uint64_t n, uint64_t base, uint32_t range,
                        const long long* __restrict__ values, long long* __restrict__ g_cnt, long long* __restrict__ g_sum,
                        long long* __restrict__ g_mn, long long* __restrict__ g_mx) {
  extern __shared__ __align__(16) unsigned char dg_raw[];
  uint32_t* s_tag = reinterpret_cast<uint32_t*>(dg_raw);                 // group number + 1, 0 = free
  uint32_t* s_cnt = s_tag + DG_SLOTS;
  long long* s_sum = reinterpret_cast<long long*>(s_cnt + DG_SLOTS);     // only the arrays that are used exist
  long long* s_mn = s_sum + (g_sum ? DG_SLOTS : 0);
  long long* s_mx = s_mn + (g_mn ? DG_SLOTS : 0);
  for (int i = threadIdx.x; i < DG_SLOTS; i += DG_THREADS) {
    s_tag[i] = 0;
    s_cnt[i] = 0;
    if (g_sum) s_sum[i] = 0;
    if (g_mn) s_mn[i] = INT64_MAX;
    if (g_mx) s_mx[i] = INT64_MIN;
  }
  __syncthreads();
  const uint64_t tiles = (n + DG_TILE - 1) / DG_TILE;
  for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const uint64_t r0 = tile * DG_TILE + threadIdx.x;
    unsigned long long code[DG_IPT];
    long long v[DG_IPT];
#pragma unroll
    for (int j = 0; j < DG_IPT; ++j) {
      const uint64_t r = r0 + (uint64_t)j * DG_THREADS;
      code[j] = r < n ? load_encode(keys, r, kt, false) : 0ull;
      v[j] = (values != nullptr && r < n) ? values[r] : 0ll;
    }
#pragma unroll
    for (int j = 0; j < DG_IPT; ++j) {
      const uint64_t r = r0 + (uint64_t)j * DG_THREADS;
      if (r >= n) continue;
      const uint32_t gidx = (uint32_t)(code[j] - base);
      if (gidx >= range) continue;                       // cannot happen when base / range come from the column itself
      uint32_t slot = (gidx * 2654435761u) >> 21;        // 11 bits
      bool placed = false;
#pragma unroll
      for (int probe = 0; probe < 2 && !placed; ++probe) {
        uint32_t t = s_tag[slot];
        if (t == 0) t = atomicCAS(&s_tag[slot], 0u, gidx + 1u);
        if (t == 0 || t == gidx + 1u) {
          atomicAdd(&s_cnt[slot], 1u);
          if (g_sum) atomicAdd(reinterpret_cast<unsigned long long*>(&s_sum[slot]), (unsigned long long)v[j]);
          if (g_mn) atomicMin(&s_mn[slot], v[j]);
          if (g_mx) atomicMax(&s_mx[slot], v[j]);
          placed = true;
        }
        slot = (slot + 1) & (DG_SLOTS - 1);
      }
      if (!placed) {                                     // both slots taken by other keys of this tile
        if (g_cnt) atomicAdd(reinterpret_cast<unsigned long long*>(&g_cnt[gidx]), 1ull);
        if (g_sum) atomicAdd(reinterpret_cast<unsigned long long*>(&g_sum[gidx]), (unsigned long long)v[j]);
        if (g_mn) atomicMin(&g_mn[gidx], v[j]);
        if (g_mx) atomicMax(&g_mx[gidx], v[j]);
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < DG_SLOTS; i += DG_THREADS) {   // flush the tile's table and empty it
      const uint32_t t = s_tag[i];
      if (t != 0) {
        const uint32_t gidx = t - 1u;
        if (g_cnt) atomicAdd(reinterpret_cast<unsigned long long*>(&g_cnt[gidx]), (unsigned long long)s_cnt[i]);
        if (g_sum) atomicAdd(reinterpret_cast<unsigned long long*>(&g_sum[gidx]), (unsigned long long)s_sum[i]);
        if (g_mn) atomicMin(&g_mn[gidx], s_mn[i]);
        if (g_mx) atomicMax(&g_mx[gidx], s_mx[i]);
        s_tag[i] = 0;
        s_cnt[i] = 0;
        if (g_sum) s_sum[i] = 0;
        if (g_mn) s_mn[i] = INT64_MAX;
        if (g_mx) s_mx[i] = INT64_MIN;
      }
    }
    __syncthreads();
  }
}

// Compaction of the non-empty table entries, in index (= key) order: per-block counts, one-block scan, write-out.
constexpr int DC_THREADS = 256;
constexpr int DC_IPT = 8;
constexpr int DC_TILE = DC_THREADS * DC_IPT;
constexpr int DC_MAX_TABS = 8;
struct DenseTabs {
  const long long* in[DC_MAX_TABS];
  long long* out[DC_MAX_TABS];
  int n;
};

__global__ void __launch_bounds__(DC_THREADS)
dense_count_kernel(const long long* __restrict__ cnt, uint64_t range, uint32_t* __restrict__ block_counts) {
  __shared__ uint32_t s_sum;
  if (threadIdx.x == 0) s_sum = 0;
  __syncthreads();
  uint32_t c = 0;
  const uint64_t base = (uint64_t)blockIdx.x * DC_TILE;
#pragma unroll
  for (int j = 0; j < DC_IPT; ++j) {
    const uint64_t i = base + (uint64_t)j * DC_THREADS + threadIdx.x;
    c += (i < range && cnt[i] != 0) ? 1u : 0u;
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_sum, c);
  __syncthreads();
  if (threadIdx.x == 0) block_counts[blockIdx.x] = s_sum;
}

__global__ void __launch_bounds__(SCAN_THREADS)
dense_scan_kernel(uint32_t* __restrict__ block_counts, uint32_t blocks, uint32_t* __restrict__ num_groups_out) {
  __shared__ uint32_t s_warp[32];
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t per = (blocks + SCAN_THREADS - 1) / SCAN_THREADS;
  const uint32_t lo = min(blocks, threadIdx.x * per), hi = min(blocks, lo + per);
  uint32_t sum = 0;
  for (uint32_t t = lo; t < hi; ++t) sum += block_counts[t];
  uint32_t x = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  uint32_t off = 0;
  for (unsigned k = 0; k < w; ++k) off += s_warp[k];
  uint32_t run = off + x - sum;
  for (uint32_t t = lo; t < hi; ++t) {
    const uint32_t c = block_counts[t];
    block_counts[t] = run;
    run += c;
  }
  if (threadIdx.x == SCAN_THREADS - 1) *num_groups_out = off + x;
}

__global__ void __launch_bounds__(DC_THREADS)
dense_write_kernel(const long long* __restrict__ cnt, uint64_t range, uint64_t base_code, int kt,
                   const uint32_t* __restrict__ block_base, void* __restrict__ keys_out, long long* __restrict__ count_out,
                   const DenseTabs tabs) {
  // entries of a tile in index order: thread t owns DC_IPT consecutive entries
  __shared__ uint32_t s_warp[DC_THREADS / 32];
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint64_t first = (uint64_t)blockIdx.x * DC_TILE + (uint64_t)threadIdx.x * DC_IPT;
  uint32_t mine = 0;
  bool live[DC_IPT];
#pragma unroll
  for (int j = 0; j < DC_IPT; ++j) {
    live[j] = first + j < range && cnt[first + j] != 0;
    mine += live[j] ? 1u : 0u;
  }
  uint32_t x = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  uint32_t off = block_base[blockIdx.x];
  for (unsigned k = 0; k < w; ++k) off += s_warp[k];
  uint32_t g = off + x - mine;
#pragma unroll
  for (int j = 0; j < DC_IPT; ++j) {
    if (!live[j]) continue;
    const uint64_t i = first + j;
    store_decoded_key(keys_out, g, base_code + i, kt);
    if (count_out) count_out[g] = cnt[i];
#pragma unroll
    for (int t = 0; t < DC_MAX_TABS; ++t)
      if (t < tabs.n) tabs.out[t][g] = tabs.in[t][i];
    ++g;
  }
}

// ---------------------------------------------------------------------------
// Nulls in value columns: the reference's null policy (null_aggregate.py:11-17)
// ---------------------------------------------------------------------------
// A nullable column is its values plus a validity byte per row. The reductions above have no notion of validity, so a
// null row is given the identity of the reduction it enters - 0 for sums, the type's largest / smallest value for
// min / max - and the number of valid rows per group is reduced like any other int64 column (out_valid64).
template <typename T>
__global__ void __launch_bounds__(256)
mask_values_kernel(const T* __restrict__ values, const unsigned char* __restrict__ valid, uint64_t n, T fill,
                   T* __restrict__ out_values, long long* __restrict__ out_valid64) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const bool ok = valid[i] != 0;
    if (out_values) out_values[i] = ok ? values[i] : fill;
    if (out_valid64) out_valid64[i] = ok ? 1ll : 0ll;
  }
}

// validity of a group's result: ignore_nulls: at least one valid row; else: no null row at all (and at least one row)
__global__ void __launch_bounds__(256)
group_validity_kernel(const long long* __restrict__ nvalid, const long long* __restrict__ rows, uint64_t g, int ignore_nulls,
                      unsigned char* __restrict__ out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = (nvalid[i] > 0 && (ignore_nulls || nvalid[i] == rows[i])) ? 1 : 0;
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
  B200X_ON_DEVICE_OF(group_starts);
  const uint64_t* one[1] = {codes};
  return b200x_group_boundaries_multi(one, 1, n, group_starts, num_groups_out, temp, temp_bytes, stream_);
}

extern "C" int b200x_group_boundaries_multi(const uint64_t* const* codes_host, int num_cols, size_t n,
                                            uint32_t* group_starts, uint32_t* num_groups_out, void* temp,
                                            size_t* temp_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(group_starts);
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
  B200X_ON_DEVICE_OF(values);
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
  B200X_ON_DEVICE_OF(counts_out);
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
  B200X_ON_DEVICE_OF(out);
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
  B200X_ON_DEVICE_OF(out);
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
  B200X_ON_DEVICE_OF(out);
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

extern "C" int b200x_key_code_range(const void* keys, int key_type, size_t n, uint64_t* range_out, void* stream_) {
  B200X_ON_DEVICE_OF(range_out);
  cudaStream_t st = (cudaStream_t)stream_;
  if (range_out == nullptr || key_type < 0 || key_type >= KEY_TYPE_COUNT || (n > 0 && keys == nullptr)) {
    set_error("b200x_key_code_range: bad arguments");
    return B200X_ERR_ARG;
  }
  const unsigned long long init[2] = {~0ull, 0ull};
  B200X_CUDA(cudaMemcpyAsync(range_out, init, sizeof(init), cudaMemcpyHostToDevice, st));
  if (n == 0) return B200X_OK;
  uint64_t blocks = (n + 255) / 256 / 4 + 1;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  key_code_range_kernel<<<(uint32_t)blocks, 256, 0, st>>>(keys, key_type, n, (unsigned long long*)range_out);
  B200X_LAUNCH_CHECK("key_code_range_kernel");
  return B200X_OK;
}

extern "C" int b200x_dense_group_accumulate(const void* keys, int key_type, size_t n, uint64_t base_code, size_t range,
                                            const int64_t* values, int init, int64_t* count_tab, int64_t* sum_tab,
                                            int64_t* min_tab, int64_t* max_tab, void* stream_) {
  B200X_ON_DEVICE_OF((keys ? (const void*)keys : (const void*)count_tab));
  cudaStream_t st = (cudaStream_t)stream_;
  if (key_type < 0 || key_type >= KEY_TYPE_COUNT || key_type == KEY_F64 || key_type == KEY_F32 || range == 0 ||
      range > ((size_t)1 << 30) || (n > 0 && keys == nullptr) ||
      (values == nullptr && (sum_tab != nullptr || min_tab != nullptr || max_tab != nullptr)) ||
      (count_tab == nullptr && sum_tab == nullptr && min_tab == nullptr && max_tab == nullptr)) {
    set_error("b200x_dense_group_accumulate: bad arguments (integer key types, 1 <= range <= 2^30, values needed for sum/min/max)");
    return B200X_ERR_ARG;
  }
  if (init) {
    uint64_t blocks = (range + 255) / 256;
    const uint64_t cap = (uint64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    dense_init_kernel<<<(uint32_t)blocks, 256, 0, st>>>(range, (long long*)count_tab, (long long*)sum_tab, (long long*)min_tab,
                                                        (long long*)max_tab);
    B200X_LAUNCH_CHECK("dense_init_kernel");
  }
  if (n == 0) return B200X_OK;
  const size_t smem = (size_t)DG_SLOTS * (8 + 8 * ((sum_tab ? 1 : 0) + (min_tab ? 1 : 0) + (max_tab ? 1 : 0)));
  static thread_local int attr_dev = -1;
  int dev = 0;
  B200X_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200X_CUDA(cudaFuncSetAttribute(dense_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SLOTS * 32));
    attr_dev = dev;
  }
  uint64_t tiles = (n + DG_TILE - 1) / DG_TILE;
  const uint64_t cap = (uint64_t)sm_count() * 6;
  if (tiles > cap) tiles = cap;
  prof_begin(st, PROF_SEGRED);
  dense_accumulate_kernel<<<(uint32_t)tiles, DG_THREADS, smem, st>>>(keys, key_type, n, base_code, (uint32_t)range,
                                                                    (const long long*)values, (long long*)count_tab,
                                                                    (long long*)sum_tab, (long long*)min_tab,
                                                                    (long long*)max_tab);
  prof_end(st);
  B200X_LAUNCH_CHECK("dense_accumulate_kernel");
  return B200X_OK;
}

extern "C" int b200x_dense_group_compact(uint64_t base_code, size_t range, int key_type, const int64_t* count_tab,
                                         int num_tabs, const int64_t* const* tabs_host, void* keys_out, int64_t* count_out,
                                         int64_t* const* tabs_out_host, uint32_t* num_groups_out, void* temp,
                                         size_t* temp_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(count_tab);
  cudaStream_t st = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_dense_group_compact: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  const uint32_t blocks = (uint32_t)((range + DC_TILE - 1) / DC_TILE);
  const size_t need = align_up((size_t)(blocks + 1) * sizeof(uint32_t), 256);
  if (temp == nullptr) {
    *temp_bytes = need;
    return B200X_OK;
  }
  if (*temp_bytes < need || range == 0 || range > ((size_t)1 << 30) || count_tab == nullptr || keys_out == nullptr ||
      num_groups_out == nullptr || num_tabs < 0 || num_tabs > DC_MAX_TABS || key_type < 0 || key_type >= KEY_TYPE_COUNT ||
      (num_tabs > 0 && (tabs_host == nullptr || tabs_out_host == nullptr))) {
    set_error("b200x_dense_group_compact: bad arguments or temp too small (%zu < %zu)", *temp_bytes, need);
    return B200X_ERR_ARG;
  }
  DenseTabs tabs;
  tabs.n = num_tabs;
  for (int t = 0; t < DC_MAX_TABS; ++t) {
    tabs.in[t] = t < num_tabs ? (const long long*)tabs_host[t] : nullptr;
    tabs.out[t] = t < num_tabs ? (long long*)tabs_out_host[t] : nullptr;
  }
  uint32_t* block_counts = (uint32_t*)temp;
  dense_count_kernel<<<blocks, DC_THREADS, 0, st>>>((const long long*)count_tab, range, block_counts);
  B200X_LAUNCH_CHECK("dense_count_kernel");
  dense_scan_kernel<<<1, SCAN_THREADS, 0, st>>>(block_counts, blocks, num_groups_out);
  B200X_LAUNCH_CHECK("dense_scan_kernel");
  dense_write_kernel<<<blocks, DC_THREADS, 0, st>>>((const long long*)count_tab, range, base_code, key_type, block_counts,
                                                    keys_out, (long long*)count_out, tabs);
  B200X_LAUNCH_CHECK("dense_write_kernel");
  return B200X_OK;
}

extern "C" int b200x_mask_values(const void* values, int value_type, const uint8_t* valid, size_t n, int fill_mode,
                                 void* out_values, int64_t* out_valid64, void* stream_) {
  B200X_ON_DEVICE_OF(valid);
  cudaStream_t st = (cudaStream_t)stream_;
  if (n == 0) return B200X_OK;
  if (valid == nullptr || (out_values == nullptr && out_valid64 == nullptr) || (out_values != nullptr && values == nullptr) ||
      (value_type != 0 && value_type != 1) || fill_mode < 0 || fill_mode > 2) {
    set_error("b200x_mask_values: bad arguments (int64 / float64 values, fill_mode 0 zero, 1 largest, 2 smallest)");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (value_type == 0) {   // B200X_VAL_I64
    const int64_t fill = fill_mode == 0 ? 0 : (fill_mode == 1 ? INT64_MAX : INT64_MIN);
    mask_values_kernel<int64_t><<<(uint32_t)blocks, 256, 0, st>>>((const int64_t*)values, valid, n, fill, (int64_t*)out_values,
                                                                  (long long*)out_valid64);
  } else {
    const double fill = fill_mode == 0 ? 0.0 : (fill_mode == 1 ? (double)INFINITY : -(double)INFINITY);
    mask_values_kernel<double><<<(uint32_t)blocks, 256, 0, st>>>((const double*)values, valid, n, fill, (double*)out_values,
                                                                 (long long*)out_valid64);
  }
  B200X_LAUNCH_CHECK("mask_values_kernel");
  return B200X_OK;
}

extern "C" int b200x_group_validity(const int64_t* nvalid, const int64_t* rows, size_t num_groups, int ignore_nulls,
                                    uint8_t* out, void* stream_) {
  B200X_ON_DEVICE_OF(out);
  if (num_groups == 0) return B200X_OK;
  if (nvalid == nullptr || rows == nullptr || out == nullptr) {
    set_error("b200x_group_validity: NULL pointer");
    return B200X_ERR_ARG;
  }
  uint64_t blocks = (num_groups + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  group_validity_kernel<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream_>>>((const long long*)nvalid, (const long long*)rows,
                                                                           num_groups, ignore_nulls, out);
  B200X_LAUNCH_CHECK("group_validity_kernel");
  return B200X_OK;
}
