// Stable LSD radix argsort ("onesweep": one histogram sweep up front, then one
// read + one write of the (code, row-index) pairs per 8-bit digit, chained
// between thread blocks with decoupled look-back).
//
// Replaces, for a device-resident block, the reference's
//   pyarrow.compute.sort_indices(table, sort_keys=...)
// call in transform_pyarrow.sort / concat_and_sort
// (python/ray/data/_internal/arrow_ops/transform_pyarrow.py:17-21, :269-276),
// reached from ArrowBlockAccessor.sort_and_partition / merge_sorted_blocks
// (python/ray/data/_internal/arrow_block.py:449-464, :554-567).
//
// Properties the parity contract needs:
//   * stable (equal keys keep input order) in ascending AND descending order
//     (descending = ascending sort of complemented codes, see common.cuh);
//   * -0.0 ties with +0.0, NaN sorts last in both directions;
//   * multi-column keys = chained calls, last key column first, feeding
//     perm_out of one call as perm_in of the next.
//
// Everything is decided on the device (which digits are constant and can be
// skipped, which buffer a pass reads / writes), so the entry point never
// synchronises with the host.
#include "common.cuh"
#include "../../include/b200x.h"

namespace b200x {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 8;

// look-back word: [flag:2][count:30]
constexpr uint32_t LB_FLAG_AGG = 1u << 30;   // tile-local count only
constexpr uint32_t LB_FLAG_INCL = 2u << 30;  // inclusive prefix over tiles 0..t
constexpr uint32_t LB_VALUE_MASK = (1u << 30) - 1;

struct SortPlan {
  uint32_t active[MAX_PASSES];   // 1 if the digit varies over the input
  uint32_t src_sel[MAX_PASSES];  // ping-pong buffer the pass reads from (ignored for the first active pass)
  int32_t first;                 // first active pass (-1: none)
  int32_t last;                  // last active pass (-1: none)
  uint32_t tile_counter[MAX_PASSES];
  uint32_t shift[MAX_PASSES];    // bit position of the 8-bit digit of each pass
  uint32_t pad[6];
};

// Which algorithm sorts this input; decided on the device from the keys themselves.
//   LSD    : one pass per varying byte, least significant first (any input).
//   HYBRID : two passes on the top 16 varying bits (LSD order, so the result is grouped into at
//            most 65536 runs by those bits), then every run - a few thousand pairs - is finished
//            in shared memory over the remaining low bits. 3 trips through HBM instead of up to 8.
//            Chosen when there are enough rows, more than 16 varying bits and no run exceeds what
//            one block can hold. The 65536 runs are cut either
//              U: uniformly, bin = (code - smallest code) >> shiftA (random ints, hashes, ids), or
//              T: through a 65536-entry table built from a sample when U overflows: the U bins become
//                 coarse buckets; a bucket holding more than the target number of rows is split into
//                 2^s bins by its next s bits, consecutive sparse buckets are merged into one bin, so
//                 that every bin holds about the same number of rows (a piecewise-linear CDF: floats
//                 from any smooth distribution, clustered timestamps, skewed ids).
constexpr uint32_t MODE_LSD = 0, MODE_HYBRID = 1;
constexpr uint32_t JOINT_BINS = 1u << 16;
constexpr int KEY_FIELD = 48;  // T: pass A rewrites the key as [bin:16][bits below the bin, left-aligned:48]
constexpr uint32_t TAB_MERGED = 31;  // table entry = bin << 5 | s; s = 31: the bucket shares its bin with neighbours

struct SortCtl {
  unsigned long long max_code;   // largest code
  unsigned long long max_inv;    // largest ~code (smallest code = ~max_inv); both start at 0
  unsigned long long base;       // smallest code: HYBRID sorts code - base, so a narrow key range anywhere
                                 // in the 64-bit space (one rank's slice of a distributed sort, timestamps)
                                 // still spreads over all 65536 runs
  uint32_t candidate;            // hybrid is worth checking (enough rows, > 16 varying bits)
  uint32_t mode;
  uint32_t shiftA, shiftB;       // digit positions of the two global passes (low one first)
  uint32_t max_run;
  uint32_t tmode;                // 1: bins come from the table (T), 0: uniform cut (U)
  uint32_t shiftC;               // T: coarse bucket = (code - base) >> shiftC (= U's shiftA, <= 48)
  uint32_t pad[5];
};

// T: key -> [bin:16][remaining bits:48], order-preserving. A split bucket keeps the R = shiftC - s bits
// below its bin exactly (R <= 48). A merged bin spans several buckets: the field gets the key's top 48
// bits (monotone, not exact) and the finisher re-reads those few keys from the column.
__device__ __forceinline__ uint64_t remap_key(uint64_t k, const uint32_t* __restrict__ tab, uint32_t shiftC) {
  const uint32_t e = __ldg(tab + (uint32_t)(k >> shiftC));
  const uint32_t sb = e & 31u;
  const uint64_t hi = (uint64_t)(e >> 5) << KEY_FIELD;
  if (sb == TAB_MERGED) return hi | (k >> 16);
  const uint64_t rem = k & ((1ull << shiftC) - 1ull);
  return hi + (rem << (KEY_FIELD - ((int)shiftC - (int)sb)));
}

// ---------------------------------------------------------------------------
// Kernel 1: all-digit histogram of the encoded keys (one read of the key column)
// ---------------------------------------------------------------------------
constexpr int HIST_THREADS = 512;
constexpr int HIST_IPT = 8;

__global__ void __launch_bounds__(HIST_THREADS)
hist_kernel(const void* __restrict__ raw, const uint32_t* __restrict__ perm_in, int kt, int desc, uint32_t n,
            int npass, uint32_t* __restrict__ ghist /*[MAX_PASSES][RADIX]*/, const SortCtl* __restrict__ ctl) {
  if (ctl->mode == MODE_HYBRID) return;
  __shared__ uint32_t sh[MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < MAX_PASSES * RADIX; i += HIST_THREADS) sh[i] = 0;
  __syncthreads();

  const uint32_t chunk = HIST_THREADS * HIST_IPT;
  const unsigned lane = threadIdx.x & 31;
  for (uint64_t base = (uint64_t)blockIdx.x * chunk; base < n; base += (uint64_t)gridDim.x * chunk) {
    uint64_t code[HIST_IPT];
    bool valid[HIST_IPT];
#pragma unroll
    for (int j = 0; j < HIST_IPT; ++j) {
      uint64_t i = base + (uint64_t)j * HIST_THREADS + threadIdx.x;
      valid[j] = i < n;
      code[j] = 0;
      if (valid[j]) {
        size_t src = perm_in ? (size_t)perm_in[i] : (size_t)i;
        code[j] = load_encode(raw, src, kt, desc != 0);
      }
    }
#pragma unroll
    for (int j = 0; j < HIST_IPT; ++j) {
      const unsigned act = __ballot_sync(0xffffffffu, valid[j]);
      if (!valid[j]) continue;
      for (int p = 0; p < npass; ++p) {
        const uint32_t d = (uint32_t)(code[j] >> (p * RADIX_BITS)) & (RADIX - 1);
        int same;
        __match_all_sync(act, d, &same);
        if (same) {
          // whole warp hits one bin (constant digit): one add instead of a 32-way conflict
          if (lane == (unsigned)(__ffs(act) - 1)) atomicAdd(&sh[p * RADIX + d], (uint32_t)__popc(act));
        } else {
          atomicAdd(&sh[p * RADIX + d], 1u);
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npass * RADIX; i += HIST_THREADS) {
    uint32_t v = sh[i];
    if (v) atomicAdd(&ghist[i], v);
  }
}

// ---------------------------------------------------------------------------
// Kernel 2: per-digit exclusive scan + plan (1 block of RADIX threads)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(RADIX)
plan_kernel(uint32_t* __restrict__ ghist, SortPlan* __restrict__ plan, uint32_t n, int npass) {
  __shared__ uint32_t warp_tot[RADIX / 32];
  __shared__ uint32_t s_active[MAX_PASSES];
  const int d = threadIdx.x;
  const unsigned lane = d & 31, w = d >> 5;
  if (d < MAX_PASSES) s_active[d] = 0;
  __syncthreads();
  for (int p = 0; p < npass; ++p) {
    const uint32_t c = ghist[p * RADIX + d];
    // a digit is constant iff one bin holds all n keys
    if (c != 0 && c != n) s_active[p] = 1;  // benign race: all writers store 1
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    uint32_t off = 0;
    for (unsigned k = 0; k < w; ++k) off += warp_tot[k];
    ghist[p * RADIX + d] = off + x - c;  // exclusive
    __syncthreads();
  }
  if (d == 0) {
    int first = -1, last = -1;
    uint32_t sel = 0;
    for (int p = 0; p < MAX_PASSES; ++p) {
      const uint32_t a = (p < npass) ? s_active[p] : 0;
      plan->active[p] = a;
      plan->tile_counter[p] = 0;
      plan->src_sel[p] = 0;
      plan->shift[p] = (uint32_t)(p * RADIX_BITS);
      if (a) {
        if (first < 0) {
          first = p;  // reads the raw column, writes buffer 0
          sel = 0;
        } else {
          plan->src_sel[p] = sel;
          sel ^= 1;
        }
        last = p;
      }
    }
    plan->first = first;
    plan->last = last;
  }
}

// ---------------------------------------------------------------------------
// Kernel 3: one digit pass
// ---------------------------------------------------------------------------
struct PassArgs {
  const void* raw;          // raw key column (first active pass only)
  const uint32_t* perm_in;  // optional row indirection for the first pass
  uint64_t* kbuf[2];        // ping-pong code buffers
  uint32_t* vbuf[2];        // ping-pong row-index buffers
  uint64_t* codes_out;      // optional: sorted codes (last pass)
  void* keys_out;           // optional: sorted keys decoded to the column type (last pass, integer types)
  uint32_t* perm_out;       // sorted row indices (last pass)
  const uint32_t* gbase;    // [RADIX] exclusive digit offsets of this pass
  uint32_t* lookback;       // [tiles][RADIX]
  SortPlan* plan;
  const SortCtl* ctl;       // HYBRID passes sort code - ctl->base (NULL: plain codes)
  uint32_t* mark_runs;      // HYBRID pass B: [65536] first position of every (B, A) run, found while writing
  uint32_t n;
  int pass;
  int kt;
  int desc;
};

// Lanes of the warp that hold the same 8-bit digit. MATCH.ANY runs on the slow ADU pipe
// (it capped the pass at ~65% ADU utilisation in profiles/), so the peer mask is built
// from one VOTE.BALLOT per digit bit: test bit -> ballot -> complement if my bit is 0 -> and.
__device__ __forceinline__ unsigned match_digit(uint32_t d) {
  unsigned peers;
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " .reg .b32 m, t;\n"
      " and.b32 t, %1, 1;   setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; mov.b32 %0, m;\n"
      " and.b32 t, %1, 2;   setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 4;   setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 8;   setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 16;  setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 32;  setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 64;  setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      " and.b32 t, %1, 128; setp.ne.u32 p, t, 0; vote.sync.ballot.b32 m, p, 0xffffffff; @!p not.b32 m, m; and.b32 %0, %0, m;\n"
      "}\n"
      : "=r"(peers)
      : "r"(d));
  return peers;
}

// code -> value of the (integer) key column it came from
__device__ __forceinline__ void store_decoded(void* out, size_t i, uint64_t code, int kt, bool desc) {
  switch (kt) {
    case KEY_I64: ((uint64_t*)out)[i] = (desc ? ~code : code) ^ 0x8000000000000000ull; break;
    case KEY_U64: ((uint64_t*)out)[i] = desc ? ~code : code; break;
    case KEY_I32: ((uint32_t*)out)[i] = (uint32_t)(desc ? ~code : code) ^ 0x80000000u; break;
    case KEY_U32: ((uint32_t*)out)[i] = (uint32_t)(desc ? ~code : code); break;
    case KEY_I16: ((uint16_t*)out)[i] = (uint16_t)((uint16_t)(desc ? ~code : code) ^ 0x8000u); break;
    case KEY_U16: ((uint16_t*)out)[i] = (uint16_t)(desc ? ~code : code); break;
    case KEY_I8: ((uint8_t*)out)[i] = (uint8_t)((uint8_t)(desc ? ~code : code) ^ 0x80u); break;
    default: ((uint8_t*)out)[i] = (uint8_t)(desc ? ~code : code); break;
  }
}

template <int THREADS, int IPT>
struct PassSmem {
  static constexpr int TILE = THREADS * IPT;
  static constexpr int WARPS = THREADS / 32;
  static constexpr size_t bytes = (size_t)TILE * 8 + (size_t)TILE * 4 + (size_t)WARPS * RADIX * 2 + RADIX * 4 + 64;
};

// WALK = false: one block per tile (fastest when the pass is active: blocks in different phases overlap).
// WALK = true: as many blocks as are resident, each claiming tiles until none is left - launching it
// for a pass that turns out inactive costs nothing, whereas ~20000 blocks that return at once cost
// ~50 us per launch. The LSD passes, which idle whenever the device picks HYBRID, use WALK.
template <int THREADS, int IPT, int MIN_BLOCKS, bool WALK>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
onesweep_pass_kernel(const PassArgs a) {
  constexpr int TILE = THREADS * IPT;
  constexpr int WARPS = THREADS / 32;
  static_assert(THREADS >= RADIX, "need one thread per digit");
  static_assert(TILE < 65536, "tile-local offsets are kept in 16 bits");

  SortPlan* plan = a.plan;
  if (!plan->active[a.pass]) return;

  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* skeys = (uint64_t*)smem_raw;                       // [TILE]
  uint32_t* svals = (uint32_t*)(skeys + TILE);                 // [TILE]
  uint32_t* obase = svals + TILE;                              // [RADIX] global base - tile-local start
  uint16_t* whist = (uint16_t*)(obase + RADIX);                // [WARPS][RADIX]
  __shared__ uint32_t s_tile;
  __shared__ uint32_t swarp[RADIX / 32];

  const int tid = threadIdx.x;
  const unsigned lane = tid & 31;
  const int warp = tid >> 5;
  const int shift = (int)plan->shift[a.pass];
  const bool is_first = plan->first == a.pass;
  const bool is_last = plan->last == a.pass;

  const uint32_t n = a.n;
  const uint32_t num_tiles = (uint32_t)(((uint64_t)n + TILE - 1) / TILE);
  do {
  // Tiles are handed out in launch order so that a tile only ever waits on
  // tiles that are already resident (no look-back deadlock).
  if (tid == 0) s_tile = atomicAdd(&plan->tile_counter[a.pass], 1u);
  for (int i = tid; i < WARPS * RADIX / 2; i += THREADS) ((uint32_t*)whist)[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  if (WALK && tile >= num_tiles) break;
  const uint64_t tile_base = (uint64_t)tile * TILE;
  const uint32_t n_valid = (uint32_t)min((uint64_t)TILE, (uint64_t)n - tile_base);

  // ---- load: warp-striped, item j of lane l = tile_base + warp*32*IPT + j*32 + l
  uint64_t key[IPT];
  uint32_t val[IPT];
  const uint32_t wbase = (uint32_t)warp * 32 * IPT + lane;
  if (is_first) {
    const int width = key_type_bytes(a.kt);
    const unsigned long long base = a.ctl ? a.ctl->base : 0ull;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t t = wbase + j * 32;
      val[j] = (uint32_t)(tile_base + t);
      if (a.perm_in) val[j] = (t < n_valid) ? a.perm_in[tile_base + t] : 0u;
    }
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t t = wbase + j * 32;
      key[j] = (t < n_valid) ? load_raw_bits(a.raw, (size_t)val[j], width) : 0ull;
    }
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t t = wbase + j * 32;
      key[j] = (t < n_valid) ? encode_bits(key[j], a.kt, a.desc != 0) - base : ~0ull;
      if (t >= n_valid) val[j] = 0;
    }
  } else {
    const uint32_t sel = plan->src_sel[a.pass];
    const uint64_t* kin = (sel ? a.kbuf[1] : a.kbuf[0]) + tile_base;
    const uint32_t* vin = (sel ? a.vbuf[1] : a.vbuf[0]) + tile_base;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t t = wbase + j * 32;
      key[j] = ~0ull;
      val[j] = 0;
      if (t < n_valid) {
        key[j] = ld_stream_u64(kin + t);
        val[j] = ld_stream_u32(vin + t);
      }
    }
  }

  // ---- rank within the warp (stable: order = (item j, lane))
  uint16_t* wh = whist + warp * RADIX;
  uint32_t rank[IPT];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int j = 0; j < IPT; ++j) {
    const uint32_t d = (uint32_t)(key[j] >> shift) & (RADIX - 1);
    const unsigned peers = match_digit(d);
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if ((int)lane == leader) {
      old = wh[d];
      wh[d] = (uint16_t)(old + __popc(peers));
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[j] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();

  // ---- per-digit: scan over warps, tile histogram, look-back
  uint32_t my_count = 0, my_start = 0;
  if (tid < RADIX) {
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = whist[w * RADIX + tid];
      whist[w * RADIX + tid] = (uint16_t)sum;
      sum += c;
    }
    // padding items of a partial tile all carry digit 255 and sit at the end
    if (tid == RADIX - 1) sum -= (uint32_t)(TILE - n_valid);
    my_count = sum;
    // publish the tile-local count as early as possible
    volatile uint32_t* lb = a.lookback + (size_t)tile * RADIX + tid;
    *lb = (tile == 0 ? LB_FLAG_INCL : LB_FLAG_AGG) | my_count;

    // exclusive scan of counts over the 256 digits -> start of each digit inside the tile
    uint32_t x = my_count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) swarp[warp] = x;
    my_start = x - my_count;
  }
  __syncthreads();
  if (tid < RADIX) {
    uint32_t off = 0;
    for (int k = 0; k < warp; ++k) off += swarp[k];
    my_start += off;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) whist[w * RADIX + tid] += (uint16_t)my_start;
  }
  __syncthreads();

  // ---- scatter into shared memory in tile-sorted order
#pragma unroll
  for (int j = 0; j < IPT; ++j) {
    const uint32_t d = (uint32_t)(key[j] >> shift) & (RADIX - 1);
    const uint32_t pos = (uint32_t)wh[d] + rank[j];
    skeys[pos] = key[j];
    svals[pos] = val[j];
  }

  // ---- decoupled look-back (digit threads), overlapped with the scatter above
  if (tid < RADIX) {
    uint32_t excl = 0;
    if (tile > 0) {
      int64_t t = (int64_t)tile - 1;
      while (true) {
        volatile uint32_t* p = a.lookback + (size_t)t * RADIX + tid;
        uint32_t v = *p;
        while ((v & ~LB_VALUE_MASK) == 0) {
          __nanosleep(20);
          v = *p;
        }
        excl += v & LB_VALUE_MASK;
        if (v & LB_FLAG_INCL) break;
        --t;
      }
      volatile uint32_t* lb = a.lookback + (size_t)tile * RADIX + tid;
      *lb = LB_FLAG_INCL | (excl + my_count);
    }
    obase[tid] = a.gbase[tid] + excl - my_start;
  }
  __syncthreads();

  // ---- coalesced write-out: consecutive threads -> consecutive addresses per digit run
  if (!is_last) {
    const uint32_t dsel = is_first ? 0u : (plan->src_sel[a.pass] ^ 1u);
    uint64_t* kout = dsel ? a.kbuf[1] : a.kbuf[0];
    uint32_t* vout = dsel ? a.vbuf[1] : a.vbuf[0];
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t i = tid + j * THREADS;
      if (i < n_valid) {
        const uint64_t k = skeys[i];
        const uint32_t d = (uint32_t)(k >> shift) & (RADIX - 1);
        const uint32_t dst = obase[d] + i;
        kout[dst] = k;
        vout[dst] = svals[i];
        if (a.mark_runs != nullptr && a.pass == 1) {
          // The input of pass B is sorted by digit A, so inside a tile the pairs of one digit B are in
          // (B, A) order: a pair whose predecessor in the tile lies in another (B, A) bin is the first of
          // its bin in this tile. The run may have been begun by an earlier tile: keep the minimum.
          const uint32_t mshift = plan->shift[0];
          const uint32_t bin = (uint32_t)(k >> mshift) & (JOINT_BINS - 1);
          const uint32_t pbin = i ? (uint32_t)(skeys[i - 1] >> mshift) & (JOINT_BINS - 1) : ~0u;
          if (pbin != bin) atomicMin(&a.mark_runs[bin], dst);
        }
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      const uint32_t i = tid + j * THREADS;
      if (i < n_valid) {
        const uint64_t k = skeys[i];
        const uint32_t d = (uint32_t)(k >> shift) & (RADIX - 1);
        const uint32_t dst = obase[d] + i;
        a.perm_out[dst] = svals[i];
        if (a.codes_out) a.codes_out[dst] = k;
        if (a.keys_out) store_decoded(a.keys_out, dst, k, a.kt, a.desc != 0);
      }
    }
  }
  if (WALK) __syncthreads();  // shared memory is rewritten by the next tile
  } while (WALK);
}

// ---------------------------------------------------------------------------
// Kernel 4: nothing to sort (every digit constant): identity permutation
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
no_pass_kernel(const void* __restrict__ raw, const uint32_t* __restrict__ perm_in, int kt, int desc, uint32_t n,
               const SortPlan* __restrict__ plan, const SortCtl* __restrict__ ctl, uint64_t* __restrict__ codes_out,
               void* __restrict__ keys_out, uint32_t* __restrict__ perm_out) {
  if (plan->first >= 0 || ctl->mode == MODE_HYBRID) return;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t src = perm_in ? perm_in[i] : (uint32_t)i;
    perm_out[i] = src;
    const uint64_t c = load_encode(raw, src, kt, desc != 0);
    if (codes_out) codes_out[i] = c;
    if (keys_out) store_decoded(keys_out, i, c, kt, desc != 0);
  }
}

// ---------------------------------------------------------------------------
// HYBRID mode kernels
// ---------------------------------------------------------------------------
// Tunable from the host (b200x_set_tuning): fewest rows for which HYBRID is considered.
static unsigned long long g_hybrid_min_rows = 1ull << 22;

// H1: smallest and largest code. One read of the key column.
__global__ void __launch_bounds__(512)
analyze_kernel(const void* __restrict__ raw, const uint32_t* __restrict__ perm_in, int kt, int desc, uint32_t n,
               SortCtl* __restrict__ ctl) {
  unsigned long long o = 0, na = 0;  // running max of code / of ~code
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int width = key_type_bytes(kt);
  constexpr int U = 8;  // independent loads in flight per thread
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < n; i += U * stride) {
    uint64_t c[U];
    if (perm_in) {
#pragma unroll
      for (int u = 0; u < U; ++u) c[u] = perm_in[i + u * stride];
#pragma unroll
      for (int u = 0; u < U; ++u) c[u] = load_raw_bits(raw, (size_t)c[u], width);
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) c[u] = load_raw_bits(raw, (size_t)(i + u * stride), width);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t e = encode_bits(c[u], kt, desc != 0);
      o = max(o, (unsigned long long)e);
      na = max(na, (unsigned long long)~e);
    }
  }
  for (; i < n; i += stride) {
    const uint64_t c = load_encode(raw, perm_in ? (size_t)perm_in[i] : (size_t)i, kt, desc != 0);
    o = max(o, (unsigned long long)c);
    na = max(na, (unsigned long long)~c);
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    o = max(o, __shfl_xor_sync(0xffffffffu, o, d));
    na = max(na, __shfl_xor_sync(0xffffffffu, na, d));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&ctl->max_code, o);
    atomicMax(&ctl->max_inv, na);
  }
}

// H2: choose the two global digits = the top 16 bits of (code - smallest code).
__global__ void hybrid_plan0_kernel(SortCtl* __restrict__ ctl, uint32_t n, unsigned long long min_rows) {
  const unsigned long long lo = ~ctl->max_inv;
  const unsigned long long vary = ctl->max_code - lo;  // width of the occupied code range
  ctl->mode = MODE_LSD;
  ctl->candidate = 0;
  ctl->base = lo;
  if (vary == 0) return;
  const int top = 64 - __clzll((long long)vary);  // bits needed for code - base
  if (top > 16 && (unsigned long long)n >= min_rows) {
    ctl->candidate = 1;
    ctl->shiftB = (uint32_t)(top - 8);
    ctl->shiftA = (uint32_t)(top - 16);
  }
}

constexpr uint32_t SAMPLE_EVERY = 64;
constexpr uint32_t FIN_IPT = 8;
// finisher block sizes: runs <= 2048 / <= 4096 / <= 8192 pairs
constexpr uint32_t FIN_SMALL_THREADS = 256, FIN_BIG_THREADS = 512, FIN_HUGE_THREADS = 1024;
constexpr uint32_t FIN_MAX = FIN_HUGE_THREADS * FIN_IPT;
// a sampled count of FIN_MAX / SAMPLE_EVERY is what the longest admissible run shows on average; allowing
// 1.25 x that (2.8 sigma) keeps data that fits from being turned away
constexpr uint32_t SAMPLED_LIMIT = (FIN_MAX + FIN_MAX / 4) / SAMPLE_EVERY;
// U is kept while its longest sampled run stays within what the 512-thread finisher holds
constexpr uint32_t SAMPLED_LIMIT_U = (FIN_BIG_THREADS * FIN_IPT + FIN_BIG_THREADS * FIN_IPT / 2) / SAMPLE_EVERY;

// H2b: every 64th key counts its bin - a cheap early warning for runs far beyond what the finisher
// holds. First call (which = 0): U bins; the counts also serve as the coarse histogram the T table is
// built from if U fails. Second call (which = 1, only if T was chosen): T bins.
__global__ void __launch_bounds__(256)
sample_bins_kernel(const void* __restrict__ raw, const uint32_t* __restrict__ perm_in, int kt, int desc, uint32_t n,
                   const SortCtl* __restrict__ ctl, int which, uint32_t* __restrict__ sampled /*[JOINT_BINS]*/,
                   const uint32_t* __restrict__ tab) {
  if (!ctl->candidate) return;
  if (which == 1 && !ctl->tmode) return;
  const uint32_t shiftA = ctl->shiftA, shiftC = ctl->shiftC;
  const unsigned long long base = ctl->base;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t ns = n / SAMPLE_EVERY;
  for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < ns; s += stride) {
    const uint64_t i = s * SAMPLE_EVERY;
    uint64_t k = load_encode(raw, perm_in ? (size_t)perm_in[i] : (size_t)i, kt, desc != 0) - base;
    if (which == 1) k = remap_key(k, tab, shiftC);
    atomicAdd(&sampled[(uint32_t)(k >> shiftA) & (JOINT_BINS - 1)], 1u);
  }
}

// H2c (one block of 1024, 64 consecutive buckets per thread): keep U if its sample looks fine;
// otherwise build the T table from the sampled U bins. A bucket with more than `target` rows gets
// 2^s bins, s = ceil(log2(rows / target)); the others are merged, in order, into bins of about
// target / 2 rows (a new bin starts after a split bucket and whenever the running row count of
// sparse buckets crosses a multiple of target / 2). target = the smallest value, searched in steps
// of 2^(1/4), for which all bins together number <= 65536.
__global__ void __launch_bounds__(1024)
hybrid_planT_kernel(SortCtl* __restrict__ ctl, const uint32_t* __restrict__ sampled, uint32_t* __restrict__ tab,
                    uint32_t n) {
  if (!ctl->candidate) return;
  __shared__ uint32_t s_red[32];
  __shared__ uint32_t s_val;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  constexpr int PER = JOINT_BINS / 1024;
  constexpr uint32_t DENSE = 0x80000000u;
  const uint32_t* mine = sampled + tid * PER;
  auto block_sum = [&](uint32_t v) -> uint32_t {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_red[w] = v;
    __syncthreads();
    uint32_t t = 0;
    for (int k = 0; k < 32; ++k) t += s_red[k];
    return t;
  };
  auto block_excl_scan = [&](uint32_t v) -> uint32_t {
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (uint32_t)o) x += y;
    }
    __syncthreads();
    if (lane == 31) s_red[w] = x;
    __syncthreads();
    uint32_t off = x - v;
    for (uint32_t k = 0; k < w; ++k) off += s_red[k];
    return off;
  };
  {
    uint32_t m = 0;
#pragma unroll 8
    for (uint32_t i = tid; i < JOINT_BINS; i += 1024) m = max(m, sampled[i]);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_red[w] = m;
    __syncthreads();
    if (tid == 0) {
      uint32_t t = 0;
      for (int k = 0; k < 32; ++k) t = max(t, s_red[k]);
      s_val = t;
    }
    __syncthreads();
    if (s_val <= SAMPLED_LIMIT_U) return;  // U (tmode stays 0)
  }
  const uint32_t shiftC = ctl->shiftA;  // the U bins are the coarse buckets
  const uint32_t smax = min(16u, shiftC);
  // Bucket sizes in quarter-octaves above T0 = n / 65536: Q = ceil(4 log2(rows / T0)), 0 if rows <= T0.
  // With target = T0 * 2^(t/4) a bucket is split iff Q > t, into 2^s bins, s = ceil((Q - t) / 4).
  const float T0 = fmaxf((float)n / (float)JOINT_BINS, 64.0f);
  auto q_of = [&](uint32_t samples) -> int {
    const float r = (float)samples * (float)SAMPLE_EVERY;
    if (!(r > T0)) return 0;
    return min((int)ceilf(4.0f * log2f(r / T0)), 127);
  };
  auto bits_for = [&](int Q, int t) -> uint32_t { return min((uint32_t)((Q - t + 3) / 4), smax); };  // Q > t
  __shared__ uint32_t s_cnt[128], s_rows[128];
  __shared__ int s_t;
  if (tid < 128) s_cnt[tid] = s_rows[tid] = 0;
  if (tid == 0) s_t = 1 << 30;
  __syncthreads();
  for (int q = 0; q < PER; ++q) {
    const uint32_t c = mine[q];
    if (c) {
      const int Q = q_of(c);
      atomicAdd(&s_cnt[Q], 1u);
      atomicAdd(&s_rows[Q], c);
    }
  }
  __syncthreads();
  constexpr int T_STEPS = 48;
  if (tid < T_STEPS) {
    const int t = (int)tid;
    const float target_t = T0 * exp2f(0.25f * (float)t);
    float bins = 2.0f, sparse = 0.0f;
    for (int Q = 0; Q < 128; ++Q) {
      if (Q > t) bins += (float)s_cnt[Q] * (float)((1u << bits_for(Q, t)) + 1u);  // + 1: a split bucket ends a merged run
      else sparse += (float)s_rows[Q];
    }
    bins += sparse * (float)SAMPLE_EVERY / (0.5f * target_t);
    // a merged bin holds < 1.5 x target rows, a split one about target (the density varies little inside a bucket)
    if (bins <= (float)JOINT_BINS && target_t * 1.75f <= (float)FIN_MAX) atomicMin(&s_t, t);
  }
  __syncthreads();
  const int t_sel = s_t;
  if (t_sel >= T_STEPS) {
    if (tid == 0) ctl->candidate = 0;
    return;
  }
  const float target = T0 * exp2f(0.25f * (float)t_sel);
  const float quantum = 0.5f * target;
  // interim entries: DENSE | s for split buckets, the quantum index of the running sparse row count otherwise
  uint32_t my_sparse = 0;
  for (int q = 0; q < PER; ++q) {
    if (!(q_of(mine[q]) > t_sel)) my_sparse += mine[q];
  }
  uint32_t running = block_excl_scan(my_sparse);
  for (int q = 0; q < PER; ++q) {
    const int Q = q_of(mine[q]);
    if (Q > t_sel) {
      tab[tid * PER + q] = DENSE | bits_for(Q, t_sel);
    } else {
      tab[tid * PER + q] = (uint32_t)((float)running * (float)SAMPLE_EVERY / quantum);
      running += mine[q];
    }
  }
  __syncthreads();
  const uint32_t prev0 = tid == 0 ? DENSE : tab[tid * PER - 1];
  uint32_t cnt = 0, prev = prev0;
  for (int q = 0; q < PER; ++q) {
    const uint32_t e = tab[tid * PER + q];
    if (e & DENSE) cnt += 1u << (e & 31u);
    else cnt += ((prev & DENSE) || prev != e) ? 1u : 0u;
    prev = e;
  }
  uint32_t run = block_excl_scan(cnt);  // (its barriers also order the reads above before the writes below)
  const uint32_t total = block_sum(cnt);
  if (total > JOINT_BINS) {
    if (tid == 0) ctl->candidate = 0;
    return;
  }
  prev = prev0;
  for (int q = 0; q < PER; ++q) {
    const uint32_t e = tab[tid * PER + q];
    if (e & DENSE) {
      tab[tid * PER + q] = (run << 5) | (e & 31u);
      run += 1u << (e & 31u);
    } else {
      if ((prev & DENSE) || prev != e) ++run;
      tab[tid * PER + q] = ((run - 1u) << 5) | TAB_MERGED;
    }
    prev = e;
  }
  if (tid == 0) {
    ctl->tmode = 1;
    ctl->shiftC = shiftC;
    ctl->shiftA = (uint32_t)KEY_FIELD;
    ctl->shiftB = (uint32_t)KEY_FIELD + RADIX_BITS;
  }
}

// H2d (one block of 1024): T only - the sampled T bins must look admissible before a full pass is spent on them.
__global__ void __launch_bounds__(1024)
hybrid_checkT_kernel(SortCtl* __restrict__ ctl, const uint32_t* __restrict__ sampledT) {
  if (!ctl->candidate || !ctl->tmode) return;
  __shared__ uint32_t s_m;
  if (threadIdx.x == 0) s_m = 0;
  __syncthreads();
  uint32_t m = 0;
  for (uint32_t i = threadIdx.x; i < JOINT_BINS; i += 1024) m = max(m, sampledT[i]);
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(&s_m, m);
  __syncthreads();
  if (threadIdx.x == 0 && s_m > SAMPLED_LIMIT) ctl->candidate = 0;
}

// H3: histograms of the two global digits (A = bits [shiftA, shiftA+8), B = the 8 bits above) of
// code - base. One read of the key column, two shared-memory atomics per key at full occupancy.
// (A 65536-bin joint histogram - which would also give the run boundaries up front - cost 0.7 ms
// per 100 M keys with global atomics and 1.6 ms with a 128 KB shared-memory table; the boundaries
// are instead marked by pass B as it writes, see mark_runs.)
// TMODE: the keys go through the bucket table first, and the rewritten keys and their row numbers
// are stored (kout, vout) for pass A to read instead of the column. Both variants are launched; the
// one that does not apply returns at once.
template <bool TMODE>
__global__ void __launch_bounds__(512, 2)
digit2_hist_kernel(const void* __restrict__ raw, const uint32_t* __restrict__ perm_in, int kt, int desc, uint32_t n,
                   const SortCtl* __restrict__ ctl, uint32_t* __restrict__ hist2 /*[2][RADIX]*/,
                   const uint32_t* __restrict__ tab, uint64_t* __restrict__ kout, uint32_t* __restrict__ vout) {
  if (!ctl->candidate || (ctl->tmode != 0) != TMODE) return;
  __shared__ uint32_t sh[2 * RADIX];
  for (int i = threadIdx.x; i < 2 * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t shiftA = ctl->shiftA, shiftC = ctl->shiftC;
  const unsigned long long base = ctl->base;
  const int width = key_type_bytes(kt);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  constexpr int U = 8;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < n; i += U * stride) {
    uint64_t c[U];
    uint32_t src[U];
    if (perm_in) {
#pragma unroll
      for (int u = 0; u < U; ++u) src[u] = perm_in[i + u * stride];
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) src[u] = (uint32_t)(i + u * stride);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) c[u] = load_raw_bits(raw, (size_t)src[u], width);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      uint64_t k = encode_bits(c[u], kt, desc != 0) - base;
      if (TMODE) {
        k = remap_key(k, tab, shiftC);
        kout[i + u * stride] = k;
        vout[i + u * stride] = src[u];
      }
      const uint32_t bin = (uint32_t)(k >> shiftA) & (JOINT_BINS - 1);
      atomicAdd(&sh[bin & 255u], 1u);
      atomicAdd(&sh[RADIX + (bin >> 8)], 1u);
    }
  }
  for (; i < n; i += stride) {
    const uint32_t src = perm_in ? perm_in[i] : (uint32_t)i;
    uint64_t k = load_encode(raw, (size_t)src, kt, desc != 0) - base;
    if (TMODE) {
      k = remap_key(k, tab, shiftC);
      kout[i] = k;
      vout[i] = src;
    }
    const uint32_t bin = (uint32_t)(k >> shiftA) & (JOINT_BINS - 1);
    atomicAdd(&sh[bin & 255u], 1u);
    atomicAdd(&sh[RADIX + (bin >> 8)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * RADIX; i += blockDim.x) {
    const uint32_t v = sh[i];
    if (v) atomicAdd(&hist2[i], v);
  }
}

// H4 (one block of RADIX threads): exclusive scans of the two digit histograms; arm the two passes
// unless a digit bin is so full that some run must exceed what the finisher holds.
__global__ void __launch_bounds__(RADIX)
hybrid_plan1_kernel(SortCtl* __restrict__ ctl, const uint32_t* __restrict__ hist2, const uint32_t* __restrict__ sampledU,
                    const uint32_t* __restrict__ sampledT, uint32_t* __restrict__ hbase /*[2][RADIX]*/,
                    SortPlan* __restrict__ hplan, uint32_t n) {
  __shared__ uint32_t warp_tot[RADIX / 32];
  __shared__ uint32_t s_max, s_smax;
  const int d = threadIdx.x;
  const unsigned lane = d & 31, w = d >> 5;
  if (d == 0) {
    s_smax = 0;
    for (int p = 0; p < MAX_PASSES; ++p) {
      hplan->active[p] = 0;
      hplan->src_sel[p] = 0;
      hplan->tile_counter[p] = 0;
      hplan->shift[p] = 0;
    }
    hplan->first = -1;
    hplan->last = -1;
    s_max = 0;
  }
  if (!ctl->candidate) return;
  __syncthreads();
  {
    const uint32_t* sampled = ctl->tmode ? sampledT : sampledU;
    uint32_t m = 0;
    for (int i = d; i < (int)JOINT_BINS; i += RADIX) m = max(m, sampled[i]);
    atomicMax(&s_smax, m);
  }
  for (int p = 0; p < 2; ++p) {
    const uint32_t c = hist2[p * RADIX + d];
    atomicMax(&s_max, c);
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    uint32_t off = 0;
    for (unsigned k = 0; k < w; ++k) off += warp_tot[k];
    hbase[p * RADIX + d] = off + x - c;
    __syncthreads();
  }
  if (d == 0 && (unsigned long long)s_max <= (unsigned long long)RADIX * FIN_MAX && s_smax <= SAMPLED_LIMIT) {
    hplan->active[0] = hplan->active[1] = 1;
    hplan->shift[0] = ctl->shiftA;
    hplan->shift[1] = ctl->shiftB;
    hplan->first = ctl->tmode ? -1 : 0;  // pass A reads the raw column (T: the rewritten keys in buffer 1), writes buffer 0
    hplan->src_sel[0] = 1;
    hplan->last = -1;   // neither pass writes the outputs: the finisher does
    hplan->src_sel[1] = 0;  // pass B reads buffer 0, writes buffer 1
  }
}

// H4b (one block of 1024, after pass B): pass B left in run_start[bin] the position of the first pair of
// every non-empty (B, A) bin and 0xFFFFFFFF for empty bins. Fill the empty ones (an empty run starts
// where the next run starts), find the longest run and decide the mode.
__global__ void __launch_bounds__(1024)
runs_fill_kernel(SortCtl* __restrict__ ctl, const SortPlan* __restrict__ hplan, uint32_t* __restrict__ run_start,
                 uint32_t n) {
  if (!hplan->active[1]) return;
  __shared__ uint32_t s_wmin[32];
  __shared__ uint32_t s_max;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_max = 0;
  // warp w owns bins [w * 2048, (w + 1) * 2048), 32 consecutive bins per step (coalesced)
  constexpr uint32_t PER_WARP = JOINT_BINS / 32, STEPS = PER_WARP / 32;
  uint32_t* mine = run_start + w * PER_WARP + lane;
  uint32_t mn = 0xFFFFFFFFu;
#pragma unroll 8
  for (uint32_t it = 0; it < STEPS; ++it) mn = min(mn, mine[it * 32]);
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  if (lane == 0) s_wmin[w] = mn;
  __syncthreads();
  uint32_t carry = n;  // start of the first non-empty run after the bins still to be visited
  for (uint32_t k = w + 1; k < 32; ++k) carry = min(carry, s_wmin[k]);
  uint32_t mx = 0;
  for (int it = (int)STEPS - 1; it >= 0; --it) {
    uint32_t x = mine[it * 32];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_down_sync(0xffffffffu, x, o);
      if (lane + o < 32) x = min(x, y);
    }
    x = min(x, carry);  // an empty bin starts where the next run starts
    uint32_t nxt = __shfl_down_sync(0xffffffffu, x, 1);
    if (lane == 31) nxt = carry;
    mx = max(mx, nxt - x);
    mine[it * 32] = x;
    carry = __shfl_sync(0xffffffffu, x, 0);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) atomicMax(&s_max, mx);
  if (tid == 0) run_start[JOINT_BINS] = n;
  __syncthreads();
  if (tid == 0) {
    ctl->max_run = s_max;
    ctl->mode = s_max <= FIN_MAX ? MODE_HYBRID : MODE_LSD;
  }
}

// H5: finish one run (all pairs that agree on the top 16 varying bits) in shared memory.
struct FinArgs {
  const uint64_t* kin;   // pairs after the two global passes
  const uint32_t* vin;
  uint32_t* perm_out;
  uint64_t* codes_out;
  void* keys_out;
  const uint32_t* run_start;  // [65537]
  const SortCtl* ctl;
  const uint32_t* tab;        // T: bucket table
  const void* raw;            // T: the key column (re-read for buckets whose low bits did not fit the key field)
  int kt;
  int desc;
};

#ifndef B200X_FIN_TPS
#define B200X_FIN_TPS 1024  // resident finisher threads per SM the register budget is set for
#endif
constexpr uint32_t TIE_MAX = 48;  // longest group of equal windows the in-place tie repair handles

// window of the one-round counting sort: about one bin per pair (2048 bins for runs <= 2048 pairs, ...)
template <uint32_t THREADS>
struct FinWindow {
  static constexpr int BITS = THREADS == 256 ? 11 : THREADS == 512 ? 12 : 13;
  static constexpr uint32_t BINS = 1u << BITS;
};

template <uint32_t THREADS>
struct FinSmem {
  uint64_t keys[THREADS * FIN_IPT];
  uint32_t vals[THREADS * FIN_IPT];
  uint32_t hist[FinWindow<THREADS>::BINS + 1];  // counting sort: bin counts, then bin starts
  uint16_t ord[THREADS * FIN_IPT];              // counting sort: position of the pair in the run as loaded
  uint16_t whist[(THREADS / 32) * RADIX];
  uint32_t swarp[32];
  uint32_t slow;
};

// One stable counting-sort round on up to 8 bits at `shift` over the block's m pairs.
// Pairs live in registers in warp-striped order (item j of lane l of warp w = w*32*nit + j*32 + l,
// so register order is memory order) and come back in the same arrangement, now sorted.
// Slots beyond m hold the all-ones code and stay at the end.
template <uint32_t THREADS>
__device__ __forceinline__ void fin_round(uint64_t (&key)[FIN_IPT], uint32_t (&val)[FIN_IPT], uint32_t nit, int shift,
                                          uint32_t mask, FinSmem<THREADS>& sm, bool reload) {
  constexpr uint32_t WARPS = THREADS / 32;
  static_assert(THREADS >= RADIX, "one thread per digit");
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (uint32_t i = tid; i < WARPS * RADIX / 2; i += THREADS) ((uint32_t*)sm.whist)[i] = 0;
  __syncthreads();
  uint16_t* wh = sm.whist + warp * RADIX;
  uint32_t rank[FIN_IPT];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (uint32_t j = 0; j < FIN_IPT; ++j) {
    if (j < nit) {
      const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
      const unsigned peers = match_digit(d);
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if ((int)lane == leader) {
        old = wh[d];
        wh[d] = (uint16_t)(old + __popc(peers));
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      rank[j] = old + __popc(peers & lt);
      __syncwarp();
    }
  }
  __syncthreads();
  uint32_t my_count = 0, my_start = 0;
  if (tid < RADIX) {
    uint32_t sum = 0;
#pragma unroll
    for (uint32_t w = 0; w < WARPS; ++w) {
      const uint32_t c = sm.whist[w * RADIX + tid];
      sm.whist[w * RADIX + tid] = (uint16_t)sum;
      sum += c;
    }
    my_count = sum;
    uint32_t x = my_count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) sm.swarp[warp] = x;
    my_start = x - my_count;
  }
  __syncthreads();
  if (tid < RADIX) {
    uint32_t off = 0;
    for (uint32_t k = 0; k < warp; ++k) off += sm.swarp[k];
    my_start += off;
#pragma unroll
    for (uint32_t w = 0; w < WARPS; ++w) sm.whist[w * RADIX + tid] += (uint16_t)my_start;
  }
  __syncthreads();
#pragma unroll
  for (uint32_t j = 0; j < FIN_IPT; ++j) {
    if (j < nit) {
      const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
      const uint32_t pos = (uint32_t)wh[d] + rank[j];
      sm.keys[pos] = key[j];
      sm.vals[pos] = val[j];
    }
  }
  __syncthreads();
  if (reload) {
    const uint32_t wbase = warp * 32 * nit + lane;
#pragma unroll
    for (uint32_t j = 0; j < FIN_IPT; ++j) {
      if (j < nit) {
        key[j] = sm.keys[wbase + j * 32];
        val[j] = sm.vals[wbase + j * 32];
      }
    }
  }
}

__device__ __forceinline__ void fin_load(uint64_t (&key)[FIN_IPT], uint32_t (&val)[FIN_IPT], uint32_t nit, uint32_t m,
                                         const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t wbase = warp * 32 * nit + lane;
#pragma unroll
  for (uint32_t j = 0; j < FIN_IPT; ++j) {
    key[j] = ~0ull;
    val[j] = 0;
    const uint32_t t = wbase + j * 32;
    if (j < nit && t < m) {
      key[j] = ld_stream_u64(kin + t);
      val[j] = ld_stream_u32(vin + t);
    }
  }
}

// LSD over bits [lo, hi) in rounds of up to 8 bits; the result is left in shared memory.
template <uint32_t THREADS>
__device__ __forceinline__ void fin_lsd(uint64_t (&key)[FIN_IPT], uint32_t (&val)[FIN_IPT], uint32_t nit, int lo,
                                        int hi, FinSmem<THREADS>& sm) {
  for (int b = lo; b < hi; b += RADIX_BITS) {
    const int nb = min(RADIX_BITS, hi - b);
    fin_round<THREADS>(key, val, nit, b, (1u << nb) - 1u, sm, /*reload=*/b + RADIX_BITS < hi);
  }
}

// T: what a bin holds. j1 = the last bucket whose first bin is <= run. A split bucket: the run's keys agree
// above R = shiftC - s bits. A merged bin (buckets j0..j1): its keys are re-read from the column and taken
// relative to bucket j0.
__device__ __noinline__ void fin_locate_run(const uint32_t* __restrict__ tab, int shiftC, uint32_t run,
                                            uint32_t* s_bucket, int& R, unsigned long long& base, bool& reread) {
  if (threadIdx.x == 0) {
    uint32_t lo = 0, hi = JOINT_BINS - 1;
    while (lo < hi) {
      const uint32_t mid = (lo + hi + 1) >> 1;
      if ((tab[mid] >> 5) <= run) lo = mid;
      else hi = mid - 1;
    }
    const uint32_t j1 = lo, e = tab[j1];
    uint32_t j0 = j1;
    if ((e & 31u) == TAB_MERGED) {
      lo = 0, hi = j1;
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if ((tab[mid] >> 5) >= run) hi = mid;
        else lo = mid + 1;
      }
      j0 = lo;
    }
    s_bucket[0] = j0;
    s_bucket[1] = j1;
    s_bucket[2] = e;
  }
  __syncthreads();
  const uint32_t j0 = s_bucket[0], j1 = s_bucket[1], e = s_bucket[2];
  if ((e & 31u) == TAB_MERGED) {
    reread = true;
    R = shiftC + (j1 > j0 ? 32 - __clz((int)(j1 - j0)) : 0);
    base += (unsigned long long)j0 << shiftC;
  } else {
    R = shiftC - (int)(e & 31u);
    base += ((unsigned long long)j1 << shiftC) | ((unsigned long long)(run - (e >> 5)) << R);
  }
}

// Finish one run with a block of THREADS threads (m <= THREADS * FIN_IPT pairs starting at `start`).
template <uint32_t THREADS>
__device__ __forceinline__ void finish_one_run(const FinArgs& a, const SortCtl* ctl, FinSmem<THREADS>& sm,
                                               uint32_t* s_bucket, uint32_t run, uint32_t start, uint32_t m) {
  {
    const uint32_t tid = threadIdx.x;
    const bool tmode = ctl->tmode != 0;
    const int shiftC = (int)ctl->shiftC;
    int R = (int)ctl->shiftA;  // low bits still unsorted inside the run
    unsigned long long base = ctl->base;
    bool reread = false;
    if (tmode) fin_locate_run(a.tab, shiftC, run, s_bucket, R, base, reread);
    const uint32_t nit = (m + THREADS - 1) / THREADS;
    const uint64_t* kin = a.kin + start;
    const uint32_t* vin = a.vin + start;
    uint64_t key[FIN_IPT];
    uint32_t val[FIN_IPT];
    // the pairs of the run; T: keys reduced to the R bits that vary inside the run
    auto load_run = [&]() {
      fin_load(key, val, nit, m, kin, vin);
      if (tmode) {
        const uint32_t wb = (tid >> 5) * 32 * nit + (tid & 31);
#pragma unroll
        for (uint32_t j = 0; j < FIN_IPT; ++j) {
          if (j < nit && wb + j * 32 < m) {
            if (reread) key[j] = load_encode(a.raw, (size_t)val[j], a.kt, a.desc != 0) - base;
            else key[j] = (key[j] & ((1ull << KEY_FIELD) - 1ull)) >> (KEY_FIELD - R);
          }
        }
      }
    };
    load_run();
    if (tid == 0) sm.slow = 0;

    // One counting-sort round on a window of the top bits that remain, with about one bin per pair:
    // every pair takes a ticket in its bin (a shared-memory atomic, so the order inside a bin is
    // arbitrary), the bins are scanned, the pairs are scattered. Pairs that share a bin are rare for
    // spread-out keys and form tiny groups; each pair then finds its place inside its group by
    // comparing (bits below the window, position in the run as loaded) with the others - which also
    // restores stability. (Two ballot-ranked 8-bit rounds on a 16-bit window, the earlier scheme, cost
    // 311 instructions per pair; one 8-bit round left groups of ~m/256 pairs and was slower still.)
    // A group longer than TIE_MAX (duplicate-heavy keys) sends the run down the plain path: stable LSD
    // over every remaining bit.
    // (At most 8 bits left - narrow keys, many duplicates per value: one stable ballot-ranked round
    // sorts the run outright; ticketed bins would hold long groups of equal keys.)
    bool tie_repair = R > RADIX_BITS;
    uint32_t fpos[FIN_IPT];
    if (!tie_repair) {
      fin_lsd<THREADS>(key, val, nit, 0, max(R, 1), sm);  // (R = 0: equal keys, the round just stages them)
    } else {
      constexpr int WB = FinWindow<THREADS>::BITS;
      constexpr uint32_t BINS = FinWindow<THREADS>::BINS;
      const int wbits = min(R, WB);
      const int wshift = R - wbits;
      const uint32_t wmask = (1u << wbits) - 1u;
      const uint64_t lowmask = (1ull << wshift) - 1ull;
      for (uint32_t i = tid; i <= BINS; i += THREADS) sm.hist[i] = 0;
      __syncthreads();
      uint32_t ticket[FIN_IPT];
      {
        const uint32_t wb = (tid >> 5) * 32 * nit + (tid & 31);
#pragma unroll
        for (uint32_t j = 0; j < FIN_IPT; ++j) {
          if (j < nit && wb + j * 32 < m) ticket[j] = atomicAdd(&sm.hist[(uint32_t)(key[j] >> wshift) & wmask], 1u);
        }
      }
      __syncthreads();
      {
        // exclusive scan of the bin counts: BINS / THREADS consecutive bins per thread
        constexpr uint32_t PERT = BINS / THREADS;
        uint32_t c[PERT], tot = 0;
#pragma unroll
        for (uint32_t q = 0; q < PERT; ++q) {
          c[q] = sm.hist[tid * PERT + q];
          tot += c[q];
        }
        uint32_t x = tot;
        const uint32_t lane = tid & 31, w = tid >> 5;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
          if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31) sm.swarp[w] = x;
        __syncthreads();
        uint32_t off = x - tot;
        for (uint32_t k = 0; k < w; ++k) off += sm.swarp[k];
#pragma unroll
        for (uint32_t q = 0; q < PERT; ++q) {
          sm.hist[tid * PERT + q] = off;
          off += c[q];
        }
        if (tid == THREADS - 1) sm.hist[BINS] = off;  // = m
      }
      __syncthreads();
      {
        const uint32_t wb = (tid >> 5) * 32 * nit + (tid & 31);
#pragma unroll
        for (uint32_t j = 0; j < FIN_IPT; ++j) {
          if (j < nit && wb + j * 32 < m) {
            const uint32_t pos = sm.hist[(uint32_t)(key[j] >> wshift) & wmask] + ticket[j];
            sm.keys[pos] = key[j];
            sm.vals[pos] = val[j];
            sm.ord[pos] = (uint16_t)(wb + j * 32);
          }
        }
      }
      __syncthreads();
#pragma unroll
      for (uint32_t j = 0; j < FIN_IPT; ++j) {
        const uint32_t i = tid + j * THREADS;
        fpos[j] = i;
        if (i < m) {
          const uint64_t k = sm.keys[i];
          const uint32_t d = (uint32_t)(k >> wshift) & wmask;
          const uint32_t gs = sm.hist[d], ge = sm.hist[d + 1];
          if (ge - gs > 1) {
            if (ge - gs > TIE_MAX) {
              sm.slow = 1;
            } else {
              const uint64_t low = k & lowmask;
              const uint32_t me = sm.ord[i];
              uint32_t ahead = 0;  // pairs of my bin that sort before me: smaller low bits, or equal and loaded earlier
              for (uint32_t e = gs; e < ge; ++e) {
                const uint64_t ol = sm.keys[e] & lowmask;
                ahead += (ol < low || (ol == low && sm.ord[e] < me)) ? 1u : 0u;
              }
              fpos[j] = gs + ahead;
            }
          }
        }
      }
    }
    __syncthreads();
    if (tie_repair && sm.slow) {
      load_run();
      fin_lsd<THREADS>(key, val, nit, 0, R, sm);
      tie_repair = false;
    }
#pragma unroll
    for (uint32_t j = 0; j < FIN_IPT; ++j) {
      const uint32_t i = tid + j * THREADS;
      if (i < m) {
        const uint32_t dst = start + (tie_repair ? fpos[j] : i);
        const uint64_t k = sm.keys[i] + base;
        a.perm_out[dst] = sm.vals[i];
        if (a.codes_out) a.codes_out[dst] = k;
        if (a.keys_out) store_decoded(a.keys_out, dst, k, a.kt, a.desc != 0);
      }
    }
  }
}

// Block sizes by run length: 256 threads for runs <= 2048 pairs, 512 up to 4096, 1024 up to 8192.
// The first two launch one block per run (a block whose run belongs to another size class returns
// at once); the 1024-thread kernel, rarely needed, launches only as many blocks as are resident and
// walks over the runs, so that launching it for nothing costs nothing. (Walking was tried for the
// 512-thread kernel too: its idle launch costs 60 us, but walking - strided or with a run counter -
// made a table-binned sort of 100 M doubles, where half of the rows are in such runs, 0.13-0.25 ms slower.)
__device__ __forceinline__ uint32_t fin_class(uint32_t m) {
  return m <= FIN_SMALL_THREADS * FIN_IPT ? FIN_SMALL_THREADS : m <= FIN_BIG_THREADS * FIN_IPT ? FIN_BIG_THREADS
                                                                                                : FIN_HUGE_THREADS;
}

template <uint32_t THREADS>
__global__ void __launch_bounds__(THREADS, B200X_FIN_TPS / THREADS)
finish_runs_kernel(const FinArgs a) {
  const SortCtl* ctl = a.ctl;
  if (ctl->mode != MODE_HYBRID) return;
  if (THREADS != FIN_SMALL_THREADS && ctl->max_run <= FIN_SMALL_THREADS * FIN_IPT) return;
  const uint32_t start = a.run_start[blockIdx.x];
  const uint32_t m = a.run_start[blockIdx.x + 1] - start;
  if (m == 0 || fin_class(m) != THREADS) return;
  extern __shared__ __align__(16) unsigned char fin_raw[];
  __shared__ uint32_t s_bucket[3];
  finish_one_run<THREADS>(a, ctl, *reinterpret_cast<FinSmem<THREADS>*>(fin_raw), s_bucket, blockIdx.x, start, m);
}

template <uint32_t THREADS>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
finish_runs_walk_kernel(const FinArgs a) {
  const SortCtl* ctl = a.ctl;
  if (ctl->mode != MODE_HYBRID || ctl->max_run <= (THREADS / 2) * FIN_IPT) return;  // no run of this size class
  extern __shared__ __align__(16) unsigned char fin_raw[];
  __shared__ uint32_t s_bucket[3];
  for (uint32_t run = blockIdx.x; run < JOINT_BINS; run += gridDim.x) {
    const uint32_t start = a.run_start[run];
    const uint32_t m = a.run_start[run + 1] - start;
    if (m == 0 || fin_class(m) != THREADS) continue;
    __syncthreads();  // the previous run's shared memory is no longer read
    finish_one_run<THREADS>(a, ctl, *reinterpret_cast<FinSmem<THREADS>*>(fin_raw), s_bucket, run, start, m);
  }
}

// Tile shape (measured on B200, 100M int64 keys, 8 passes): 384x12x3 7.17 ms, 512x12x2 7.38,
// 256x16x3 7.41, 384x16x2 7.42, 256x24x2 7.69, 1024x8x1 8.63, 512x16x1 8.82, 256x32x1 9.95.
// 384 x 12 stages 4608 pairs (54 KB) in shared memory; three blocks share an SM.
#ifndef B200X_PASS_THREADS
#define B200X_PASS_THREADS 384
#define B200X_PASS_IPT 12
#define B200X_PASS_MINB 3
#endif
constexpr int PASS_THREADS = B200X_PASS_THREADS;
constexpr int PASS_IPT = B200X_PASS_IPT;
constexpr int PASS_MINB = B200X_PASS_MINB;
constexpr int PASS_TILE = PASS_THREADS * PASS_IPT;

// blocks of the finisher that are resident on the GPU at once (it walks over the runs)
template <uint32_t THREADS>
static uint32_t fin_grid(size_t smem, int sms) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, finish_runs_walk_kernel<THREADS>, (int)THREADS, smem) !=
          cudaSuccess || per_sm < 1)
    per_sm = 1;
  const uint32_t g = (uint32_t)per_sm * (uint32_t)sms;
  return g > JOINT_BINS ? JOINT_BINS : g;
}

struct TempLayout {
  size_t off_plan, off_hplan, off_ctl, off_hist, off_hbase, off_joint, off_runs, off_lookback, off_hlookback;
  size_t off_k0, off_k1, off_v0, off_v1, total;
  uint32_t tiles;
};

static TempLayout temp_layout(size_t n, int npass) {
  TempLayout L;
  L.tiles = (uint32_t)((n + PASS_TILE - 1) / PASS_TILE);
  size_t o = 0;
  L.off_plan = o;       o = align_up(o + sizeof(SortPlan), 256);
  L.off_hplan = o;      o = align_up(o + sizeof(SortPlan), 256);
  L.off_ctl = o;        o = align_up(o + sizeof(SortCtl), 256);
  L.off_hist = o;       o = align_up(o + sizeof(uint32_t) * MAX_PASSES * RADIX, 256);
  L.off_hbase = o;      o = align_up(o + sizeof(uint32_t) * 2 * RADIX, 256);
  // digit histograms, sampled U bins, sampled T bins, T table
  L.off_joint = o;      o = align_up(o + sizeof(uint32_t) * (2 * RADIX + 3 * JOINT_BINS), 256);
  L.off_runs = o;       o = align_up(o + sizeof(uint32_t) * (JOINT_BINS + 1), 256);
  L.off_lookback = o;   o = align_up(o + sizeof(uint32_t) * (size_t)npass * L.tiles * RADIX, 256);
  L.off_hlookback = o;  o = align_up(o + sizeof(uint32_t) * (size_t)2 * L.tiles * RADIX, 256);
  L.off_k0 = o;         o = align_up(o + n * 8, 256);
  L.off_k1 = o;         o = align_up(o + n * 8, 256);
  L.off_v0 = o;         o = align_up(o + n * 4, 256);
  L.off_v1 = o;         o = align_up(o + n * 4, 256);
  L.total = o;
  return L;
}

}  // namespace b200x

using namespace b200x;

namespace b200x { extern int g_scatter_blocks_per_sm; extern int g_gather_chains; extern int g_gather_blocks_per_sm; extern int g_scatter_bulk; }

extern "C" int b200x_set_tuning(int key, long long value) {
  if (key == 0 && value >= 0) {
    g_hybrid_min_rows = (unsigned long long)value;
    return B200X_OK;
  }
  if (key == 1 && value >= 0 && value <= 32) {
    b200x::g_scatter_blocks_per_sm = (int)value;
    return B200X_OK;
  }
  if (key == 4 && (value == 0 || value == 1)) {
    b200x::g_scatter_bulk = (int)value;
    return B200X_OK;
  }
  if (key == 3 && value >= 1 && value <= 32) {
    b200x::g_gather_blocks_per_sm = (int)value;
    return B200X_OK;
  }
  if (key == 2 && (value == 2 || value == 4)) {
    b200x::g_gather_chains = (int)value;
    return B200X_OK;
  }
  set_error("b200x_set_tuning: unknown key %d or bad value", key);
  return B200X_ERR_ARG;
}

extern "C" int b200x_radix_argsort(const void* keys, int key_type, int descending, const uint32_t* perm_in,
                                   uint32_t* perm_out, uint64_t* codes_out, void* keys_out, size_t n, int begin_bit,
                                   int end_bit, void* temp, size_t* temp_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(keys);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_radix_argsort: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  if (key_type < 0 || key_type >= KEY_TYPE_COUNT) {
    set_error("b200x_radix_argsort: bad key_type %d", key_type);
    return B200X_ERR_ARG;
  }
  if (keys_out != nullptr && (key_type == KEY_F64 || key_type == KEY_F32)) {
    // -0.0 and NaN payloads do not survive the code; float keys are gathered like any other column
    set_error("b200x_radix_argsort: keys_out is only available for integer key types");
    return B200X_ERR_ARG;
  }
  if (n >= ((size_t)1 << 30)) {
    set_error("b200x_radix_argsort: n=%zu exceeds the 2^30-1 rows one call can sort", n);
    return B200X_ERR_TOO_LARGE;
  }
  const int kbits = key_type_bits(key_type);
  if (begin_bit != 0) {
    set_error("b200x_radix_argsort: begin_bit must be 0");
    return B200X_ERR_ARG;
  }
  if (end_bit <= 0 || end_bit > 64) end_bit = 64;
  // NaN of a 32-bit float maps to the all-ones 64-bit code, so floats always scan all digits
  int bits = (key_type == KEY_F32 || key_type == KEY_F64) ? 64 : min(kbits, end_bit);
  const int npass = (bits + RADIX_BITS - 1) / RADIX_BITS;

  const TempLayout L = temp_layout(n, npass);
  if (temp == nullptr) {
    *temp_bytes = L.total;
    return B200X_OK;
  }
  if (*temp_bytes < L.total) {
    set_error("b200x_radix_argsort: temp buffer %zu < required %zu bytes", *temp_bytes, L.total);
    return B200X_ERR_TEMP;
  }
  if (((uintptr_t)temp & 255) != 0) {
    set_error("b200x_radix_argsort: temp must be 256-byte aligned");
    return B200X_ERR_ALIGN;
  }
  if (n == 0) return B200X_OK;
  if (keys == nullptr || perm_out == nullptr) {
    set_error("b200x_radix_argsort: keys/perm_out is NULL");
    return B200X_ERR_ARG;
  }

  unsigned char* base = (unsigned char*)temp;
  SortPlan* plan = (SortPlan*)(base + L.off_plan);
  SortPlan* hplan = (SortPlan*)(base + L.off_hplan);
  SortCtl* ctl = (SortCtl*)(base + L.off_ctl);
  uint32_t* ghist = (uint32_t*)(base + L.off_hist);
  uint32_t* hbase = (uint32_t*)(base + L.off_hbase);
  uint32_t* hist2 = (uint32_t*)(base + L.off_joint);
  uint32_t* run_start = (uint32_t*)(base + L.off_runs);
  uint32_t* lookback = (uint32_t*)(base + L.off_lookback);
  uint32_t* hlookback = (uint32_t*)(base + L.off_hlookback);

  // plans, control block, histograms and look-back words start at zero (one contiguous clear)
  B200X_CUDA(cudaMemsetAsync(base, 0, L.off_k0, stream));

  const int sms = sm_count();
  constexpr size_t pass_smem = PassSmem<PASS_THREADS, PASS_IPT>::bytes;
  constexpr size_t fin_small_smem = sizeof(FinSmem<FIN_SMALL_THREADS>), fin_big_smem = sizeof(FinSmem<FIN_BIG_THREADS>);
  constexpr size_t fin_huge_smem = sizeof(FinSmem<FIN_HUGE_THREADS>);
  static thread_local int smem_set_dev = -1;
  {
    int dev = 0;
    B200X_CUDA(cudaGetDevice(&dev));
    if (dev != smem_set_dev) {
      B200X_CUDA(cudaFuncSetAttribute(onesweep_pass_kernel<PASS_THREADS, PASS_IPT, PASS_MINB, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pass_smem));
      B200X_CUDA(cudaFuncSetAttribute(onesweep_pass_kernel<PASS_THREADS, PASS_IPT, PASS_MINB, true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pass_smem));
      B200X_CUDA(cudaFuncSetAttribute(finish_runs_kernel<FIN_SMALL_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)fin_small_smem));
      B200X_CUDA(cudaFuncSetAttribute(finish_runs_kernel<FIN_BIG_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)fin_big_smem));
      B200X_CUDA(cudaFuncSetAttribute(finish_runs_walk_kernel<FIN_HUGE_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)fin_huge_smem));
      smem_set_dev = dev;
    }
  }

  PassArgs a;
  a.raw = keys;
  a.perm_in = perm_in;
  a.kbuf[0] = (uint64_t*)(base + L.off_k0);
  a.kbuf[1] = (uint64_t*)(base + L.off_k1);
  a.vbuf[0] = (uint32_t*)(base + L.off_v0);
  a.vbuf[1] = (uint32_t*)(base + L.off_v1);
  a.codes_out = codes_out;
  a.keys_out = keys_out;
  a.perm_out = perm_out;
  a.n = (uint32_t)n;
  a.kt = key_type;
  a.desc = descending;

  // ---- analysis + HYBRID path (kernels return at once when the device picks LSD)
  const bool try_hybrid = bits > 2 * RADIX_BITS && (unsigned long long)n >= g_hybrid_min_rows;
  if (try_hybrid) {
    uint32_t blocks = (uint32_t)((n + 512 * 8 - 1) / (512 * 8));
    if (blocks > (uint32_t)sms * 4) blocks = (uint32_t)sms * 4;
    analyze_kernel<<<blocks, 512, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, ctl);
    B200X_LAUNCH_CHECK("analyze_kernel");
    hybrid_plan0_kernel<<<1, 1, 0, stream>>>(ctl, (uint32_t)n, g_hybrid_min_rows);
    B200X_LAUNCH_CHECK("hybrid_plan0_kernel");
    B200X_CUDA(cudaMemsetAsync(run_start, 0xFF, sizeof(uint32_t) * (JOINT_BINS + 1), stream));
    uint32_t* sampledU = hist2 + 2 * RADIX;
    uint32_t* sampledT = sampledU + JOINT_BINS;
    uint32_t* tab = sampledT + JOINT_BINS;
    uint32_t sblocks = (uint32_t)((n / SAMPLE_EVERY + 255) / 256);
    if (sblocks > (uint32_t)sms * 8) sblocks = (uint32_t)sms * 8;
    if (sblocks == 0) sblocks = 1;
    sample_bins_kernel<<<sblocks, 256, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, ctl, 0, sampledU,
                                                    tab);
    B200X_LAUNCH_CHECK("sample_bins_kernel");
    hybrid_planT_kernel<<<1, 1024, 0, stream>>>(ctl, sampledU, tab, (uint32_t)n);
    B200X_LAUNCH_CHECK("hybrid_planT_kernel");
    sample_bins_kernel<<<sblocks, 256, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, ctl, 1, sampledT,
                                                    tab);
    B200X_LAUNCH_CHECK("sample_bins_kernel");
    hybrid_checkT_kernel<<<1, 1024, 0, stream>>>(ctl, sampledT);
    B200X_LAUNCH_CHECK("hybrid_checkT_kernel");
    prof_begin(stream, PROF_JOINT);
    digit2_hist_kernel<false><<<blocks, 512, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, ctl, hist2,
                                                          tab, nullptr, nullptr);
    prof_end(stream);
    B200X_LAUNCH_CHECK("digit2_hist_kernel");
    prof_begin(stream, PROF_JOINT);
    digit2_hist_kernel<true><<<blocks, 512, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, ctl, hist2,
                                                         tab, a.kbuf[1], a.vbuf[1]);
    prof_end(stream);
    B200X_LAUNCH_CHECK("digit2_hist_kernel");
    hybrid_plan1_kernel<<<1, RADIX, 0, stream>>>(ctl, hist2, sampledU, sampledT, hbase, hplan, (uint32_t)n);
    B200X_LAUNCH_CHECK("hybrid_plan1_kernel");
    a.plan = hplan;
    a.ctl = ctl;
    a.mark_runs = run_start;
    for (int p = 0; p < 2; ++p) {
      a.pass = p;
      a.gbase = hbase + p * RADIX;
      a.lookback = hlookback + (size_t)p * L.tiles * RADIX;
      prof_begin(stream, PROF_PASS);
      onesweep_pass_kernel<PASS_THREADS, PASS_IPT, PASS_MINB, false><<<L.tiles, PASS_THREADS, pass_smem, stream>>>(a);
      prof_end(stream);
      B200X_LAUNCH_CHECK("onesweep_pass_kernel");
    }
    runs_fill_kernel<<<1, 1024, 0, stream>>>(ctl, hplan, run_start, (uint32_t)n);
    B200X_LAUNCH_CHECK("runs_fill_kernel");
    FinArgs f;
    f.kin = a.kbuf[1];
    f.vin = a.vbuf[1];
    f.perm_out = perm_out;
    f.codes_out = codes_out;
    f.keys_out = keys_out;
    f.run_start = run_start;
    f.ctl = ctl;
    f.tab = tab;
    f.raw = keys;
    f.kt = key_type;
    f.desc = descending;
    prof_begin(stream, PROF_FINISH);
    finish_runs_kernel<FIN_SMALL_THREADS><<<JOINT_BINS, FIN_SMALL_THREADS, fin_small_smem, stream>>>(f);
    prof_end(stream);
    B200X_LAUNCH_CHECK("finish_runs_kernel");
    prof_begin(stream, PROF_FINISH);
    finish_runs_kernel<FIN_BIG_THREADS><<<JOINT_BINS, FIN_BIG_THREADS, fin_big_smem, stream>>>(f);
    prof_end(stream);
    B200X_LAUNCH_CHECK("finish_runs_kernel");
    prof_begin(stream, PROF_FINISH);
    finish_runs_walk_kernel<FIN_HUGE_THREADS><<<fin_grid<FIN_HUGE_THREADS>(fin_huge_smem, sms), FIN_HUGE_THREADS, fin_huge_smem, stream>>>(f);
    prof_end(stream);
    B200X_LAUNCH_CHECK("finish_runs_kernel");
  }

  // ---- LSD path (kernels return at once when the device picked HYBRID)
  {
    const uint32_t chunk = HIST_THREADS * HIST_IPT;
    uint32_t blocks = (uint32_t)((n + chunk - 1) / chunk);
    const uint32_t cap = (uint32_t)sms * 4;
    if (blocks > cap) blocks = cap;
    prof_begin(stream, PROF_HIST);
    hist_kernel<<<blocks, HIST_THREADS, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, npass, ghist,
                                                     ctl);
    prof_end(stream);
    B200X_LAUNCH_CHECK("hist_kernel");
  }
  plan_kernel<<<1, RADIX, 0, stream>>>(ghist, plan, (uint32_t)n, npass);
  B200X_LAUNCH_CHECK("plan_kernel");
  a.plan = plan;
  a.ctl = nullptr;
  a.mark_runs = nullptr;
  uint32_t walk_grid = (uint32_t)sms * PASS_MINB;
  if (walk_grid > L.tiles) walk_grid = L.tiles;
  for (int p = 0; p < npass; ++p) {
    a.pass = p;
    a.gbase = ghist + p * RADIX;
    a.lookback = lookback + (size_t)p * L.tiles * RADIX;
    prof_begin(stream, PROF_PASS);
    if (try_hybrid)  // the device may well pick HYBRID: idle passes must be free
      onesweep_pass_kernel<PASS_THREADS, PASS_IPT, PASS_MINB, true><<<walk_grid, PASS_THREADS, pass_smem, stream>>>(a);
    else
      onesweep_pass_kernel<PASS_THREADS, PASS_IPT, PASS_MINB, false><<<L.tiles, PASS_THREADS, pass_smem, stream>>>(a);
    prof_end(stream);
    B200X_LAUNCH_CHECK("onesweep_pass_kernel");
  }
  {
    uint32_t blocks = (uint32_t)((n + 255) / 256);
    const uint32_t cap = (uint32_t)sms * 8;
    if (blocks > cap) blocks = cap;
    no_pass_kernel<<<blocks, 256, 0, stream>>>(keys, perm_in, key_type, descending, (uint32_t)n, plan, ctl, codes_out,
                                               keys_out, perm_out);
    B200X_LAUNCH_CHECK("no_pass_kernel");
  }
  return B200X_OK;
}

// ---------------------------------------------------------------------------
// b200x_radix_argsort_static: the same sort for callers that cannot allocate
// ---------------------------------------------------------------------------
// Some callers of this library have no way to hand over a scratch buffer: an operator invoked
// from inside a framework whose caching allocator already owns all of HBM (a raw cudaMalloc next
// to it fails or fragments the pool), a callback on a stream-ordered hot loop where cudaMalloc's
// implicit device synchronisation is not acceptable, or a thin FFI shim (cgo / JNI / ctypes
// without torch) that has no device allocator at all. For them the scratch lives inside the
// library image: one file-scope __device__ array per GPU, sized for the largest input this
// entry point accepts. The array exists on every GPU that loads libb200x.so, from module load
// until the process exits (about 27 MiB per GPU).
//
// The array's name is only a host-side stand-in; its device address is looked up once per
// device with cudaGetSymbolAddress while that device is current, and that address is what the
// kernels receive.
#include <mutex>

namespace b200x {

constexpr size_t STATIC_SORT_MAX_ROWS = (size_t)1 << 20;  // 1,048,576 rows per call
constexpr size_t STATIC_SORT_TILES = (STATIC_SORT_MAX_ROWS + PASS_TILE - 1) / PASS_TILE;
// ping-pong pairs (24 B/row) + look-back words of 8 + 2 passes + plans, histograms, run table and alignment slack
constexpr size_t STATIC_SORT_BYTES =
    24 * STATIC_SORT_MAX_ROWS + (MAX_PASSES + 2) * STATIC_SORT_TILES * RADIX * sizeof(uint32_t) + 2 * 1024 * 1024;

static __device__ __align__(256) unsigned char g_static_sort_scratch[STATIC_SORT_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static void* g_static_sort_ptr[STATIC_MAX_DEVICES];  // device address of the array on each GPU (0 = not looked up yet)
static std::mutex g_static_sort_mu;

}  // namespace b200x

extern "C" int b200x_radix_argsort_static(const void* keys, int key_type, int descending, const uint32_t* perm_in,
                                          uint32_t* perm_out, uint64_t* codes_out, void* keys_out, size_t n,
                                          void* stream) {
  if (n > STATIC_SORT_MAX_ROWS) {
    set_error("b200x_radix_argsort_static: n=%zu exceeds the %zu rows the built-in scratch holds; "
              "use b200x_radix_argsort with a caller-provided temp buffer",
              n, (size_t)STATIC_SORT_MAX_ROWS);
    return B200X_ERR_TOO_LARGE;
  }
  int dev = 0;
  B200X_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) {
    set_error("b200x_radix_argsort_static: device ordinal %d out of range", dev);
    return B200X_ERR_ARG;
  }
  void* scratch = nullptr;
  {
    // First call on this device: resolve the device address of the array (the current device
    // is the one the caller made current, i.e. the one the data lives on).
    std::lock_guard<std::mutex> lock(g_static_sort_mu);
    scratch = g_static_sort_ptr[dev];
    if (scratch == nullptr) {
      B200X_CUDA(cudaGetSymbolAddress(&scratch, g_static_sort_scratch));
      g_static_sort_ptr[dev] = scratch;
    }
  }
  // The sort clears its plan, histograms and look-back words on `stream` before it uses them, so
  // whatever the previous call left in the array does not matter; the ping-pong buffers are
  // written before they are read.
  size_t bytes = STATIC_SORT_BYTES;
  return b200x_radix_argsort(keys, key_type, descending, perm_in, perm_out, codes_out, keys_out, n, 0, 0, scratch,
                             &bytes, stream);
}
