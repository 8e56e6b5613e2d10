// Fused partition-gather + exchange over NVLink peer memory.
//
// The reference moves every map output to its reducer through the object store (plasma shared
// memory inside a node, chunked gRPC between nodes:
// python/ray/data/_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:86-124,
// push_based_shuffle_task_scheduler.py:514-568). On one NVSwitch box the same hop is ONE kernel:
// each rank walks its rows in destination order (perm = the local sort order, so the rows for
// rank d are one contiguous run), reads the row from its own HBM and stores it straight into rank
// d's receive buffer through the peer mapping, at the offset reserved for this sender. There is
// no send buffer and no separate copy: the gather that would have filled the send buffer is the
// transfer. Consecutive threads write consecutive remote addresses (rows for one peer are
// contiguous on both sides), so NVLink sees full-width stores.
//
// Ordering across ranks (all writers done before anybody reads) is the caller's job: a barrier on
// the symmetric-memory signal pads before and after this kernel (distributed.py).
#include "common.cuh"
#include "../../include/b200x.h"

namespace b200x {

constexpr int MAX_PEERS = 16;

struct PeerPlan {
  unsigned char* base[MAX_PEERS];          // receive buffer of rank d (peer-mapped address)
  unsigned long long row_off[MAX_PEERS];   // first row in rank d's buffer reserved for this sender
  unsigned long long send_off[MAX_PEERS + 1];  // rows [send_off[d], send_off[d+1]) of the send order go to rank d
  int world;
};

template <typename V>
__global__ void __launch_bounds__(256)
scatter_rows_to_peers_kernel(const V* __restrict__ src, const uint32_t* __restrict__ perm, const PeerPlan p, uint64_t n,
                             uint32_t vec_per_row) {
  constexpr int U = 8;  // stores in flight per thread: NVLink latency is ~2 us, keep the links busy from few blocks
  const uint64_t total = n * vec_per_row;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += U * stride) {
    V v[U];
    V* dst[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t t = t0 + u * stride;
      dst[u] = nullptr;
      if (t < total) {
        const uint64_t i = t / vec_per_row;
        const uint32_t c = (uint32_t)(t - i * vec_per_row);
        int d = 0;
#pragma unroll
        for (int k = 1; k < MAX_PEERS; ++k)
          if (k < p.world && i >= p.send_off[k]) d = k;
        const uint64_t srow = perm ? (uint64_t)perm[i] : i;
        v[u] = src[srow * vec_per_row + c];
        dst[u] = reinterpret_cast<V*>(p.base[d]) + (p.row_off[d] + (i - p.send_off[d])) * vec_per_row + c;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (dst[u]) *dst[u] = v[u];
  }
}

}  // namespace b200x

using namespace b200x;

extern "C" int b200x_scatter_rows_to_peers(const void* src, size_t row_bytes, const uint32_t* perm, size_t n,
                                           const unsigned long long* send_off_host, void* const* peer_base_host,
                                           const unsigned long long* peer_row_off_host, int world, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world <= 0 || world > MAX_PEERS) {
    set_error("b200x_scatter_rows_to_peers: world=%d out of range [1, %d]", world, MAX_PEERS);
    return B200X_ERR_ARG;
  }
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || send_off_host == nullptr || peer_base_host == nullptr || peer_row_off_host == nullptr) {
    set_error("b200x_scatter_rows_to_peers: NULL pointer");
    return B200X_ERR_ARG;
  }
  PeerPlan p;
  p.world = world;
  uintptr_t al = (uintptr_t)src | (uintptr_t)row_bytes;
  for (int d = 0; d < MAX_PEERS; ++d) {
    p.base[d] = d < world ? (unsigned char*)peer_base_host[d] : nullptr;
    p.row_off[d] = d < world ? peer_row_off_host[d] : 0;
    if (d < world) al |= (uintptr_t)p.base[d];
  }
  for (int d = 0; d <= MAX_PEERS; ++d) p.send_off[d] = d <= world ? send_off_host[d] : send_off_host[world];
  if (p.send_off[0] != 0 || p.send_off[world] != n) {
    set_error("b200x_scatter_rows_to_peers: send offsets must start at 0 and end at n");
    return B200X_ERR_ARG;
  }
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  const uint64_t total = (uint64_t)n * vpr;
  // A quarter of the resident-block slots: the transfer is link-bound, and the sort of the already
  // received key column runs next to it on another stream.
  uint64_t blocks = (total / 8 + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 4;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  const uint32_t g = (uint32_t)blocks;
  prof_begin(stream, PROF_SCATTER);
  switch (v) {
    case 16: scatter_rows_to_peers_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, perm, p, n, vpr); break;
    case 8: scatter_rows_to_peers_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, perm, p, n, vpr); break;
    case 4: scatter_rows_to_peers_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t*)src, perm, p, n, vpr); break;
    case 2: scatter_rows_to_peers_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t*)src, perm, p, n, vpr); break;
    default: scatter_rows_to_peers_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t*)src, perm, p, n, vpr); break;
  }
  prof_end(stream);
  B200X_LAUNCH_CHECK("scatter_rows_to_peers_kernel");
  return B200X_OK;
}

// ---------------------------------------------------------------------------
// Partition-first exchange: where does every row go, without sorting first
// ---------------------------------------------------------------------------
// b200x_partition_targets: for every row, the rank that owns its key range (found by comparing the
// key's code with the W-1 splitter codes) and the row's slot among the rows this rank sends there,
// in row order (a stable W-way split, so ties keep their order through the exchange).
// b200x_scatter_rows_by_target: stores row r (read sequentially, full-width) at that slot of the
// owner's receive buffer. Together they replace "sort locally, gather into send order, copy":
// the payload is read once, in order, and lands remotely in one kernel.
namespace b200x {

constexpr int PT_THREADS = 256;
constexpr int PT_IPT = 8;
constexpr int PT_TILE = PT_THREADS * PT_IPT;
constexpr int PT_WARPS = PT_THREADS / 32;
constexpr uint32_t SLOT_BITS = 28;  // target = owner << 28 | slot

__device__ __forceinline__ uint32_t owner_of(uint64_t code, const uint64_t* __restrict__ s_split, int nsplit) {
  uint32_t d = 0;
  for (int k = 0; k < nsplit; ++k) d += (s_split[k] <= code) ? 1u : 0u;  // rows equal to a splitter go right
  return d;
}

// lanes of the warp with the same 4-bit value
__device__ __forceinline__ unsigned match4(uint32_t d) {
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const bool bit = (d >> b) & 1u;
    const unsigned m = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// One pass over the keys (decoupled look-back over tiles, as in the radix passes): every tile counts its rows per
// owner, publishes the counts, adds up what the tiles before it published and writes its rows' targets. Tiles are
// handed out in launch order (ticket counter), so a tile only ever waits on tiles that are already running.
// Rows are warp-striped (item j of lane l of warp w = tile row w*256 + j*32 + l), so that slots follow row order.
// tile_base[tile][owner] (the slots handed out before the tile) stays behind for the caller.
constexpr uint32_t PT_FLAG_AGG = 1u << 30, PT_FLAG_INCL = 2u << 30, PT_VALUE_MASK = (1u << 30) - 1u;

__global__ void __launch_bounds__(PT_THREADS)
pt_fused_kernel(const void* __restrict__ keys, int kt, int desc, uint64_t n, const uint64_t* __restrict__ split, int nsplit,
                uint32_t* __restrict__ tile_base /*[tiles][MAX_PEERS]*/, uint32_t* __restrict__ lookback /*[tiles][MAX_PEERS]*/,
                uint32_t* __restrict__ ticket, uint32_t tiles, uint32_t* __restrict__ target,
                unsigned long long* __restrict__ counts_out) {
  __shared__ uint64_t s_split[MAX_PEERS];
  __shared__ uint32_t s_wcnt[PT_WARPS][MAX_PEERS];
  __shared__ uint32_t s_tile;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  if (threadIdx.x < MAX_PEERS) s_split[threadIdx.x] = threadIdx.x < (unsigned)nsplit ? split[threadIdx.x] : ~0ull;
  if (threadIdx.x < PT_WARPS * MAX_PEERS) (&s_wcnt[0][0])[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  if (tile >= tiles) return;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = lanemask_lt();
  const uint64_t base = (uint64_t)tile * PT_TILE + (uint64_t)warp * 32 * PT_IPT + lane;
  uint32_t dest[PT_IPT], rank[PT_IPT];
  const int width = key_type_bytes(kt);
  uint64_t raw[PT_IPT];
#pragma unroll
  for (int j = 0; j < PT_IPT; ++j) {
    const uint64_t r = base + (uint64_t)j * 32;
    raw[j] = r < n ? load_raw_bits(keys, r, width) : 0ull;
  }
#pragma unroll
  for (int j = 0; j < PT_IPT; ++j) {
    const uint64_t r = base + (uint64_t)j * 32;
    const bool valid = r < n;
    dest[j] = valid ? owner_of(encode_bits(raw[j], kt, desc != 0), s_split, nsplit) : (uint32_t)(MAX_PEERS - 1);
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    const unsigned peers = match4(dest[j]) & act;
    uint32_t old = 0;
    const int leader = __ffs(peers) - 1;
    if (valid && (int)lane == leader) {
      old = s_wcnt[warp][dest[j]];
      s_wcnt[warp][dest[j]] = old + __popc(peers);
    }
    old = __shfl_sync(0xffffffffu, old, leader < 0 ? 0 : leader);
    rank[j] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  // one thread per owner: scan over the warps, publish the tile's count, look back, publish the inclusive count
  if (threadIdx.x < MAX_PEERS) {
    const unsigned d = threadIdx.x;
    uint32_t run = 0;
    for (int w = 0; w < PT_WARPS; ++w) {
      const uint32_t c = s_wcnt[w][d];
      s_wcnt[w][d] = run;
      run += c;
    }
    volatile uint32_t* lb = lookback + (size_t)tile * MAX_PEERS + d;
    *lb = (tile == 0 ? PT_FLAG_INCL : PT_FLAG_AGG) | run;
    uint32_t excl = 0;
    if (tile > 0) {
      int64_t t = (int64_t)tile - 1;
      while (true) {
        volatile uint32_t* p = lookback + (size_t)t * MAX_PEERS + d;
        uint32_t v = *p;
        while ((v & ~PT_VALUE_MASK) == 0) {
          __nanosleep(20);
          v = *p;
        }
        excl += v & PT_VALUE_MASK;
        if (v & PT_FLAG_INCL) break;
        --t;
      }
      *lb = PT_FLAG_INCL | (excl + run);
    }
    tile_base[(size_t)tile * MAX_PEERS + d] = excl;
    if (tile == tiles - 1) counts_out[d] = (unsigned long long)excl + run;
    for (int w = 0; w < PT_WARPS; ++w) s_wcnt[w][d] += excl;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < PT_IPT; ++j) {
    const uint64_t r = base + (uint64_t)j * 32;
    if (r < n) target[r] = (dest[j] << SLOT_BITS) | (s_wcnt[warp][dest[j]] + rank[j]);
  }
}

template <typename V>
__global__ void __launch_bounds__(256)
scatter_rows_by_target_kernel(const V* __restrict__ src, const uint32_t* __restrict__ target, const PeerPlan p, uint64_t n,
                              uint32_t vec_per_row) {
  constexpr int U = 4;
  const uint64_t total = n * vec_per_row;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += U * stride) {
    V v[U];
    V* dst[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t t = t0 + u * stride;
      dst[u] = nullptr;
      if (t < total) {
        const uint64_t r = t / vec_per_row;
        const uint32_t c = (uint32_t)(t - r * vec_per_row);
        const uint32_t tg = target[r];
        const uint32_t d = tg >> SLOT_BITS;
        const uint64_t slot = tg & ((1u << SLOT_BITS) - 1u);
        v[u] = src[t];
        dst[u] = reinterpret_cast<V*>(p.base[d]) + (p.row_off[d] + slot) * vec_per_row + c;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (dst[u]) *dst[u] = v[u];
  }
}

// Staged variant for rows up to 128 bytes: a block takes a tile of consecutive rows, reorders them in
// shared memory by (owner, slot) - the slots of one owner inside a tile are consecutive - and then writes
// each owner's rows as ONE contiguous run, consecutive threads storing consecutive 16-byte pieces. Remote
// stores therefore leave the SM as full 128-byte requests instead of one 8..64-byte packet per row
// (8 GPUs, 100 M rows: key column 3.45 ms unstaged).
// b200x_set_tuning(1, k): staged scatter launches at most k blocks per SM (0 = one block per tile)
int g_scatter_blocks_per_sm = 0;
// b200x_set_tuning(2, k): loads in flight per thread of the record gather (2 or 4)
int g_gather_chains = 2;
// b200x_set_tuning(3, k): most blocks per SM the record gather launches (it shares the SMs with the transfer)
int g_gather_blocks_per_sm = 16;
// b200x_set_tuning(4, 0 | 1): the record scatter writes an owner's rows with one bulk copy (1, default) or with 16-byte stores (0)
int g_scatter_bulk = 1;

constexpr int ST_THREADS = 256;
constexpr int ST_MAX_TILE = 1024;
constexpr int ST_DATA_BYTES = 32 * 1024;

template <typename V>
__global__ void __launch_bounds__(ST_THREADS)
scatter_rows_staged_kernel(const V* __restrict__ src, const uint32_t* __restrict__ target, const PeerPlan p, uint64_t n,
                           uint32_t vec_per_row, uint32_t tile_rows) {
  extern __shared__ __align__(16) unsigned char st_raw[];
  V* sdata = reinterpret_cast<V*>(st_raw);                                            // [tile_rows * vec_per_row]
  unsigned long long* sdst = reinterpret_cast<unsigned long long*>(st_raw + ST_DATA_BYTES);  // [tile_rows] destination of staged row i
  uint16_t* slp = reinterpret_cast<uint16_t*>(sdst + ST_MAX_TILE);                     // [tile_rows] staged position of tile row
  __shared__ uint32_t s_cnt[MAX_PEERS], s_min[MAX_PEERS], s_off[MAX_PEERS];
  const uint64_t num_tiles = (n + tile_rows - 1) / tile_rows;
  // grid-stride over tiles: the host can launch one block per tile, or only a few blocks per SM so that
  // the transfer leaves most of every SM to work running next to it on another stream
  for (uint64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
  const uint64_t r0 = tile * tile_rows;
  const uint32_t rows = (uint32_t)min((uint64_t)tile_rows, n - r0);
  if (threadIdx.x < MAX_PEERS) {
    s_cnt[threadIdx.x] = 0;
    s_min[threadIdx.x] = 0xFFFFFFFFu;
  }
  __syncthreads();
  constexpr int RPT = ST_MAX_TILE / ST_THREADS;
  uint32_t tg[RPT];
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const uint32_t lr = threadIdx.x + k * ST_THREADS;
    tg[k] = 0;
    if (lr < rows) {
      tg[k] = target[r0 + lr];
      const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
      atomicAdd(&s_cnt[d], 1u);
      atomicMin(&s_min[d], slot);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int d = 0; d < MAX_PEERS; ++d) {
      s_off[d] = run;
      run += s_cnt[d];
    }
  }
  __syncthreads();
  const size_t row_bytes = (size_t)vec_per_row * sizeof(V);
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const uint32_t lr = threadIdx.x + k * ST_THREADS;
    if (lr < rows) {
      const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
      const uint32_t lp = s_off[d] + (slot - s_min[d]);
      slp[lr] = (uint16_t)lp;
      sdst[lp] = (unsigned long long)(p.base[d] + (p.row_off[d] + slot) * row_bytes);
    }
  }
  __syncthreads();
  const uint32_t chunks = rows * vec_per_row;
  const V* tsrc = src + r0 * vec_per_row;
  for (uint32_t t = threadIdx.x; t < chunks; t += ST_THREADS) {  // sequential, coalesced read of the tile
    const uint32_t lr = t / vec_per_row, c = t - lr * vec_per_row;
    sdata[(uint32_t)slp[lr] * vec_per_row + c] = tsrc[t];
  }
  __syncthreads();
  for (uint32_t t = threadIdx.x; t < chunks; t += ST_THREADS) {  // owner runs leave as contiguous stores
    const uint32_t i = t / vec_per_row, c = t - i * vec_per_row;
    reinterpret_cast<V*>(sdst[i])[c] = sdata[t];
  }
  __syncthreads();
  }
}


// Realigning variant for rows whose width is not a multiple of 16 bytes (gensort's 90-byte payload, 10- or
// 2-byte key pieces): moving such rows chunk by chunk means 2-byte (or 1-byte) loads and stores - 45 remote
// stores per 90-byte row. Here the tile of consecutive source rows is copied into shared memory with 16-byte
// loads (the tile starts 16-byte aligned: tile_rows is a multiple of 16), and every owner's rows - one
// contiguous byte range at the destination - leave as ALIGNED 16-byte stores: each thread assembles one
// destination vector from the shared-memory image (five word loads and funnel shifts when its 16 bytes lie in
// one row, a byte loop when they span two rows). The last vector of a range is usually shared with the next
// rows of the same owner, which lie in a later tile: the block looks ahead in `target` for that owner's next
// row, takes the missing bytes from it and stores the whole vector, and the later tile skips its first,
// partial vector (single-byte stores are separate NVLink packets: with them a 90-byte payload moved at half
// the link rate). Only the first vector of a rank's whole range for an owner, and the last one when no further
// row of that owner exists, are written byte by byte.
constexpr int RA_THREADS = 256;
constexpr int RA_MAX_TILE = 1024;
constexpr int RA_DATA_BYTES = 32 * 1024;

__global__ void __launch_bounds__(RA_THREADS)
scatter_rows_realign_kernel(const unsigned char* __restrict__ src, const uint32_t* __restrict__ target, const PeerPlan p,
                            uint64_t n, uint32_t W, uint32_t magic /* ceil(2^32 / W) */, uint32_t tile_rows) {
  extern __shared__ __align__(16) unsigned char ra_raw[];
  unsigned char* img = ra_raw;                                                 // [tile_rows * W] (+ 16 bytes of slack)
  uint16_t* srow = reinterpret_cast<uint16_t*>(ra_raw + RA_DATA_BYTES + 32);   // [tile_rows] tile row at staged position
  __shared__ uint32_t s_cnt[MAX_PEERS], s_min[MAX_PEERS], s_off[MAX_PEERS], s_vec0[MAX_PEERS + 1];
  __shared__ unsigned long long s_dst[MAX_PEERS], s_next[MAX_PEERS];
  __shared__ uint32_t s_missing;
  constexpr unsigned long long NONE = ~0ull;
  const uint64_t num_tiles = (n + tile_rows - 1) / tile_rows;
  for (uint64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const uint64_t r0 = tile * tile_rows;
    const uint32_t rows = (uint32_t)min((uint64_t)tile_rows, n - r0);
    if (threadIdx.x < MAX_PEERS) {
      s_cnt[threadIdx.x] = 0;
      s_min[threadIdx.x] = 0xFFFFFFFFu;
      s_next[threadIdx.x] = NONE;
    }
    __syncthreads();
    constexpr int RPT = RA_MAX_TILE / RA_THREADS;
    uint32_t tg[RPT];
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const uint32_t lr = threadIdx.x + k * RA_THREADS;
      tg[k] = 0;
      if (lr < rows) {
        tg[k] = target[r0 + lr];
        const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
        atomicAdd(&s_cnt[d], 1u);
        atomicMin(&s_min[d], slot);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t run = 0, vec = 0;
      for (int d = 0; d < MAX_PEERS; ++d) {
        s_off[d] = run;
        s_vec0[d] = vec;
        if (s_cnt[d]) {
          const unsigned long long g = (unsigned long long)(p.base[d]) + (p.row_off[d] + s_min[d]) * (unsigned long long)W;
          s_dst[d] = g;
          // vectors are counted from the 128-byte line the range starts in, and every owner's count is padded to
          // a multiple of 32: a warp then always stores four whole 128-byte lines (full-line NVLink packets)
          vec += (((uint32_t)(g & 127ull) + s_cnt[d] * W + 15u) / 16u + 31u) & ~31u;
        }
        run += s_cnt[d];
      }
      s_vec0[MAX_PEERS] = vec;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const uint32_t lr = threadIdx.x + k * RA_THREADS;
      if (lr < rows) {
        const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
        srow[s_off[d] + (slot - s_min[d])] = (uint16_t)lr;  // the slots of one owner inside a tile are consecutive
      }
    }
    // the tile as it lies in the source: sequential 16-byte loads
    const uint32_t tbytes = rows * W;
    const unsigned char* tsrc = src + r0 * W;
    const uint32_t nvec = tbytes / 16;
    for (uint32_t i = threadIdx.x; i < nvec; i += RA_THREADS)
      reinterpret_cast<uint4*>(img)[i] = reinterpret_cast<const uint4*>(tsrc)[i];
    for (uint32_t i = nvec * 16 + threadIdx.x; i < tbytes; i += RA_THREADS) img[i] = tsrc[i];
    // look ahead: the next row of every owner whose range ends inside a vector
    for (uint64_t base = r0 + rows; base < n; base += RA_THREADS) {
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t missing = 0;
        for (int d = 0; d < MAX_PEERS; ++d)
          missing += (s_cnt[d] && ((s_dst[d] + (unsigned long long)s_cnt[d] * W) & 15ull) && s_next[d] == NONE) ? 1u : 0u;
        s_missing = missing;
      }
      __syncthreads();
      if (s_missing == 0) break;
      const uint64_t i = base + threadIdx.x;
      if (i < n) atomicMin(&s_next[target[i] >> SLOT_BITS], (unsigned long long)i);
    }
    __syncthreads();
    const uint32_t total = s_vec0[MAX_PEERS];
    const uint32_t* img32 = reinterpret_cast<const uint32_t*>(img);
    for (uint32_t t = threadIdx.x; t < total; t += RA_THREADS) {
      uint32_t d = 0;
      while (t >= s_vec0[d + 1]) ++d;
      const uint32_t k = t - s_vec0[d];
      const unsigned long long g = s_dst[d];
      const uint32_t a = (uint32_t)(g & 127ull), L = s_cnt[d] * W, first = s_off[d];
      const int lo = (int)(16u * k) - (int)a;  // position of the vector's first byte inside the owner's byte range
      if (lo + 16 <= 0 || lo >= (int)L) continue;  // before the range (line padding) or after it (warp padding)
      unsigned char* gp = reinterpret_cast<unsigned char*>(g - a) + 16ull * k;
      if (lo >= 0 && (uint32_t)lo + 16u <= L) {
        const uint32_t q = __umulhi((uint32_t)lo, magic), w = (uint32_t)lo - q * W;
        const uint32_t rowA = (uint32_t)srow[first + q] * W;
        uint4 v;
        if (w + 16u <= W) {
          const uint32_t A0 = rowA + w, wi = A0 >> 2, sh = (A0 & 3u) * 8u;
          const uint32_t x0 = img32[wi], x1 = img32[wi + 1], x2 = img32[wi + 2], x3 = img32[wi + 3], x4 = img32[wi + 4];
          v.x = __funnelshift_r(x0, x1, sh);
          v.y = __funnelshift_r(x1, x2, sh);
          v.z = __funnelshift_r(x2, x3, sh);
          v.w = __funnelshift_r(x3, x4, sh);
        } else {
          const uint32_t rowB = (uint32_t)srow[first + q + 1] * W;
          uint32_t out[4] = {0u, 0u, 0u, 0u};
#pragma unroll
          for (uint32_t j = 0; j < 16; ++j) {
            const uint32_t b = (w + j < W) ? img[rowA + w + j] : img[rowB + (w + j - W)];
            out[j >> 2] |= b << ((j & 3u) * 8u);
          }
          v = make_uint4(out[0], out[1], out[2], out[3]);
        }
        *reinterpret_cast<uint4*>(gp) = v;
      } else if (lo < 0) {
        // first vector of the range: an earlier tile of this rank completes it unless these are the rank's first rows
        if (s_min[d] == 0) {
          for (int o = 0; o < lo + 16; ++o) {
            const uint32_t q = __umulhi((uint32_t)o, magic), w = (uint32_t)o - q * W;
            gp[o - lo] = img[(uint32_t)srow[first + q] * W + w];
          }
        }
      } else {
        // last vector: mine are the first L - lo bytes; the rest is the beginning of the owner's next row
        const uint32_t mine = L - (uint32_t)lo;
        const unsigned long long nxt = s_next[d];
        if (nxt != NONE) {
          const unsigned char* nrow = src + nxt * W;
          uint32_t out[4] = {0u, 0u, 0u, 0u};
#pragma unroll
          for (uint32_t j = 0; j < 16; ++j) {
            uint32_t b;
            if (j < mine) {
              const uint32_t o = (uint32_t)lo + j, q = __umulhi(o, magic), w = o - q * W;
              b = img[(uint32_t)srow[first + q] * W + w];
            } else {
              b = nrow[j - mine];
            }
            out[j >> 2] |= b << ((j & 3u) * 8u);
          }
          *reinterpret_cast<uint4*>(gp) = make_uint4(out[0], out[1], out[2], out[3]);
        } else {
          for (uint32_t j = 0; j < mine; ++j) {
            const uint32_t o = (uint32_t)lo + j, q = __umulhi(o, magic), w = o - q * W;
            gp[j] = img[(uint32_t)srow[first + q] * W + w];
          }
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace b200x

extern "C" int b200x_partition_targets(const void* keys, int key_type, int descending, size_t n,
                                       const uint64_t* splitter_codes, int num_splitters, uint32_t* target_out,
                                       unsigned long long* counts_out, void* temp, size_t* temp_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(counts_out);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_partition_targets: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  if (num_splitters < 0 || num_splitters >= MAX_PEERS || key_type < 0 || key_type >= KEY_TYPE_COUNT) {
    set_error("b200x_partition_targets: bad key_type or more than %d splitters", MAX_PEERS - 1);
    return B200X_ERR_ARG;
  }
  if (n >= ((size_t)1 << SLOT_BITS)) {
    set_error("b200x_partition_targets: n=%zu exceeds 2^28-1 rows per call", n);
    return B200X_ERR_TOO_LARGE;
  }
  const uint32_t tiles = (uint32_t)((n + PT_TILE - 1) / PT_TILE);
  const size_t table_bytes = align_up((size_t)(tiles + 1) * MAX_PEERS * sizeof(uint32_t), 256);
  const size_t need = 2 * table_bytes + 256;   // tile_base | look-back words | ticket
  if (temp == nullptr) {
    *temp_bytes = need;
    return B200X_OK;
  }
  if (*temp_bytes < need) {
    set_error("b200x_partition_targets: temp buffer %zu < required %zu bytes", *temp_bytes, need);
    return B200X_ERR_TEMP;
  }
  if (counts_out == nullptr) {
    set_error("b200x_partition_targets: counts_out is NULL");
    return B200X_ERR_ARG;
  }
  if (n == 0) {
    B200X_CUDA(cudaMemsetAsync(counts_out, 0, sizeof(unsigned long long) * MAX_PEERS, stream));
    return B200X_OK;
  }
  if (keys == nullptr || target_out == nullptr || (num_splitters > 0 && splitter_codes == nullptr)) {
    set_error("b200x_partition_targets: NULL pointer");
    return B200X_ERR_ARG;
  }
  uint32_t* tile_base = (uint32_t*)temp;
  uint32_t* lookback = (uint32_t*)((unsigned char*)temp + table_bytes);
  uint32_t* ticket = (uint32_t*)((unsigned char*)temp + 2 * table_bytes);
  B200X_CUDA(cudaMemsetAsync(lookback, 0, table_bytes + 256, stream));   // look-back flags and the ticket start at zero
  pt_fused_kernel<<<tiles, PT_THREADS, 0, stream>>>(keys, key_type, descending, n, splitter_codes, num_splitters, tile_base,
                                                    lookback, ticket, tiles, target_out, counts_out);
  B200X_LAUNCH_CHECK("pt_fused_kernel");
  return B200X_OK;
}

extern "C" int b200x_scatter_rows_by_target(const void* src, size_t row_bytes, const uint32_t* target, size_t n,
                                            void* const* peer_base_host, const unsigned long long* peer_row_off_host,
                                            int world, void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world <= 0 || world > MAX_PEERS) {
    set_error("b200x_scatter_rows_by_target: world=%d out of range [1, %d]", world, MAX_PEERS);
    return B200X_ERR_ARG;
  }
  if (n == 0 || row_bytes == 0) return B200X_OK;
  if (src == nullptr || target == nullptr || peer_base_host == nullptr || peer_row_off_host == nullptr) {
    set_error("b200x_scatter_rows_by_target: NULL pointer");
    return B200X_ERR_ARG;
  }
  PeerPlan p;
  p.world = world;
  uintptr_t al = (uintptr_t)src | (uintptr_t)row_bytes;
  for (int d = 0; d < MAX_PEERS; ++d) {
    p.base[d] = d < world ? (unsigned char*)peer_base_host[d] : nullptr;
    p.row_off[d] = d < world ? peer_row_off_host[d] : 0;
    if (d < world) al |= (uintptr_t)p.base[d];
  }
  for (int d = 0; d <= MAX_PEERS; ++d) p.send_off[d] = 0;
  int v = 16;
  while (v > 1 && (al & (uintptr_t)(v - 1)) != 0) v >>= 1;
  const uint32_t vpr = (uint32_t)(row_bytes / v);
  const uint64_t total = (uint64_t)n * vpr;
  if (v < 16 && row_bytes >= 24 && row_bytes <= 1024 && ((uintptr_t)src & 15) == 0) {
    // rows that do not split into 16-byte pieces: realign in shared memory, store aligned vectors
    uint32_t tile_rows = (uint32_t)(RA_DATA_BYTES / row_bytes) / 16 * 16;
    if (tile_rows > (uint32_t)RA_MAX_TILE) tile_rows = RA_MAX_TILE;
    const size_t smem = RA_DATA_BYTES + 32 + RA_MAX_TILE * sizeof(uint16_t);
    uint64_t gtiles = (n + tile_rows - 1) / tile_rows;
    if (g_scatter_blocks_per_sm > 0 && gtiles > (uint64_t)sm_count() * g_scatter_blocks_per_sm)
      gtiles = (uint64_t)sm_count() * g_scatter_blocks_per_sm;
    const uint32_t magic = (uint32_t)((0x100000000ull + row_bytes - 1) / row_bytes);
    prof_begin(stream, PROF_SCATTER);
    scatter_rows_realign_kernel<<<(uint32_t)gtiles, RA_THREADS, smem, stream>>>((const unsigned char*)src, target, p, n,
                                                                                (uint32_t)row_bytes, magic, tile_rows);
    prof_end(stream);
    B200X_LAUNCH_CHECK("scatter_rows_realign_kernel");
    return B200X_OK;
  }
  if (row_bytes <= 128) {
    uint32_t tile_rows = (uint32_t)(ST_DATA_BYTES / row_bytes) / ST_THREADS * ST_THREADS;
    if (tile_rows > (uint32_t)ST_MAX_TILE) tile_rows = ST_MAX_TILE;
    const size_t smem = ST_DATA_BYTES + ST_MAX_TILE * sizeof(unsigned long long) + ST_MAX_TILE * sizeof(uint16_t);
    uint64_t gtiles = (n + tile_rows - 1) / tile_rows;
    if (g_scatter_blocks_per_sm > 0 && gtiles > (uint64_t)sm_count() * g_scatter_blocks_per_sm)
      gtiles = (uint64_t)sm_count() * g_scatter_blocks_per_sm;
    const uint32_t g = (uint32_t)gtiles;
    prof_begin(stream, PROF_SCATTER);
    switch (v) {
      case 16: scatter_rows_staged_kernel<uint4><<<g, ST_THREADS, smem, stream>>>((const uint4*)src, target, p, n, vpr, tile_rows); break;
      case 8: scatter_rows_staged_kernel<uint64_t><<<g, ST_THREADS, smem, stream>>>((const uint64_t*)src, target, p, n, vpr, tile_rows); break;
      case 4: scatter_rows_staged_kernel<uint32_t><<<g, ST_THREADS, smem, stream>>>((const uint32_t*)src, target, p, n, vpr, tile_rows); break;
      case 2: scatter_rows_staged_kernel<uint16_t><<<g, ST_THREADS, smem, stream>>>((const uint16_t*)src, target, p, n, vpr, tile_rows); break;
      default: scatter_rows_staged_kernel<uint8_t><<<g, ST_THREADS, smem, stream>>>((const uint8_t*)src, target, p, n, vpr, tile_rows); break;
    }
    prof_end(stream);
    B200X_LAUNCH_CHECK("scatter_rows_staged_kernel");
    return B200X_OK;
  }
  uint64_t blocks = (total / 4 + 255) / 256;
  const uint64_t cap = (uint64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  const uint32_t g = (uint32_t)blocks;
  prof_begin(stream, PROF_SCATTER);
  switch (v) {
    case 16: scatter_rows_by_target_kernel<uint4><<<g, 256, 0, stream>>>((const uint4*)src, target, p, n, vpr); break;
    case 8: scatter_rows_by_target_kernel<uint64_t><<<g, 256, 0, stream>>>((const uint64_t*)src, target, p, n, vpr); break;
    case 4: scatter_rows_by_target_kernel<uint32_t><<<g, 256, 0, stream>>>((const uint32_t*)src, target, p, n, vpr); break;
    case 2: scatter_rows_by_target_kernel<uint16_t><<<g, 256, 0, stream>>>((const uint16_t*)src, target, p, n, vpr); break;
    default: scatter_rows_by_target_kernel<uint8_t><<<g, 256, 0, stream>>>((const uint8_t*)src, target, p, n, vpr); break;
  }
  prof_end(stream);
  B200X_LAUNCH_CHECK("scatter_rows_by_target_kernel");
  return B200X_OK;
}

// ---------------------------------------------------------------------------
// Record exchange: rows kept as records at a pitch (b200x_gather_rows_pitched)
// ---------------------------------------------------------------------------
// b200x_scatter_records_by_target: the staged scatter for records. Of every source record (src_pitch bytes apart)
// only copy_bytes bytes travel - the record minus padding, and minus its first 8 bytes when those are the sort key
// that was exchanged ahead as a column of its own - and land at the start of the owner's row slot (dst_pitch
// apart). With copy_bytes a multiple of 32 and dst_pitch a multiple of 128 every row leaves as whole 32-byte
// sectors of one 128-byte line: a gensort record (10-byte key + 90-byte payload in a 128-byte record) crosses
// NVLink as 96 bytes, the 8-byte key prefix having gone ahead with the key exchange.
// b200x_gather_records_prefixed: the receiver's last step. Output record i = [prefix(sorted_keys[i]) | the received
// row idx[i]], i.e. the elided key bytes are put back from the sorted key column (which the sort produced anyway)
// while the rows are gathered into sorted order. The gather can be restricted to rows whose receive slot lies in
// given ranges (rows that have already landed when the payload arrives in chunks), and to / from positions whose
// key equals a neighbour's (their order is only final after b200x_fix_ties).
namespace b200x {

// BULK (dst_pitch == copy_bytes: an owner's rows of a tile are one contiguous byte range at the destination): the
// write-out is ONE bulk copy per owner, shared memory -> the peer's buffer, issued by one thread and carried out by
// the SM's copy unit (cp.async.bulk, SASS UBLKCP) instead of a loop of 16-byte stores through the registers of all
// threads: the block's warps are free again as soon as the copies are queued, and the stores leave in large bursts.
__device__ __forceinline__ void bulk_store_to_global(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"((uint32_t)__cvta_generic_to_shared(ssrc)), "r"(bytes)
               : "memory");
}

template <typename LV, bool BULK>
__global__ void __launch_bounds__(ST_THREADS)
scatter_records_staged_kernel(const unsigned char* __restrict__ src, uint64_t src_pitch, uint32_t copy_bytes,
                              const uint32_t* __restrict__ target, const PeerPlan p, uint64_t dst_pitch, uint64_t n,
                              uint32_t tile_rows) {
  extern __shared__ __align__(16) unsigned char st_raw[];
  unsigned char* sdata = st_raw;                                                              // [tile_rows * copy_bytes]
  unsigned long long* sdst = reinterpret_cast<unsigned long long*>(st_raw + ST_DATA_BYTES);   // [tile_rows]
  uint16_t* slp = reinterpret_cast<uint16_t*>(sdst + ST_MAX_TILE);                            // [tile_rows]
  __shared__ uint32_t s_cnt[MAX_PEERS], s_min[MAX_PEERS], s_off[MAX_PEERS];
  const uint64_t num_tiles = (n + tile_rows - 1) / tile_rows;
  const uint32_t lpr = copy_bytes / (uint32_t)sizeof(LV);   // loads per row
  const uint32_t vpr = copy_bytes / 16u;                    // 16-byte stores per row
  for (uint64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const uint64_t r0 = tile * tile_rows;
    const uint32_t rows = (uint32_t)min((uint64_t)tile_rows, n - r0);
    if (threadIdx.x < MAX_PEERS) {
      s_cnt[threadIdx.x] = 0;
      s_min[threadIdx.x] = 0xFFFFFFFFu;
    }
    __syncthreads();
    constexpr int RPT = ST_MAX_TILE / ST_THREADS;
    uint32_t tg[RPT];
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const uint32_t lr = threadIdx.x + k * ST_THREADS;
      tg[k] = 0;
      if (lr < rows) {
        tg[k] = target[r0 + lr];
        const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
        atomicAdd(&s_cnt[d], 1u);
        atomicMin(&s_min[d], slot);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t run = 0;
      for (int d = 0; d < MAX_PEERS; ++d) {
        s_off[d] = run;
        run += s_cnt[d];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const uint32_t lr = threadIdx.x + k * ST_THREADS;
      if (lr < rows) {
        const uint32_t d = tg[k] >> SLOT_BITS, slot = tg[k] & ((1u << SLOT_BITS) - 1u);
        const uint32_t lp = s_off[d] + (slot - s_min[d]);
        slp[lr] = (uint16_t)lp;
        sdst[lp] = (unsigned long long)(p.base[d] + (p.row_off[d] + slot) * dst_pitch);
      }
    }
    __syncthreads();
    const unsigned char* tsrc = src + r0 * src_pitch;
    const uint32_t loads = rows * lpr;
    for (uint32_t t = threadIdx.x; t < loads; t += ST_THREADS) {   // the tile's rows, in order
      const uint32_t lr = t / lpr, c = t - lr * lpr;
      *reinterpret_cast<LV*>(sdata + (uint32_t)slp[lr] * copy_bytes + c * (uint32_t)sizeof(LV)) =
          *reinterpret_cast<const LV*>(tsrc + (uint64_t)lr * src_pitch + c * sizeof(LV));
    }
    if (BULK) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the staged rows become visible to the copy unit
      __syncthreads();
      if (threadIdx.x < MAX_PEERS && s_cnt[threadIdx.x] != 0) {
        const uint32_t d = threadIdx.x;
        bulk_store_to_global(p.base[d] + (p.row_off[d] + s_min[d]) * dst_pitch, sdata + s_off[d] * copy_bytes,
                             s_cnt[d] * copy_bytes);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory may be rewritten
      }
      __syncthreads();
    } else {
      __syncthreads();
      const uint32_t stores = rows * vpr;
      for (uint32_t t = threadIdx.x; t < stores; t += ST_THREADS) {  // owner by owner, consecutive slots
        const uint32_t i = t / vpr, c = t - i * vpr;
        reinterpret_cast<uint4*>(sdst[i])[c] = reinterpret_cast<const uint4*>(sdata + i * copy_bytes)[c];
      }
      __syncthreads();
    }
  }
  if (BULK) {
    if (threadIdx.x < MAX_PEERS) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // all copies have been performed
    __threadfence_system();
  }
}

__device__ __forceinline__ uint64_t bswap64(uint64_t v) {
  const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  return ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | (uint64_t)__byte_perm(hi, 0, 0x0123);
}

// keep the bytes of an 8-byte word whose record position (pos0 .. pos0+7) lies below `used`
__device__ __forceinline__ uint64_t keep_below(uint64_t w, uint32_t pos0, uint32_t used) {
  if (pos0 + 8u <= used) return w;
  if (pos0 >= used) return 0ull;
  return w & ((1ull << (8u * (used - pos0))) - 1ull);
}

// 16 output bytes [16j, 16j+16) of a record = received bytes [16j - shift, 16j - shift + 16) (shift 0 or 8), the
// first 8 bytes being the key prefix when shift == 8, zero from `used` on
__device__ __forceinline__ uint4 prefixed_vector(const unsigned char* __restrict__ row, uint32_t j, uint32_t shift,
                                                 uint32_t copy_bytes, uint32_t used, uint64_t key, int prefix_be) {
  const uint32_t o0 = 16u * j, o1 = o0 + 8u;
  uint64_t w0 = 0, w1 = 0;
  if (o0 >= shift && o0 - shift + 8u <= copy_bytes) w0 = *reinterpret_cast<const uint64_t*>(row + (o0 - shift));
  if (o1 >= shift && o1 - shift + 8u <= copy_bytes) w1 = *reinterpret_cast<const uint64_t*>(row + (o1 - shift));
  if (shift == 8u && j == 0) w0 = prefix_be ? bswap64(key) : key;
  w0 = keep_below(w0, o0, used);
  w1 = keep_below(w1, o1, used);
  uint4 v;
  v.x = (uint32_t)w0;
  v.y = (uint32_t)(w0 >> 32);
  v.z = (uint32_t)w1;
  v.w = (uint32_t)(w1 >> 32);
  return v;
}

// Output positions pos[0..m) (or 0..m-1 when pos is NULL): one thread per 16-byte piece, U pieces in flight.
// A random gather is latency-bound: the kernel is kept at 32 registers so that 2048 threads are resident per SM.
// LANES: dst_vec is a power of two <= 32 and copy_bytes a multiple of 16, so the pieces of a row sit in neighbouring
// lanes (a thread's piece number never changes): every lane loads ONE aligned 16-byte vector of the received row
// and, when the row is shifted by the 8-byte key prefix, takes the other half of its output vector from the lane
// below with a shuffle.
template <bool LANES, int U>
__global__ void __launch_bounds__(256, LANES ? 8 : 4)
gather_records_list_kernel(const unsigned char* __restrict__ recv, uint64_t recv_pitch, uint32_t copy_bytes,
                           const uint32_t* __restrict__ idx, const uint32_t* __restrict__ pos, uint64_t m,
                           const uint64_t* __restrict__ keys, int prefix_be, uint32_t shift, uint32_t used,
                           unsigned char* __restrict__ dst, uint32_t dst_vec) {
  const uint64_t total = m * dst_vec;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (LANES) {
    const uint32_t c = (uint32_t)t & (dst_vec - 1u);        // stride is a multiple of 32, hence of dst_vec
    const uint32_t sh = 31u - (uint32_t)__clz(dst_vec);
    const bool has = 16u * c + 16u <= copy_bytes;
    // (all lanes of a warp leave the loop together: the trip count depends on the warp's first element only)
    for (; t - (threadIdx.x & 31) < total; t += U * stride) {
      uint32_t orow[U], s[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint64_t e = t + u * stride;
        orow[u] = e < total ? (uint32_t)(e >> sh) : 0xFFFFFFFFu;
        if (pos != nullptr && e < total) orow[u] = pos[orow[u]];
      }
#pragma unroll
      for (int u = 0; u < U; ++u) s[u] = orow[u] != 0xFFFFFFFFu ? idx[orow[u]] : 0u;
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        v[u] = make_uint4(0u, 0u, 0u, 0u);
        if (has && orow[u] != 0xFFFFFFFFu) v[u] = *reinterpret_cast<const uint4*>(recv + (uint64_t)s[u] * recv_pitch + 16u * c);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        uint64_t w0 = (uint64_t)v[u].x | ((uint64_t)v[u].y << 32), w1 = (uint64_t)v[u].z | ((uint64_t)v[u].w << 32);
        if (shift == 8u) {
          const uint64_t below = __shfl_up_sync(0xffffffffu, w1, 1);   // upper half of the vector before mine
          w1 = w0;
          w0 = below;
          if (c == 0 && orow[u] != 0xFFFFFFFFu) {
            const uint64_t key = keys[orow[u]];
            w0 = prefix_be ? bswap64(key) : key;
          }
        }
        w0 = keep_below(w0, 16u * c, used);
        w1 = keep_below(w1, 16u * c + 8u, used);
        if (orow[u] != 0xFFFFFFFFu)
          reinterpret_cast<uint4*>(dst)[(uint64_t)orow[u] * dst_vec + c] =
              make_uint4((uint32_t)w0, (uint32_t)(w0 >> 32), (uint32_t)w1, (uint32_t)(w1 >> 32));
      }
    }
  } else {
    const uint64_t q = stride / dst_vec;
    const uint32_t rem = (uint32_t)(stride - q * dst_vec);
    uint64_t r = t / dst_vec;
    uint32_t c = (uint32_t)(t - r * dst_vec);
    for (; t < total; t += stride) {
      const uint64_t orow = pos ? (uint64_t)pos[r] : r;
      const uint64_t key = (shift == 8u && c == 0) ? keys[orow] : 0ull;
      reinterpret_cast<uint4*>(dst)[orow * dst_vec + c] =
          prefixed_vector(recv + (uint64_t)idx[orow] * recv_pitch, c, shift, copy_bytes, used, key, prefix_be);
      c += rem;
      r += q;
      if (c >= dst_vec) {
        c -= dst_vec;
        ++r;
      }
    }
  }
}

// Only (tie_mode 2) or all but (tie_mode 1) the positions whose key equals a neighbour's. A warp takes 32
// consecutive output rows, each lane decides for one of them, then groups of lanes copy the selected rows.
__global__ void __launch_bounds__(256)
gather_records_ties_kernel(const unsigned char* __restrict__ recv, uint64_t recv_pitch, uint32_t copy_bytes,
                           const uint32_t* __restrict__ idx, uint64_t n, const uint64_t* __restrict__ keys, int prefix_be,
                           uint32_t shift, uint32_t used, unsigned char* __restrict__ dst, uint32_t dst_vec,
                           uint32_t lanes_per_row, int tie_mode) {
  const unsigned lane = threadIdx.x & 31;
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t groups = 32 / lanes_per_row;        // rows copied at once by one warp
  const uint32_t g = lane / lanes_per_row, c0 = lane % lanes_per_row;
  for (uint64_t r0 = warp * 32; r0 < n; r0 += nwarps * 32) {
    const uint64_t i = r0 + lane;
    const uint64_t key = i < n ? keys[i] : 0ull;
    uint64_t prev = __shfl_up_sync(0xffffffffu, key, 1), next = __shfl_down_sync(0xffffffffu, key, 1);
    if (lane == 0 && i > 0 && i < n) prev = keys[i - 1];
    if (lane == 31 && i + 1 < n) next = keys[i + 1];
    const bool tied = i < n && ((i > 0 && prev == key) || (i + 1 < n && next == key));
    const bool go = i < n && (tied == (tie_mode == 2));
    const unsigned any = __ballot_sync(0xffffffffu, go);
    if (any == 0) continue;
    const uint32_t s = go ? idx[i] : 0u;
    for (uint32_t b = 0; b < 32; b += groups) {
      const uint32_t ri = b + g;                        // row of this lane group within the warp's 32
      const uint32_t srow = __shfl_sync(0xffffffffu, s, ri);
      const uint64_t krow = __shfl_sync(0xffffffffu, key, ri);
      if (!((any >> ri) & 1u)) continue;
      const unsigned char* row = recv + (uint64_t)srow * recv_pitch;
      uint4* drow = reinterpret_cast<uint4*>(dst) + (r0 + ri) * dst_vec;
      for (uint32_t j = c0; j < dst_vec; j += lanes_per_row) drow[j] = prefixed_vector(row, j, shift, copy_bytes, used, krow, prefix_be);
    }
  }
}

// Output positions grouped by the chunk their row arrived in. starts[r * chunks + c] = first receive slot of chunk
// c of source r (ascending over the whole array). The order inside a chunk's list does not matter.
constexpr int SP_MAX_STARTS = 128;
struct ChunkStarts {
  uint32_t start[SP_MAX_STARTS];
  uint32_t list_off[MAX_PEERS];   // where every chunk's list begins in positions_out
  int total, chunks;
};

// Every warp owns a contiguous range of output positions; pass 1 counts its positions per chunk, a scan turns the
// counts into list offsets, pass 2 writes the positions - so every list comes out in ascending order, which keeps
// the later reads of idx[] and sorted_keys[] through a list close to sequential. No global atomics.
__device__ __forceinline__ int chunk_of_slot(const uint32_t* s_start, int total, int chunks, uint32_t s) {
  int lo = 0, hi = total;                     // number of starts <= s, by bisection
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (s_start[mid] <= s) lo = mid + 1; else hi = mid;
  }
  return lo > 0 ? (lo - 1) % chunks : 0;
}

template <bool WRITE>
__global__ void __launch_bounds__(256)
split_positions_kernel(const uint32_t* __restrict__ idx, uint64_t n, uint64_t rows_per_warp, const ChunkStarts cs,
                       uint32_t* __restrict__ warp_counts /*[warps][MAX_PEERS]: counts (pass 1) / exclusive offsets (pass 2)*/,
                       uint32_t* __restrict__ positions_out) {
  __shared__ uint32_t s_start[SP_MAX_STARTS];
  for (int i = threadIdx.x; i < SP_MAX_STARTS; i += blockDim.x) s_start[i] = i < cs.total ? cs.start[i] : 0xFFFFFFFFu;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, lt = lanemask_lt();
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t lo = warp * rows_per_warp, hi = min(n, lo + rows_per_warp);
  // lane c keeps chunk c's running count (pass 1) or next list position (pass 2)
  uint32_t mine = 0;
  if (WRITE && lane < (unsigned)cs.chunks) mine = cs.list_off[lane] + warp_counts[warp * MAX_PEERS + lane];
  for (uint64_t r0 = lo; r0 < hi; r0 += 32) {
    const uint64_t i = r0 + lane;
    const int chunk = i < hi ? chunk_of_slot(s_start, cs.total, cs.chunks, idx[i]) : -1;
    for (int c = 0; c < cs.chunks; ++c) {
      const unsigned m = __ballot_sync(0xffffffffu, chunk == c);
      if (m == 0) continue;
      if (WRITE) {
        const uint32_t base = __shfl_sync(0xffffffffu, mine, c);
        if (chunk == c) positions_out[base + __popc(m & lt)] = (uint32_t)i;
      }
      if ((int)lane == c) mine += (uint32_t)__popc(m);
    }
  }
  if (!WRITE && lane < (unsigned)MAX_PEERS) warp_counts[warp * MAX_PEERS + lane] = lane < (unsigned)cs.chunks ? mine : 0u;
}

// exclusive scan over the warps of every chunk's counts (one block per chunk)
__global__ void __launch_bounds__(1024)
split_scan_kernel(uint32_t* __restrict__ warp_counts, uint32_t warps) {
  __shared__ uint32_t s_warp[32];
  const int c = blockIdx.x;
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t per = (warps + 1023) / 1024;
  const uint32_t lo = min(warps, threadIdx.x * per), hi = min(warps, lo + per);
  uint32_t sum = 0;
  for (uint32_t t = lo; t < hi; ++t) sum += warp_counts[(size_t)t * MAX_PEERS + c];
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
    const uint32_t v = warp_counts[(size_t)t * MAX_PEERS + c];
    warp_counts[(size_t)t * MAX_PEERS + c] = run;
    run += v;
  }
}

}  // namespace b200x

extern "C" int b200x_scatter_records_by_target(const void* src, size_t src_pitch, size_t copy_bytes, const uint32_t* target,
                                               size_t n, void* const* peer_base_host,
                                               const unsigned long long* peer_row_off_host, size_t dst_pitch, int world,
                                               void* stream_) {
  B200X_ON_DEVICE_OF(src);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world <= 0 || world > MAX_PEERS) {
    set_error("b200x_scatter_records_by_target: world=%d out of range [1, %d]", world, MAX_PEERS);
    return B200X_ERR_ARG;
  }
  if (n == 0 || copy_bytes == 0) return B200X_OK;
  if (src == nullptr || target == nullptr || peer_base_host == nullptr || peer_row_off_host == nullptr) {
    set_error("b200x_scatter_records_by_target: NULL pointer");
    return B200X_ERR_ARG;
  }
  if ((copy_bytes & 15) != 0 || copy_bytes > 1024 || copy_bytes > src_pitch || copy_bytes > dst_pitch || (dst_pitch & 15) != 0 ||
      (src_pitch & 7) != 0 || ((uintptr_t)src & 7) != 0) {
    set_error("b200x_scatter_records_by_target: copy_bytes=%zu must be a multiple of 16, <= 1024 and <= both pitches (%zu, %zu); "
              "dst_pitch a multiple of 16; src and src_pitch multiples of 8", copy_bytes, src_pitch, dst_pitch);
    return B200X_ERR_ALIGN;
  }
  PeerPlan p;
  p.world = world;
  for (int d = 0; d < MAX_PEERS; ++d) {
    p.base[d] = d < world ? (unsigned char*)peer_base_host[d] : nullptr;
    p.row_off[d] = d < world ? peer_row_off_host[d] : 0;
    if (d < world && ((uintptr_t)p.base[d] & 15) != 0) {
      set_error("b200x_scatter_records_by_target: receive buffer %d is not 16-byte aligned", d);
      return B200X_ERR_ALIGN;
    }
  }
  for (int d = 0; d <= MAX_PEERS; ++d) p.send_off[d] = 0;
  uint32_t tile_rows = (uint32_t)(ST_DATA_BYTES / copy_bytes) / 32 * 32;
  if (tile_rows > (uint32_t)ST_MAX_TILE) tile_rows = ST_MAX_TILE;
  const size_t smem = ST_DATA_BYTES + ST_MAX_TILE * sizeof(unsigned long long) + ST_MAX_TILE * sizeof(uint16_t);
  uint64_t gtiles = (n + tile_rows - 1) / tile_rows;
  if (g_scatter_blocks_per_sm > 0 && gtiles > (uint64_t)sm_count() * g_scatter_blocks_per_sm)
    gtiles = (uint64_t)sm_count() * g_scatter_blocks_per_sm;
  const bool wide = (((uintptr_t)src | src_pitch) & 15) == 0;
  const bool bulk = dst_pitch == copy_bytes && g_scatter_bulk != 0;
  const unsigned char* s8 = (const unsigned char*)src;
  const uint32_t cb = (uint32_t)copy_bytes, g = (uint32_t)gtiles;
  prof_begin(stream, PROF_SCATTER);
  if (wide && bulk)
    scatter_records_staged_kernel<uint4, true><<<g, ST_THREADS, smem, stream>>>(s8, src_pitch, cb, target, p, dst_pitch, n, tile_rows);
  else if (wide)
    scatter_records_staged_kernel<uint4, false><<<g, ST_THREADS, smem, stream>>>(s8, src_pitch, cb, target, p, dst_pitch, n, tile_rows);
  else if (bulk)
    scatter_records_staged_kernel<uint64_t, true><<<g, ST_THREADS, smem, stream>>>(s8, src_pitch, cb, target, p, dst_pitch, n, tile_rows);
  else
    scatter_records_staged_kernel<uint64_t, false><<<g, ST_THREADS, smem, stream>>>(s8, src_pitch, cb, target, p, dst_pitch, n, tile_rows);
  prof_end(stream);
  B200X_LAUNCH_CHECK("scatter_records_staged_kernel");
  return B200X_OK;
}

extern "C" int b200x_gather_records_prefixed(const void* recv, size_t recv_pitch, size_t copy_bytes, const uint32_t* idx,
                                             size_t n, const uint64_t* sorted_keys, int prefix_be, size_t shift,
                                             size_t used_bytes, void* dst, size_t dst_pitch, int tie_mode,
                                             const uint32_t* positions, size_t num_positions, void* stream_) {
  B200X_ON_DEVICE_OF(idx);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n == 0 || (positions != nullptr && num_positions == 0)) return B200X_OK;
  if (recv == nullptr || idx == nullptr || dst == nullptr || (shift != 0 && shift != 8) || (shift == 8 && sorted_keys == nullptr) ||
      (tie_mode != 0 && sorted_keys == nullptr) || tie_mode < 0 || tie_mode > 2 || (positions != nullptr && tie_mode != 0) ||
      num_positions > n) {
    set_error("b200x_gather_records_prefixed: bad arguments (shift 0 or 8; sorted_keys needed for shift 8 and the tie modes; "
              "a position list excludes the tie modes)");
    return B200X_ERR_ARG;
  }
  if ((copy_bytes & 7) != 0 || copy_bytes > recv_pitch || (recv_pitch & 7) != 0 || ((uintptr_t)recv & 7) != 0 ||
      (dst_pitch & 15) != 0 || ((uintptr_t)dst & 15) != 0 || used_bytes > dst_pitch || dst_pitch > (1u << 16)) {
    set_error("b200x_gather_records_prefixed: copy_bytes / recv_pitch / recv must be multiples of 8, dst and dst_pitch of 16, "
              "used_bytes <= dst_pitch <= 65536");
    return B200X_ERR_ALIGN;
  }
  const uint32_t dst_vec = (uint32_t)(dst_pitch / 16);
  const uint64_t cap = (uint64_t)sm_count() * g_gather_blocks_per_sm;
  prof_begin(stream, PROF_GATHER);
  if (tie_mode == 0) {
    const uint64_t m = positions != nullptr ? num_positions : n;
    uint64_t blocks = (m * dst_vec + 255) / 256;
    if (blocks > cap) blocks = cap;
    const bool lanes = dst_vec <= 32 && (dst_vec & (dst_vec - 1)) == 0 && (copy_bytes & 15) == 0 && ((uintptr_t)recv & 15) == 0 &&
                       (recv_pitch & 15) == 0;
    if (lanes && g_gather_chains == 4)
      gather_records_list_kernel<true, 4><<<(uint32_t)blocks, 256, 0, stream>>>(
          (const unsigned char*)recv, recv_pitch, (uint32_t)copy_bytes, idx, positions, m, sorted_keys, prefix_be, (uint32_t)shift,
          (uint32_t)used_bytes, (unsigned char*)dst, dst_vec);
    else if (lanes)
      gather_records_list_kernel<true, 2><<<(uint32_t)blocks, 256, 0, stream>>>(
          (const unsigned char*)recv, recv_pitch, (uint32_t)copy_bytes, idx, positions, m, sorted_keys, prefix_be, (uint32_t)shift,
          (uint32_t)used_bytes, (unsigned char*)dst, dst_vec);
    else
      gather_records_list_kernel<false, 1><<<(uint32_t)blocks, 256, 0, stream>>>(
          (const unsigned char*)recv, recv_pitch, (uint32_t)copy_bytes, idx, positions, m, sorted_keys, prefix_be, (uint32_t)shift,
          (uint32_t)used_bytes, (unsigned char*)dst, dst_vec);
  } else {
    uint32_t lpr = 1;
    while (lpr < dst_vec && lpr < 32) lpr <<= 1;
    uint64_t blocks = ((uint64_t)n + 255) / 256;
    if (blocks > cap) blocks = cap;
    gather_records_ties_kernel<<<(uint32_t)blocks, 256, 0, stream>>>(
        (const unsigned char*)recv, recv_pitch, (uint32_t)copy_bytes, idx, n, sorted_keys, prefix_be, (uint32_t)shift,
        (uint32_t)used_bytes, (unsigned char*)dst, dst_vec, lpr, tie_mode);
  }
  prof_end(stream);
  B200X_LAUNCH_CHECK("gather_records_prefixed");
  return B200X_OK;
}

extern "C" int b200x_split_positions_by_chunk(const uint32_t* idx, size_t n, const uint32_t* chunk_starts_host, int num_sources,
                                              int num_chunks, const uint32_t* list_offsets_host, uint32_t* positions_out,
                                              void* temp, size_t* temp_bytes, void* stream_) {
  B200X_ON_DEVICE_OF(positions_out);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (temp_bytes == nullptr) {
    set_error("b200x_split_positions_by_chunk: temp_bytes is NULL");
    return B200X_ERR_ARG;
  }
  // every warp takes a contiguous range of positions (a multiple of 32 rows)
  uint64_t warps = (uint64_t)sm_count() * 8 * 8;
  uint64_t rows_per_warp = ((n + warps - 1) / warps + 31) / 32 * 32;
  if (rows_per_warp == 0) rows_per_warp = 32;
  warps = (n + rows_per_warp - 1) / rows_per_warp;
  if (warps == 0) warps = 1;
  warps = (warps + 7) / 8 * 8;   // whole blocks of 8 warps
  const size_t need = align_up(warps * MAX_PEERS * sizeof(uint32_t), 256);
  if (temp == nullptr) {
    *temp_bytes = need;
    return B200X_OK;
  }
  if (*temp_bytes < need || num_sources < 1 || num_chunks < 1 || num_chunks > MAX_PEERS ||
      num_sources * num_chunks > SP_MAX_STARTS || chunk_starts_host == nullptr || list_offsets_host == nullptr ||
      (n > 0 && (idx == nullptr || positions_out == nullptr))) {
    set_error("b200x_split_positions_by_chunk: bad arguments (sources x chunks <= %d, chunks <= %d) or temp too small",
              SP_MAX_STARTS, MAX_PEERS);
    return B200X_ERR_ARG;
  }
  ChunkStarts cs;
  cs.total = num_sources * num_chunks;
  cs.chunks = num_chunks;
  for (int i = 0; i < SP_MAX_STARTS; ++i) cs.start[i] = i < cs.total ? chunk_starts_host[i] : 0xFFFFFFFFu;
  for (int i = 1; i < cs.total; ++i)
    if (cs.start[i] < cs.start[i - 1]) {
      set_error("b200x_split_positions_by_chunk: chunk starts must ascend");
      return B200X_ERR_ARG;
    }
  for (int c = 0; c < MAX_PEERS; ++c) cs.list_off[c] = c < num_chunks ? list_offsets_host[c] : 0u;
  if (n == 0) return B200X_OK;
  uint32_t* warp_counts = (uint32_t*)temp;
  const uint32_t blocks = (uint32_t)(warps / 8);
  split_positions_kernel<false><<<blocks, 256, 0, stream>>>(idx, n, rows_per_warp, cs, warp_counts, positions_out);
  B200X_LAUNCH_CHECK("split_positions_kernel<count>");
  split_scan_kernel<<<num_chunks, 1024, 0, stream>>>(warp_counts, (uint32_t)warps);
  B200X_LAUNCH_CHECK("split_scan_kernel");
  split_positions_kernel<true><<<blocks, 256, 0, stream>>>(idx, n, rows_per_warp, cs, warp_counts, positions_out);
  B200X_LAUNCH_CHECK("split_positions_kernel<write>");
  return B200X_OK;
}
