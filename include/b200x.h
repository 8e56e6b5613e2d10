/* b200x — C ABI of the B200-native all-to-all hot path (sort / random_shuffle /
 * sort-based groupby-aggregate) behind the ray.data block interface.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / Arrow types.
 * The reference has no native interface for this path (it is Python over
 * pyarrow/numpy); each entry point below cites the reference call site whose
 * arithmetic it replaces, relative to /root/reference/python/ray/data/.
 *
 * Conventions shared by every entry point
 *   - Every pointer is a DEVICE pointer unless its name ends in _host.
 *   - The caller owns every buffer. Entry points that need scratch take
 *     `void* temp, size_t* temp_bytes`: call once with temp == NULL to get the
 *     required size in *temp_bytes, then again with a 256-byte aligned buffer.
 *   - `stream` is a cudaStream_t (NULL = legacy default stream). Calls are
 *     asynchronous: they enqueue work and return without synchronising.
 *   - Every entry point that launches work makes the GPU that owns its first device buffer current for the
 *     duration of the call and restores the caller's current device before it returns, so it may be called from
 *     any thread (the reference's executor runs the bulk function on helper threads). `stream` must be a stream of
 *     that GPU (or NULL).
 *   - Return value: 0 = OK; > 0 = argument error (B200X_ERR_*);
 *     < 0 = -(cudaError_t). b200x_last_error_string() gives the message of the
 *     last failure on the calling thread. Nothing throws across the ABI.
 */
#ifndef B200X_H
#define B200X_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200X_VERSION 100

/* key column element types (enum KeyType in csrc/common.cuh) */
#define B200X_KEY_I64 0
#define B200X_KEY_U64 1
#define B200X_KEY_F64 2
#define B200X_KEY_I32 3
#define B200X_KEY_U32 4
#define B200X_KEY_F32 5
#define B200X_KEY_I16 6
#define B200X_KEY_U16 7
#define B200X_KEY_I8 8
#define B200X_KEY_U8 9

/* value column element types for the aggregation kernels */
#define B200X_VAL_I64 0
#define B200X_VAL_F64 1

int b200x_version(void);
const char* b200x_last_error_string(void);
int b200x_device_count(void);

/* Measurement hooks (bench.py): number of kernels the library has launched in this process,
 * and CUDA-event timing of the main kernels on their launching stream: enable(1) clears the
 * record and starts recording one event pair per launch on the calling thread; read(kernel, ...)
 * synchronises and returns the summed duration and the launch count of one kernel class
 * (0 radix pass, 1 row gather, 2 run finisher, 3 digit histogram, 4 row permutation, 5 segmented reduce,
 * 6 LSD histogram, 7 peer scatter). Launches that
 * return at once because the device chose the other sort path are not counted. */
unsigned long long b200x_launch_count(void);
int b200x_profile_enable(int on);
int b200x_profile_read(int kernel, double* total_ms, int* launches);

/* Stable argsort of one key column. Radix sort on order-preserving 64-bit codes with 8-bit digits
 * and decoupled look-back ("onesweep" passes). The device inspects the keys and picks
 *   - LSD: one pass per varying byte (constant bytes are skipped), or
 *   - HYBRID: two passes group the pairs into <= 65536 runs by a monotone function of the key - the top
 *     16 bits of (code - smallest code), or, for skewed distributions, a sampled bucket table that
 *     equalises the runs - then each run is finished in shared memory (3 trips through HBM instead of
 *     up to 8). Chosen when no run exceeds 8192 pairs; duplicate-heavy columns fall back to LSD.
 *   perm_out[i] = row that lands at sorted position i
 *   codes_out   = (optional, may be NULL) the sorted order-preserving 64-bit key codes;
 *                 ascending for both sort directions (see b200x_encode_keys)
 *   keys_out    = (optional, integer key types only) the sorted key column itself, decoded from
 *                 the codes by the last pass (saves a random gather of the key column)
 *   perm_in     = (optional) rows are visited in the order perm_in[0..n): the key of
 *                 position i is keys[perm_in[i]]. Chaining calls from the last key
 *                 column to the first gives a stable multi-column sort.
 *   n < 2^30. begin_bit must be 0; end_bit (0 = all) bounds the significant code bits.
 * Replaces pyarrow.compute.sort_indices in transform_pyarrow.sort / concat_and_sort:
 *   _internal/arrow_ops/transform_pyarrow.py:17-21, :269-276
 *   (via ArrowBlockAccessor.sort_and_partition / merge_sorted_blocks, _internal/arrow_block.py:449-464, :554-567) */
int b200x_radix_argsort(const void* keys, int key_type, int descending, const uint32_t* perm_in, uint32_t* perm_out,
                        uint64_t* codes_out, void* keys_out, size_t n, int begin_bit, int end_bit, void* temp,
                        size_t* temp_bytes, void* stream);

/* Tuning knobs (process-wide). key 0: fewest rows for which b200x_radix_argsort considers its
 * HYBRID mode (two global passes on the top 16 varying bits, then every run is finished in shared
 * memory); default 2^22. key 1: most blocks per SM the staged peer scatter launches (0 = one block per tile);
 * a small value leaves room for work running next to the transfer on another stream. key 2: loads in flight per
 * thread of b200x_gather_records_prefixed (2 or 4); key 3: most blocks per SM it launches (default 16; fewer leave
 * room for a transfer running next to it); key 4: b200x_scatter_records_by_target writes an owner's rows of a tile
 * with one bulk copy from shared memory (1, default) or with 16-byte stores (0). Results never depend on any of them. */
int b200x_set_tuning(int key, long long value);

/* b200x_radix_argsort for callers that cannot allocate device memory. (No counterpart in the
 * reference: ray.data always allocates through Arrow / the object store.)
 * Who needs it: code that runs where cudaMalloc is off limits or simply not available - an
 * operator called from a process whose caching allocator already owns the GPU's memory, a step
 * of a latency-critical stream loop that must not take cudaMalloc's implicit synchronisation,
 * or a foreign-function shim with no device allocator of its own. Such a caller still passes its
 * inputs and outputs; only the sort's scratch comes from a fixed array inside the library
 * (about 27 MiB that every GPU which loads libb200x.so keeps until the process exits).
 *   - n is limited to 1,048,576 rows (2^20); larger n returns B200X_ERR_TOO_LARGE with a message.
 *   - Arguments are those of b200x_radix_argsort minus begin_bit/end_bit/temp/temp_bytes; results are identical.
 *   - The first call on each device looks up the array's device address and sets a kernel
 *     attribute; make that first call outside any stream capture, with the device current.
 *   - No cudaMalloc / cudaFree / allocating library routine is called (leave b200x_profile_enable
 *     off on such a path: timing events are created on first use).
 *   - All calls on one GPU share the one array: calls whose kernels may overlap on the GPU (different
 *     streams, different threads) race and corrupt each other; serialise them on one stream. */
int b200x_radix_argsort_static(const void* keys, int key_type, int descending, const uint32_t* perm_in,
                               uint32_t* perm_out, uint64_t* codes_out, void* keys_out, size_t n, void* stream);

/* dst[i, :] = src[idx[i], :] for rows of row_bytes bytes (a fixed-width column or a
 * packed payload). Replaces Table.take: _internal/arrow_ops/transform_pyarrow.py:24-50,
 * _internal/arrow_block.py:346-355. */
int b200x_gather_rows(const void* src, size_t row_bytes, const uint32_t* idx, size_t n, void* dst, void* stream);

/* Record layout: a block may keep several fixed-width columns of a row side by side in ONE record per row,
 * records lying at a pitch (bytes from one row to the next) that is a multiple of 16, so that a row moves in one
 * pass of aligned 16-byte accesses whatever its columns' widths (gensort: 10-byte key + 90-byte payload in a
 * 128-byte record = one DRAM line). dst[i, 0:row_bytes) = src[idx[i], 0:row_bytes), rows src_pitch / dst_pitch
 * bytes apart; bases, pitches and row_bytes are multiples of 16, row_bytes <= either pitch. Same call sites as
 * b200x_gather_rows (Table.take moves all columns of a row: transform_pyarrow.py:24-50). */
int b200x_gather_rows_pitched(const void* src, size_t src_pitch, const uint32_t* idx, size_t n, void* dst,
                              size_t dst_pitch, size_t row_bytes, void* stream);

/* dst[r, 0:row_bytes) = src[r, 0:row_bytes) for n rows lying src_pitch / dst_pitch bytes apart (any alignment; the
 * widest access that divides all of them is used): columns <-> records at block ingest / egress, the step
 * BlockAccessor.for_block / to_arrow perform when a block changes format (block.py:380-407). */
int b200x_copy_rows_2d(const void* src, size_t src_pitch, void* dst, size_t dst_pitch, size_t row_bytes, size_t n,
                       void* stream);

/* Same gather for ncols 8-byte columns sharing one index vector; srcs_host/dsts_host are HOST
 * arrays of device pointers. */
int b200x_gather_columns8(const void* const* srcs_host, void* const* dsts_host, int ncols, const uint32_t* idx, size_t n,
                          void* stream);

/* out[i] = start + i */
int b200x_iota_u32(uint32_t* out, size_t n, uint32_t start, void* stream);

/* codes_out[i] = order-preserving 64-bit code of keys[i]: integers in numeric order,
 * floats with -0.0 == +0.0 and NaN last; descending complements every non-NaN code.
 * This is the comparison pyarrow's sort applies (call sites as for b200x_radix_argsort). */
int b200x_encode_keys(const void* keys, int key_type, int descending, size_t n, uint64_t* codes_out, void* stream);

/* *count_out = number of positions i >= 1 with keys[i] == keys[i-1] in a sorted key column (compared as
 * codes: -0.0 == +0.0). Zero = all keys distinct: a multi-column sort (transform_pyarrow.py:17-21 with several
 * sort keys) then needs only the sort by its first column. No counterpart in the reference. */
int b200x_count_adjacent_equal(const void* keys, int key_type, size_t n, uint64_t* count_out, void* stream);

/* Multi-column sort whose first column is (almost) distinct. After a stable sort by the first key column alone
 * (keys_sorted = that column in sorted order, or its codes with key_type U64; perm = the permutation, updated in
 * place) every run of equal first keys is ordered by the remaining num_cols (1..3) key columns with a stable
 * insertion sort of the run's permutation entries. The *_host arguments are HOST arrays over those columns:
 * cols_host[c] = device address of row 0's value, types_host[c] = its key type, strides_host[c] = bytes from one
 * row's value to the next (NULL or 0 = dense), be_bytes_host[c] > 0 = the value is that many big-endian bytes (the
 * next piece of a byte-string key, read where it lies inside the row's record; types_host[c] is then ignored;
 * NULL = none). Values are compared as codes in the given direction. stats_out[0] (device) = rows that lie in such
 * runs, stats_out[1] = longest run; runs longer than 64 rows are left untouched, so a caller that reads
 * stats_out[1] > 64 must sort by all columns instead. Gives the order of pyarrow.compute.sort_indices with several
 * sort keys (transform_pyarrow.py:17-21) for one sort pass instead of one per column. */
int b200x_fix_ties(const void* keys_sorted, int key_type, size_t n, uint32_t* perm, const void* const* cols_host,
                   const int* types_host, const unsigned long long* strides_host, const int* be_bytes_host, int num_cols,
                   int descending, unsigned long long* stats_out, void* stream);

/* out[i] = bytes [offset, offset + width) of row i of a fixed-width byte-string column (row_bytes per row)
 * as a big-endian unsigned integer, left-aligned in an out_bytes-wide (2, 4 or 8) unsigned integer and
 * zero-padded; 1 <= width <= out_bytes. Sorting a w-byte key = chained b200x_radix_argsort calls (key types
 * U64 / U32 / U16) over its chunks, last chunk first: the lexicographic order pyarrow.compute.sort_indices
 * gives binary / fixed_size_binary keys (transform_pyarrow.py:17-21; gensort-style 10-byte keys). */
int b200x_pack_key_bytes(const void* col, size_t row_bytes, size_t offset, int width, size_t n, void* out,
                         int out_bytes, void* stream);

/* All ceil(width / 8) chunks of b200x_pack_key_bytes in one pass over a key column whose rows lie row_pitch bytes
 * apart, row_pitch a multiple of 16 (a key held inside a record, see b200x_gather_rows_pitched): each row's key
 * is read with the aligned 16-byte loads that cover it. outs_host / out_bytes_host are HOST arrays: chunk k goes
 * to outs_host[k] as an unsigned integer of out_bytes_host[k] bytes (8 for full chunks; 2, 4 or 8 for the last
 * one, at least its byte count). width <= 32. Same reference call site as b200x_pack_key_bytes. */
int b200x_pack_key_chunks(const void* col, size_t row_pitch, int width, size_t n, void* const* outs_host,
                          const int* out_bytes_host, void* stream);

/* The inverse of b200x_pack_key_bytes: writes the first `width` big-endian bytes of chunk[i] to bytes
 * [offset, offset + width) of row i (the key column is rebuilt from its sorted chunks instead of being
 * carried through the exchange). No counterpart in the reference. */
int b200x_unpack_key_bytes(const void* chunk, int chunk_bytes, int width, void* col, size_t row_bytes, size_t offset,
                           size_t n, void* stream);

/* out_index[b] = lower bound (upper == 0) or upper bound (upper != 0) of bound_codes[b] in the
 * ascending array codes[0..n). Replaces find_partition_index / find_partitions:
 * _internal/util.py:719-758, :761-779 (numpy.searchsorted per boundary). */
int b200x_search_sorted_codes(const uint64_t* codes, size_t n, const uint64_t* bound_codes, int num_bounds, int upper,
                              uint64_t* out_index, void* stream);

/* The same search on a key column that is already SORTED in the given direction (codes are computed
 * on the fly), so cutting a sorted block at boundaries needs no code array. */
int b200x_search_sorted_keys(const void* keys_sorted, int key_type, int descending, size_t n,
                             const uint64_t* bound_codes, int num_bounds, int upper, uint64_t* out_index, void* stream);

/* out_index[i] (i < parts-1) = the cut that splits a SORTED key column into `parts` near-equal blocks
 * without separating equal keys: the first row whose key equals the key at row floor((i+1)*n/parts).
 * (The reference's rule that equal keys land in one partition: _internal/util.py:756-758.) */
int b200x_even_cuts(const void* keys_sorted, int key_type, int descending, size_t n, int parts, uint64_t* out_index,
                    void* stream);

/* Fused partition-gather + exchange over NVLink peer memory (one GPU per process, all on one box).
 * Rows are visited in send order i = 0..n-1 (row perm[i] of src, or row i if perm is NULL); rows
 * [send_off_host[d], send_off_host[d+1]) go to rank d and are stored directly into that rank's receive buffer
 * peer_base_host[d] (a peer-mapped device address, e.g. from torch symmetric memory), starting at row
 * peer_row_off_host[d] of it. The *_host arrays are HOST arrays of length world (+1 for send_off). No send
 * buffer exists; the caller orders ranks with a barrier before and after. world <= 16.
 * Replaces the object-store hop between map and reduce tasks
 * (_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:86-124,
 *  push_based_shuffle_task_scheduler.py:514-568). */
int b200x_scatter_rows_to_peers(const void* src, size_t row_bytes, const uint32_t* perm, size_t n,
                                const unsigned long long* send_off_host, void* const* peer_base_host,
                                const unsigned long long* peer_row_off_host, int world, void* stream);

/* Partition-first variant of the same exchange (no local sort before the transfer).
 * b200x_partition_targets: target_out[r] = owner << 28 | slot for every row r, where owner = number of
 * splitter codes <= code(keys[r]) (so equal keys share an owner, _internal/util.py:756-758) and slot = the row's
 * position among the rows this rank sends to that owner, in row order (a stable split: ties keep their
 * order through the exchange). counts_out[16] (device) = rows per owner. n < 2^28, <= 15 splitters.
 * After the call `temp` holds, as uint32 [ceil(n / 2048)][16], the slots handed out per owner before every tile of
 * 2048 rows (row t = rows per owner among the first 2048 * t rows): what a sender needs to say which slots a
 * chunk of its rows fills.
 * b200x_scatter_rows_by_target: row r of src (read in order) is stored at row peer_row_off_host[owner] + slot
 * of peer_base_host[owner]. */
int b200x_partition_targets(const void* keys, int key_type, int descending, size_t n, const uint64_t* splitter_codes,
                            int num_splitters, uint32_t* target_out, unsigned long long* counts_out, void* temp,
                            size_t* temp_bytes, void* stream);
int b200x_scatter_rows_by_target(const void* src, size_t row_bytes, const uint32_t* target, size_t n,
                                 void* const* peer_base_host, const unsigned long long* peer_row_off_host, int world,
                                 void* stream);

/* The same exchange for rows kept as records (b200x_gather_rows_pitched). b200x_scatter_records_by_target: of
 * every source record (src_pitch bytes apart; src may point 8 bytes into the records) copy_bytes bytes (a multiple
 * of 16) are stored at the start of row peer_row_off_host[owner] + slot of peer_base_host[owner], rows dst_pitch
 * apart. A record thus travels without its padding - and without its first 8 bytes when those are the sort key that
 * went ahead with the key column: a gensort record in a 128-byte slot crosses NVLink as 96 bytes = three whole
 * sectors of one line. Passing src + r0 * src_pitch, target + r0 and a row count sends a chunk of the rows.
 * b200x_gather_records_prefixed: the receiver's last step: dst record i (dst_pitch apart) = [8-byte key prefix |
 * received row idx[i]] when shift == 8 - the prefix taken from sorted_keys[i] (big-endian bytes if prefix_be, else
 * the integer as it is), i.e. the elided key bytes are put back from the sorted key column while the rows are
 * gathered into sorted order - or just the received row when shift == 0; bytes from used_bytes on are zero.
 * positions (device, num_positions entries; NULL = all n) lists the output positions to write: the rows of one
 * chunk that has landed; tie_mode 1 skips, 2 selects the positions whose key equals a neighbour's (their order is
 * final only after b200x_fix_ties), 0 takes all; a position list goes with tie_mode 0 only.
 * b200x_split_positions_by_chunk builds those lists: chunk_starts_host[r * num_chunks + c] = first receive slot of
 * chunk c of source rank r (ascending; HOST array); output position i (whose row sits in slot idx[i]) is appended
 * to the list of its chunk, the lists lying in positions_out from list_offsets_host[c] on (sizes are known from
 * the count tables); every list comes out in ascending order. Two-phase temp query as elsewhere.
 * Replaces the same object-store hop as b200x_scatter_rows_to_peers, and the reducer's concat + take
 * (_internal/arrow_block.py:554-567, transform_pyarrow.py:269-276). */
int b200x_scatter_records_by_target(const void* src, size_t src_pitch, size_t copy_bytes, const uint32_t* target, size_t n,
                                    void* const* peer_base_host, const unsigned long long* peer_row_off_host,
                                    size_t dst_pitch, int world, void* stream);
int b200x_gather_records_prefixed(const void* recv, size_t recv_pitch, size_t copy_bytes, const uint32_t* idx, size_t n,
                                  const uint64_t* sorted_keys, int prefix_be, size_t shift, size_t used_bytes, void* dst,
                                  size_t dst_pitch, int tie_mode, const uint32_t* positions, size_t num_positions,
                                  void* stream);
int b200x_split_positions_by_chunk(const uint32_t* idx, size_t n, const uint32_t* chunk_starts_host, int num_sources,
                                   int num_chunks, const uint32_t* list_offsets_host, uint32_t* positions_out, void* temp,
                                   size_t* temp_bytes, void* stream);

/* Seeded pseudo-random bijection pi on [0, n) (6 alternating Feistel rounds over the 2^ceil(log2 n)
 * domain, cycle-walked into range). Deterministic in (seed, n).
 *   inverse != 0 evaluates pi^-1.
 * b200x_permute_rows:   dst[j, :] = src[pi(j), :]      (permutation + take fused, no index array)
 * b200x_perm_indices:   out[t] = pi(start + t), t in [0, count)
 * b200x_perm_indices_host: same values computed on the host (out_host is a HOST pointer)
 * Replaces numpy.random.RandomState(seed).permutation(n) + take in
 * ArrowBlockAccessor.random_shuffle (_internal/arrow_block.py:249-255), reached from
 * ShuffleTaskSpec.map / reduce (_internal/planner/exchange/shuffle_task_spec.py:81-84, :119-123).
 * The MT19937 stream itself is not reproduced: the contract is an exact multiset
 * permutation that is deterministic per seed. */
int b200x_permute_rows(const void* src, size_t row_bytes, size_t n, uint64_t seed, int inverse, void* dst, void* stream);
int b200x_permute_rows_pitched(const void* src, size_t src_pitch, size_t n, uint64_t seed, int inverse, void* dst,
                               size_t dst_pitch, size_t row_bytes, void* stream); /* records at a pitch, see b200x_gather_rows_pitched */
int b200x_perm_indices(size_t n, uint64_t seed, int inverse, size_t start, size_t count, uint32_t* out, void* stream);

/* random_shuffle across the GPUs of one box as ONE scatter over peer memory: this rank holds rows [first_global,
 * first_global + n_local) of a dataset of total_rows rows (rows src_pitch bytes apart, row_bytes bytes each); row G is
 * stored at position q = pi^-1(G) of the shuffled dataset, which lies on rank q / rows_per_rank at row
 * q % rows_per_rank of peer_base_host[rank] (rows dst_pitch apart; HOST array of peer-mapped device addresses). The
 * concatenation of the ranks' buffers equals b200x_permute_rows of the concatenated dataset with the same seed. No
 * counts are exchanged (every rank gets rows_per_rank rows, the last one the rest), no send buffer, no second local
 * shuffle. Replaces ShuffleTaskSpec.map + the object-store hop + ShuffleTaskSpec.reduce
 * (_internal/planner/exchange/shuffle_task_spec.py:81-132). */
int b200x_permute_scatter_to_peers(const void* src, size_t src_pitch, size_t row_bytes, size_t n_local, size_t first_global,
                                   size_t total_rows, uint64_t seed, size_t rows_per_rank, void* const* peer_base_host,
                                   size_t dst_pitch, int world, void* stream);
int b200x_perm_indices_host(size_t n, uint64_t seed, int inverse, size_t start, size_t count, uint32_t* out_host);

/* Group boundaries of an ascending code array: group_starts[g] = first position of group g,
 * *num_groups_out (device) = number of distinct codes. group_starts needs room for n entries.
 * Replaces the per-row key comparison loop of ArrowBlockAccessor.combine
 * (_internal/arrow_block.py:490-515) and GroupedData.map_groups' boundary finder
 * (grouped_data.py:181-200). */
int b200x_group_boundaries(const uint64_t* codes, size_t n, uint32_t* group_starts, uint32_t* num_groups_out, void* temp,
                           size_t* temp_bytes, void* stream);

/* The same for a multi-column key: codes_host is a HOST array of num_cols (<= 4) device code arrays of the
 * rows in sorted order; a group starts where any column changes (groupby(["a", "b"]), dataset.py:1788-1836). */
int b200x_group_boundaries_multi(const uint64_t* const* codes_host, int num_cols, size_t n, uint32_t* group_starts,
                                 uint32_t* num_groups_out, void* temp, size_t* temp_bytes, void* stream);

/* Segmented reduction of one value column over the groups found above (group_starts and the
 * device scalar num_groups as written by b200x_group_boundaries). Row i of the sorted order
 * holds values[perm[i]] (perm may be NULL = already in sorted order). For every group g
 * writes, where the output pointer is not NULL:
 *   sum_out[g] (int64 wrap-around or float64), min_out[g], max_out[g]  (same type as values)
 * If center (float64 per group) is given, the reduced quantity is (v - center[g])^2 as float64
 * (sum of squared differences for Std) and all outputs are float64; weights (int64 per input
 * row, optional, only with center) multiplies each squared difference and addend (float64 per
 * input row, optional, only with center) is added to it: together they merge Std partials,
 * M2 = sum_p M2_p + count_p * (mean_p - mean)^2 (aggregate/_aggregate.py:264-283).
 * Deterministic for a given n (fixed reduction tree).
 * Replaces pyarrow.compute.sum/min_max on each group's slice and the accumulator merges:
 * _internal/arrow_block.py:370-447 (sum/min/max/mean/sum_of_squared_diffs_from_mean),
 * aggregate/_aggregate.py:105-319, _internal/arrow_block.py:569-680 (aggregate_combined_blocks). */
int b200x_segmented_reduce(const uint32_t* group_starts, const uint32_t* num_groups, const uint32_t* perm,
                           const void* values, int value_type, size_t n, const double* center, const int64_t* weights,
                           const double* addend, void* sum_out, void* min_out, void* max_out, void* temp,
                           size_t* temp_bytes, void* stream);

/* Dense-range aggregation: grouping by an integer key column whose order-preserving codes span a small range needs
 * no sort - the group of a row is code - min(code), and the non-empty entries of a table indexed by that number,
 * read in index order, are the groups in ascending key order (ids, categories, dates; BASELINE configs[2]).
 * Integer sums / counts / minima / maxima do not depend on the order of the additions, so the result equals the
 * sorted path's bit for bit; float columns stay on the sorted path (fixed reduction order).
 *   b200x_key_code_range: range_out[0] = smallest, range_out[1] = largest code of the column (device, 2 x uint64).
 *   b200x_dense_group_accumulate: for every row r: g = code(keys[r]) - base_code (< range): count_tab[g] += 1,
 *     sum_tab[g] += values[r] (int64 wrap-around), min_tab / max_tab[g] = min / max(.., values[r]); table pointers
 *     that are NULL are skipped, values (int64) may be NULL when only counts are wanted; init != 0 first sets the
 *     given tables to 0 / 0 / INT64_MAX / INT64_MIN. Heavy keys are combined per tile in shared memory first.
 *   b200x_dense_group_compact: the entries with count_tab[g] != 0 in index order: keys_out[j] = the key (decoded
 *     to key_type), count_out[j] (may be NULL), tabs_out_host[t][j] = tabs_host[t][g] for num_tabs (<= 8) int64
 *     tables (HOST arrays of device pointers); *num_groups_out (device) = how many. Outputs need room for
 *     min(range, n) entries. Two-phase temp query as elsewhere.
 * Replaces the same arithmetic as b200x_segmented_reduce + b200x_group_counts + the key sort in front of them
 * (_internal/arrow_block.py:466-548, :569-680; aggregate/_aggregate.py:90-241). */
int b200x_key_code_range(const void* keys, int key_type, size_t n, uint64_t* range_out, void* stream);
int b200x_dense_group_accumulate(const void* keys, int key_type, size_t n, uint64_t base_code, size_t range,
                                 const int64_t* values, int init, int64_t* count_tab, int64_t* sum_tab, int64_t* min_tab,
                                 int64_t* max_tab, void* stream);
int b200x_dense_group_compact(uint64_t base_code, size_t range, int key_type, const int64_t* count_tab, int num_tabs,
                              const int64_t* const* tabs_host, void* keys_out, int64_t* count_out,
                              int64_t* const* tabs_out_host, uint32_t* num_groups_out, void* temp, size_t* temp_bytes,
                              void* stream);

/* counts_out[g] = rows in group g (int64), g < *num_groups; max_groups bounds the launch.
 * Replaces Count's accumulate_block (num_rows of the group slice): aggregate/_aggregate.py:90-102. */
int b200x_group_counts(const uint32_t* group_starts, const uint32_t* num_groups, size_t max_groups, size_t n,
                       int64_t* counts_out, void* stream);

/* Nulls in value columns (the reference's null policy, _internal/null_aggregate.py:11-17, applied by the wrappers of
 * :28-250 and by pyarrow's skip_nulls in _internal/arrow_block.py:387-425). A nullable column is its values plus one
 * validity byte per row (non-zero = present).
 * b200x_mask_values: out_values[i] = valid[i] ? values[i] : the identity of the reduction the column enters next
 *   (fill_mode 0: 0 for sums; 1: the largest int64 / +inf for minima; 2: the smallest / -inf for maxima), and
 *   out_valid64[i] = valid[i] ? 1 : 0 as int64 (reduced per group like any other column: the number of non-null rows,
 *   which is also Mean's and Std's denominator). Either output may be NULL.
 * b200x_group_validity: out[g] = 1 if group g's result exists: ignore_nulls != 0: nvalid[g] > 0 (nulls are skipped, an
 *   all-null group has no result); ignore_nulls == 0: nvalid[g] == rows[g] > 0 (one null makes the result null). */
int b200x_mask_values(const void* values, int value_type, const uint8_t* valid, size_t n, int fill_mode, void* out_values,
                      int64_t* out_valid64, void* stream);
int b200x_group_validity(const int64_t* nvalid, const int64_t* rows, size_t num_groups, int ignore_nulls, uint8_t* out,
                         void* stream);

/* mode 0: out[g] = sum[g] / counts[g]                         (Mean.finalize, aggregate/_aggregate.py:238)
 * mode 1: out[g] = counts[g] < 2 ? 0 : sqrt(sum[g] / (counts[g] - ddof))   (Std.finalize, :296-302) */
int b200x_finalize_groups(const void* sum, int value_type, const int64_t* counts, size_t num_groups, int mode, int ddof,
                          double* out, void* stream);

/* out[g] = max(|min_in[g]|, |max_in[g]|), same type as the inputs (AbsMax, aggregate/_aggregate.py:323-357). */
int b200x_absmax_groups(const void* min_in, const void* max_in, int value_type, size_t num_groups, void* out, void* stream);

/* out[g] = q-quantile of group g's values by linear interpolation between the two nearest ranks,
 * rounded to 5 decimals when it falls between two rows, exactly as the reference's `percentile`
 * (Quantile, aggregate/_aggregate.py:360-431, finalize :406-417). `perm` must visit the rows ordered
 * by (group key, value); group_starts are positions in that order. 0 <= q <= 1. out is float64. */
int b200x_group_quantile(const uint32_t* group_starts, const uint32_t* num_groups, size_t max_groups,
                         const uint32_t* perm, const void* values, int value_type, size_t n, double q, double* out,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200X_H */
