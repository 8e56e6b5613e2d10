// Shared helpers for the b200x kernels (sm_100a only).
//
// Error convention of every extern "C" entry point (see include/b200x.h):
//   0            success
//   > 0          argument error (B200X_ERR_*)
//   < 0          -(cudaError_t) of the failing CUDA runtime call
// The message of the last failure on the calling host thread is kept in a
// thread-local buffer and returned by b200x_last_error_string().
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define B200X_OK 0
#define B200X_ERR_ARG 1        // null pointer / bad enum / bad size
#define B200X_ERR_TOO_LARGE 2  // n exceeds the limit of the entry point
#define B200X_ERR_TEMP 3       // temp buffer too small / misaligned
#define B200X_ERR_ALIGN 4      // pointer alignment requirement violated

namespace b200x {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define B200X_CUDA(call)                                         \
  do {                                                           \
    cudaError_t _e = (call);                                     \
    if (_e != cudaSuccess) return ::b200x::cuda_fail(_e, #call); \
  } while (0)

#define B200X_LAUNCH_CHECK(name)                                  \
  do {                                                            \
    ::b200x::count_launch();                                      \
    cudaError_t _e = cudaGetLastError();                          \
    if (_e != cudaSuccess) return ::b200x::cuda_fail(_e, name);   \
  } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Number of SMs of the current device (cached per device).
int sm_count();

// Every entry point runs on the GPU that owns its (first local) buffer, whatever device is current on the calling
// thread - the reference's executor calls the bulk function from helper threads, which start out on device 0 - and
// leaves the caller's current device as it found it. A NULL or host pointer leaves everything alone.
struct DeviceGuard {
  int prev = -1;
  bool switched = false;
  explicit DeviceGuard(const void* ptr);
  ~DeviceGuard();
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;
};
#define B200X_ON_DEVICE_OF(ptr) ::b200x::DeviceGuard _b200x_device_guard((const void*)(ptr))

// Every kernel launch of the library is counted (b200x_launch_count).
void count_launch();

// Optional CUDA-event bracket around kernel launches, recorded on the launching stream while
// b200x_profile_enable(1) is in effect (bench.py's live per-kernel timing).
enum ProfKernel : int { PROF_PASS = 0, PROF_GATHER = 1, PROF_FINISH = 2, PROF_JOINT = 3, PROF_PERMUTE = 4, PROF_SEGRED = 5,
                        PROF_HIST = 6, PROF_SCATTER = 7, PROF_COUNT = 8 };
void prof_begin(cudaStream_t s, int kernel);
void prof_end(cudaStream_t s);

// ---- key encoding: raw column value -> order-preserving 64-bit code -------
// Ascending order of codes == the order pyarrow.compute.sort_indices uses
// (reference call site: python/ray/data/_internal/arrow_ops/transform_pyarrow.py:17-21):
//   * integers in numeric order
//   * floats in numeric order with -0.0 == +0.0 (a tie), every NaN (either
//     sign) after all numbers
// Descending = bitwise complement of the code of every non-NaN value, NaN still
// maps to the largest code, so that a *stable ascending* radix sort of the codes
// gives Arrow's stable descending order (ties keep input order, NaN last).
enum KeyType : int {
  KEY_I64 = 0,
  KEY_U64 = 1,
  KEY_F64 = 2,
  KEY_I32 = 3,
  KEY_U32 = 4,
  KEY_F32 = 5,
  KEY_I16 = 6,
  KEY_U16 = 7,
  KEY_I8 = 8,
  KEY_U8 = 9,
  KEY_TYPE_COUNT = 10
};

__host__ __device__ __forceinline__ int key_type_bytes(int kt) {
  switch (kt) {
    case KEY_I64: case KEY_U64: case KEY_F64: return 8;
    case KEY_I32: case KEY_U32: case KEY_F32: return 4;
    case KEY_I16: case KEY_U16: return 2;
    default: return 1;
  }
}

// Significant bits of the code for a key type (upper bits are constant).
__host__ __device__ __forceinline__ int key_type_bits(int kt) { return key_type_bytes(kt) * 8; }

__device__ __forceinline__ uint64_t encode_f64_bits(uint64_t b, bool desc) {
  const uint64_t abs = b & 0x7FFFFFFFFFFFFFFFull;
  if (abs > 0x7FF0000000000000ull) return ~0ull;           // NaN -> last, both directions
  if (abs == 0) b = 0;                                     // -0.0 ties with +0.0
  uint64_t c = (b >> 63) ? ~b : (b | 0x8000000000000000ull);
  return desc ? ~c : c;
}

__device__ __forceinline__ uint64_t encode_f32_bits(uint32_t b, bool desc) {
  const uint32_t abs = b & 0x7FFFFFFFu;
  if (abs > 0x7F800000u) return ~0ull;                     // NaN
  if (abs == 0) b = 0;
  uint32_t c = (b >> 31) ? ~b : (b | 0x80000000u);
  // 32 significant bits; descending complements only those bits so that the
  // upper 32 stay constant (zero) and their radix passes are skipped.
  return desc ? (uint64_t)(uint32_t)(~c) : (uint64_t)c;
}

// Raw bits of element i (zero-extended) and their code, separately: kernels load a batch of raw
// bits first (one plain load per element, so the loads overlap) and encode afterwards.
__device__ __forceinline__ uint64_t load_raw_bits(const void* __restrict__ col, size_t i, int width) {
  switch (width) {
    case 8: return ((const uint64_t*)col)[i];
    case 4: return ((const uint32_t*)col)[i];
    case 2: return ((const uint16_t*)col)[i];
    default: return ((const uint8_t*)col)[i];
  }
}

__device__ __forceinline__ uint64_t encode_bits(uint64_t b, int kt, bool desc) {
  switch (kt) {
    case KEY_I64: { const uint64_t c = b ^ 0x8000000000000000ull; return desc ? ~c : c; }
    case KEY_U64: return desc ? ~b : b;
    case KEY_F64: return encode_f64_bits(b, desc);
    case KEY_I32: { const uint32_t c = (uint32_t)b ^ 0x80000000u; return desc ? (uint32_t)~c : c; }
    case KEY_U32: { const uint32_t c = (uint32_t)b; return desc ? (uint32_t)~c : c; }
    case KEY_F32: return encode_f32_bits((uint32_t)b, desc);
    case KEY_I16: { const uint16_t c = (uint16_t)b ^ 0x8000u; return desc ? (uint16_t)~c : c; }
    case KEY_U16: { const uint16_t c = (uint16_t)b; return desc ? (uint16_t)~c : c; }
    case KEY_I8: { const uint8_t c = (uint8_t)b ^ 0x80u; return desc ? (uint8_t)~c : c; }
    default: { const uint8_t c = (uint8_t)b; return desc ? (uint8_t)~c : c; }
  }
}

// Load element i of a raw key column and encode it.
__device__ __forceinline__ uint64_t load_encode(const void* __restrict__ col, size_t i, int kt, bool desc) {
  switch (kt) {
    case KEY_I64: {
      uint64_t c = ((const uint64_t*)col)[i] ^ 0x8000000000000000ull;
      return desc ? ~c : c;
    }
    case KEY_U64: {
      uint64_t c = ((const uint64_t*)col)[i];
      return desc ? ~c : c;
    }
    case KEY_F64:
      return encode_f64_bits(((const uint64_t*)col)[i], desc);
    case KEY_I32: {
      uint32_t c = ((const uint32_t*)col)[i] ^ 0x80000000u;
      return desc ? (uint32_t)~c : c;
    }
    case KEY_U32: {
      uint32_t c = ((const uint32_t*)col)[i];
      return desc ? (uint32_t)~c : c;
    }
    case KEY_F32:
      return encode_f32_bits(((const uint32_t*)col)[i], desc);
    case KEY_I16: {
      uint16_t c = ((const uint16_t*)col)[i] ^ 0x8000u;
      return desc ? (uint16_t)~c : c;
    }
    case KEY_U16: {
      uint16_t c = ((const uint16_t*)col)[i];
      return desc ? (uint16_t)~c : c;
    }
    case KEY_I8: {
      uint8_t c = ((const uint8_t*)col)[i] ^ 0x80u;
      return desc ? (uint8_t)~c : c;
    }
    default: {
      uint8_t c = ((const uint8_t*)col)[i];
      return desc ? (uint8_t)~c : c;
    }
  }
}

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Streaming (read-once) loads / (write-once) stores: keep them out of L1.
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

}  // namespace b200x
