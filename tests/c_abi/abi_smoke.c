/* A plain C caller of the b200x C ABI (no Python, no torch): what a cgo / JNI / FFI shim does.
 *
 * Sorts n random int64 keys with the two-phase scratch query of b200x_radix_argsort, gathers a
 * 64-byte payload through the permutation with b200x_gather_rows, and checks both on the host.
 * Build (tests/test_c_abi_gpu.py does this):
 *   nvcc -x c -o abi_smoke tests/c_abi/abi_smoke.c -Iinclude -L<pkg> -lb200x -Xlinker -rpath=<pkg>
 * Exit code 0 = everything checked out. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "b200x.h"

#define CHECK_CUDA(x)                                                                    \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) {                                                             \
      fprintf(stderr, "%s:%d CUDA error %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return 2;                                                                          \
    }                                                                                    \
  } while (0)

#define CHECK_B200X(x)                                                                   \
  do {                                                                                   \
    int rc_ = (x);                                                                       \
    if (rc_ != 0) {                                                                      \
      fprintf(stderr, "%s:%d b200x status %d: %s\n", __FILE__, __LINE__, rc_, b200x_last_error_string()); \
      return 3;                                                                          \
    }                                                                                    \
  } while (0)

static uint64_t splitmix(uint64_t* s) {
  uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

int main(int argc, char** argv) {
  const size_t n = argc > 1 ? (size_t)strtoull(argv[1], NULL, 10) : 5000000;
  const size_t row_bytes = 64;
  if (b200x_device_count() < 1) {
    fprintf(stderr, "no CUDA device: the library has no CPU path\n");
    return 4;
  }
  int64_t* h_keys = (int64_t*)malloc(n * sizeof(int64_t));
  unsigned char* h_pay = (unsigned char*)malloc(n * row_bytes);
  uint64_t seed = 42;
  for (size_t i = 0; i < n; ++i) {
    h_keys[i] = (int64_t)(splitmix(&seed) >> 1) - (int64_t)(1ull << 61);  /* negative and positive keys */
    if (i % 7 == 0 && i) h_keys[i] = h_keys[i - 1];                       /* ties: stability is checked */
    memset(h_pay + i * row_bytes, 0, row_bytes);
    memcpy(h_pay + i * row_bytes, &i, sizeof(i));                         /* payload carries its row number */
  }
  int64_t *d_keys, *d_sorted;
  uint32_t* d_perm;
  unsigned char *d_pay, *d_out, *d_temp;
  cudaStream_t stream;
  CHECK_CUDA(cudaStreamCreate(&stream));
  CHECK_CUDA(cudaMalloc((void**)&d_keys, n * sizeof(int64_t)));
  CHECK_CUDA(cudaMalloc((void**)&d_sorted, n * sizeof(int64_t)));
  CHECK_CUDA(cudaMalloc((void**)&d_perm, n * sizeof(uint32_t)));
  CHECK_CUDA(cudaMalloc((void**)&d_pay, n * row_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_out, n * row_bytes));
  CHECK_CUDA(cudaMemcpyAsync(d_keys, h_keys, n * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
  CHECK_CUDA(cudaMemcpyAsync(d_pay, h_pay, n * row_bytes, cudaMemcpyHostToDevice, stream));

  size_t temp_bytes = 0; /* phase 1: temp == NULL -> required scratch size */
  CHECK_B200X(b200x_radix_argsort(d_keys, B200X_KEY_I64, 0, NULL, d_perm, NULL, d_sorted, n, 0, 0, NULL, &temp_bytes, stream));
  CHECK_CUDA(cudaMalloc((void**)&d_temp, temp_bytes));
  const unsigned long long launches0 = b200x_launch_count();
  CHECK_B200X(b200x_radix_argsort(d_keys, B200X_KEY_I64, 0, NULL, d_perm, NULL, d_sorted, n, 0, 0, d_temp, &temp_bytes, stream));
  CHECK_B200X(b200x_gather_rows(d_pay, row_bytes, d_perm, n, d_out, stream));
  CHECK_CUDA(cudaStreamSynchronize(stream));
  const unsigned long long launches = b200x_launch_count() - launches0;

  int64_t* s_keys = (int64_t*)malloc(n * sizeof(int64_t));
  uint32_t* s_perm = (uint32_t*)malloc(n * sizeof(uint32_t));
  unsigned char* s_pay = (unsigned char*)malloc(n * row_bytes);
  CHECK_CUDA(cudaMemcpy(s_keys, d_sorted, n * sizeof(int64_t), cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(s_perm, d_perm, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(s_pay, d_out, n * row_bytes, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < n; ++i) {
    size_t src;
    memcpy(&src, s_pay + i * row_bytes, sizeof(src));
    if (src != s_perm[i] || h_keys[src] != s_keys[i]) {
      fprintf(stderr, "row %zu: payload / key does not follow the permutation\n", i);
      return 5;
    }
    if (i && (s_keys[i - 1] > s_keys[i] || (s_keys[i - 1] == s_keys[i] && s_perm[i - 1] > s_perm[i]))) {
      fprintf(stderr, "row %zu: order or stability violated\n", i);
      return 6;
    }
  }
  /* error convention: a bad argument returns a non-zero status and a message, nothing is thrown */
  if (b200x_radix_argsort(d_keys, 999, 0, NULL, d_perm, NULL, NULL, n, 0, 0, d_temp, &temp_bytes, stream) == 0 ||
      strlen(b200x_last_error_string()) == 0) {
    fprintf(stderr, "bad key_type was not rejected\n");
    return 7;
  }
  printf("abi_smoke ok: %zu rows sorted and gathered through the C ABI, %llu kernel launches, version %d\n", n, launches,
         b200x_version());
  return 0;
}
