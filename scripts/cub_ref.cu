// Developer yardstick (NOT part of the product): cub::DeviceRadixSort::SortPairs timing on the same shapes.
#include <cub/cub.cuh>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <random>
template <typename K>
void run(size_t n, const char* name, int end_bit) {
  K *k_in, *k_out; uint32_t *v_in, *v_out;
  cudaMalloc(&k_in, n * sizeof(K)); cudaMalloc(&k_out, n * sizeof(K));
  cudaMalloc(&v_in, n * 4); cudaMalloc(&v_out, n * 4);
  std::vector<K> h(n); std::mt19937_64 g(1); for (auto& x : h) x = (K)g();
  if (end_bit < (int)sizeof(K) * 8) for (auto& x : h) x &= (((K)1) << end_bit) - 1;
  cudaMemcpy(k_in, h.data(), n * sizeof(K), cudaMemcpyHostToDevice);
  void* tmp = nullptr; size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(tmp, tb, k_in, k_out, v_in, v_out, n, 0, end_bit);
  cudaMalloc(&tmp, tb);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9;
  for (int it = 0; it < 6; ++it) {
    cudaEventRecord(a);
    cub::DeviceRadixSort::SortPairs(tmp, tb, k_in, k_out, v_in, v_out, n, 0, end_bit);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (it > 1 && ms < best) best = ms;
  }
  printf("cub SortPairs %-28s n=%zu bits=%d: %.3f ms (temp %.1f MB)\n", name, n, end_bit, best, tb / 1e6);
  cudaFree(k_in); cudaFree(k_out); cudaFree(v_in); cudaFree(v_out); cudaFree(tmp);
}
int main() {
  size_t n = 100000000;
  run<uint64_t>(n, "u64 key + u32 value", 64);
  run<uint64_t>(n, "u64 key + u32 value", 32);
  run<uint32_t>(n, "u32 key + u32 value", 32);
  return 0;
}
