// Stage ablation of the onesweep pass (developer tool, not part of the product):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DSOAR_SORT_PROBE tools/sort_probe.cu \
//        soar_b200/csrc/prof.cu soar_b200/csrc/api_stub... -> see tools/build_probe.sh
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cub/device/device_radix_sort.cuh>   // the bar to beat only: the product never includes CUB
#include "../soar_b200/csrc/binning.cu"

namespace soar {
void set_error(const char* fmt, ...) { (void)fmt; }
KernelScope::KernelScope(int, cudaStream_t) : id_(0), stream_(0), e1_(nullptr) {}
KernelScope::~KernelScope() {}
void KernelScope::count_extra_launch() {}
}  // namespace soar
using namespace soar;

int main(int argc, char** argv) {
  const uint32_t n = argc > 1 ? (uint32_t)atoll(argv[1]) : 8900000u;
  const int bits = argc > 2 ? atoi(argv[2]) : 13;
  std::vector<uint32_t> hk(n), hv(n);
  uint32_t x = 12345;
  for (uint32_t i = 0; i < n; ++i) { x = x * 1664525u + 1013904223u; hk[i] = (x >> 8) & ((1u << bits) - 1); hv[i] = i; }
  uint32_t *k[2], *v[2];
  for (int i = 0; i < 2; ++i) { cudaMalloc(&k[i], n * 4); cudaMalloc(&v[i], n * 4); }
  const int passes = sort_passes(bits);
  const int tiles = sort_tiles_for(n);
  uint32_t *ghist, *status;
  cudaMalloc(&ghist, 4 * 256 * 4);
  const size_t status_bytes = ((size_t)passes * tiles * 256 + 8) * 4;
  cudaMalloc(&status, status_bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 4; ++mode) {
    cudaMemcpyToSymbol(g_probe, &mode, sizeof(int));
    float best = 1e9f;
    for (int it = 0; it < 5; ++it) {
      cudaMemcpy(k[0], hk.data(), n * 4, cudaMemcpyHostToDevice);
      cudaMemcpy(v[0], hv.data(), n * 4, cudaMemcpyHostToDevice);
      cudaMemset(ghist, 0, 4 * 256 * 4);
      cudaMemset(status, 0, status_bytes);
      int res = 0;
      cudaEventRecord(e0);
      int rc = launch_radix_sort(k, v, n, bits, false, false, false, ghist, status, tiles, &res, 0);
      cudaEventRecord(e1);
      cudaError_t err = cudaDeviceSynchronize();
      if (rc || err) { printf("mode %d: rc %d err %s\n", mode, rc, cudaGetErrorString(err)); return 1; }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      best = ms < best ? ms : best;
    }
    printf("n=%u bits=%d passes=%d tiles=%d probe=%d (bit0 no look-back, bit1 no ranking): %.1f us total (hist+scan+passes)\n",
           n, bits, passes, tiles, mode, best * 1e3f);
  }
  // ---- the bar: cub::DeviceRadixSort::SortPairs on the same (u32, u32) input, same significant bits ----
  {
    void* temp = nullptr; size_t temp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(temp, temp_bytes, k[0], k[1], v[0], v[1], (int)n, 0, bits);
    cudaMalloc(&temp, temp_bytes);
    float best = 1e9f;
    for (int it = 0; it < 5; ++it) {
      cudaMemcpy(k[0], hk.data(), n * 4, cudaMemcpyHostToDevice);
      cudaMemcpy(v[0], hv.data(), n * 4, cudaMemcpyHostToDevice);
      cudaEventRecord(e0);
      cub::DeviceRadixSort::SortPairs(temp, temp_bytes, k[0], k[1], v[0], v[1], (int)n, 0, bits);
      cudaEventRecord(e1);
      cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      best = ms < best ? ms : best;
    }
    printf("n=%u bits=%d cub::DeviceRadixSort::SortPairs (CUDA 12.9): %.1f us, temp %zu bytes\n", n, bits, best * 1e3f, temp_bytes);
  }
  return 0;
}
