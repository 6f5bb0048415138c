// oracle/cub_sort.cu -- TEST / BENCH INFRASTRUCTURE, never linked into the product.
// cub::DeviceRadixSort::SortPairs<uint32, uint32> behind a C ABI: the library call the reference makes
// (DGR/cuda_rasterizer/rasterizer_impl.cu:300-308, there on 64-bit keys) on the depth keys our sort handles, so
// that tools/sort_bench.py can time both on the same keys on the same GPU.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

extern "C" {

size_t cub_sort_temp_bytes(int n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, n);
  return bytes;
}

// keys_in / vals_in -> keys_out / vals_out, all 32 bits, on `stream`
int cub_sort_pairs(void* temp, size_t temp_bytes, const uint32_t* keys_in, uint32_t* keys_out,
                   const uint32_t* vals_in, uint32_t* vals_out, int n, int end_bit, void* stream) {
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit,
                                              (cudaStream_t)stream);
}
}
