// csrc/compact.cuh -- row compaction for the densify / clone / split / prune step.
//
// The reference rebuilds all six parameter tensors, their two Adam moments and three statistics tensors with
// boolean-mask indexing and torch.cat (scene/gaussian_model.py:273-327: `tensor[mask]`, `torch.cat((a, b))` per
// tensor and per optimizer state: ~40 nonzero / index / cat launches and as many temporaries per call).  Here:
//   compact_count / compact_offsets / compact_write  mask (P bytes) -> ascending indices of the set entries + their
//                                                    number (the same order as tensor[mask])
//   gather_rows_kernel                               ONE launch moves the rows of up to 24 tensors:
//                                                    dst[j, :] = src[idx[j], :] (idx < 0: zeros), any row width
// Pure data movement: results are bit-identical to the torch statements.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int CP_THREADS = 256;
constexpr int CP_ITEMS = 16;                      // mask bytes per thread
constexpr int CP_BLOCK = CP_THREADS * CP_ITEMS;   // 4096 mask entries per CTA

__global__ void __launch_bounds__(CP_THREADS) compact_count_kernel(int P, const uint8_t* __restrict__ mask,
                                                                  uint32_t* __restrict__ block_count) {
  __shared__ uint32_t s_w[CP_THREADS / 32];
  const int base = blockIdx.x * CP_BLOCK + threadIdx.x * CP_ITEMS;
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < CP_ITEMS; k++)
    if (base + k < P) c += mask[base + k] != 0;
  c = __reduce_add_sync(0xffffffffu, c);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int k = 0; k < CP_THREADS / 32; k++) t += s_w[k];
    block_count[blockIdx.x] = t;
  }
}

// one CTA: exclusive scan of the block counts (in place), total -> total_out[0] (64-bit)
__global__ void __launch_bounds__(1024) compact_offsets_kernel(int n_blocks, uint32_t* __restrict__ block_count,
                                                              unsigned long long* __restrict__ total_out) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_carry = 0u;
  __syncthreads();
  for (int b0 = 0; b0 < n_blocks; b0 += 1024) {
    const int i = b0 + tid;
    const uint32_t c = i < n_blocks ? block_count[i] : 0u;
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    if (lane == 31) s_w[w] = v;
    __syncthreads();
    if (w == 0) {
      uint32_t x = s_w[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += t;
      }
      s_w[lane] = x;
    }
    __syncthreads();
    const uint32_t incl = s_carry + v + (w > 0 ? s_w[w - 1] : 0u);
    if (i < n_blocks) block_count[i] = incl - c;
    __syncthreads();
    if (tid == 1023) s_carry = incl;
    __syncthreads();
  }
  if (tid == 0) total_out[0] = s_carry;
}

// idx_out[offset + rank] = i for every set mask entry i, in ascending order; `add` is added to every index written
// (lets a caller build index lists into a concatenation)
__global__ void __launch_bounds__(CP_THREADS) compact_write_kernel(int P, const uint8_t* __restrict__ mask,
                                                                  const uint32_t* __restrict__ block_offset,
                                                                  int32_t* __restrict__ idx_out, int32_t add) {
  __shared__ uint32_t s_w[CP_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int base = blockIdx.x * CP_BLOCK + tid * CP_ITEMS;
  uint32_t bits = 0;
#pragma unroll
  for (int k = 0; k < CP_ITEMS; k++)
    if (base + k < P && mask[base + k] != 0) bits |= 1u << k;
  const uint32_t c = __popc(bits);
  uint32_t v = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) s_w[w] = v;
  __syncthreads();
  uint32_t pre = block_offset[blockIdx.x];
  for (int k = 0; k < w; k++) pre += s_w[k];
  uint32_t pos = pre + v - c;
#pragma unroll
  for (int k = 0; k < CP_ITEMS; k++)
    if (bits & (1u << k)) idx_out[pos++] = base + k + add;
}

constexpr int GATHER_MAX_TABLES = 24;
struct GatherTable {
  const float* src;
  float* dst;
  const int32_t* idx;
  int row_floats;
  int first_block;  // first CTA (of blockIdx.x) working on this table
};
struct GatherArgs {
  GatherTable t[GATHER_MAX_TABLES];
  int n_tables;
};
constexpr int GATHER_THREADS = 256;
constexpr int GATHER_ELEMS = 4096;  // output floats per CTA

// dst[j * w + c] = idx[j] >= 0 ? src[idx[j] * w + c] : 0 for all tables; a CTA owns GATHER_ELEMS consecutive output
// floats of one table (consecutive threads -> consecutive output floats: coalesced stores, row-contiguous loads).
__global__ void __launch_bounds__(GATHER_THREADS) gather_rows_kernel(const __grid_constant__ GatherArgs a,
                                                                    long long n_rows,
                                                                    const int32_t* __restrict__ idx) {
  int k = 0;
#pragma unroll 1
  while (k + 1 < a.n_tables && (int)blockIdx.x >= a.t[k + 1].first_block) k++;
  const GatherTable& t = a.t[k];
  const long long total = n_rows * t.row_floats;
  const long long e0 = (long long)(blockIdx.x - t.first_block) * GATHER_ELEMS;
#pragma unroll 4
  for (int u = threadIdx.x; u < GATHER_ELEMS; u += GATHER_THREADS) {
    const long long e = e0 + u;
    if (e >= total) break;
    const long long row = e / t.row_floats;
    const int col = (int)(e - row * t.row_floats);
    const int32_t s = __ldg((t.idx ? t.idx : idx) + row);
    GSR_CHECK(s < 0x7fffffff);
    t.dst[e] = s >= 0 ? __ldg(t.src + (long long)s * t.row_floats + col) : 0.f;
  }
}

}  // namespace gsr
