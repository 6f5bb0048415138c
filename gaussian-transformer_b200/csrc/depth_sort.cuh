// csrc/depth_sort.cuh -- stable LSD radix sort of (u32 key, u32 value) pairs, 8-bit digits.
//
// Role on the path: the reference sorts R (tile|depth, id) 64-bit pairs with
// cub::DeviceRadixSort (DGR/cuda_rasterizer/rasterizer_impl.cu:300-308), R being
// ~14x the number of visible Gaussians at 1080p.  The depth part of that order
// does not depend on the tile, so here the P Gaussians are sorted by depth ONCE
// (this file: keys = depth bits, values = Gaussian ids, stable => equal depths
// stay in ascending id order exactly like the reference's stable sort), and the
// tile part is resolved afterwards by an order-preserving binning pass
// (binning.cuh) that needs no keys at all.
//
// Per pass: (A) per-CTA digit histogram, (B) one-CTA exclusive scan over the
// digit-major histogram matrix, (C) stable scatter.  In (C) every warp owns a
// contiguous 512-key segment of the CTA's 4096-key tile, keeps it in registers,
// ranks keys with match.any against per-warp digit counters in shared memory
// and writes them to their final position.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ROUNDS = 16;                         // keys per thread
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;       // 4096 keys per CTA
constexpr int RS_SEG = 32 * RS_ROUNDS;                // 512 keys per warp

__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift,
                                                                int n_cta, uint32_t* __restrict__ cta_hist) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0u;
  __syncthreads();
  const int base = blockIdx.x * RS_TILE;
#pragma unroll 4
  for (int k = 0; k < RS_ROUNDS; k++) {
    const int i = base + k * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[(__ldg(keys + i) >> shift) & 255u], 1u);
  }
  __syncthreads();
  cta_hist[(size_t)threadIdx.x * n_cta + blockIdx.x] = h[threadIdx.x];
}

// In-place exclusive scan of n u32 values by ONE CTA of 1024 threads (n up to a few million).
// If total_out != nullptr the grand total (64-bit) is written there as two u32 words.
__global__ void __launch_bounds__(1024) exclusive_scan_kernel(uint32_t* __restrict__ data, int n,
                                                              uint32_t* __restrict__ total_out) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0ull;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    const uint32_t c = i < n ? data[i] : 0u;
    unsigned long long v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    if (wid == 0) {
      unsigned long long w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_tot[lane] = w;
    }
    __syncthreads();
    const unsigned long long incl = carry_s + v + (wid > 0 ? warp_tot[wid - 1] : 0ull);
    if (i < n) data[i] = (uint32_t)(incl - c);
    __syncthreads();
    if (tid == 1023) carry_s = incl;
    __syncthreads();
  }
  if (tid == 0 && total_out) {
    total_out[0] = (uint32_t)(carry_s & 0xffffffffull);
    total_out[1] = (uint32_t)(carry_s >> 32);
  }
}

// vals_in == nullptr means "identity" (value = index), used by the first pass.
__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                   const uint32_t* __restrict__ vals_in, int n,
                                                                   int shift, int n_cta,
                                                                   const uint32_t* __restrict__ cta_offs,
                                                                   uint32_t* __restrict__ keys_out,
                                                                   uint32_t* __restrict__ vals_out) {
  __shared__ uint32_t cnt[RS_WARPS][256];  // per-warp digit counts, then per-warp running output positions
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int k = tid; k < RS_WARPS * 256; k += RS_THREADS) (&cnt[0][0])[k] = 0u;
  __syncthreads();
  const int seg = blockIdx.x * RS_TILE + w * RS_SEG;
  uint32_t key[RS_ROUNDS], val[RS_ROUNDS];
  const unsigned lt = (1u << lane) - 1u;
  // phase 1: load the warp's segment (in order) and count digits
#pragma unroll
  for (int k = 0; k < RS_ROUNDS; k++) {
    const int i = seg + k * 32 + lane;
    const bool ok = i < n;
    key[k] = ok ? __ldg(keys_in + i) : 0xffffffffu;
    val[k] = ok ? (vals_in ? __ldg(vals_in + i) : (uint32_t)i) : 0u;
    const uint32_t d = (key[k] >> shift) & 255u;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const unsigned m = __match_any_sync(act, d);
      if ((m & lt) == 0u) cnt[w][d] += __popc(m);  // leader of its digit group; digits are distinct across leaders
    }
    __syncwarp();
  }
  __syncthreads();
  // per digit: exclusive prefix over the warps + this CTA's global offset
  {
    const int d = tid;  // RS_THREADS == 256 digits
    uint32_t run = __ldg(cta_offs + (size_t)d * n_cta + blockIdx.x);
#pragma unroll
    for (int ww = 0; ww < RS_WARPS; ww++) {
      const uint32_t c = cnt[ww][d];
      cnt[ww][d] = run;
      run += c;
    }
  }
  __syncthreads();
  // phase 2: stable ranks and scatter
#pragma unroll
  for (int k = 0; k < RS_ROUNDS; k++) {
    const int i = seg + k * 32 + lane;
    const bool ok = i < n;
    const uint32_t d = (key[k] >> shift) & 255u;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const unsigned m = __match_any_sync(act, d);
      const uint32_t pos = cnt[w][d] + __popc(m & lt);
      keys_out[pos] = key[k];
      vals_out[pos] = val[k];
      __syncwarp(m);
      if ((m & lt) == 0u) cnt[w][d] += __popc(m);
    }
    __syncwarp();
  }
}

}  // namespace gsr
