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
// Per pass: (A) per-CTA digit histogram, (B) per-digit row scan of the
// digit-major histogram matrix (256 CTAs), (C) stable scatter.  In (C) every warp owns a
// contiguous 512-key segment of the CTA's 4096-key tile, keeps it in registers,
// ranks keys with match.any against per-warp digit counters in shared memory
// and writes them to their final position.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ROUNDS = GSR_SORT_ROUNDS;             // keys per thread
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;       // keys per CTA
static_assert(RS_TILE == SORT_TILE_KEYS, "scratch sizing and the sort disagree");
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

// One CTA per digit d: in-place exclusive scan of row d of the digit-major matrix cta_hist[256][n_cta]
// (the position, among keys with digit d, of the first such key of each CTA tile); digit_total[d] = row sum.
__global__ void __launch_bounds__(256) radix_rowscan_kernel(uint32_t* __restrict__ cta_hist, int n_cta,
                                                            uint32_t* __restrict__ digit_total) {
  __shared__ uint32_t warp_tot[8];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint32_t* row = cta_hist + (size_t)blockIdx.x * n_cta;
  if (tid == 0) carry_s = 0u;
  __syncthreads();
  for (int base = 0; base < n_cta; base += 256) {
    const int i = base + tid;
    const uint32_t c = i < n_cta ? row[i] : 0u;
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    uint32_t pre = carry_s;
#pragma unroll
    for (int k = 0; k < 8; k++)
      if (k < wid) pre += warp_tot[k];
    if (i < n_cta) row[i] = pre + v - c;
    __syncthreads();
    if (tid == 255) carry_s = pre + v;
    __syncthreads();
  }
  if (tid == 0) digit_total[blockIdx.x] = carry_s;
}

// vals_in == nullptr means "identity" (value = index), used by the first pass.
__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                   const uint32_t* __restrict__ vals_in, int n,
                                                                   int shift, int n_cta,
                                                                   const uint32_t* __restrict__ cta_offs,
                                                                   const uint32_t* __restrict__ digit_total,
                                                                   uint32_t* __restrict__ keys_out,
                                                                   uint32_t* __restrict__ vals_out) {
  __shared__ uint32_t cnt[RS_WARPS][256];  // per-warp digit counts, then per-warp running output positions
  __shared__ uint32_t dbase[256];          // exclusive scan of digit_total
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int k = tid; k < RS_WARPS * 256; k += RS_THREADS) (&cnt[0][0])[k] = 0u;
  __syncthreads();
  const int seg = blockIdx.x * RS_TILE + w * RS_SEG;
  uint32_t key[RS_ROUNDS], val[RS_ROUNDS];
  unsigned grp[RS_ROUNDS];  // lanes of this round that share the lane's digit (match.any is slow: done once)
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
    grp[k] = 0u;
    if (ok) {
      const unsigned m = __match_any_sync(act, d);
      grp[k] = m;
      if ((m & lt) == 0u) cnt[w][d] += __popc(m);  // leader of its digit group; digits are distinct across leaders
    }
    __syncwarp();
  }
  __syncthreads();
  // exclusive scan of the 256 digit totals (every CTA redoes it: 256 values)
  {
    const uint32_t c = __ldg(digit_total + tid);
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    dbase[tid] = v;  // inclusive within the warp's 32 digits
    __syncthreads();
    uint32_t pre = 0;
    for (int ww = 0; ww < w; ww++) pre += dbase[ww * 32 + 31];
    __syncthreads();
    dbase[tid] = pre + v - c;
    __syncthreads();
  }
  // per digit: exclusive prefix over the warps + this CTA's global offset
  {
    const int d = tid;  // RS_THREADS == 256 digits
    uint32_t run = dbase[d] + __ldg(cta_offs + (size_t)d * n_cta + blockIdx.x);
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
    if (ok) {
      const unsigned m = grp[k];
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
