// csrc/binning.cuh -- from per-Gaussian rectangles to per-tile depth-sorted lists.
//
// Replaces, with identical observable results (ranges, point_list):
//   cub::DeviceScan::InclusiveSum + D2H     DGR/cuda_rasterizer/rasterizer_impl.cu:277-281
//   duplicateWithKeys                       :70-111
//   cub::DeviceRadixSort::SortPairs (64-bit tile|depth keys, stable)  :300-308
//   cudaMemset(ranges) + identifyTileRanges :310-318, :116-138
//
// B200 design.  The reference globally sorts R (tile|depth) keys with a 6-pass
// LSD radix sort (~150 B of HBM traffic per instance).  Here the tile part of
// the key is resolved by *bucketing*: preprocess counts instances per tile, a
// one-CTA scan turns counts into tile ranges (so `ranges` needs no extra pass),
// `scatter_instances` drops (depth|id) keys into their tile's bucket with an
// atomic cursor, and each bucket is sorted on-chip in shared memory.  The
// reference's stable sort orders equal (tile, depth) keys by emission order,
// which is ascending Gaussian id (each Gaussian appears at most once per tile),
// so sorting the 64-bit composite (depth_bits << 32 | id) reproduces its order
// exactly and needs no stability.  HBM traffic: 8 B written + 8 B read + 4 B
// written per instance.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int SORT_SMALL_CAP = 2048;   // keys per tile handled by the 256-thread kernel (16 KB smem)
constexpr int SORT_BIG_CAP = 24576;    // keys per tile handled by the 1024-thread kernel (192 KB smem)

// One CTA: exclusive scan of tile_count -> ranges, zero cursors, collect tiles
// that need the big / global sort paths, publish R.
__global__ void __launch_bounds__(1024) tile_scan_kernel(int T, const uint32_t* __restrict__ tile_count,
                                                         uint2* __restrict__ ranges,
                                                         uint32_t* __restrict__ tile_cursor,
                                                         uint32_t* __restrict__ big_tiles,
                                                         uint32_t* __restrict__ ctrl) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry_s;
  __shared__ uint32_t nbig_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) { carry_s = 0ull; nbig_s = 0u; }
  __syncthreads();
  for (int base = 0; base < T; base += 1024) {
    const int t = base + tid;
    const uint32_t c = t < T ? tile_count[t] : 0u;
    unsigned long long v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long n = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += n;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    if (wid == 0) {
      unsigned long long w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        unsigned long long n = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += n;
      }
      warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const unsigned long long carry = carry_s;
    const unsigned long long incl = carry + v + (wid > 0 ? warp_tot[wid - 1] : 0ull);
    if (t < T) {
      const unsigned long long excl = incl - c;
      // saturate: an overflowing R is reported through ctrl and never rendered
      const uint32_t lo = excl > 0xffffffffull ? 0xffffffffu : (uint32_t)excl;
      const uint32_t hi = incl > 0xffffffffull ? 0xffffffffu : (uint32_t)incl;
      ranges[t] = c ? make_uint2(lo, hi) : make_uint2(0u, 0u);
      tile_cursor[t] = 0u;
      if (c > (uint32_t)SORT_SMALL_CAP) big_tiles[atomicAdd(&nbig_s, 1u)] = (uint32_t)t;
    }
    __syncthreads();
    if (tid == 1023) carry_s = incl;
    __syncthreads();
  }
  if (tid == 0) {
    const unsigned long long R = carry_s;
    ctrl[CTRL_R_LO] = (uint32_t)(R & 0xffffffffull);
    ctrl[CTRL_R_HI] = (uint32_t)(R >> 32);
    ctrl[CTRL_NBIG] = nbig_s;
  }
}

// Per visible Gaussian: recompute its rectangle from (mean2D, radius) exactly as
// duplicateWithKeys does (rasterizer_impl.cu:85-90) and append (depth|id) to
// each covered tile's bucket.
__global__ void __launch_bounds__(256) scatter_instances_kernel(int P, const float4* __restrict__ g0, int gx, int gy,
                                                                const uint2* __restrict__ ranges,
                                                                uint32_t* __restrict__ tile_cursor,
                                                                unsigned long long* __restrict__ keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float4 a = __ldg(g0 + i);
  const int radius = __float_as_int(a.w);
  if (radius <= 0) return;
  int minx, miny, maxx, maxy;
  get_rect(a.x, a.y, radius, gx, gy, minx, miny, maxx, maxy);
  const unsigned long long key = ((unsigned long long)__float_as_uint(a.z) << 32) | (uint32_t)i;
  for (int y = miny; y < maxy; y++)
    for (int x = minx; x < maxx; x++) {
      const int t = y * gx + x;
      const uint32_t slot = __ldg(&ranges[t].x) + atomicAdd(&tile_cursor[t], 1u);
      keys[slot] = key;
    }
}

// Bitonic network with all comparators ascending (mirror step + half cleaners):
// positions >= n behave as +inf and never move, so any n works without padding.
template <int THREADS>
__device__ __forceinline__ void bitonic_sort_shared(unsigned long long* s, int n) {
  int N = 1;
  while (N < n) N <<= 1;
  const int half = N >> 1;
  for (int k = 2; k <= N; k <<= 1) {
    // mirror step: i in lower half of each k-block, partner = i ^ (k-1)
    {
      const int hk = k >> 1;
      for (int c = threadIdx.x; c < half; c += THREADS) {
        const int i = (c / hk) * k + (c % hk);
        const int p = i ^ (k - 1);
        if (p < n) {
          unsigned long long a = s[i], b = s[p];
          if (a > b) { s[i] = b; s[p] = a; }
        }
      }
      __syncthreads();
    }
    for (int j = k >> 2; j > 0; j >>= 1) {
      for (int c = threadIdx.x; c < half; c += THREADS) {
        const int i = (c / j) * (2 * j) + (c % j);
        const int p = i + j;
        if (p < n) {
          unsigned long long a = s[i], b = s[p];
          if (a > b) { s[i] = b; s[p] = a; }
        }
      }
      __syncthreads();
    }
  }
}

template <int THREADS>
__device__ __forceinline__ void sort_one_tile_shared(unsigned long long* s, const unsigned long long* __restrict__ keys,
                                                     uint32_t* __restrict__ point_list, uint32_t start, int n) {
  for (int i = threadIdx.x; i < n; i += THREADS) s[i] = keys[start + i];
  __syncthreads();
  bitonic_sort_shared<THREADS>(s, n);
  for (int i = threadIdx.x; i < n; i += THREADS) point_list[start + i] = (uint32_t)s[i];
  __syncthreads();
}

// One CTA per tile, lists up to SORT_SMALL_CAP.
__global__ void __launch_bounds__(256) tile_sort_small_kernel(const uint2* __restrict__ ranges,
                                                              const unsigned long long* __restrict__ keys,
                                                              uint32_t* __restrict__ point_list) {
  __shared__ unsigned long long s[SORT_SMALL_CAP];
  const uint2 r = ranges[blockIdx.x];
  const int n = (int)(r.y - r.x);
  if (n <= 0 || n > SORT_SMALL_CAP) return;
  if (n == 1) {
    if (threadIdx.x == 0) point_list[r.x] = (uint32_t)keys[r.x];
    return;
  }
  sort_one_tile_shared<256>(s, keys, point_list, r.x, n);
}

// Persistent grid over the (rare) tiles with long lists: <= SORT_BIG_CAP keys are
// sorted in 192 KB of shared memory; longer lists fall back to the same network
// run in place in global memory (correct for any n, slow, practically unused).
__global__ void __launch_bounds__(1024) tile_sort_big_kernel(const uint32_t* __restrict__ big_tiles,
                                                             const uint32_t* __restrict__ ctrl,
                                                             const uint2* __restrict__ ranges,
                                                             unsigned long long* __restrict__ keys,
                                                             uint32_t* __restrict__ point_list) {
  extern __shared__ __align__(16) unsigned long long sbig[];
  const uint32_t nbig = ctrl[CTRL_NBIG];
  for (uint32_t b = blockIdx.x; b < nbig; b += gridDim.x) {
    const uint2 r = ranges[big_tiles[b]];
    const int n = (int)(r.y - r.x);
    if (n <= SORT_BIG_CAP) {
      sort_one_tile_shared<1024>(sbig, keys, point_list, r.x, n);
    } else {
      unsigned long long* g = keys + r.x;
      long long N = 1;
      while (N < n) N <<= 1;
      const long long half = N >> 1;
      for (long long k = 2; k <= N; k <<= 1) {
        const long long hk = k >> 1;
        for (long long c = threadIdx.x; c < half; c += 1024) {
          const long long i = (c / hk) * k + (c % hk);
          const long long p = i ^ (k - 1);
          if (p < n) {
            unsigned long long a = g[i], q = g[p];
            if (a > q) { g[i] = q; g[p] = a; }
          }
        }
        __syncthreads();
        for (long long j = k >> 2; j > 0; j >>= 1) {
          for (long long c = threadIdx.x; c < half; c += 1024) {
            const long long i = (c / j) * (2 * j) + (c % j);
            const long long p = i + j;
            if (p < n) {
              unsigned long long a = g[i], q = g[p];
              if (a > q) { g[i] = q; g[p] = a; }
            }
          }
          __syncthreads();
        }
      }
      for (int i = threadIdx.x; i < n; i += 1024) point_list[r.x + i] = (uint32_t)g[i];
      __syncthreads();
    }
  }
}

}  // namespace gsr
