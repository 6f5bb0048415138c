// csrc/onesweep.cuh -- stable LSD radix sort of (u32 key, u32 value) pairs: ONE kernel per 8-bit pass
// ("onesweep": per-tile digit histogram, chained look-back over the earlier tiles' histograms, rank and scatter
// in the same kernel) plus one up-front kernel that builds the digit histograms of all four passes.
//
// Role on the path: the reference sorts R (tile|depth, id) 64-bit pairs with cub::DeviceRadixSort::SortPairs
// (DGR/cuda_rasterizer/rasterizer_impl.cu:300-308; CUB picks its onesweep policy on this architecture).  Here the P
// Gaussians are sorted by their 32-bit depth key once (see depth_sort.cuh for why that is enough) -- 5 launches
// instead of the 12 of the histogram / row-scan / scatter scheme of depth_sort.cuh, which stays in the tree for the
// smallest inputs and as the A/B reference (gsr_set_option("sort", 0)).
//
// B200 notes.  The whole working set (16 B per Gaussian) lives in the L2, so a pass is bound by latency and by the
// look-back chain, not by bandwidth: all tiles of a pass are resident at once (<= 2 tiles per SM at 1 M keys), so a
// tile cannot expect its predecessor's inclusive prefix to be there; it sums the predecessors' tile histograms
// instead (16 independent loads in flight per thread, one thread per digit) and stops at the first inclusive prefix
// it meets; the walk has two levels (tiles of the own group, then whole groups of 16 tiles).  Tile numbers are handed out by an atomic counter in arrival order, so every
// predecessor a tile waits for is running or finished (forward progress without relying on block scheduling order).
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int OS_THREADS = 256;
constexpr int OS_WARPS = OS_THREADS / 32;
#ifndef GSR_OS_MATCH
#define GSR_OS_MATCH 0
#endif
#ifndef GSR_OS_LOOKBACK
#define GSR_OS_LOOKBACK 16
#endif
constexpr int OS_LOOKBACK = GSR_OS_LOOKBACK;      // earlier tiles read per look-back step (independent loads in flight)
constexpr int OS_GROUP = OS_GROUP_TILES;           // tiles per look-back group (common.cuh sizes the scratch)
constexpr uint32_t OS_FLAG_AGG = 1u << 30;        // status word = count | flag: this tile's own histogram entry
constexpr uint32_t OS_FLAG_INC = 2u << 30;        //                             inclusive prefix over tiles 0..t
constexpr uint32_t OS_COUNT_MASK = (1u << 30) - 1u;


// lanes of `act` whose 8-bit digit equals this lane's: eight ballots (ALU pipe) instead of one MATCH.ANY (MIO pipe,
// low throughput)
__device__ __forceinline__ unsigned match_digit_all(unsigned act, uint32_t d) {
#if GSR_OS_MATCH
  return __match_any_sync(0xffffffffu, d) & act;
#else
  unsigned m = act;  // every lane votes (full mask: no convergence bookkeeping); lanes outside act are masked out
#pragma unroll
  for (int b = 0; b < 8; b++) {
    const bool bit = (d >> b) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, bit);
    m &= bit ? bal : ~bal;
  }
  return m;
#endif
}

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Sums the published counts of entries hi, hi-1, ..., lo of one digit's column (stride 256 words), waiting for each
// to be published, until one of them is an inclusive prefix.  Returns true iff it stopped at an inclusive prefix.
__device__ __forceinline__ bool os_walk(const uint32_t* col, int hi, int lo, uint32_t& excl) {
  int t = hi;
  while (t >= lo) {
    uint32_t s[OS_LOOKBACK];
#pragma unroll
    for (int j = 0; j < OS_LOOKBACK; j++) s[j] = (t - j >= lo) ? ld_relaxed(col + (size_t)(t - j) * 256) : 0u;
#pragma unroll
    for (int j = 0; j < OS_LOOKBACK; j++) {
      if (t < lo || (s[j] >> 30) == 0u) break;  // past the end, or not published yet: re-read from here
      excl += s[j] & OS_COUNT_MASK;
      t--;
      if (s[j] & OS_FLAG_INC) return true;
    }
  }
  return false;
}

// Digit histograms of all four passes in one read of the keys: hist[pass][digit].
__global__ void __launch_bounds__(OS_THREADS) os_hist_kernel(const uint32_t* __restrict__ keys, int n,
                                                            uint32_t* __restrict__ hist) {
  __shared__ uint32_t h[4][256];
  for (int k = threadIdx.x; k < 4 * 256; k += OS_THREADS) (&h[0][0])[k] = 0u;
  __syncthreads();
  const int n4 = n >> 2;
  const uint4* __restrict__ k4 = reinterpret_cast<const uint4*>(keys);
  for (int i = blockIdx.x * OS_THREADS + threadIdx.x; i < n4; i += gridDim.x * OS_THREADS) {
    const uint4 q = __ldg(k4 + i);
    const uint32_t v[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      atomicAdd(&h[0][v[j] & 255u], 1u);
      atomicAdd(&h[1][(v[j] >> 8) & 255u], 1u);
      atomicAdd(&h[2][(v[j] >> 16) & 255u], 1u);
      atomicAdd(&h[3][v[j] >> 24], 1u);
    }
  }
  if (blockIdx.x == 0)
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += OS_THREADS) {
      const uint32_t v = __ldg(keys + i);
      atomicAdd(&h[0][v & 255u], 1u);
      atomicAdd(&h[1][(v >> 8) & 255u], 1u);
      atomicAdd(&h[2][(v >> 16) & 255u], 1u);
      atomicAdd(&h[3][v >> 24], 1u);
    }
  __syncthreads();
  for (int k = threadIdx.x; k < 4 * 256; k += OS_THREADS) {
    const uint32_t c = (&h[0][0])[k];
    if (c) atomicAdd(hist + k, c);
  }
}

// One pass.  vals_in == nullptr: value = index (first pass).  hist = this pass's 256 digit totals, counter = this
// pass's tile counter (zeroed), status / gstatus = this pass's [tiles][256] / [groups][256] look-back words (zeroed).
// OS_KPT keys per thread: a tile is 256 * OS_KPT keys, a warp owns 32 * OS_KPT consecutive ones.
template <int OS_KPT>
__global__ void __launch_bounds__(OS_THREADS) os_pass_kernel(const uint32_t* __restrict__ keys_in,
                                                            const uint32_t* __restrict__ vals_in, int n, int shift,
                                                            const uint32_t* __restrict__ hist,
                                                            uint32_t* __restrict__ counter,
                                                            uint32_t* __restrict__ status,
                                                            uint32_t* __restrict__ gstatus,
                                                            uint32_t* __restrict__ keys_out,
                                                            uint32_t* __restrict__ vals_out) {
  constexpr int OS_TILE = OS_THREADS * OS_KPT, OS_SEG = 32 * OS_KPT;
  __shared__ uint32_t cnt[OS_WARPS][256];  // per-warp digit counts, then per-warp running positions in the tile
  __shared__ uint32_t s_scan[16], s_delta[256];
  __shared__ uint32_t s_keys[OS_TILE], s_vals[OS_TILE];
  __shared__ uint32_t s_tile;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(counter, 1u);
  for (int k = tid; k < OS_WARPS * 256; k += OS_THREADS) (&cnt[0][0])[k] = 0u;
  __syncthreads();
  const int tile = (int)s_tile;
  const int seg = tile * OS_TILE + w * OS_SEG;
  uint32_t key[OS_KPT], val[OS_KPT];
  unsigned grp[OS_KPT];  // lanes of the round that share this lane's digit
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int k = 0; k < OS_KPT; k++) {
    const int i = seg + k * 32 + lane;
    key[k] = i < n ? __ldg(keys_in + i) : 0xffffffffu;
  }
#pragma unroll
  for (int k = 0; k < OS_KPT; k++) {
    const int i = seg + k * 32 + lane;
    val[k] = vals_in ? (i < n ? __ldg(vals_in + i) : 0u) : (uint32_t)i;
  }
  // phase 1: per-warp digit counts (rounds in order: what makes the ranks below stable)
#pragma unroll
  for (int k = 0; k < OS_KPT; k++) {
    const int i = seg + k * 32 + lane;
    const bool ok = i < n;
    const uint32_t d = (key[k] >> shift) & 255u;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    const unsigned m = ok ? match_digit_all(act, d) : (match_digit_all(act, d), 0u);
    grp[k] = m;
    if (m && (m & lt) == 0u) cnt[w][d] += __popc(m);  // the group's leader; digits differ between leaders
    __syncwarp();
  }
  __syncthreads();
  // phase 2a (thread = digit): tile histogram, published at once so that later tiles can walk over it; first slot of
  // the digit inside the tile (local) and in the output (global, from the up-front histogram); per-warp local bases
  const int d_ = tid;
  uint32_t tot = 0, lstart, gstart;
  uint32_t* const mine = status + (size_t)tile * 256 + d_;
  {
#pragma unroll
    for (int ww = 0; ww < OS_WARPS; ww++) tot += cnt[ww][d_];
    st_relaxed(mine, tot | (tile == 0 ? OS_FLAG_INC : OS_FLAG_AGG));
    const uint32_t c = __ldg(hist + d_);
    uint32_t v = c, u = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o), t2 = __shfl_up_sync(0xffffffffu, u, o);
      if (lane >= o) { v += t; u += t2; }
    }
    if (lane == 31) { s_scan[w] = v; s_scan[8 + w] = u; }
    __syncthreads();
    uint32_t pre = 0, pre2 = 0;
    for (int ww = 0; ww < w; ww++) { pre += s_scan[ww]; pre2 += s_scan[8 + ww]; }
    gstart = pre + v - c;
    lstart = pre2 + u - tot;
    uint32_t run = lstart;
#pragma unroll
    for (int ww = 0; ww < OS_WARPS; ww++) {
      const uint32_t cc = cnt[ww][d_];
      cnt[ww][d_] = run;
      run += cc;
    }
  }
  __syncthreads();
  // phase 3: stable ranks; the tile is put in digit order in shared memory (needs no look-back result)
#pragma unroll
  for (int k = 0; k < OS_KPT; k++) {
    const uint32_t d = (key[k] >> shift) & 255u;
    const unsigned m = grp[k];
    if (m) {
      const uint32_t lp = cnt[w][d] + __popc(m & lt);
      GSR_CHECK(lp < (uint32_t)OS_TILE);
      s_keys[lp] = key[k];
      s_vals[lp] = val[k];
    }
    __syncwarp();
    if (m && (m & lt) == 0u) cnt[w][d] += __popc(m);
    __syncwarp();
  }
  // phase 2b (thread = digit): look back, two levels.  (A) the earlier tiles of this tile's group of OS_GROUP tiles;
  // (B) the earlier groups, whose sums are published by each group's last tile.  Either walk stops at the first
  // inclusive prefix.  All tiles of the first wave start together and find no inclusive prefixes yet, so a one-level
  // walk is quadratic in the number of resident tiles (245 tiles at 1 M keys: 28 % of the kernel's instructions).
  {
    uint32_t excl = 0;
    if (tile > 0) {
      const int g = tile / OS_GROUP;
      const bool done = os_walk(status + (size_t)d_, tile - 1, g * OS_GROUP, excl);
      const bool last = (tile % OS_GROUP) == OS_GROUP - 1;
      uint32_t* const gmine = gstatus + (size_t)g * 256 + d_;
      if (last && !done) st_relaxed(gmine, (excl + tot) | OS_FLAG_AGG);  // the group's own sum
      if (!done && g > 0) os_walk(gstatus + (size_t)d_, g - 1, 0, excl);
      st_relaxed(mine, (excl + tot) | OS_FLAG_INC);
      if (last) st_relaxed(gmine, (excl + tot) | OS_FLAG_INC);
    }
    s_delta[d_] = gstart + excl - lstart;  // output position = position in the sorted tile + delta[digit]
  }
  __syncthreads();
  // phase 4: coalesced scatter -- consecutive threads write consecutive slots of a digit's run (a scatter straight
  // from registers costs one 32-byte store wavefront per key and value: 8192 per tile and pass, the bottleneck
  // of the first version of this kernel)
  const int n_tile = min(OS_TILE, n - tile * OS_TILE);
#pragma unroll
  for (int k = 0; k < OS_KPT; k++) {
    const int idx = k * OS_THREADS + tid;
    if (idx < n_tile) {
      const uint32_t kk = s_keys[idx];
      const uint32_t pos = (uint32_t)idx + s_delta[(kk >> shift) & 255u];
      GSR_CHECK(pos < (uint32_t)n);
      keys_out[pos] = kk;
      vals_out[pos] = s_vals[idx];
    }
  }
}

// Host side: sorts n pairs, result in keys[0] / vals[0] (vals of the first pass = index).  scratch:
// os_scratch_words(n) words.  Returns the number of kernels launched.
inline int os_sort_launch(uint32_t* const keys[2], uint32_t* const vals[2], int n, uint32_t* scratch,
                          cudaStream_t stream) {
  const int tiles = (int)os_tiles((size_t)n);
  uint32_t* hist = scratch;
  uint32_t* counters = scratch + 4 * 256;
  uint32_t* status = scratch + 4 * 256 + 64;
  const size_t per_pass = (size_t)(tiles + os_groups((size_t)n)) * 256;
  cudaMemsetAsync(scratch, 0, 4 * os_scratch_words((size_t)n), stream);
  const int hist_ctas = tiles < 592 ? (tiles < 1 ? 1 : tiles) : 592;
  os_hist_kernel<<<hist_ctas, OS_THREADS, 0, stream>>>(keys[0], n, hist);
  for (int pass = 0; pass < 4; pass++) {
    const int src = pass & 1, dst = src ^ 1;
    const uint32_t* vin = pass == 0 ? nullptr : vals[src];
    uint32_t* st = status + pass * per_pass;
    if (os_kpt((size_t)n) == 4)
      os_pass_kernel<4><<<tiles, OS_THREADS, 0, stream>>>(keys[src], vin, n, 8 * pass, hist + 256 * pass,
                                                          counters + pass, st, st + (size_t)tiles * 256, keys[dst],
                                                          vals[dst]);
    else
      os_pass_kernel<16><<<tiles, OS_THREADS, 0, stream>>>(keys[src], vin, n, 8 * pass, hist + 256 * pass,
                                                           counters + pass, st, st + (size_t)tiles * 256, keys[dst],
                                                           vals[dst]);
  }
  return 5;
}

}  // namespace gsr
