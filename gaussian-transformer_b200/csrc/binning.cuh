// csrc/binning.cuh -- from depth-ordered Gaussians to per-tile depth-sorted lists.
//
// Replaces, with identical observable results (ranges, point_list, num_rendered):
//   cub::DeviceScan::InclusiveSum + D2H     DGR/cuda_rasterizer/rasterizer_impl.cu:277-281
//   duplicateWithKeys                       :70-111
//   cub::DeviceRadixSort::SortPairs (64-bit tile|depth keys, stable)  :300-308
//   cudaMemset(ranges) + identifyTileRanges :310-318, :116-138
//
// B200 design.  The reference materialises R = sum(tiles_touched) 12-byte
// (key, id) pairs and radix-sorts them in 6 passes (~150 B of HBM traffic per
// instance; R = 12 M at 1 M Gaussians / 1080p).  Here the Gaussians are already
// in depth order (depth_sort.cuh), so the tile lists can be produced by an
// ORDER-PRESERVING binning that never forms a key:
//   1. the depth-ordered sequence is cut into chunks; a warp walks one chunk
//      sequentially and counts, per tile, how many of its Gaussians cover it
//      (`bin_pass<false>`): row `chunk` of a chunks x tiles matrix;
//   2. `column_scan_kernel` turns every column into exclusive prefixes (the
//      position, inside the tile's list, of the chunk's first instance) and
//      yields the per-tile totals; `tile_scan_kernel` turns those into `ranges`
//      and num_rendered -- `ranges` needs no boundary-detection pass;
//   3. the same walk again (`bin_pass<true>`) with the per-tile cursors
//      initialised to ranges[tile].x + prefix writes each Gaussian id straight
//      to its final slot of point_list.
// A warp handles its chunk strictly in order (one Gaussian at a time, lanes over
// the tiles of its rectangle), which is what makes the lists come out sorted
// by (depth, id) without a per-tile sort.  To bound the per-warp cursor table
// (shared memory) the tile grid is cut into bands of whole tile rows; a warp
// task is (chunk, band).  HBM traffic: 4 B per instance + the small matrix.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int BIN_WARPS = 4;           // warps per CTA in bin_pass
constexpr int BAND_TILES_MAX = 2048;   // cursor-table entries per warp (8 KB)
#ifndef GSR_BAND_ROWS
#define GSR_BAND_ROWS 4
#endif
constexpr int BAND_ROWS = GSR_BAND_ROWS;  // tile rows per band

struct BinPlan {
  int chunk;       // Gaussians per chunk (a power of two >= 128)
  int chunk_shift; // log2(chunk)
  int n_chunks;
  int band_major;  // task order of bin_pass_kernel: 1 = all chunks of band 0 first (keeps the lists being written small)
  int rows_per_band, n_bands, band_tiles;
};

inline BinPlan make_bin_plan(int P, int gx, int gy) {
  BinPlan p;
  const int shift = bin_chunk_shift((size_t)P, (size_t)gx * gy);  // common.cuh (also sizes the count matrix)
  const int chunk = 1 << shift;
  p.chunk_shift = shift;
  p.chunk = chunk;
  p.n_chunks = P > 0 ? (P + chunk - 1) / chunk : 0;
  p.band_major = 1;  // measured at 1 M Gaussians / 1080p: 151 -> 135 us for the scatter walk
  p.rows_per_band = BAND_TILES_MAX / gx;
  if (p.rows_per_band < 1) p.rows_per_band = 0;  // image wider than 32768 px: unsupported, caller reports it
  // (measured at 1M Gaussians / 1080p with the rectangles precomputed: 4, 8, 17 tile rows per band give 148, 149,
  // 152 us -- the walk is bound by its 12 M scattered 4-byte writes, not by the scan or the serial chain)
  if (p.rows_per_band > BAND_ROWS) p.rows_per_band = BAND_ROWS;
  if (p.rows_per_band > gy) p.rows_per_band = gy;
  if (p.rows_per_band > 1023) p.rows_per_band = 1023;  // row offsets travel in 10 bits of a packed shuffle word
  p.n_bands = p.rows_per_band ? (gy + p.rows_per_band - 1) / p.rows_per_band : 0;
  p.band_tiles = p.rows_per_band * gx;
  return p;
}

// The ordered walk of Gaussians [begin, end) of the depth order for tile rows [by0, by1): strictly one Gaussian at a
// time, lanes over the tiles of its rectangle.  tab[] = this warp's per-tile cursors (band-local tile index);
// SCATTER = false only counts (tab[idx]++), SCATTER = true also writes the id at the cursor.
template <bool SCATTER>
__device__ __forceinline__ void bin_walk(int begin, int end, int lane, uint32_t* tab, uint4* desc,
                                         const uint32_t* __restrict__ order, const uint2* __restrict__ rects,
                                         const float4* __restrict__ g0, int gx, int gy, int by0, int by1,
                                         uint32_t* __restrict__ point_list, uint32_t list_cap = 0xffffffffu) {
  (void)list_cap;
  const unsigned lt = (1u << lane) - 1u;
  for (int r0 = begin; r0 < end; r0 += 32) {
    const int k = r0 + lane;
    // descriptor: x = id, y = first table index | width << 16, z = tiles | magic << 16
    uint4 d = make_uint4(0u, 0u, 0u, 0u);
    bool has = false;
    if (k < end) {
      d.x = __ldg(order + k);
      int minx = 0, miny = 0, maxx = 0, maxy = 0;
      if (rects != nullptr) {  // clipped rectangles in depth order, left by bin_count_cta_kernel
        const uint2 r = __ldg(rects + k);
        minx = r.x & 0xffff; maxx = r.x >> 16; miny = r.y & 0xffff; maxy = r.y >> 16;
      } else {
        const float4 a = __ldg(g0 + d.x);
        const int radius = __float_as_int(a.w);
        // same rectangle as duplicateWithKeys recomputes (rasterizer_impl.cu:85-90)
        if (radius > 0) get_rect(a.x, a.y, radius, gx, gy, minx, miny, maxx, maxy);
      }
      const int y0 = max(miny, by0), y1 = min(maxy, by1);
      has = (y1 > y0) && (maxx > minx);
      if (has) {
        const int wdt = maxx - minx, nt = wdt * (y1 - y0);
        // floor(t / wdt) == (t * magic) >> 10 for t < 32, wdt <= 32 (only used when the rect has <= 32 tiles)
        const uint32_t magic = wdt <= 32 ? (1024u + wdt - 1u) / wdt : 0u;
        d.y = (uint32_t)((y0 - by0) * gx + minx) | ((uint32_t)wdt << 16);  // both <= 2048
        d.z = (uint32_t)nt | (magic << 16);
      }
    }
    const unsigned live = __ballot_sync(0xffffffffu, has);
    if (has) desc[__popc(live & lt)] = d;
    __syncwarp();
    const int n_live = __popc(live);
    // strictly in order, one Gaussian at a time, lanes over the tiles of its rectangle
    for (int s_ = 0; s_ < n_live; s_++) {
      const uint4 e = desc[s_];  // broadcast read
      const int rowbase = e.y & 0xffff, wdt = e.y >> 16, nt = e.z & 0xffff;
      if (nt <= 32) {
        if (lane < nt) {
          const int ty = (int)(((uint32_t)lane * (e.z >> 16)) >> 10);
          const int idx = rowbase + ty * gx + (lane - ty * wdt);
          GSR_CHECK(idx >= 0 && idx < (by1 - by0) * gx);
          const uint32_t v = tab[idx];
          tab[idx] = v + 1u;
          GSR_CHECK(!SCATTER || v < list_cap);
          if (SCATTER) point_list[v] = e.x;
        }
      } else {
        for (int t = lane; t < nt; t += 32) {
          const int ty = t / wdt;
          const int idx = rowbase + ty * gx + (t - ty * wdt);
          GSR_CHECK(idx >= 0 && idx < (by1 - by0) * gx);
          const uint32_t v = tab[idx];
          tab[idx] = v + 1u;
          GSR_CHECK(!SCATTER || v < list_cap);
          if (SCATTER) point_list[v] = e.x;
        }
      }
      __syncwarp();
    }
  }
}

// One warp per (chunk, band).  SCATTER = false: count into matrix[chunk][tile];
// SCATTER = true: cursors start at ranges[tile].x + matrix[chunk][tile] and ids are written to point_list.
template <bool SCATTER>
__global__ void __launch_bounds__(BIN_WARPS * 32) bin_pass_kernel(int P, const uint32_t* __restrict__ order,
                                                                  const float4* __restrict__ g0, int gx, int gy,
                                                                  BinPlan plan, uint32_t* __restrict__ matrix,
                                                                  const uint2* __restrict__ ranges,
                                                                  uint32_t* __restrict__ point_list,
                                                                  const uint2* __restrict__ rects,
                                                                  const uint32_t* __restrict__ guard, uint32_t cap) {
  extern __shared__ uint32_t s_tab[];
  __shared__ uint4 s_desc[BIN_WARPS][32];  // per warp: the step's live Gaussians, compacted in order
  if (guard != nullptr && (guard[CTRL_R_HI] != 0u || guard[CTRL_R_LO] > cap)) return;  // see render_fwd_kernel
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int task = blockIdx.x * BIN_WARPS + w;
  if (task >= plan.n_chunks * plan.n_bands) return;
  int chunk, band;
  if (plan.band_major) { band = task / plan.n_chunks; chunk = task - band * plan.n_chunks; }
  else { chunk = task / plan.n_bands; band = task - chunk * plan.n_bands; }
  const int by0 = band * plan.rows_per_band, by1 = min(gy, by0 + plan.rows_per_band);
  const int ntab = (by1 - by0) * gx;
  const int tile0 = by0 * gx;
  uint32_t* tab = s_tab + w * plan.band_tiles;
  uint4* desc = s_desc[w];
  uint32_t* row = matrix + (size_t)chunk * gx * gy + tile0;
  for (int t = lane; t < ntab; t += 32) tab[t] = SCATTER ? (__ldg(&ranges[tile0 + t].x) + row[t]) : 0u;
  __syncwarp();
  const int begin = chunk * plan.chunk, end = min(P, begin + plan.chunk);
  bin_walk<SCATTER>(begin, end, lane, tab, desc, order, rects, g0, gx, gy, by0, by1, point_list,
                    cap);
  if (!SCATTER) {
    __syncwarp();
    for (int t = lane; t < ntab; t += 32) row[t] = tab[t];
  }
}


// CTA-cooperative scatter for long lists.  One CTA per (band, chunk), band-major: its 8 warps split the chunk into 8
// consecutive sub-chunks, (1) count their per-tile hits into their own tables, (2) turn the tables into cursors
// (tile's list start + instances of earlier chunks + of earlier sub-chunks), (3) walk again and place.  Twice the
// scanning of bin_pass<true>, but 8x shorter serial chains and -- decisive once point_list outgrows the L2 -- the
// CTAs in flight at any time write the lists of one or two bands only, so every 32-byte sector of a list is completed
// while it is still L2-resident.  (6 M Gaussians at 3840x2160, 228 M instances: bin_pass<true> wrote each sector to
// DRAM ~7 times, 12.4 GB of traffic for a 0.9 GB list, 8.3 ms.)
__global__ void __launch_bounds__(BIN_WARPS * 32) bin_scatter_cta_kernel(
    int P, const uint32_t* __restrict__ order, const float4* __restrict__ g0, int gx, int gy, BinPlan plan,
    const uint32_t* __restrict__ matrix, const uint2* __restrict__ ranges, uint32_t* __restrict__ point_list,
    const uint2* __restrict__ rects, const uint32_t* __restrict__ guard, uint32_t cap) {
  extern __shared__ uint32_t s_tab[];      // [BIN_WARPS][band_tiles]
  __shared__ uint4 s_desc[BIN_WARPS][32];
  if (guard != nullptr && (guard[CTRL_R_HI] != 0u || guard[CTRL_R_LO] > cap)) return;  // see render_fwd_kernel
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int band = blockIdx.x / plan.n_chunks, chunk = blockIdx.x - band * plan.n_chunks;
  const int by0 = band * plan.rows_per_band, by1 = min(gy, by0 + plan.rows_per_band);
  const int ntab = (by1 - by0) * gx, tile0 = by0 * gx;
  uint32_t* tab = s_tab + w * plan.band_tiles;
  const int sub = plan.chunk / BIN_WARPS;  // a multiple of 32 (chunk >= 256)
  const int begin = min(P, chunk * plan.chunk + w * sub), end = min(P, begin + sub);
  for (int t = lane; t < ntab; t += 32) tab[t] = 0u;
  __syncwarp();
  bin_walk<false>(begin, end, lane, tab, s_desc[w], order, rects, g0, gx, gy, by0, by1, nullptr);
  __syncthreads();
  const uint32_t* row = matrix + (size_t)chunk * gx * gy + tile0;
  for (int t = tid; t < ntab; t += BIN_WARPS * 32) {
    uint32_t run = __ldg(&ranges[tile0 + t].x) + row[t];
#pragma unroll
    for (int ww = 0; ww < BIN_WARPS; ww++) {
      const uint32_t c = s_tab[ww * plan.band_tiles + t];
      s_tab[ww * plan.band_tiles + t] = run;
      run += c;
    }
  }
  __syncthreads();
  bin_walk<true>(begin, end, lane, tab, s_desc[w], order, rects, g0, gx, gy, by0, by1, point_list,
                 cap);
}

// ---------------------------------------------------------------------------------------------------------------
// Counting needs no order, so the counting half is fully parallel whenever a table of all T tiles fits in shared
// memory: bin_count_cta<<<chunk>>> = one CTA per chunk, every thread adds its own Gaussians' rectangles with
// shared-memory atomics (as per-row difference arrays, see the kernel).  It also leaves each Gaussian's clipped tile rectangle in depth order (rects[k]), so the
// ordered scatter walk (bin_pass<true>) reads 8 coalesced bytes per Gaussian instead of gathering g0 again.
// (Measured alternative, removed: an unordered atomic scatter of sorted positions followed by a per-tile fix-up
// that ranks the 1.4-element (chunk, tile) cells -- bit-exact, but 12 M scattered 4-byte writes + 12 M gathers cost
// 97 + 95 us against 150 us for the ordered walk, whose floor is the same 12 M scattered writes.)
constexpr int BINC_THREADS = 256;
constexpr size_t BINC_SMEM_MAX = 200 * 1024;
#ifndef GSR_BINC_BANDS
#define GSR_BINC_BANDS 1
#endif
constexpr int BINC_BANDS = GSR_BINC_BANDS;  // minimum number of bands of tile rows per chunk
constexpr size_t BINC_TABLE_TARGET = 48 * 1024;  // per-CTA table size that still lets ~4 CTAs share an SM

// Bands of tile rows per chunk in bin_count_cta_kernel (its blockIdx.y): 1 while the table of all tiles is small
// (1080p: 32 KB); larger grids are cut so that several CTAs fit an SM (3840x2160: 130 KB -> 3 bands of 43 KB; one
// 130 KB table per SM left 8 warps per SM for 228 M shared-memory atomics, 0.68 ms).
inline int binc_bands(int gx, int gy) {
  const size_t bytes = (size_t)4 * gx * gy;
  int bands = (int)((bytes + BINC_TABLE_TARGET - 1) / BINC_TABLE_TARGET);
  if (bands < BINC_BANDS) bands = BINC_BANDS;
  if (bands > gy) bands = gy;
  return bands < 1 ? 1 : bands;
}
inline size_t binc_smem(int gx, int gy) {  // table of one band
  const int bands = binc_bands(gx, gy), rows = (gy + bands - 1) / bands;
  return (size_t)4 * rows * (gx + 1);  // difference array: one extra column per row
}
// the shared-memory table must fit (tile coordinates always fit the 16-bit rectangle fields: W, H < 2^20)
inline bool bin_parallel(int gx, int gy, const BinPlan& plan) {
  return binc_smem(gx, gy) <= BINC_SMEM_MAX && binc_bands(gx, gy) <= 65535 && plan.n_bands > 0;
}

__device__ __forceinline__ uint2 pack_rect(int minx, int miny, int maxx, int maxy) {
  return make_uint2((uint32_t)minx | ((uint32_t)maxx << 16), (uint32_t)miny | ((uint32_t)maxy << 16));
}

__global__ void __launch_bounds__(BINC_THREADS) bin_count_cta_kernel(int P, const uint32_t* __restrict__ order,
                                                                     const float4* __restrict__ g0, int gx, int gy,
                                                                     int chunk_size, uint2* __restrict__ rects,
                                                                     uint32_t* __restrict__ matrix) {
  extern __shared__ uint32_t s_tab[];
  const int tid = threadIdx.x, lane = tid & 31;
  // blockIdx.y = band of tile rows: a chunk's work is split over several CTAs (near chunks hold many more
  // instances than far ones); every band CTA scans the chunk and keeps what reaches its rows
  const int rows = (gy + gridDim.y - 1) / gridDim.y, by0 = blockIdx.y * rows, by1 = min(gy, by0 + rows);
  const int tile0 = by0 * gx;
  for (int t = tid; t < max(0, by1 - by0) * (gx + 1); t += BINC_THREADS) s_tab[t] = 0u;
  __syncthreads();
  const int begin = blockIdx.x * chunk_size, end = min(P, begin + chunk_size);
  for (int k0 = begin; k0 < end; k0 += BINC_THREADS) {
    const int k = k0 + tid;
    uint2 r = make_uint2(0u, 0u);
    if (k < end) {
      const float4 a = __ldg(g0 + __ldg(order + k));
      const int radius = __float_as_int(a.w);
      if (radius > 0) {
        int minx, miny, maxx, maxy;
        get_rect(a.x, a.y, radius, gx, gy, minx, miny, maxx, maxy);  // as duplicateWithKeys recomputes it
        if (maxx > minx && maxy > miny) r = pack_rect(minx, miny, maxx, maxy);
      }
      if (blockIdx.y == 0) rects[k] = r;
    }
    // Difference array per tile row: +1 at the rectangle's first column, -1 behind its last one (the table has
    // gx + 1 columns); a row prefix sum at the end turns them into per-tile counts.  2 atomics per tile row instead
    // of one per tile, and -- lanes run in lockstep -- the loop length is the tallest rectangle of the warp in ROWS.
    {
      const int minx = r.x & 0xffff, maxx = r.x >> 16;
      const int y0 = max((int)(r.y & 0xffff), by0) - by0, y1 = min((int)(r.y >> 16), by1) - by0;
      if (maxx > minx)
        for (int y = y0; y < y1; y++) {
          GSR_CHECK(y >= 0 && y < by1 - by0 && minx >= 0 && maxx <= gx);
          atomicAdd(&s_tab[y * (gx + 1) + minx], 1u);
          atomicAdd(&s_tab[y * (gx + 1) + maxx], 0xffffffffu);
        }
    }
  }
  __syncthreads();
  // row prefix sums (one warp per row, 32 columns at a time with a carry) straight into matrix[chunk][tile]
  uint32_t* out = matrix + (size_t)blockIdx.x * gx * gy + tile0;
  const int n_rows = max(0, by1 - by0);
  for (int y = tid >> 5; y < n_rows; y += BINC_THREADS / 32) {
    uint32_t carry = 0;
    for (int x0 = 0; x0 < gx; x0 += 32) {
      const int x = x0 + lane;
      uint32_t v = x < gx ? s_tab[y * (gx + 1) + x] : 0u;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      v += carry;
      if (x < gx) out[y * gx + x] = v;
      carry = __shfl_sync(0xffffffffu, v, 31);
    }
  }
}

// matrix[chunk][tile] -> exclusive prefix over chunks (in place); tile_count[tile] = column total.
// CTA = 32 tiles x 32 chunk groups.
__global__ void __launch_bounds__(1024) column_scan_kernel(uint32_t* __restrict__ matrix, int n_chunks, int T,
                                                          uint32_t* __restrict__ tile_count) {
  __shared__ uint32_t part[32][33];
  const int tx = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int tile = blockIdx.x * 32 + tx;
  const int per = (n_chunks + 31) / 32;
  const int c0 = min(n_chunks, g * per), c1 = min(n_chunks, c0 + per);
  uint32_t sum = 0;
  if (tile < T)
    for (int c = c0; c < c1; c++) sum += matrix[(size_t)c * T + tile];
  part[g][tx] = sum;
  __syncthreads();
  uint32_t run = 0;
  for (int gg = 0; gg < g; gg++) run += part[gg][tx];
  if (tile < T) {
    for (int c = c0; c < c1; c++) {
      const size_t i = (size_t)c * T + tile;
      const uint32_t v = matrix[i];
      matrix[i] = run;
      run += v;
    }
    if (g == 31) tile_count[tile] = run;
  }
}

// One CTA: exclusive scan of tile_count -> ranges ((0,0) for empty tiles, like the reference's memset +
// identifyTileRanges), publish R = num_rendered.
__global__ void __launch_bounds__(1024) tile_scan_kernel(int T, const uint32_t* __restrict__ tile_count,
                                                         uint2* __restrict__ ranges, uint32_t* __restrict__ ctrl) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0ull;
  __syncthreads();
  for (int base = 0; base < T; base += 1024) {
    const int t = base + tid;
    const uint32_t c = t < T ? tile_count[t] : 0u;
    unsigned long long v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long nb = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += nb;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    if (wid == 0) {
      unsigned long long ws = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long nb = __shfl_up_sync(0xffffffffu, ws, o);
        if (lane >= o) ws += nb;
      }
      warp_tot[lane] = ws;
    }
    __syncthreads();
    const unsigned long long incl = carry_s + v + (wid > 0 ? warp_tot[wid - 1] : 0ull);
    if (t < T) {
      const unsigned long long excl = incl - c;
      // saturate: an overflowing R is reported through ctrl and never rendered
      const uint32_t lo = excl > 0xffffffffull ? 0xffffffffu : (uint32_t)excl;
      const uint32_t hi = incl > 0xffffffffull ? 0xffffffffu : (uint32_t)incl;
      ranges[t] = c ? make_uint2(lo, hi) : make_uint2(0u, 0u);
    }
    __syncthreads();
    if (tid == 1023) carry_s = incl;
    __syncthreads();
  }
  if (tid == 0) {
    ctrl[CTRL_R_LO] = (uint32_t)(carry_s & 0xffffffffull);
    ctrl[CTRL_R_HI] = (uint32_t)(carry_s >> 32);
  }
}

// Sync-free forward (gsr_forward_capacity): when num_rendered does not fit the caller's list capacity the scatter and
// the blend do nothing (their guard); this empties every tile's range as well, so that a backward launched without
// the host having looked at num_rendered walks no list at all instead of reading past the buffer.
__global__ void __launch_bounds__(256) void_ranges_on_overflow_kernel(int T, uint2* __restrict__ ranges,
                                                                      const uint32_t* __restrict__ ctrl, uint32_t cap,
                                                                      long long* __restrict__ result) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // {num_rendered, flags}: device memory or device-mapped pinned host memory
    result[0] = (long long)(((unsigned long long)ctrl[CTRL_R_HI] << 32) | ctrl[CTRL_R_LO]);
    result[1] = (long long)ctrl[CTRL_FLAGS];
    __threadfence_system();
  }
  if (ctrl[CTRL_R_HI] == 0u && ctrl[CTRL_R_LO] <= cap) return;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t < T) ranges[t] = make_uint2(0u, 0u);
}

}  // namespace gsr
