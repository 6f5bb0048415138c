// csrc/render_bwd.cuh -- backward of the alpha compositing, one CTA (8 warps) per tile.
// Behaviour: DGR/cuda_rasterizer/backward.cu:399-557 (renderCUDA): back-to-front
// replay from final_T / n_contrib, same power / alpha tests as the forward.
//
// B200 design.  The reference issues 9 float atomics per contributing
// (pixel, Gaussian) pair and that is where its time goes.  Here
//   * the same two-level culling as the forward (cull.cuh) removes instances
//     that contribute to no pixel of the tile / of the warp's 8x4 block;
//   * the 9 partial gradients of a Gaussian are summed across the warp with a
//     reduce-scatter butterfly: 12 shuffles instead of 45, after which 9 lanes
//     each hold one finished sum;
//   * those 9 lanes issue ONE predicated RED instruction into the per-Gaussian
//     accumulators (common.cuh BwdView: two float4 arrays + one float array),
//     i.e. one atomic instruction per (warp, Gaussian) instead of 9 per
//     (pixel, Gaussian).
#pragma once
#include "common.cuh"
#include "cull.cuh"
#include "render_fwd.cuh"

namespace gsr {

// Reduce-scatter butterfly over the warp for 9 values.  One level: lanes whose bit `xorm` is clear keep the first
// value of each pair and receive the partner's copy of it, the others the second; the adds of two pairs are issued
// as one packed FADD2.  After the call, lane l with (l & 3) == 0 holds the warp-wide sum of
// v[4*b2 + 2*b3 + b4] (b2 = bit 2 of l, ...), and lanes with (l & 2) hold the sum of v8.
__device__ __forceinline__ float2 rs_pair2(float a0, float b0, float a1, float b1, bool hi, int xorm) {
  const float2 send = make_float2(hi ? a0 : b0, hi ? a1 : b1);
  const float2 keep = make_float2(hi ? b0 : a0, hi ? b1 : a1);
  return __fadd2_rn(keep, make_float2(__shfl_xor_sync(0xffffffffu, send.x, xorm),
                                      __shfl_xor_sync(0xffffffffu, send.y, xorm)));
}

__device__ __forceinline__ float warp_reduce_scatter9(float v0, float v1, float v2, float v3, float v4, float v5,
                                                      float v6, float v7, float v8, int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  const float2 r01 = rs_pair2(v0, v1, v2, v3, b4, 16), r23 = rs_pair2(v4, v5, v6, v7, b4, 16);
  float s = v8 + __shfl_xor_sync(0xffffffffu, v8, 16);
  const float2 t01 = rs_pair2(r01.x, r01.y, r23.x, r23.y, b3, 8);
  s += __shfl_xor_sync(0xffffffffu, s, 8);
  // level xor 4: the pair (t01.x, t01.y) and the running s advance together
  const float send = b2 ? t01.x : t01.y, keep = b2 ? t01.y : t01.x;
  const float2 us = __fadd2_rn(make_float2(keep, s), make_float2(__shfl_xor_sync(0xffffffffu, send, 4),
                                                                 __shfl_xor_sync(0xffffffffu, s, 4)));
  const float send2 = b1 ? us.x : us.y, keep2 = b1 ? us.y : us.x;
  float w = keep2 + __shfl_xor_sync(0xffffffffu, send2, 2);
  w += __shfl_xor_sync(0xffffffffu, w, 1);
  return w;
}

__global__ void __launch_bounds__(256) render_bwd_kernel(const uint2* __restrict__ ranges,
                                                         const uint32_t* __restrict__ point_list, int W, int H,
                                                         const float* __restrict__ bg, const float4* __restrict__ g0,
                                                         const float4* __restrict__ g1, const float4* __restrict__ g2,
                                                         const float* __restrict__ colors_precomp,
                                                         const float* __restrict__ final_T,
                                                         const uint32_t* __restrict__ n_contrib,
                                                         const float* __restrict__ dL_dpix, BwdView acc) {
  __shared__ StageRec s_rec[RENDER_BATCH];
  __shared__ float s_rba[RENDER_BATCH];
  __shared__ uint32_t s_id[RENDER_BATCH];
  __shared__ int s_warp[8];
  __shared__ uint32_t s_maxc;

  const int tid = threadIdx.x, lane = tid & 31;
  int lx, ly;
  warp_pixel(tid, lx, ly);
  const int pix_x = blockIdx.x * TILE + lx, pix_y = blockIdx.y * TILE + ly;
  const bool inside = pix_x < W && pix_y < H;
  const float2 npix = make_float2(-(float)pix_x, -(float)pix_y);
  const float tx0 = (float)(blockIdx.x * TILE), ty0 = (float)(blockIdx.y * TILE);
  const float wx0 = tx0 + (float)(((tid >> 5) & 1) << 3), wy0 = ty0 + (float)(((tid >> 5) >> 1) << 2);
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  if (range.y == range.x) return;
  const size_t N = (size_t)W * H, pix = (size_t)pix_y * W + pix_x;
  const uint32_t rec_addr = smem_addr(s_rec);
  const uint32_t id_addr = smem_addr(s_id);

  const float T_final = inside ? final_T[pix] : 0.f;
  float T = T_final;
  const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
  float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f;
  if (inside) {
    dLp0 = dL_dpix[pix];
    dLp1 = dL_dpix[N + pix];
    dLp2 = dL_dpix[2 * N + pix];
  }
  const float bg_dot = __ldg(bg) * dLp0 + __ldg(bg + 1) * dLp1 + __ldg(bg + 2) * dLp2;
  const float tfinal_bg = T_final * bg_dot;
  float S = 0.f, lcd = 0.f, last_alpha = 0.f;

  // list entries behind every pixel's last contributor are never touched
  uint32_t wmax = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
  if (tid == 0) s_maxc = 0u;
  __syncthreads();
  if (lane == 0) atomicMax(&s_maxc, wmax);
  __syncthreads();
  const int n_eff = (int)s_maxc;
  // which accumulator slot this lane commits after the reduce-scatter (-1: none)
  const int commit = ((lane & 3) == 0) ? (((lane >> 2) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 4) & 1))
                                       : (lane == 2 ? 8 : -1);
  float* const commit_base = commit < 0 ? nullptr
                             : commit < 4 ? reinterpret_cast<float*>(acc.acc0) + commit
                             : commit < 8 ? reinterpret_cast<float*>(acc.acc1) + (commit - 4)
                                          : acc.acc2;
  const uint32_t commit_shift = commit == 8 ? 0u : 2u;  // element offset of Gaussian id = id << shift (P < 2^29)

  for (int start = 0; start < n_eff; start += RENDER_BATCH) {
    __syncthreads();  // previous batch fully consumed before it is overwritten
    const int n_batch = min(RENDER_BATCH, n_eff - start);
    const int first = n_eff - 1 - start;
    const int cnt = stage_batch(tid, n_batch, point_list + range.x + first, -1, first, g0, g1, g2, colors_precomp,
                                tx0, ty0, s_rec, s_rba, s_id, s_warp);
    for (int c0 = 0; c0 < cnt; c0 += 32) {
      bool hit = false;
      if (c0 + lane < cnt && __float_as_uint(s_rec[c0 + lane].c.w) < wmax) {
        const float4 a = s_rec[c0 + lane].a, b = s_rec[c0 + lane].b;
        CullPre pre;
        pre.thr = b.z; pre.rBC = b.w; pre.rBA = s_rba[c0 + lane];
        hit = !cull_rect(a.x, a.y, a.z, b.x, a.w, pre, wx0, wx0 + 7.f, wy0, wy0 + 3.f);
      }
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      const uint32_t chunk_addr = rec_addr + (uint32_t)c0 * (uint32_t)sizeof(StageRec);
      const uint32_t chunk_id_addr = id_addr + (uint32_t)c0 * 4u;
      while (mask) {
        const uint32_t bit = (uint32_t)(__ffs(mask) - 1);
        mask &= mask - 1;
        const uint32_t addr = chunk_addr + bit * (uint32_t)sizeof(StageRec);
        const float4 a = lds128(addr);
        const float2 b = lds64(addr + 16);
        const float4 col = lds128(addr + 32);
        float2 d, pp;  // (dx, dy), (dx*A, dy*C)
        const float power = gauss_power2(make_float2(a.x, a.y), npix, make_float2(a.z, a.w), b.x, d, pp);
        // exp(power) as ex2.approx(power * log2 e): 2 instructions instead of expf's 10.  Relative error ~1e-6
        // (|power| <= ~10 times one rounding of the product, plus 2 ulp of ex2.approx): far inside the gradient
        // tolerance; an alpha within 1e-6 of 1/255 may be classified differently from the forward (a few dozen
        // pixel-Gaussian pairs per frame, each worth < 0.4 % of one pixel's contribution).  The forward keeps expf.
        float G;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(power * 1.4426950408889634f));
        const float alpha = fminf(0.99f, __fmul_rn(b.y, G));
        // same acceptance as the forward (backward.cu:488-497): behind the last contributor, power > 0 and
        // alpha < 1/255 are skipped
        const bool valid = (__float_as_uint(col.w) < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, valid)) continue;
        // Predicated body.  Algebra relative to backward.cu:499-554 (results identical up to rounding):
        //  * the pixel's dL/dpixel is constant over the list, so the reference's three accumulated colours enter
        //    only through their dot product with it: S = <accum_rec, dLp> obeys the same linear recurrence;
        //  * the per-Gaussian constant factors of the 2-D gradients (opacity, -1, -0.5, 0.5*W, 0.5*H) are applied
        //    once per Gaussian in preprocess_bwd instead of once per (pixel, Gaussian) pair.
        const float om = __fsub_rn(1.0f, alpha);  // alpha <= 0.99 => om in [0.01, 1]
        float rinv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rinv) : "f"(om));  // 1 ulp; T picks up <= ~1e-7 per list entry
        const float Tn = T * rinv;
        const float cd = fmaf(col.z, dLp2, fmaf(col.y, dLp1, col.x * dLp0));  // <colour, dLp>
        const float Sn = fmaf(last_alpha, lcd, (1.f - last_alpha) * S);
        const float dL_dalpha = fmaf(cd - Sn, Tn, -(tfinal_bg * rinv));
        const float w = valid ? alpha * Tn : 0.f;          // dchannel_dcolor
        const float q = valid ? G * dL_dalpha : 0.f;       // dL_dopacity; common factor of the 2-D gradients
        T = valid ? Tn : T;
        S = valid ? Sn : S;
        lcd = valid ? cd : lcd;
        last_alpha = valid ? alpha : last_alpha;
        const float2 qq = make_float2(q, q);
        const float2 v01 = __fmul2_rn(make_float2(w, w), make_float2(dLp0, dLp1));
        const float v2 = w * dLp2;
        const float2 qd = __fmul2_rn(qq, d);   // q dx, q dy  -> dL_dmean2D (conic and constants in preprocess_bwd)
        const float2 v68 = __fmul2_rn(qd, d);  // q dx dx, q dy dy  -> dL_dconic.xx, .yy (* -0.5 opacity)
        const float v7 = qd.x * d.y;           //                   -> dL_dconic.xy
        const float sum = warp_reduce_scatter9(v01.x, v01.y, v2, q, qd.x, qd.y, v68.x, v7, v68.y, lane);
        const uint32_t off = lds32(chunk_id_addr + bit * 4u) << commit_shift;
        if (commit >= 0) atomicAdd(commit_base + off, sum);
      }
    }
  }
}


// ------------------------------------------------------------------------------------------------------------------
// render_bwd2: TWO pixels per thread.  4 warps per tile, warp w owns the 8x8 block at (8*(w&1), 8*(w>>1)); lane l
// holds pixels (l&7, l>>3) and (l&7, (l>>3)+4) of it, i.e. two pixels of one column: dx is shared, dy differs by 4.
// Why: the hit loop is issue bound and every hit pays a fixed price that does not depend on the number of pixels
// behind it (mask walk, 3 shared loads, the 9-value reduce-scatter butterfly, the commit).  With two pixels per lane
//   * that fixed price is paid once per 64 pixels instead of once per 32, and an 8x8 block is hit by ~0.6x as many
//     Gaussians as its two 8x4 halves together;
//   * the per-pixel arithmetic of the two pixels pairs up in Blackwell's packed fp32 instructions (FADD2 / FMUL2 /
//     FFMA2: two IEEE fp32 results per issue slot);
//   * a skipped pixel is handled arithmetically instead of by selects: its alpha is forced to 0, which makes
//     1/(1-alpha) = 1, the T update a multiplication by one and the colour recurrence the identity.
// State per pixel: T and U = alpha_j c_j + (1-alpha_j) U  (c_j = <colour_j, dL/dpixel>), which is the reference's
// accum_rec (backward.cu:515-519) dotted with dL/dpixel and shifted by one list entry: the reference's value at entry j
// is U after entry j+1 (the previous one in back-to-front order).
constexpr int RB2_THREADS = 128;
#ifndef GSR_RB2_CTAS
#define GSR_RB2_CTAS 8
#endif


// Gather + tile-level cull + order-preserving compaction of up to 256 list entries by 128 threads (entries tid and
// tid + 128).  Entry k of the batch is list position first - k (back to front).  s_warp: 8 ints.
__device__ __forceinline__ int stage_batch2(int tid, int n_batch, const uint32_t* __restrict__ list_ptr, int first,
                                            const float4* __restrict__ g0, const float4* __restrict__ g1,
                                            const float4* __restrict__ g2, const float* __restrict__ colors_precomp,
                                            float tx0, float ty0, StageRec* s_rec, float* s_rba, uint32_t* s_id,
                                            int* s_warp) {
  const int lane = tid & 31, w = tid >> 5;
  bool keep[2] = {false, false};
  uint32_t id[2] = {0u, 0u};
  float4 a[2], co[2];
  CullPre pre[2];
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int e = k * RB2_THREADS + tid;
    if (e < n_batch) id[k] = __ldg(list_ptr - e);
  }
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int e = k * RB2_THREADS + tid;
    if (e < n_batch) {
      a[k] = __ldg(g0 + id[k]);
      co[k] = __ldg(g1 + id[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int e = k * RB2_THREADS + tid;
    if (e < n_batch) {
      keep[k] = cull_prepare(a[k].x, a[k].y, co[k].x, co[k].y, co[k].z, co[k].w, tx0, tx0 + 15.f, ty0, ty0 + 15.f,
                             pre[k]);
      if (keep[k])
        keep[k] = !cull_rect(a[k].x, a[k].y, co[k].x, co[k].y, co[k].z, pre[k], tx0, tx0 + 15.f, ty0, ty0 + 15.f);
    }
  }
  const unsigned m0 = __ballot_sync(0xffffffffu, keep[0]), m1 = __ballot_sync(0xffffffffu, keep[1]);
  if (lane == 0) {
    s_warp[w] = __popc(m0);
    s_warp[4 + w] = __popc(m1);
  }
  __syncthreads();
  int base[2] = {0, 0}, total = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int c = s_warp[k];
    if (k < w) base[0] += c;
    if (k < 4 + w) base[1] += c;
    total += c;
  }
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (keep[k]) {
      const int slot = base[k] + __popc((k ? m1 : m0) & ((1u << lane) - 1u));
      const float posf = __uint_as_float((uint32_t)(first - (k * RB2_THREADS + tid)));
      float4 c;
      if (colors_precomp != nullptr)
        c = make_float4(__ldg(colors_precomp + 3 * (size_t)id[k]), __ldg(colors_precomp + 3 * (size_t)id[k] + 1),
                        __ldg(colors_precomp + 3 * (size_t)id[k] + 2), posf);
      else {
        c = __ldg(g2 + id[k]);
        c.w = posf;
      }
      GSR_CHECK(slot >= 0 && slot < RENDER_BATCH);
      s_rec[slot].a = make_float4(a[k].x, a[k].y, co[k].x, co[k].z);
      s_rec[slot].b = make_float4(co[k].y, co[k].w, pre[k].thr, pre[k].rBC);
      s_rec[slot].c = c;
      s_rba[slot] = pre[k].rBA;
      s_id[slot] = id[k];
    }
  }
  __syncthreads();
  return total;
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// COUNT (test instantiation, option render_bwd = 3): additionally evaluates the forward's exact alpha (expf) and
// counts, in counters[0..2], the accepted (pixel, Gaussian) pairs, the pairs the forward accepted but this kernel's
// ex2.approx alpha rejects, and the reverse (defined deviation 4).
// ---- helpers of the PAIR instantiation (small scenes, see render_fwd2_kernel<true>): a list entry for the lane's two
// pixels, split into (1) what depends on neither T nor U, (2) the recurrence on T and U with the nine sums it feeds
// into the reduction, so that two entries' shared-memory loads, exponentials and reduce-scatter butterflies overlap.
// Same operations as the loop body of the default instantiation.
struct BwdHit2 {
  float dx;
  float2 dy, G, power;
  float al0, al1;
  float4 col;
};
__device__ __forceinline__ BwdHit2 bwd2_eval(uint32_t addr, float npx, float2 npy) {
  BwdHit2 h;
  const float4 a = lds128(addr);      // mean.x, mean.y, conic A, conic C
  const float2 b = lds64(addr + 16);  // conic B, opacity
  h.col = lds128(addr + 32);
  h.dx = a.x + npx;
  h.dy = __fadd2_rn(make_float2(a.y, a.y), npy);
  const float dxA = h.dx * a.z, ndxB = -h.dx * b.x;
  const float2 t1 = __fmul2_rn(h.dy, __fmul2_rn(h.dy, make_float2(a.w, a.w)));
  const float2 s = __ffma2_rn(make_float2(h.dx, h.dx), make_float2(dxA, dxA), t1);
  const float2 nv = __fmul2_rn(h.dy, make_float2(ndxB, ndxB));
  h.power = __ffma2_rn(s, make_float2(-0.5f, -0.5f), nv);
  const float2 pl = __fmul2_rn(h.power, make_float2(1.4426950408889634f, 1.4426950408889634f));
  h.G = make_float2(ex2_approx(pl.x), ex2_approx(pl.y));
  const float2 og = __fmul2_rn(make_float2(b.y, b.y), h.G);
  h.al0 = fminf(0.99f, og.x);
  h.al1 = fminf(0.99f, og.y);
  return h;
}
struct BwdSums9 {
  float v[9];
};
__device__ __forceinline__ BwdSums9 bwd2_step(const BwdHit2& h, bool v0, bool v1, float2& T, float2& U, float2 dLr,
                                              float2 dLg, float2 dLb, float2 ntfbg) {
  const float2 alpha = make_float2(v0 ? h.al0 : 0.f, v1 ? h.al1 : 0.f);
  const float2 Ge = make_float2(v0 ? h.G.x : 0.f, v1 ? h.G.y : 0.f);
  const float2 om = __ffma2_rn(alpha, make_float2(-1.f, -1.f), make_float2(1.f, 1.f));
  const float2 rinv = make_float2(rcp_approx(om.x), rcp_approx(om.y));
  T = __fmul2_rn(T, rinv);
  const float2 cr = make_float2(h.col.x, h.col.x), cg = make_float2(h.col.y, h.col.y), cb = make_float2(h.col.z, h.col.z);
  const float2 cd = __ffma2_rn(cb, dLb, __ffma2_rn(cg, dLg, __fmul2_rn(cr, dLr)));
  const float2 cdU = __fadd2_rn(cd, make_float2(-U.x, -U.y));
  const float2 dLda = __ffma2_rn(cdU, T, __fmul2_rn(ntfbg, rinv));
  U = __ffma2_rn(alpha, cdU, U);
  const float2 wgt = __fmul2_rn(alpha, T);
  const float2 q = __fmul2_rn(Ge, dLda);
  const float2 wr = __fmul2_rn(wgt, dLr), wg = __fmul2_rn(wgt, dLg), wb = __fmul2_rn(wgt, dLb);
  const float2 qdy = __fmul2_rn(q, h.dy);
  const float2 qdy2 = __fmul2_rn(qdy, h.dy);
  const float qs = q.x + q.y, sy = qdy.x + qdy.y;
  const float sx = h.dx * qs;
  BwdSums9 r;
  r.v[0] = wr.x + wr.y; r.v[1] = wg.x + wg.y; r.v[2] = wb.x + wb.y;
  r.v[3] = qs; r.v[4] = sx; r.v[5] = sy; r.v[6] = h.dx * sx; r.v[7] = h.dx * sy; r.v[8] = qdy2.x + qdy2.y;
  return r;
}

// PAIR (with COUNT = false only): the small-scene instantiation, two list entries per step of the hit loop.
template <bool COUNT, bool PAIR = false>
__global__ void __launch_bounds__(RB2_THREADS, (COUNT || PAIR) ? 4 : GSR_RB2_CTAS) render_bwd2_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
    const float* __restrict__ bg, const float4* __restrict__ g0, const float4* __restrict__ g1,
    const float4* __restrict__ g2, const float* __restrict__ colors_precomp, const float* __restrict__ final_T,
    const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix, BwdView acc,
    unsigned long long* __restrict__ counters) {
  __shared__ StageRec s_rec[RENDER_BATCH];
  __shared__ float s_rba[RENDER_BATCH];
  __shared__ uint32_t s_id[RENDER_BATCH];
  __shared__ int s_warp[8];
  __shared__ uint32_t s_maxc;

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  if (range.y == range.x) return;
  const int pix_x = blockIdx.x * TILE + ((w & 1) << 3) + (lane & 7);
  const int pix_y0 = blockIdx.y * TILE + ((w >> 1) << 3) + (lane >> 3), pix_y1 = pix_y0 + 4;
  const bool in0 = pix_x < W && pix_y0 < H, in1 = pix_x < W && pix_y1 < H;
  const float npx = -(float)pix_x;
  const float2 npy = make_float2(-(float)pix_y0, -(float)pix_y1);
  const float tx0 = (float)(blockIdx.x * TILE), ty0 = (float)(blockIdx.y * TILE);
  const float wx0 = tx0 + (float)((w & 1) << 3), wy0 = ty0 + (float)((w >> 1) << 3);
  const size_t N = (size_t)W * H, p0 = (size_t)pix_y0 * W + pix_x, p1 = (size_t)pix_y1 * W + pix_x;
  const uint32_t rec_addr = smem_addr(s_rec);
  const uint32_t id_addr = smem_addr(s_id);

  float2 T = make_float2(in0 ? final_T[p0] : 0.f, in1 ? final_T[p1] : 0.f);
  const uint32_t lc0 = in0 ? n_contrib[p0] : 0u, lc1 = in1 ? n_contrib[p1] : 0u;
  float2 dLr = make_float2(0.f, 0.f), dLg = dLr, dLb = dLr;
  if (in0) { dLr.x = dL_dpix[p0]; dLg.x = dL_dpix[N + p0]; dLb.x = dL_dpix[2 * N + p0]; }
  if (in1) { dLr.y = dL_dpix[p1]; dLg.y = dL_dpix[N + p1]; dLb.y = dL_dpix[2 * N + p1]; }
  const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
  // -T_final * <bg, dL/dpixel>  (backward.cu:531-534), negated once here
  const float2 ntfbg = make_float2(-T.x * (bg0 * dLr.x + bg1 * dLg.x + bg2 * dLb.x),
                                   -T.y * (bg0 * dLr.y + bg1 * dLg.y + bg2 * dLb.y));
  float2 U = make_float2(0.f, 0.f);

  // list entries behind every pixel's last contributor are never touched
  uint32_t wmax = max(lc0, lc1);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
  if (tid == 0) s_maxc = 0u;
  __syncthreads();
  if (lane == 0) atomicMax(&s_maxc, wmax);
  __syncthreads();
  const int n_eff = (int)s_maxc;
  const int commit = ((lane & 3) == 0) ? (((lane >> 2) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 4) & 1))
                                       : (lane == 2 ? 8 : -1);
  float* const commit_base = commit < 0 ? nullptr
                             : commit < 4 ? reinterpret_cast<float*>(acc.acc0) + commit
                             : commit < 8 ? reinterpret_cast<float*>(acc.acc1) + (commit - 4)
                                          : acc.acc2;
  const uint32_t commit_shift = commit == 8 ? 0u : 2u;
  const float2 one2 = make_float2(1.f, 1.f);

  for (int start = 0; start < n_eff; start += RENDER_BATCH) {
    __syncthreads();  // previous batch fully consumed before it is overwritten
    const int n_batch = min(RENDER_BATCH, n_eff - start);
    const int first = n_eff - 1 - start;
    const int cnt = stage_batch2(tid, n_batch, point_list + range.x + first, first, g0, g1, g2, colors_precomp, tx0,
                                 ty0, s_rec, s_rba, s_id, s_warp);
    for (int c0 = 0; c0 < cnt; c0 += 32) {
      bool hit = false;
      if (c0 + lane < cnt && __float_as_uint(s_rec[c0 + lane].c.w) < wmax) {
        const float4 a = s_rec[c0 + lane].a, b = s_rec[c0 + lane].b;
        CullPre pre;
        pre.thr = b.z; pre.rBC = b.w; pre.rBA = s_rba[c0 + lane];
        hit = !cull_rect(a.x, a.y, a.z, b.x, a.w, pre, wx0, wx0 + 7.f, wy0, wy0 + 7.f);
      }
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      const uint32_t chunk_addr = rec_addr + (uint32_t)c0 * (uint32_t)sizeof(StageRec);
      const uint32_t chunk_id_addr = id_addr + (uint32_t)c0 * 4u;
      if constexpr (PAIR) {
        while (mask) {
          const uint32_t bitA = (uint32_t)(__ffs(mask) - 1);
          mask &= mask - 1;
          const bool two = mask != 0u;  // warp-uniform
          const uint32_t bitB = two ? (uint32_t)(__ffs(mask) - 1) : bitA;
          mask &= mask - 1;             // no-op on 0
          const BwdHit2 hA = bwd2_eval(chunk_addr + bitA * (uint32_t)sizeof(StageRec), npx, npy);
          const BwdHit2 hB = bwd2_eval(chunk_addr + bitB * (uint32_t)sizeof(StageRec), npx, npy);
          const uint32_t posA = __float_as_uint(hA.col.w), posB = __float_as_uint(hB.col.w);
          const bool vA0 = (posA < lc0) && !(hA.power.x > 0.0f) && !(hA.al0 < 1.0f / 255.0f);
          const bool vA1 = (posA < lc1) && !(hA.power.y > 0.0f) && !(hA.al1 < 1.0f / 255.0f);
          const bool vB0 = two && (posB < lc0) && !(hB.power.x > 0.0f) && !(hB.al0 < 1.0f / 255.0f);
          const bool vB1 = two && (posB < lc1) && !(hB.power.y > 0.0f) && !(hB.al1 < 1.0f / 255.0f);
          const bool anyA = __any_sync(0xffffffffu, vA0 || vA1), anyB = __any_sync(0xffffffffu, vB0 || vB1);
          if (!(anyA || anyB)) continue;
          // an entry no pixel accepts has alpha = 0 everywhere: T and U pass through unchanged, its sums are 0
          const BwdSums9 sA = bwd2_step(hA, vA0, vA1, T, U, dLr, dLg, dLb, ntfbg);
          const BwdSums9 sB = bwd2_step(hB, vB0, vB1, T, U, dLr, dLg, dLb, ntfbg);
          if (anyA && anyB) {
            const float sumA = warp_reduce_scatter9(sA.v[0], sA.v[1], sA.v[2], sA.v[3], sA.v[4], sA.v[5], sA.v[6],
                                                    sA.v[7], sA.v[8], lane);
            const float sumB = warp_reduce_scatter9(sB.v[0], sB.v[1], sB.v[2], sB.v[3], sB.v[4], sB.v[5], sB.v[6],
                                                    sB.v[7], sB.v[8], lane);
            const uint32_t offA = lds32(chunk_id_addr + bitA * 4u) << commit_shift;
            const uint32_t offB = lds32(chunk_id_addr + bitB * 4u) << commit_shift;
            if (commit >= 0) {
              atomicAdd(commit_base + offA, sumA);
              atomicAdd(commit_base + offB, sumB);
            }
          } else {
            const BwdSums9& s1 = anyA ? sA : sB;
            const float sum = warp_reduce_scatter9(s1.v[0], s1.v[1], s1.v[2], s1.v[3], s1.v[4], s1.v[5], s1.v[6],
                                                   s1.v[7], s1.v[8], lane);
            const uint32_t off = lds32(chunk_id_addr + (anyA ? bitA : bitB) * 4u) << commit_shift;
            if (commit >= 0) atomicAdd(commit_base + off, sum);
          }
        }
      } else
      while (mask) {
        const uint32_t bit = (uint32_t)(__ffs(mask) - 1);
        mask &= mask - 1;
        const uint32_t addr = chunk_addr + bit * (uint32_t)sizeof(StageRec);
        const float4 a = lds128(addr);     // mean.x, mean.y, conic A, conic C
        const float2 b = lds64(addr + 16);  // conic B, opacity
        const float4 col = lds128(addr + 32);
        // power for both pixels (same expression as forward.cu:337-339; dx is common to the two)
        const float dx = a.x + npx;
        const float2 dy = __fadd2_rn(make_float2(a.y, a.y), npy);
        const float dxA = dx * a.z, ndxB = -dx * b.x;
        const float2 t1 = __fmul2_rn(dy, __fmul2_rn(dy, make_float2(a.w, a.w)));
        const float2 s = __ffma2_rn(make_float2(dx, dx), make_float2(dxA, dxA), t1);
        const float2 nv = __fmul2_rn(dy, make_float2(ndxB, ndxB));
        const float2 power = __ffma2_rn(s, make_float2(-0.5f, -0.5f), nv);
        // exp(power) as ex2.approx(power * log2 e), see render_bwd_kernel
        const float2 pl = __fmul2_rn(power, make_float2(1.4426950408889634f, 1.4426950408889634f));
        const float2 G = make_float2(ex2_approx(pl.x), ex2_approx(pl.y));
        const float2 og = __fmul2_rn(make_float2(b.y, b.y), G);
        const float al0 = fminf(0.99f, og.x), al1 = fminf(0.99f, og.y);
        const uint32_t pos = __float_as_uint(col.w);
        // same acceptance as the forward (backward.cu:488-497)
        const bool v0 = (pos < lc0) && !(power.x > 0.0f) && !(al0 < 1.0f / 255.0f);
        const bool v1 = (pos < lc1) && !(power.y > 0.0f) && !(al1 < 1.0f / 255.0f);
        if (COUNT) {
          const float e0 = fminf(0.99f, __fmul_rn(b.y, expf(power.x))), e1 = fminf(0.99f, __fmul_rn(b.y, expf(power.y)));
          const bool f0 = (pos < lc0) && !(power.x > 0.0f) && !(e0 < 1.0f / 255.0f);
          const bool f1 = (pos < lc1) && !(power.y > 0.0f) && !(e1 < 1.0f / 255.0f);
          const unsigned n_acc = __popc(__ballot_sync(0xffffffffu, f0)) + __popc(__ballot_sync(0xffffffffu, f1));
          const unsigned n_lost = __popc(__ballot_sync(0xffffffffu, f0 && !v0)) +
                                  __popc(__ballot_sync(0xffffffffu, f1 && !v1));
          const unsigned n_extra = __popc(__ballot_sync(0xffffffffu, v0 && !f0)) +
                                   __popc(__ballot_sync(0xffffffffu, v1 && !f1));
          if (lane == 0) {
            if (n_acc) atomicAdd(counters, (unsigned long long)n_acc);
            if (n_lost) atomicAdd(counters + 1, (unsigned long long)n_lost);
            if (n_extra) atomicAdd(counters + 2, (unsigned long long)n_extra);
          }
        }
        if (!__any_sync(0xffffffffu, v0 || v1)) continue;
        const float2 alpha = make_float2(v0 ? al0 : 0.f, v1 ? al1 : 0.f);
        const float2 Ge = make_float2(v0 ? G.x : 0.f, v1 ? G.y : 0.f);
        const float2 om = __ffma2_rn(alpha, make_float2(-1.f, -1.f), one2);  // in [0.01, 1]; exactly 1 when skipped
        const float2 rinv = make_float2(rcp_approx(om.x), rcp_approx(om.y));
        T = __fmul2_rn(T, rinv);
        const float2 cr = make_float2(col.x, col.x), cg = make_float2(col.y, col.y), cb = make_float2(col.z, col.z);
        const float2 cd = __ffma2_rn(cb, dLb, __ffma2_rn(cg, dLg, __fmul2_rn(cr, dLr)));  // <colour, dLp>
        const float2 cdU = __fadd2_rn(cd, make_float2(-U.x, -U.y));
        // dL/dalpha = (c - accum_rec) * T - T_final/(1-alpha) * <bg, dLp>
        const float2 dLda = __ffma2_rn(cdU, T, __fmul2_rn(ntfbg, rinv));
        U = __ffma2_rn(alpha, cdU, U);
        const float2 wgt = __fmul2_rn(alpha, T);   // dchannel/dcolour
        const float2 q = __fmul2_rn(Ge, dLda);     // dL/dopacity; common factor of the 2-D gradients
        // sums over the lane's two pixels
        const float2 wr = __fmul2_rn(wgt, dLr), wg = __fmul2_rn(wgt, dLg), wb = __fmul2_rn(wgt, dLb);
        const float2 qdy = __fmul2_rn(q, dy);
        const float2 qdy2 = __fmul2_rn(qdy, dy);
        const float qs = q.x + q.y, sy = qdy.x + qdy.y;
        const float v0_ = wr.x + wr.y, v1_ = wg.x + wg.y, v2_ = wb.x + wb.y;
        const float sx = dx * qs;
        const float sum = warp_reduce_scatter9(v0_, v1_, v2_, qs, sx, sy, dx * sx, dx * sy, qdy2.x + qdy2.y, lane);
        const uint32_t off = lds32(chunk_id_addr + bit * 4u) << commit_shift;
        if (commit >= 0) atomicAdd(commit_base + off, sum);
      }
    }
  }
}

}  // namespace gsr
