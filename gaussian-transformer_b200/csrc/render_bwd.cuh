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
//   * those 9 lanes issue ONE predicated RED instruction into a 48-byte
//     per-Gaussian accumulator row (9 floats used), i.e. one atomic instruction
//     per (warp, Gaussian) instead of 9 per (pixel, Gaussian).
#pragma once
#include "common.cuh"
#include "cull.cuh"
#include "render_fwd.cuh"

namespace gsr {

constexpr int ACC_STRIDE = 12;  // floats per Gaussian in the accumulator array (9 used)

// pair step of the reduce-scatter: lanes with `hi` clear end up with sum(a), lanes with it set with sum(b)
__device__ __forceinline__ float rs_pair(float a, float b, bool hi, int xorm) {
  const float send = hi ? a : b;
  const float keep = hi ? b : a;
  return keep + __shfl_xor_sync(0xffffffffu, send, xorm);
}

// After the call, lane l with (l & 3) == 0 holds the warp-wide sum of v[4*b2 + 2*b3 + b4]
// (b2 = bit 2 of l, ...), and lanes with (l & 2) hold the sum of v8.
__device__ __forceinline__ float warp_reduce_scatter9(float v0, float v1, float v2, float v3, float v4, float v5,
                                                      float v6, float v7, float v8, int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  const float r0 = rs_pair(v0, v1, b4, 16), r1 = rs_pair(v2, v3, b4, 16), r2 = rs_pair(v4, v5, b4, 16),
              r3 = rs_pair(v6, v7, b4, 16);
  float s = v8 + __shfl_xor_sync(0xffffffffu, v8, 16);
  const float t0 = rs_pair(r0, r1, b3, 8), t1 = rs_pair(r2, r3, b3, 8);
  s += __shfl_xor_sync(0xffffffffu, s, 8);
  const float u0 = rs_pair(t0, t1, b2, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  float w = rs_pair(u0, s, b1, 2);
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
                                                         const float* __restrict__ dL_dpix, float* __restrict__ acc) {
  __shared__ StageRec s_rec[RENDER_BATCH];
  __shared__ uint32_t s_pos[RENDER_BATCH];
  __shared__ uint32_t s_id[RENDER_BATCH];
  __shared__ int s_warp[8];
  __shared__ uint32_t s_maxc;

  const int tid = threadIdx.x, lane = tid & 31;
  int lx, ly;
  warp_pixel(tid, lx, ly);
  const int pix_x = blockIdx.x * TILE + lx, pix_y = blockIdx.y * TILE + ly;
  const bool inside = pix_x < W && pix_y < H;
  const float pfx = (float)pix_x, pfy = (float)pix_y;
  const float tx0 = (float)(blockIdx.x * TILE), ty0 = (float)(blockIdx.y * TILE);
  const float wx0 = tx0 + (float)(((tid >> 5) & 1) << 3), wy0 = ty0 + (float)(((tid >> 5) >> 1) << 2);
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  if (range.y == range.x) return;
  const size_t N = (size_t)W * H, pix = (size_t)pix_y * W + pix_x;

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
  float acr0 = 0.f, acr1 = 0.f, acr2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

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

  for (int start = 0; start < n_eff; start += RENDER_BATCH) {
    __syncthreads();  // previous batch fully consumed before it is overwritten
    const int n_batch = min(RENDER_BATCH, n_eff - start);
    const int first = n_eff - 1 - start;
    const int cnt = stage_batch(tid, n_batch, point_list + range.x + first, -1, first, g0, g1, g2, colors_precomp,
                                tx0, ty0, s_rec, s_pos, s_id, s_warp);
    for (int c0 = 0; c0 < cnt; c0 += 32) {
      bool hit = false;
      if (c0 + lane < cnt && s_pos[c0 + lane] < wmax) {
        const float4 a = s_rec[c0 + lane].a, b = s_rec[c0 + lane].b;
        CullPre pre;
        pre.thr = b.z; pre.rBC = b.w; pre.rBA = s_rec[c0 + lane].c.w;
        hit = !cull_rect(a.x, a.y, a.z, a.w, b.x, pre, wx0, wx0 + 7.f, wy0, wy0 + 3.f);
      }
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int j = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const float4 a = s_rec[j].a, b = s_rec[j].b;
        bool valid = s_pos[j] < last_contributor;
        float G = 0.f, alpha = 0.f;
        const float dx = __fsub_rn(a.x, pfx), dy = __fsub_rn(a.y, pfy);
        if (valid) {
          const float power = gauss_power(dx, dy, a.z, a.w, b.x);
          valid = !(power > 0.0f);
          if (valid) {
            G = expf(power);
            alpha = fminf(0.99f, __fmul_rn(b.y, G));
            valid = !(alpha < 1.0f / 255.0f);
          }
        }
        if (!__any_sync(0xffffffffu, valid)) continue;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f;
        if (valid) {
          T = T / (1.f - alpha);
          const float dchannel_dcolor = alpha * T;
          const float4 col = s_rec[j].c;
          acr0 = last_alpha * lc0 + (1.f - last_alpha) * acr0;
          acr1 = last_alpha * lc1 + (1.f - last_alpha) * acr1;
          acr2 = last_alpha * lc2 + (1.f - last_alpha) * acr2;
          lc0 = col.x; lc1 = col.y; lc2 = col.z;
          float dL_dalpha = (col.x - acr0) * dLp0 + (col.y - acr1) * dLp1 + (col.z - acr2) * dLp2;
          v0 = dchannel_dcolor * dLp0;
          v1 = dchannel_dcolor * dLp1;
          v2 = dchannel_dcolor * dLp2;
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
          const float dL_dG = b.y * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * a.z - gdy * a.w;
          const float dG_ddely = -gdy * b.x - gdx * a.w;
          v3 = G * dL_dalpha;                // dL_dopacity
          v4 = dL_dG * dG_ddelx * ddelx_dx;  // dL_dmean2D.x
          v5 = dL_dG * dG_ddely * ddely_dy;  // dL_dmean2D.y
          v6 = -0.5f * gdx * dx * dL_dG;     // dL_dconic.xx
          v7 = -0.5f * gdx * dy * dL_dG;     // dL_dconic.xy
          v8 = -0.5f * gdy * dy * dL_dG;     // dL_dconic.yy
        }
        const float sum = warp_reduce_scatter9(v0, v1, v2, v3, v4, v5, v6, v7, v8, lane);
        if (commit >= 0) atomicAdd(acc + (size_t)s_id[j] * ACC_STRIDE + commit, sum);
      }
    }
  }
}

}  // namespace gsr
