// csrc/render_bwd.cuh -- backward of the alpha compositing, one CTA per tile.
// Behaviour: DGR/cuda_rasterizer/backward.cu:399-557 (renderCUDA): back-to-front
// replay from final_T / n_contrib, same power / alpha tests as the forward.
//
// B200 design: the reference issues 9 float atomics per contributing
// (pixel, Gaussian) pair.  Here the 9 partial gradients are first summed across
// the warp with shuffles (only when some lane contributes), and one lane issues
// two 128-bit vector REDs + one scalar RED per (warp, Gaussian) into 36-byte
// per-Gaussian accumulators: 32x fewer atomic operations, 3 instead of 9
// instructions each.
#pragma once
#include "common.cuh"

namespace gsr {

__global__ void __launch_bounds__(256) render_bwd_kernel(const uint2* __restrict__ ranges,
                                                         const uint32_t* __restrict__ point_list, int W, int H,
                                                         const float* __restrict__ bg, const float4* __restrict__ g0,
                                                         const float4* __restrict__ g1, const float4* __restrict__ g2,
                                                         const float* __restrict__ colors_precomp,
                                                         const float* __restrict__ final_T,
                                                         const uint32_t* __restrict__ n_contrib,
                                                         const float* __restrict__ dL_dpix, float4* __restrict__ accA,
                                                         float4* __restrict__ accB, float* __restrict__ accC) {
  __shared__ uint32_t s_id[256];
  __shared__ float2 s_xy[256];
  __shared__ float4 s_co[256];
  __shared__ float4 s_rgb[256];

  const int tid = threadIdx.y * TILE + threadIdx.x;
  const int lane = tid & 31;
  const int pix_x = blockIdx.x * TILE + threadIdx.x, pix_y = blockIdx.y * TILE + threadIdx.y;
  const bool inside = pix_x < W && pix_y < H;
  const float pfx = (float)pix_x, pfy = (float)pix_y;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const int n = (int)(range.y - range.x);
  if (n == 0) return;
  const int rounds = (n + 255) / 256;
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

  // the whole CTA can skip list entries behind every pixel's last contributor
  __shared__ uint32_t s_maxc;
  if (tid == 0) s_maxc = 0u;
  __syncthreads();
  atomicMax(&s_maxc, last_contributor);
  __syncthreads();
  const uint32_t max_contrib = s_maxc;
  if (max_contrib == 0u) return;

  uint32_t contributor = (uint32_t)n;
  int todo = n;
  for (int rd = 0; rd < rounds; rd++, todo -= 256) {
    __syncthreads();
    // entries of this round have list index in (contributor-256, contributor-1]
    const int progress = rd * 256 + tid;
    const bool round_live = (uint32_t)(todo - min(256, todo)) < max_contrib;
    if (round_live && progress < n) {
      const uint32_t id = __ldg(point_list + range.y - progress - 1);
      s_id[tid] = id;
      const float4 a = __ldg(g0 + id);
      s_xy[tid] = make_float2(a.x, a.y);
      s_co[tid] = __ldg(g1 + id);
      if (colors_precomp != nullptr)
        s_rgb[tid] = make_float4(__ldg(colors_precomp + 3 * (size_t)id), __ldg(colors_precomp + 3 * (size_t)id + 1),
                                 __ldg(colors_precomp + 3 * (size_t)id + 2), 0.f);
      else
        s_rgb[tid] = __ldg(g2 + id);
    }
    __syncthreads();
    const int cnt = min(256, todo);
    if (!round_live) {
      contributor -= cnt;
      continue;
    }
    for (int j = 0; j < cnt; j++) {
      contributor--;
      bool valid = contributor < last_contributor;
      float G = 0.f, alpha = 0.f, dx = 0.f, dy = 0.f;
      float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) {
        const float2 xy = s_xy[j];
        co = s_co[j];
        dx = __fsub_rn(xy.x, pfx);
        dy = __fsub_rn(xy.y, pfy);
        const float power = gauss_power(dx, dy, co.x, co.y, co.z);
        valid = !(power > 0.0f);
        if (valid) {
          G = expf(power);
          alpha = fminf(0.99f, __fmul_rn(co.w, G));
          valid = !(alpha < 1.0f / 255.0f);
        }
      }
      if (!__any_sync(0xffffffffu, valid)) continue;
      float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, v4 = 0.f, v5 = 0.f, v6 = 0.f, v7 = 0.f, v8 = 0.f;
      if (valid) {
        T = T / (1.f - alpha);
        const float dchannel_dcolor = alpha * T;
        const float4 col = s_rgb[j];
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
        const float dL_dG = co.w * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * co.x - gdy * co.y;
        const float dG_ddely = -gdy * co.z - gdx * co.y;
        v3 = G * dL_dalpha;                  // dL_dopacity
        v4 = dL_dG * dG_ddelx * ddelx_dx;    // dL_dmean2D.x
        v5 = dL_dG * dG_ddely * ddely_dy;    // dL_dmean2D.y
        v6 = -0.5f * gdx * dx * dL_dG;       // dL_dconic.xx
        v7 = -0.5f * gdx * dy * dL_dG;       // dL_dconic.xy
        v8 = -0.5f * gdy * dy * dL_dG;       // dL_dconic.yy
      }
      v0 = warp_sum(v0); v1 = warp_sum(v1); v2 = warp_sum(v2);
      v3 = warp_sum(v3); v4 = warp_sum(v4); v5 = warp_sum(v5);
      v6 = warp_sum(v6); v7 = warp_sum(v7); v8 = warp_sum(v8);
      if (lane == 0) {
        const uint32_t id = s_id[j];
        red_add_v4(accA + id, v0, v1, v2, v3);
        red_add_v4(accB + id, v4, v5, v6, v7);
        atomicAdd(accC + id, v8);
      }
    }
  }
}

}  // namespace gsr
