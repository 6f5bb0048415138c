// csrc/sh_exchange.cuh -- the two kernels behind the data-parallel SH gradient exchange (gsr_dp.py).
//
// For ONE view the SH gradient of a Gaussian is a rank-1 product (DGR/cuda_rasterizer/backward.cu:20-139):
//     dL/dsh(i, k, c) = basis_k(dir_i) * f(i, c),   f = dL/dcolor with the forward's clamped channels zeroed,
// dir_i = normalize(mean_i - campos).  Gaussians are replicated on every rank, so a rank can rebuild the SH
// gradient of ANY view from that view's camera position and its 3 floats per Gaussian.  Data-parallel ranks
// therefore all-gather the factors (12 B per Gaussian per view) instead of all-reducing 12*M B per Gaussian
// (192 B at M = 16) and sum the outer products locally; preprocess_bwd then does not write dL/dsh at all.
#pragma once
#include "common.cuh"
#include "preprocess_bwd.cuh"
#include "preprocess_fwd.cuh"

namespace gsr {

// factors[i] = dL_dcolors[i] with the channels the forward clamped (g2 flags bits 0..2) set to zero.
__global__ void __launch_bounds__(256) sh_color_factors_kernel(int P, const float4* __restrict__ g2,
                                                               const float* __restrict__ dL_dcolors,
                                                               float* __restrict__ factors) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const uint32_t flags = __float_as_uint(__ldg(&g2[i].w));
  const size_t I = (size_t)i;
  factors[3 * I] = (flags & 1u) ? 0.f : dL_dcolors[3 * I];
  factors[3 * I + 1] = (flags & 2u) ? 0.f : dL_dcolors[3 * I + 1];
  factors[3 * I + 2] = (flags & 4u) ? 0.f : dL_dcolors[3 * I + 2];
}

// The same factors read from the blend accumulators (acc0 = dL_dcolor.rgb, dL_dopacity) between the two backward
// stages, so that their exchange can start before preprocess_bwd runs.
__global__ void __launch_bounds__(256) sh_acc_factors_kernel(int P, const float4* __restrict__ g2,
                                                             const float4* __restrict__ acc0,
                                                             float* __restrict__ factors) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const uint32_t flags = __float_as_uint(__ldg(&g2[i].w));
  const float4 a = __ldg(acc0 + i);
  const size_t I = (size_t)i;
  factors[3 * I] = (flags & 1u) ? 0.f : a.x;
  factors[3 * I + 1] = (flags & 2u) ? 0.f : a.y;
  factors[3 * I + 2] = (flags & 4u) ? 0.f : a.z;
}

// dL_dsh[i] (+)= sum_v basis(normalize(mean_i - campos_v)) (x) factors[v][i].  One thread per Gaussian, 48
// accumulators in registers; STAGED (M == 16, 16-byte aligned output): rows leave through shared memory with
// coalesced 128-bit stores like preprocess_bwd_staged_kernel.
template <bool STAGED>
__global__ void __launch_bounds__(PB_THREADS) sh_grad_from_factors_kernel(
    int P, int D, int M, int V, const float* __restrict__ means3D, const float* __restrict__ campos,
    const float* __restrict__ factors, float* __restrict__ dL_dsh, int accumulate) {
  extern __shared__ float4 s_tile[];  // STAGED: [PB_THREADS][13]
  const int tid = threadIdx.x;
  const int base = blockIdx.x * PB_THREADS;
  const int i = base + tid;
  const int nG = min(PB_THREADS, P - base);
  const size_t I = (size_t)i;
  float v[48];
#pragma unroll
  for (int k = 0; k < 48; k++) v[k] = 0.f;
  if (i < P) {
    const float p0 = __ldg(means3D + 3 * I), p1 = __ldg(means3D + 3 * I + 1), p2 = __ldg(means3D + 3 * I + 2);
    for (int w = 0; w < V; w++) {
      const float* f = factors + ((size_t)w * P + I) * 3;
      const float f0 = __ldg(f), f1 = __ldg(f + 1), f2 = __ldg(f + 2);
      if (f0 == 0.f && f1 == 0.f && f2 == 0.f) continue;  // invisible in this view (or fully clamped)
      const float dox = p0 - __ldg(campos + 3 * w), doy = p1 - __ldg(campos + 3 * w + 1),
                  doz = p2 - __ldg(campos + 3 * w + 2);
      const float len = sqrtf(dox * dox + doy * doy + doz * doz);
      float basis[16];
#pragma unroll
      for (int k = 0; k < 16; k++) basis[k] = 0.f;
      sh_basis(D, dox / len, doy / len, doz / len, basis);
#pragma unroll
      for (int k = 0; k < 16; k++) {
        v[3 * k] = fmaf(basis[k], f0, v[3 * k]);
        v[3 * k + 1] = fmaf(basis[k], f1, v[3 * k + 1]);
        v[3 * k + 2] = fmaf(basis[k], f2, v[3 * k + 2]);
      }
    }
  }
  if (STAGED) {  // M == 16
    float4* row4 = s_tile + tid * 13;
    if (i < P) {
#pragma unroll
      for (int k = 0; k < 12; k++) row4[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
    __syncthreads();
    float4* __restrict__ dst4 = reinterpret_cast<float4*>(dL_dsh + (size_t)base * 48);
#pragma unroll 4
    for (int j = tid; j < nG * 12; j += PB_THREADS) {
      const int row = j / 12, col = j - row * 12;
      float4 q = s_tile[row * 13 + col];
      if (accumulate) {
        const float4 o = dst4[j];
        q.x += o.x; q.y += o.y; q.z += o.z; q.w += o.w;
      }
      __stcs(dst4 + j, q);
    }
  } else if (i < P) {
    float* dst = dL_dsh + I * M * 3;
    const int nf = 3 * M;
#pragma unroll
    for (int k = 0; k < 48; k++)
      if (k < nf) dst[k] = accumulate ? dst[k] + v[k] : v[k];
    if (!accumulate)
      for (int k = 48; k < nf; k++) dst[k] = 0.f;
  }
}

}  // namespace gsr
