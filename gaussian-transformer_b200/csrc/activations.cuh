// csrc/activations.cuh -- the parameter activations in front of the rasterizer:
// scene/gaussian_model.py:95-106 get_scaling = exp(_scaling), get_rotation = normalize(_rotation),
// get_opacity = sigmoid(_opacity), as set up at :43-51.  The reference runs them as ~8 elementwise / reduce kernels
// forward and ~17 backward on every iteration; here each direction is one pass over 32 B per Gaussian.
// The expressions are the ones torch's CUDA kernels evaluate (sigmoid: 1 / (1 + expf(-x)), exp: expf,
// F.normalize: x / max(||x||_2, 1e-12)), so that the rasterizer sees the same inputs.
#pragma once
#include "common.cuh"

namespace gsr {

// (quat_norm / act_sigmoid / act_normalize live in common.cuh: preprocess_fwd / preprocess_bwd apply the same
// expressions in place when the rasterizer is handed the raw parameters, gsr_inputs.raw_params.)
__global__ void __launch_bounds__(256) activate_kernel(int P, const float* __restrict__ opacity_raw,
                                                       const float* __restrict__ scaling_raw,
                                                       const float4* __restrict__ rotation_raw,
                                                       float* __restrict__ opacities, float* __restrict__ scales,
                                                       float4* __restrict__ rotations) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const size_t I = (size_t)i;
  opacities[I] = act_sigmoid(__ldg(opacity_raw + I));
#pragma unroll
  for (int k = 0; k < 3; k++) scales[3 * I + k] = expf(__ldg(scaling_raw + 3 * I + k));
  rotations[I] = act_normalize(__ldg(rotation_raw + I));
}

// Chain rule of the three activations; the activated values are recomputed from the raw parameters.
__global__ void __launch_bounds__(256) activate_bwd_kernel(
    int P, const float* __restrict__ opacity_raw, const float* __restrict__ scaling_raw,
    const float4* __restrict__ rotation_raw, const float* __restrict__ g_opac, const float* __restrict__ g_scales,
    const float4* __restrict__ g_rot, float* __restrict__ d_opacity_raw, float* __restrict__ d_scaling_raw,
    float4* __restrict__ d_rotation_raw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const size_t I = (size_t)i;
  const float o = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-__ldg(opacity_raw + I))));
  d_opacity_raw[I] = __ldg(g_opac + I) * (1.0f - o) * o;  // sigmoid_backward: grad * (1 - y) * y
#pragma unroll
  for (int k = 0; k < 3; k++) d_scaling_raw[3 * I + k] = __ldg(g_scales + 3 * I + k) * expf(__ldg(scaling_raw + 3 * I + k));
  const float4 q = __ldg(rotation_raw + I), g = __ldg(g_rot + I);
  const float nr = quat_norm(q);
  const float n = fmaxf(nr, 1e-12f);
  const float inv = 1.0f / n;
  const float4 u = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
  // y = q / n: dq = g / n - (norm > eps ? u * <u, g> / n : 0)
  const float dot = nr > 1e-12f ? (u.x * g.x + u.y * g.y + u.z * g.z + u.w * g.w) : 0.f;
  d_rotation_raw[I] = make_float4((g.x - u.x * dot) * inv, (g.y - u.y * dot) * inv, (g.z - u.z * dot) * inv,
                                  (g.w - u.w * dot) * inv);
}

}  // namespace gsr

namespace gsr {

// Per-iteration densification statistics (train.py:113-116, scene/gaussian_model.py:405-407), for the Gaussians the
// view saw (radii > 0):  max_radii2D = max(max_radii2D, radii);  xyz_gradient_accum += ||dL/dmeans2D.xy||;
// denom += 1.  The reference does this with boolean-mask indexing (a nonzero() host synchronisation, gathers and
// index_puts); here it is one pass over 28 B per Gaussian with no synchronisation.  The 2-norm is torch's
// sqrt(x*x + y*y) with both products rounded.
__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const int* __restrict__ radii,
                                                            const float* __restrict__ dL_dmeans2D,
                                                            float* __restrict__ max_radii2D,
                                                            float* __restrict__ xyz_gradient_accum,
                                                            float* __restrict__ denom) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const int r = __ldg(radii + i);
  if (r <= 0) return;
  const size_t I = (size_t)i;
  max_radii2D[I] = fmaxf(max_radii2D[I], (float)r);
  const float gx = __ldg(dL_dmeans2D + 3 * I), gy = __ldg(dL_dmeans2D + 3 * I + 1);
  xyz_gradient_accum[I] = __fadd_rn(xyz_gradient_accum[I], __fsqrt_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy))));
  denom[I] = __fadd_rn(denom[I], 1.0f);
}

}  // namespace gsr
