// csrc/adam.cuh -- the optimiser step that consumes the path's gradients:
// torch.optim.Adam(l, lr=0.0, eps=1e-15) over the six parameter groups of scene/gaussian_model.py:149-167
// (xyz, f_dc, f_rest, opacity, scaling, rotation; per-group lr, betas (0.9, 0.999), no weight decay, no amsgrad).
//
// The reference's default (foreach) implementation runs 8 multi-tensor kernels per step and re-reads every array
// several times (~76 B/element); this is ONE launch for all groups at the compulsory 28 B/element
// (read p, g, m, v; write p, m, v).  Per-element arithmetic follows torch's op order
// (torch/optim/adam.py _multi_tensor_adam: lerp, mul, addcmul, sqrt, div, add, addcdiv) so results agree to
// rounding (tolerance in tests/test_gpu_adam.py); the bias corrections are formed on the host in double
// precision exactly as torch does.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int ADAM_MAX_GROUPS = 8;
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_CHUNK = ADAM_THREADS * 16;  // elements per CTA

struct AdamGroup {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  float neg_step_size;  // -(lr / (1 - beta1^t))
  float bc2_sqrt;       // sqrt(1 - beta2^t)
  int first_block;      // prefix sum of chunk counts
  int vec;              // all four arrays 16-byte aligned
};

struct AdamArgs {
  AdamGroup grp[ADAM_MAX_GROUPS];
  int n_groups;
  float w1;     // 1 - beta1
  float beta2;
  float w2;     // 1 - beta2
  float eps;
};

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const AdamArgs& a,
                                          const AdamGroup& G) {
  m = fmaf(a.w1, g - m, m);                           // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(__fmul_rn(a.w2, g), g, __fmul_rn(v, a.beta2));  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = __fadd_rn(__fdiv_rn(sqrtf(v), G.bc2_sqrt), a.eps);
  p = fmaf(G.neg_step_size, __fdiv_rn(m, denom), p);  // param.addcdiv_(exp_avg, denom, value = -step_size)
}

__global__ void __launch_bounds__(ADAM_THREADS) adam_step_kernel(const __grid_constant__ AdamArgs a) {
  int gi = 0;
#pragma unroll
  for (int k = 1; k < ADAM_MAX_GROUPS; k++)
    if (k < a.n_groups && (int)blockIdx.x >= a.grp[k].first_block) gi = k;
  const AdamGroup& G = a.grp[gi];
  const long long base = (long long)((int)blockIdx.x - G.first_block) * ADAM_CHUNK;
  const long long end = min(G.n, base + ADAM_CHUNK);
  if (G.vec && base + ADAM_CHUNK <= G.n) {
#pragma unroll
    for (int r = 0; r < ADAM_CHUNK / (ADAM_THREADS * 4); r++) {
      const long long i = base + (long long)(r * ADAM_THREADS + threadIdx.x) * 4;
      float4 p = *reinterpret_cast<const float4*>(G.p + i);
      const float4 g = __ldcs(reinterpret_cast<const float4*>(G.g + i));
      float4 m = *reinterpret_cast<const float4*>(G.m + i);
      float4 v = *reinterpret_cast<const float4*>(G.v + i);
      adam_elem(p.x, g.x, m.x, v.x, a, G);
      adam_elem(p.y, g.y, m.y, v.y, a, G);
      adam_elem(p.z, g.z, m.z, v.z, a, G);
      adam_elem(p.w, g.w, m.w, v.w, a, G);
      *reinterpret_cast<float4*>(G.p + i) = p;
      *reinterpret_cast<float4*>(G.m + i) = m;
      *reinterpret_cast<float4*>(G.v + i) = v;
    }
  } else {
    for (long long i = base + threadIdx.x; i < end; i += ADAM_THREADS) {
      float p = G.p[i], m = G.m[i], v = G.v[i];
      adam_elem(p, G.g[i], m, v, a, G);
      G.p[i] = p;
      G.m[i] = m;
      G.v[i] = v;
    }
  }
}

}  // namespace gsr
