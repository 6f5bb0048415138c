// csrc/preprocess_bwd.cuh -- per-Gaussian backward stage, ONE fused kernel.
// Behaviour: DGR/cuda_rasterizer/backward.cu:144-274 (computeCov2DCUDA) followed
// by :346-396 (preprocessCUDA) with computeColorFromSH :20-139, computeCov3D
// :278-341 and dnormvdv (auxiliary.h:107-117).
//
// B200 design: the reference runs two kernels that re-read the same parameters
// and relies on nine memset gradient tensors.  This kernel reads the 36-byte
// accumulators written by render_bwd once, recomputes cov3D from scale/rotation
// instead of round-tripping it through HBM, and writes every output gradient in
// full (zeros for Gaussians with radii <= 0), so no output memset pass exists.
// It is HBM-bound (about 0.5 KB per Gaussian), so all array-of-struct traffic
// goes through shared memory: a CTA owns 128 consecutive Gaussians, whose SH
// rows (128 x 192 B at M = 16) and 12/24-byte gradient rows are contiguous in
// global memory and are moved with coalesced 128-bit loads / stores; each
// thread then works on its own (bank-conflict-free, padded) row.
// Gradients are tolerance-checked (1e-3 relative), so plain fp32 expressions.
#pragma once
#include "common.cuh"
#include "preprocess_fwd.cuh"

namespace gsr {

struct BwdOut {
  float* dL_dmeans3D;  // (P,3)
  float* dL_dmeans2D;  // (P,3)
  float* dL_dsh;       // (P,M,3) or NULL; split SH storage: the DC rows (P,1,3)
  float* dL_dsh_rest;  // split SH storage: (P,M-1,3), else unused
  float* dL_dcolors;   // (P,3)
  float* dL_dopacity;  // (P)
  float* dL_dscales;   // (P,3)
  float* dL_drot;      // (P,4)
  float* dL_dcov3D;    // (P,6)
  float* dL_dconic;    // (P,4)
};

constexpr int PB_THREADS = 128;

// SH backward for one Gaussian (backward.cu:20-139), in place on a register
// array: on entry v[3k+c] = coefficient (k, c) for k < K; on exit v[3k+c] =
// dL/dsh(k, c) for k < K and 0 for the other rows.  Returns dL/d(dir).
template <int NF>
__device__ __forceinline__ void sh_backward_inplace(int deg, float* v, float x, float y, float z, const float* dRGB,
                                                    float* ddir) {
  float dx_[3] = {0.f, 0.f, 0.f}, dy_[3] = {0.f, 0.f, 0.f}, dz_[3] = {0.f, 0.f, 0.f};
#define SHC(k, c) ((3 * (k) + (c)) < NF ? v[(3 * (k) + (c)) < NF ? (3 * (k) + (c)) : 0] : 0.f)
#define DSH(k, c, val) do { if ((3 * (k) + (c)) < NF) v[(3 * (k) + (c)) < NF ? (3 * (k) + (c)) : 0] = (val); } while (0)
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float g = dRGB[c];
    if (deg > 0) {
      dx_[c] = -SH_C1 * SHC(3, c);
      dy_[c] = -SH_C1 * SHC(1, c);
      dz_[c] = SH_C1 * SHC(2, c);
      if (deg > 1) {
        const float s4 = SHC(4, c), s5 = SHC(5, c), s6 = SHC(6, c), s7 = SHC(7, c), s8 = SHC(8, c);
        dx_[c] += SH_C2_0 * y * s4 + SH_C2_2 * 2.f * -x * s6 + SH_C2_3 * z * s7 + SH_C2_4 * 2.f * x * s8;
        dy_[c] += SH_C2_0 * x * s4 + SH_C2_1 * z * s5 + SH_C2_2 * 2.f * -y * s6 + SH_C2_4 * 2.f * -y * s8;
        dz_[c] += SH_C2_1 * y * s5 + SH_C2_2 * 2.f * 2.f * z * s6 + SH_C2_3 * x * s7;
        if (deg > 2) {
          const float s9 = SHC(9, c), s10 = SHC(10, c), s11 = SHC(11, c), s12 = SHC(12, c), s13 = SHC(13, c),
                      s14 = SHC(14, c), s15 = SHC(15, c);
          dx_[c] += SH_C3_0 * s9 * 3.f * 2.f * xy + SH_C3_1 * s10 * yz + SH_C3_2 * s11 * -2.f * xy +
                    SH_C3_3 * s12 * -3.f * 2.f * xz + SH_C3_4 * s13 * (-3.f * xx + 4.f * zz - yy) +
                    SH_C3_5 * s14 * 2.f * xz + SH_C3_6 * s15 * 3.f * (xx - yy);
          dy_[c] += SH_C3_0 * s9 * 3.f * (xx - yy) + SH_C3_1 * s10 * xz +
                    SH_C3_2 * s11 * (-3.f * yy + 4.f * zz - xx) + SH_C3_3 * s12 * -3.f * 2.f * yz +
                    SH_C3_4 * s13 * -2.f * xy + SH_C3_5 * s14 * -2.f * yz + SH_C3_6 * s15 * -3.f * 2.f * xy;
          dz_[c] += SH_C3_1 * s10 * xy + SH_C3_2 * s11 * 4.f * 2.f * yz +
                    SH_C3_3 * s12 * 3.f * (2.f * zz - xx - yy) + SH_C3_4 * s13 * 4.f * 2.f * xz +
                    SH_C3_5 * s14 * (xx - yy);
        }
      }
    }
    // all coefficient reads of this channel are done: overwrite with the gradients
    DSH(0, c, SH_C0 * g);
    DSH(1, c, deg > 0 ? -SH_C1 * y * g : 0.f);
    DSH(2, c, deg > 0 ? SH_C1 * z * g : 0.f);
    DSH(3, c, deg > 0 ? -SH_C1 * x * g : 0.f);
    DSH(4, c, deg > 1 ? SH_C2_0 * xy * g : 0.f);
    DSH(5, c, deg > 1 ? SH_C2_1 * yz * g : 0.f);
    DSH(6, c, deg > 1 ? SH_C2_2 * (2.f * zz - xx - yy) * g : 0.f);
    DSH(7, c, deg > 1 ? SH_C2_3 * xz * g : 0.f);
    DSH(8, c, deg > 1 ? SH_C2_4 * (xx - yy) * g : 0.f);
    DSH(9, c, deg > 2 ? SH_C3_0 * y * (3.f * xx - yy) * g : 0.f);
    DSH(10, c, deg > 2 ? SH_C3_1 * xy * z * g : 0.f);
    DSH(11, c, deg > 2 ? SH_C3_2 * y * (4.f * zz - xx - yy) * g : 0.f);
    DSH(12, c, deg > 2 ? SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy) * g : 0.f);
    DSH(13, c, deg > 2 ? SH_C3_4 * x * (4.f * zz - xx - yy) * g : 0.f);
    DSH(14, c, deg > 2 ? SH_C3_5 * z * (xx - yy) * g : 0.f);
    DSH(15, c, deg > 2 ? SH_C3_6 * x * (xx - 3.f * yy) * g : 0.f);
  }
#undef SHC
#undef DSH
  ddir[0] = dx_[0] * dRGB[0] + dx_[1] * dRGB[1] + dx_[2] * dRGB[2];
  ddir[1] = dy_[0] * dRGB[0] + dy_[1] * dRGB[1] + dy_[2] * dRGB[2];
  ddir[2] = dz_[0] * dRGB[0] + dz_[1] * dRGB[1] + dz_[2] * dRGB[2];
}

struct GaussGrad {
  float dm[3], dcov[6], dscale[3], dq[4];
};

// Everything except the SH part for one visible Gaussian: computeCov2DCUDA (backward.cu:144-274), the projection
// part of preprocessCUDA (:366-385) and computeCov3D backward (:278-341).  dmx/dmy = dL/dmean2D, (dcx,dcy,dcz) =
// dL/dconic (xx, xy, yy).
__device__ __forceinline__ void geom_backward(size_t I, const float* __restrict__ scales,
                                              const float* __restrict__ rotations,
                                              const float* __restrict__ cov3D_precomp, const Cam& cam, float p0,
                                              float p1, float p2, float dmx, float dmy, float dcx, float dcy,
                                              float dcz, GaussGrad& o) {
  const float* __restrict__ view = cam.view;
  const float* __restrict__ proj = cam.proj;
  float c3[6];
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f;
  float4 q_raw = make_float4(0.f, 0.f, 0.f, 0.f);
  if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; k++) c3[k] = __ldg(cov3D_precomp + 6 * I + k);
  } else {
    s0 = __ldg(scales + 3 * I); s1 = __ldg(scales + 3 * I + 1); s2 = __ldg(scales + 3 * I + 2);
    float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + I);
    q_raw = q;
    if (cam.raw_params) {  // raw parameters in: the same activated values as the forward computed
      s0 = expf(s0); s1 = expf(s1); s2 = expf(s2);
      q = act_normalize(q);
    }
    q0 = q.x; q1 = q.y; q2 = q.z; q3 = q.w;
    cov3d_from_scale_rot(s0, s1, s2, cam.scale_modifier, q0, q1, q2, q3, c3);
  }
  // ---- computeCov2DCUDA (backward.cu:144-274)
  float tx = view[0] * p0 + view[4] * p1 + view[8] * p2 + view[12];
  float ty = view[1] * p0 + view[5] * p1 + view[9] * p2 + view[13];
  const float tz = view[2] * p0 + view[6] * p1 + view[10] * p2 + view[14];
  const float limx = 1.3f * cam.tanfovx, limy = 1.3f * cam.tanfovy;
  const float txtz = tx / tz, tytz = ty / tz;
  tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
  ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
  const float xgm = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  const float ygm = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  const float hx_ = cam.focal_x, hy_ = cam.focal_y;
  const float Jx = hx_ / tz, Jy = hy_ / tz;
  const float J02 = -(hx_ * tx) / (tz * tz), J12 = -(hy_ * ty) / (tz * tz);
  float T0[3], T1[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    T0[k] = view[4 * k] * Jx + view[4 * k + 2] * J02;
    T1[k] = view[4 * k + 1] * Jy + view[4 * k + 2] * J12;
  }
  const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
  float TV0[3], TV1[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    TV0[k] = T0[0] * V[k][0] + T0[1] * V[k][1] + T0[2] * V[k][2];
    TV1[k] = T1[0] * V[k][0] + T1[1] * V[k][1] + T1[2] * V[k][2];
  }
  const float a = T0[0] * TV0[0] + T0[1] * TV0[1] + T0[2] * TV0[2] + 0.3f;
  const float b = T0[0] * TV1[0] + T0[1] * TV1[1] + T0[2] * TV1[2];
  const float c = T1[0] * TV1[0] + T1[1] * TV1[1] + T1[2] * TV1[2] + 0.3f;
  const float denom = a * c - b * b;
  float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float* dcov = o.dcov;
  if (denom2inv != 0.f) {
    dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
    dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
    dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
    dcov[0] = T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc;
    dcov[3] = T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc;
    dcov[5] = T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc;
    dcov[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
    dcov[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
    dcov[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
  }
  float dT0[3], dT1[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    dT0[k] = 2 * TV0[k] * dL_da + TV1[k] * dL_db;
    dT1[k] = 2 * TV1[k] * dL_dc + TV0[k] * dL_db;
  }
  const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
  const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
  const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
  const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
  const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
  const float dtx = xgm * -hx_ * itz2 * dJ02;
  const float dty = ygm * -hy_ * itz2 * dJ12;
  const float dtz = -hx_ * itz2 * dJ00 - hy_ * itz2 * dJ11 + (2 * hx_ * tx) * itz3 * dJ02 + (2 * hy_ * ty) * itz3 * dJ12;
  float* dm = o.dm;
  dm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
  dm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
  dm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
  // ---- preprocessCUDA (backward.cu:366-385): projection part
  const float hx = proj[0] * p0 + proj[4] * p1 + proj[8] * p2 + proj[12];
  const float hy = proj[1] * p0 + proj[5] * p1 + proj[9] * p2 + proj[13];
  const float hw = proj[3] * p0 + proj[7] * p1 + proj[11] * p2 + proj[15];
  const float m_w = 1.0f / (hw + 0.0000001f);
  const float mul1 = hx * m_w * m_w, mul2 = hy * m_w * m_w;
  dm[0] += (proj[0] * m_w - proj[3] * mul1) * dmx + (proj[1] * m_w - proj[3] * mul2) * dmy;
  dm[1] += (proj[4] * m_w - proj[7] * mul1) * dmx + (proj[5] * m_w - proj[7] * mul2) * dmy;
  dm[2] += (proj[8] * m_w - proj[11] * mul1) * dmx + (proj[9] * m_w - proj[11] * mul2) * dmy;
  // ---- computeCov3D backward (backward.cu:278-341)
  if (scales != nullptr) {
    const float r = q0, x = q1, y = q2, z = q3;
    const float Rg[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                            {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                            {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float sv[3] = {cam.scale_modifier * s0, cam.scale_modifier * s1, cam.scale_modifier * s2};
    const float Sg[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                            {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                            {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
    float dMt[3][3];
#pragma unroll
    for (int a_ = 0; a_ < 3; a_++)
#pragma unroll
      for (int b_ = 0; b_ < 3; b_++) {
        float acc_ = 0.f;
#pragma unroll
        for (int k = 0; k < 3; k++) acc_ += (sv[a_] * Rg[k][a_]) * Sg[b_][k];
        dMt[a_][b_] = 2.0f * acc_;
      }
#pragma unroll
    for (int a_ = 0; a_ < 3; a_++) {
      o.dscale[a_] = Rg[0][a_] * dMt[a_][0] + Rg[1][a_] * dMt[a_][1] + Rg[2][a_] * dMt[a_][2];
#pragma unroll
      for (int b_ = 0; b_ < 3; b_++) dMt[a_][b_] *= sv[a_];
    }
    float* dq = o.dq;
    dq[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    dq[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) -
            4 * x * (dMt[2][2] + dMt[1][1]);
    dq[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) -
            4 * y * (dMt[2][2] + dMt[0][0]);
    dq[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) -
            4 * z * (dMt[1][1] + dMt[0][0]);
    if (cam.raw_params) {
      // chain rule of exp and F.normalize (the autograd replay of gaussian_model.py:95-102, activations.cuh):
      // d/ds_raw = g * exp(s_raw);  y = q / n: dq_raw = g / n - (n > eps ? y <y, g> / n : 0)
      o.dscale[0] *= s0; o.dscale[1] *= s1; o.dscale[2] *= s2;
      const float nr = quat_norm(q_raw);
      const float inv = 1.0f / fmaxf(nr, 1e-12f);
      const float dot = nr > 1e-12f ? (r * dq[0] + x * dq[1] + y * dq[2] + z * dq[3]) : 0.f;
      dq[0] = (dq[0] - r * dot) * inv; dq[1] = (dq[1] - x * dot) * inv;
      dq[2] = (dq[2] - y * dot) * inv; dq[3] = (dq[3] - z * dot) * inv;
    }
  }
}

// dnormvdv (auxiliary.h:107-117): dL/d(dir) of the normalised view direction -> dL/dmean
__device__ __forceinline__ void dir_backward(float dox, float doy, float doz, const float* ddir, float* dm) {
  const float sum2 = dox * dox + doy * doy + doz * doz;
  const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  dm[0] += ((sum2 - dox * dox) * ddir[0] - doy * dox * ddir[1] - doz * dox * ddir[2]) * inv32;
  dm[1] += (-dox * doy * ddir[0] + (sum2 - doy * doy) * ddir[1] - doz * doy * ddir[2]) * inv32;
  dm[2] += (-dox * doz * ddir[0] - doy * doz * ddir[1] + (sum2 - doz * doz) * ddir[2]) * inv32;
}

// coalesced copy of `nfloats` staged floats to global memory (dst 16-byte aligned)
__device__ __forceinline__ void store_staged(float* __restrict__ dst, const float* s, int nfloats, int tid) {
  const int n4 = nfloats >> 2;
#pragma unroll 1
  for (int j = tid; j < n4; j += PB_THREADS)
    __stcs(reinterpret_cast<float4*>(dst) + j, reinterpret_cast<const float4*>(s)[j]);
#pragma unroll 1
  for (int j = (n4 << 2) + tid; j < nfloats; j += PB_THREADS) dst[j] = s[j];
}

// Staged variant: SH rows of MT coefficients (3*MT floats, a multiple of 4) with 16-byte aligned sh / dL_dsh and
// 16-byte aligned (P,3)/(P,6) gradient outputs.  MT = 16 (degree <= 3 storage) or 4 (degree <= 1 storage).
// SPLIT: split SH storage (gsr_inputs.sh_rest): `shs` = DC rows (P,1,3), sh_rest = (P,MT-1,3).  The CTA's 128 rest
// rows are ONE contiguous block of 128 * 3 * (MT-1) floats, staged flat (row stride 3*(MT-1) words, odd =>
// conflict-free) and streamed back the same way; the DC row is three plain loads / a staged (P,3)-style store.
template <int MT, bool SPLIT>
__global__ void __launch_bounds__(PB_THREADS, 5) preprocess_bwd_staged_kernel(
    int P, int D, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float4* __restrict__ g1, const float* __restrict__ shs,
    const float* __restrict__ sh_rest, const float* __restrict__ cov3D_precomp, Cam cam,
    const int* __restrict__ radii, const float4* __restrict__ g2, BwdView acc, BwdOut out) {
  constexpr int NF = 3 * MT, F4 = NF / 4, RS4 = F4 | 1;  // row stride in float4, odd => conflict-free 128-bit rows
  constexpr int NR = NF - 3;                             // SPLIT: floats per rest row
  extern __shared__ float4 s_tile[];                     // [PB_THREADS][RS4], or SPLIT: float [PB_THREADS][NR]
  float* const s_rest = reinterpret_cast<float*>(s_tile);
  // dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dscales (, SPLIT: dL_dsh DC rows)
  __shared__ __align__(16) float s_v3[SPLIT ? 5 : 4][PB_THREADS * 3];
  __shared__ __align__(16) float s_cov[PB_THREADS * 6];
  __shared__ uint32_t s_vis[PB_THREADS / 32];

  const int tid = threadIdx.x;
  const int base = blockIdx.x * PB_THREADS;
  const int i = base + tid;
  const int nG = min(PB_THREADS, P - base);
  const size_t I = (size_t)i;
  const bool in_range = i < P;
  const bool visible = in_range && radii[i] > 0;
  const unsigned vm = __ballot_sync(0xffffffffu, visible);
  if ((tid & 31) == 0) s_vis[tid >> 5] = vm;
  float4 A = make_float4(0.f, 0.f, 0.f, 0.f), B = A;
  float Cyy = 0.f;
  if (in_range) {
    A = __ldcs(acc.acc0 + i);
    B = __ldcs(acc.acc1 + i);
    Cyy = __ldcs(acc.acc2 + i);
    finish_2d_grads(__ldg(g1 + i), cam.W, cam.H, B, Cyy);  // opacity as the forward saw it
  }
  __syncthreads();
  if (D > 0) {  // coefficient rows of the visible Gaussians, coalesced (degree 0 needs no coefficient)
    // asynchronous global->shared copies (LDGSTS): no register staging, and the copy overlaps the geometry part
    // below, which does not need the coefficients
    const uint32_t tile_addr = (uint32_t)__cvta_generic_to_shared(s_tile);
    if (SPLIT) {
      const float* __restrict__ src = sh_rest + (size_t)base * NR;
      const int nfl = nG * NR, n4 = nfl >> 2;
#pragma unroll 4
      for (int j = tid; j < n4; j += PB_THREADS) {
        const int r0 = (4 * j) / NR, r1 = (4 * j + 3) / NR;  // a 16-byte chunk touches one or two rows
        if (((s_vis[r0 >> 5] >> (r0 & 31)) | (s_vis[r1 >> 5] >> (r1 & 31))) & 1u)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile_addr + (uint32_t)j * 16u),
                       "l"(reinterpret_cast<const float4*>(src) + j)
                       : "memory");
      }
      for (int j = (n4 << 2) + tid; j < nfl; j += PB_THREADS) s_rest[j] = __ldg(src + j);  // last CTA's tail
    } else {
      const float4* __restrict__ src4 = reinterpret_cast<const float4*>(shs + (size_t)base * NF);
#pragma unroll 4
      for (int j = tid; j < nG * F4; j += PB_THREADS) {
        const int row = j / F4, col = j - row * F4;
        if ((s_vis[row >> 5] >> (row & 31)) & 1u)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile_addr + (uint32_t)(row * RS4 + col) * 16u),
                       "l"(src4 + j)
                       : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  GaussGrad o;
#pragma unroll
  for (int k = 0; k < 3; k++) { o.dm[k] = 0.f; o.dscale[k] = 0.f; }
#pragma unroll
  for (int k = 0; k < 6; k++) o.dcov[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 4; k++) o.dq[k] = 0.f;
  float4* row4 = s_tile + tid * RS4;
  float p0 = 0.f, p1 = 0.f, p2 = 0.f;
  if (visible) {
    p0 = __ldg(means3D + 3 * I); p1 = __ldg(means3D + 3 * I + 1); p2 = __ldg(means3D + 3 * I + 2);
    geom_backward(I, scales, rotations, cov3D_precomp, cam, p0, p1, p2, B.x, B.y, B.z, B.w, Cyy, o);
  }
  if (D > 0) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
  }
  if (visible) {
    const float dox = p0 - __ldg(cam.campos), doy = p1 - __ldg(cam.campos + 1), doz = p2 - __ldg(cam.campos + 2);
    const float len = sqrtf(dox * dox + doy * doy + doz * doz);
    const uint32_t flags = __float_as_uint(__ldg(&g2[i].w));
    const float dRGB[3] = {(flags & 1u) ? 0.f : A.x, (flags & 2u) ? 0.f : A.y, (flags & 4u) ? 0.f : A.z};
    float v[NF];
    if (D > 0 && SPLIT) {
      v[0] = __ldg(shs + 3 * I); v[1] = __ldg(shs + 3 * I + 1); v[2] = __ldg(shs + 3 * I + 2);
#pragma unroll
      for (int k = 0; k < NR; k++) v[3 + k] = s_rest[tid * NR + k];
    } else if (D > 0) {
#pragma unroll
      for (int k = 0; k < F4; k++) {
        const float4 q = row4[k];
        v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
      }
    } else {
#pragma unroll
      for (int k = 0; k < NF; k++) v[k] = 0.f;
    }
    float ddir[3];
    sh_backward_inplace<NF>(D, v, dox / len, doy / len, doz / len, dRGB, ddir);
    dir_backward(dox, doy, doz, ddir, o.dm);
    if (SPLIT) {
      s_v3[SPLIT ? 4 : 0][3 * tid] = v[0]; s_v3[SPLIT ? 4 : 0][3 * tid + 1] = v[1]; s_v3[SPLIT ? 4 : 0][3 * tid + 2] = v[2];
#pragma unroll
      for (int k = 0; k < NR; k++) s_rest[tid * NR + k] = v[3 + k];
    } else {
#pragma unroll
      for (int k = 0; k < F4; k++) row4[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
  } else if (in_range) {
    if (SPLIT) {
      s_v3[SPLIT ? 4 : 0][3 * tid] = 0.f; s_v3[SPLIT ? 4 : 0][3 * tid + 1] = 0.f; s_v3[SPLIT ? 4 : 0][3 * tid + 2] = 0.f;
#pragma unroll
      for (int k = 0; k < NR; k++) s_rest[tid * NR + k] = 0.f;
    } else {
#pragma unroll
      for (int k = 0; k < F4; k++) row4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  if (in_range) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      s_v3[0][3 * tid + k] = o.dm[k];
      s_v3[3][3 * tid + k] = o.dscale[k];
    }
    s_v3[1][3 * tid] = B.x; s_v3[1][3 * tid + 1] = B.y; s_v3[1][3 * tid + 2] = 0.f;
    s_v3[2][3 * tid] = A.x; s_v3[2][3 * tid + 1] = A.y; s_v3[2][3 * tid + 2] = A.z;
#pragma unroll
    for (int k = 0; k < 6; k++) s_cov[6 * tid + k] = o.dcov[k];
    // raw parameters: sigmoid backward, grad * (1 - y) * y with y = the opacity the forward stored
    const float op_ = __ldg(&g1[i].w);
    __stcs(out.dL_dopacity + I, cam.raw_params ? A.w * (1.0f - op_) * op_ : A.w);
    __stcs(reinterpret_cast<float4*>(out.dL_dconic) + I, make_float4(B.z, B.w, 0.f, Cyy));
    __stcs(reinterpret_cast<float4*>(out.dL_drot) + I, make_float4(o.dq[0], o.dq[1], o.dq[2], o.dq[3]));
  }
  __syncthreads();
  if (SPLIT) {
    if (out.dL_dsh != nullptr) {
      store_staged(out.dL_dsh_rest + (size_t)base * NR, s_rest, nG * NR, tid);
      store_staged(out.dL_dsh + (size_t)base * 3, s_v3[SPLIT ? 4 : 0], nG * 3, tid);
    }
  } else if (out.dL_dsh != nullptr) {  // dL_dsh rows: nG * F4 contiguous float4 (NULL: the caller exchanges SH factors)
    float4* __restrict__ dst4 = reinterpret_cast<float4*>(out.dL_dsh + (size_t)base * NF);
    for (int j = tid; j < nG * F4; j += PB_THREADS) {
      const int row = j / F4, col = j - row * F4;
      __stcs(dst4 + j, s_tile[row * RS4 + col]);
    }
  }
  store_staged(out.dL_dmeans3D + (size_t)base * 3, s_v3[0], nG * 3, tid);
  store_staged(out.dL_dmeans2D + (size_t)base * 3, s_v3[1], nG * 3, tid);
  store_staged(out.dL_dcolors + (size_t)base * 3, s_v3[2], nG * 3, tid);
  store_staged(out.dL_dscales + (size_t)base * 3, s_v3[3], nG * 3, tid);
  store_staged(out.dL_dcov3D + (size_t)base * 6, s_cov, nG * 6, tid);
}

// Generic variant: any M, any alignment, no SH at all (colors_precomp); direct global accesses.
__global__ void __launch_bounds__(PB_THREADS) preprocess_bwd_kernel(
    int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float4* __restrict__ g1, const float* __restrict__ shs,
    const float* __restrict__ cov3D_precomp, Cam cam, const int* __restrict__ radii, const float4* __restrict__ g2,
    BwdView acc, BwdOut out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const size_t I = (size_t)i;
  const float4 A = acc.acc0[i];
  float4 B = acc.acc1[i];
  float Cyy = acc.acc2[i];
  finish_2d_grads(__ldg(g1 + i), cam.W, cam.H, B, Cyy);  // opacity as the forward saw it
  out.dL_dcolors[3 * I] = A.x; out.dL_dcolors[3 * I + 1] = A.y; out.dL_dcolors[3 * I + 2] = A.z;
  {
    const float op_ = __ldg(&g1[i].w);
    out.dL_dopacity[I] = cam.raw_params ? A.w * (1.0f - op_) * op_ : A.w;
  }
  out.dL_dmeans2D[3 * I] = B.x; out.dL_dmeans2D[3 * I + 1] = B.y; out.dL_dmeans2D[3 * I + 2] = 0.f;
  out.dL_dconic[4 * I] = B.z; out.dL_dconic[4 * I + 1] = B.w; out.dL_dconic[4 * I + 2] = 0.f;
  out.dL_dconic[4 * I + 3] = Cyy;
  GaussGrad o;
#pragma unroll
  for (int k = 0; k < 3; k++) { o.dm[k] = 0.f; o.dscale[k] = 0.f; }
#pragma unroll
  for (int k = 0; k < 6; k++) o.dcov[k] = 0.f;
#pragma unroll
  for (int k = 0; k < 4; k++) o.dq[k] = 0.f;
  const bool visible = radii[i] > 0;
  float* dsh = out.dL_dsh ? out.dL_dsh + I * M * 3 : nullptr;
  const int nf = 3 * M;
  if (visible) {
    const float p0 = __ldg(means3D + 3 * I), p1 = __ldg(means3D + 3 * I + 1), p2 = __ldg(means3D + 3 * I + 2);
    geom_backward(I, scales, rotations, cov3D_precomp, cam, p0, p1, p2, B.x, B.y, B.z, B.w, Cyy, o);
    if (shs != nullptr) {
      const float dox = p0 - __ldg(cam.campos), doy = p1 - __ldg(cam.campos + 1), doz = p2 - __ldg(cam.campos + 2);
      const float len = sqrtf(dox * dox + doy * doy + doz * doz);
      const uint32_t flags = __float_as_uint(__ldg(&g2[i].w));
      const float dRGB[3] = {(flags & 1u) ? 0.f : A.x, (flags & 2u) ? 0.f : A.y, (flags & 4u) ? 0.f : A.z};
      const float* __restrict__ sh = shs + I * M * 3;
      float v[48];
#pragma unroll
      for (int k = 0; k < 48; k++) v[k] = (D > 0 && k < nf) ? __ldg(sh + k) : 0.f;
      float ddir[3];
      sh_backward_inplace<48>(D, v, dox / len, doy / len, doz / len, dRGB, ddir);
      dir_backward(dox, doy, doz, ddir, o.dm);
      if (dsh) {
#pragma unroll
        for (int k = 0; k < 48; k++)
          if (k < nf) dsh[k] = v[k];
        for (int k = 48; k < nf; k++) dsh[k] = 0.f;
      }
    } else if (dsh) {
      for (int k = 0; k < nf; k++) dsh[k] = 0.f;
    }
  } else if (dsh) {
    for (int k = 0; k < nf; k++) dsh[k] = 0.f;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) out.dL_dmeans3D[3 * I + k] = o.dm[k];
#pragma unroll
  for (int k = 0; k < 6; k++) out.dL_dcov3D[6 * I + k] = o.dcov[k];
#pragma unroll
  for (int k = 0; k < 3; k++) out.dL_dscales[3 * I + k] = o.dscale[k];
#pragma unroll
  for (int k = 0; k < 4; k++) out.dL_drot[4 * I + k] = o.dq[k];
}

}  // namespace gsr
