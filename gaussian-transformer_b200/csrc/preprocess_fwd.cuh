// csrc/preprocess_fwd.cuh -- per-Gaussian forward stage.
// Behaviour: DGR/cuda_rasterizer/forward.cu:155-256 (preprocessCUDA) with
// computeCov3D :118-152, computeCov2D :74-113, computeColorFromSH :20-71,
// in_frustum auxiliary.h:139-164.  Integer results (radii, tile rectangles)
// must be bit-exact, so every fp32 operation on the path to them is an explicit
// round-to-nearest intrinsic placed where the reference's compiled code rounds.
//
// B200 design: one thread per Gaussian, float4 stores of three SoA records, SH
// coefficients fetched with 128-bit read-only loads; also emits the 32-bit
// depth key the Gaussian-level sort consumes (depth_sort.cuh) -- the reference's
// per-Gaussian prefix sum + 64-bit per-instance key sort has no counterpart.
#pragma once
#include "common.cuh"

namespace gsr {

// cov3D from scale / unnormalised quaternion (forward.cu:118-152).  M = S*R
// keeps the products with the zero entries of S: they turn non-finite rotation
// entries into NaN exactly like the reference.
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float r, float x,
                                                     float y, float z, float* out) {
  float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
  float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
  float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
  float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
  float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
  float xx_p_yy = __fmaf_rn(x, x, yy);
  float yy_p_zz = __fadd_rn(yy, zz);
  float xx_p_zz = __fmaf_rn(x, x, zz);
  float R00 = __fsub_rn(1.0f, __fadd_rn(yy_p_zz, yy_p_zz)), R01 = __fadd_rn(xy_m_rz, xy_m_rz),
        R02 = __fadd_rn(xz_p_ry, xz_p_ry);
  float R10 = __fadd_rn(xy_p_rz, xy_p_rz), R11 = __fsub_rn(1.0f, __fadd_rn(xx_p_zz, xx_p_zz)),
        R12 = __fadd_rn(yz_m_rx, yz_m_rx);
  float R20 = __fadd_rn(xz_m_ry, xz_m_ry), R21 = __fadd_rn(yz_p_rx, yz_p_rx),
        R22 = __fsub_rn(1.0f, __fadd_rn(xx_p_yy, xx_p_yy));
  float sx = __fmul_rn(s0, mod), sy = __fmul_rn(s1, mod), sz = __fmul_rn(s2, mod);
  float A00 = __fmaf_rn(0.0f, R02, __fmaf_rn(0.0f, R01, __fmul_rn(sx, R00)));
  float A01 = __fmaf_rn(0.0f, R02, __fmaf_rn(sy, R01, __fmul_rn(0.0f, R00)));
  float A02 = __fmaf_rn(sz, R02, __fmaf_rn(0.0f, R01, __fmul_rn(0.0f, R00)));
  float A10 = __fmaf_rn(0.0f, R12, __fmaf_rn(sx, R10, __fmul_rn(0.0f, R11)));
  float A11 = __fmaf_rn(0.0f, R12, __fmaf_rn(0.0f, R10, __fmul_rn(sy, R11)));
  float A12 = __fmaf_rn(sz, R12, __fmaf_rn(0.0f, R10, __fmul_rn(0.0f, R11)));
  float A20 = __fmaf_rn(0.0f, R22, __fmaf_rn(sx, R20, __fmul_rn(0.0f, R21)));
  float A21 = __fmaf_rn(0.0f, R22, __fmaf_rn(0.0f, R20, __fmul_rn(sy, R21)));
  float A22 = __fmaf_rn(sz, R22, __fmaf_rn(0.0f, R20, __fmul_rn(0.0f, R21)));
  out[0] = dot3c(A00, A00, A01, A01, A02, A02);
  out[1] = dot3c(A00, A10, A01, A11, A02, A12);
  out[2] = dot3c(A00, A20, A01, A21, A02, A22);
  out[3] = dot3c(A10, A10, A11, A11, A12, A12);
  out[4] = dot3c(A10, A20, A11, A21, A12, A22);
  out[5] = dot3c(A20, A20, A21, A21, A22, A22);
}

// EWA 2-D covariance (forward.cu:74-113); returns (a, b, c) with the 0.3
// low-pass added.  t = view-space position (tx, ty, tz).
__device__ __forceinline__ void cov2d(float tx, float ty, float tz, const Cam& cam, const float* c3,
                                      const float* __restrict__ v, float& a, float& b, float& c) {
  float limx = __fmul_rn(1.3f, cam.tanfovx), limy = __fmul_rn(1.3f, cam.tanfovy);
  float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
  float cx = fminf(limx, fmaxf(-limx, txtz));
  float cy = fminf(limy, fmaxf(-limy, tytz));
  float Jx = __fdiv_rn(cam.focal_x, tz);
  float tz2 = __fmul_rn(tz, tz);
  float J02 = __fdiv_rn(__fmul_rn(__fmul_rn(tz, -cx), cam.focal_x), tz2);
  float Jy = __fdiv_rn(cam.focal_y, tz);
  float J12 = __fdiv_rn(__fmul_rn(__fmul_rn(tz, -cy), cam.focal_y), tz2);
  float T0[3], T1[3];
#pragma unroll
  for (int r = 0; r < 3; r++) {
    float w0 = v[4 * r + 0], w1 = v[4 * r + 1], w2 = v[4 * r + 2];
    T0[r] = __fmaf_rn(w2, J02, __fmaf_rn(w0, Jx, __fmul_rn(0.0f, w1)));
    T1[r] = __fmaf_rn(w2, J12, __fmaf_rn(0.0f, w0, __fmul_rn(w1, Jy)));
  }
  const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
  float B0[3], B1[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    B0[k] = dot3c(T0[0], V[0][k], T0[1], V[1][k], T0[2], V[2][k]);
    B1[k] = dot3c(T1[0], V[0][k], T1[1], V[1][k], T1[2], V[2][k]);
  }
  float c00 = dot3c(T0[0], B0[0], T0[1], B0[1], T0[2], B0[2]);
  float c01 = dot3c(T0[0], B1[0], T0[1], B1[1], T0[2], B1[2]);
  float c11 = dot3c(T1[0], B1[0], T1[1], B1[1], T1[2], B1[2]);
  a = __fadd_rn(c00, 0.3f);
  b = c01;
  c = __fadd_rn(c11, 0.3f);
}

// SH basis values for a unit direction, degree <= 3 (forward.cu:30-61).
// basis[k] multiplies coefficient k; basis[0] = C0.
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float* basis) {
  basis[0] = SH_C0;
  if (deg > 0) {
    basis[1] = -SH_C1 * y;
    basis[2] = SH_C1 * z;
    basis[3] = -SH_C1 * x;
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      basis[4] = SH_C2_0 * xy;
      basis[5] = SH_C2_1 * yz;
      basis[6] = SH_C2_2 * (2.0f * zz - xx - yy);
      basis[7] = SH_C2_3 * xz;
      basis[8] = SH_C2_4 * (xx - yy);
      if (deg > 2) {
        basis[9] = SH_C3_0 * y * (3.0f * xx - yy);
        basis[10] = SH_C3_1 * xy * z;
        basis[11] = SH_C3_2 * y * (4.0f * zz - xx - yy);
        basis[12] = SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        basis[13] = SH_C3_4 * x * (4.0f * zz - xx - yy);
        basis[14] = SH_C3_5 * z * (xx - yy);
        basis[15] = SH_C3_6 * x * (xx - 3.0f * yy);
      }
    }
  }
}

// rgb = max(sum_k basis[k]*sh[k] + 0.5, 0), clamped flags (forward.cu:63-70).
// VEC: coefficients fetched as float4 (requires 16-B aligned rows: M*12 % 16 == 0
// and an aligned base); else scalar read-only loads.
// sh_rest != nullptr: split storage, sh_row = the Gaussian's DC coefficient (3 floats), sh_rest = its M-1 others.
template <bool VEC>
__device__ __forceinline__ void sh_to_rgb(int deg, int M, const float* __restrict__ sh_row,
                                          const float* __restrict__ sh_rest, float dx, float dy,
                                          float dz, float& r, float& g, float& b, uint32_t& clamped) {
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float x = dx / len, y = dy / len, z = dz / len;
  float basis[16];
  sh_basis(deg, x, y, z, basis);
  const int K = (deg + 1) * (deg + 1);
  float acc[3] = {0.f, 0.f, 0.f};
  if (VEC) {
    // K*3 floats; K in {1,4,9,16} -> 3, 12, 27, 48 floats.  Full float4 chunks
    // then a scalar tail (K=1: 3 floats, K=9: 27 floats).
    const float4* p4 = reinterpret_cast<const float4*>(sh_row);
    const int nf = K * 3;
    const int n4 = nf >> 2;
    float buf[48];
#pragma unroll
    for (int j = 0; j < 12; j++) {
      if (j < n4) {
        float4 q = __ldg(p4 + j);
        buf[4 * j] = q.x; buf[4 * j + 1] = q.y; buf[4 * j + 2] = q.z; buf[4 * j + 3] = q.w;
      }
    }
    // (the host only selects VEC when K*3 is a multiple of 4: K = 4 or 16)
#pragma unroll
    for (int k = 0; k < 16; k++) {
      if (k < K) {
        acc[0] = fmaf(basis[k], buf[3 * k], acc[0]);
        acc[1] = fmaf(basis[k], buf[3 * k + 1], acc[1]);
        acc[2] = fmaf(basis[k], buf[3 * k + 2], acc[2]);
      }
    }
  } else {
    for (int k = 0; k < K; k++) {
      const float* c = (sh_rest != nullptr && k > 0) ? sh_rest + 3 * (k - 1) : sh_row + 3 * k;
      acc[0] = fmaf(basis[k], __ldg(c), acc[0]);
      acc[1] = fmaf(basis[k], __ldg(c + 1), acc[1]);
      acc[2] = fmaf(basis[k], __ldg(c + 2), acc[2]);
    }
  }
  (void)M;
  float rr = acc[0] + 0.5f, gg = acc[1] + 0.5f, bb = acc[2] + 0.5f;
  clamped = (rr < 0.f ? 1u : 0u) | (gg < 0.f ? 2u : 0u) | (bb < 0.f ? 4u : 0u);
  // glm::max(x, 0) is (x < 0) ? 0 : x, so a NaN colour stays NaN (forward.cu:70)
  r = rr < 0.f ? 0.f : rr;
  g = gg < 0.f ? 0.f : gg;
  b = bb < 0.f ? 0.f : bb;
}

// Geometry half of preprocessCUDA for one Gaussian (forward.cu:182-237 without the colour): cull, projection, cov3D,
// EWA cov2D, conic, radius, tile rectangle.  Returns false when the Gaussian is culled; otherwise o0 = (mean2D.x,
// mean2D.y, depth, radius bits), conic = (A, B, C), tiles = rectangle area.  Every fp32 operation on the way to an
// integer result is an explicit round-to-nearest intrinsic placed where the reference's compiled code rounds.
__device__ __forceinline__ bool pre_geometry(int i, float px3, float py3, float pz3, float s0, float s1, float s2,
                                             float4 q, const float* __restrict__ cov3D_precomp, const Cam& cam,
                                             int prefiltered, uint32_t* __restrict__ ctrl, float4& o0, float& conx,
                                             float& cony, float& conz, int& radius, int& tiles) {
  const float* __restrict__ view = cam.view;
  const float* __restrict__ proj = cam.proj;
  // in_frustum: cull iff p_view.z <= 0.2 (NaN passes, like the reference)
  const float vz = xf_row(view, 2, px3, py3, pz3);
  if (vz <= 0.2f) {
    if (prefiltered) atomicOr(&ctrl[CTRL_FLAGS], 1u);
    return false;
  }
  const float hw = xf_row(proj, 3, px3, py3, pz3);
  const float hx = xf_row(proj, 0, px3, py3, pz3);
  const float hy = xf_row(proj, 1, px3, py3, pz3);
  const float pw = __frcp_rn(__fadd_rn(hw, 0.0000001f));
  const float projx = __fmul_rn(hx, pw), projy = __fmul_rn(hy, pw);
  float c3[6];
  if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; k++) c3[k] = __ldg(cov3D_precomp + 6 * (size_t)i + k);
  } else {
    cov3d_from_scale_rot(s0, s1, s2, cam.scale_modifier, q.x, q.y, q.z, q.w, c3);
  }
  const float tx = xf_row(view, 0, px3, py3, pz3);
  const float ty = xf_row(view, 1, px3, py3, pz3);
  float a, b, c;
  cov2d(tx, ty, vz, cam, c3, view, a, b, c);
  const float det = __fmaf_rn(a, c, -__fmul_rn(b, b));
  if (det == 0.0f) return false;
  const float det_inv = __frcp_rn(det);
  conx = __fmul_rn(c, det_inv); cony = __fmul_rn(b, -det_inv); conz = __fmul_rn(a, det_inv);
  const float mid = __fmul_rn(__fadd_rn(a, c), 0.5f);
  const float disc = fmaxf(__fmaf_rn(mid, mid, -det), 0.1f);
  const float sq = __fsqrt_rn(disc);
  const float lmax = fmaxf(__fadd_rn(mid, sq), __fsub_rn(mid, sq));
  const int my_radius = __float2int_rz(ceilf(__fmul_rn(__fsqrt_rn(lmax), 3.0f)));
  // lambda_max >= 0.3 for finite input, so radius >= 2; radius 0 only arises from a NaN covariance.  The
  // reference then still counts the Gaussian's tile but never writes its sort key (duplicateWithKeys skips
  // radii <= 0, rasterizer_impl.cu:84), i.e. it renders an uninitialised key.  Defined behaviour here: culled.
  if (my_radius <= 0) return false;
  const float mx = ndc2pix(projx, cam.W), my = ndc2pix(projy, cam.H);
  int minx, miny, maxx, maxy;
  get_rect(mx, my, my_radius, cam.gx, cam.gy, minx, miny, maxx, maxy);
  tiles = (maxx - minx) * (maxy - miny);
  if (tiles == 0) return false;
  radius = my_radius;
  o0 = make_float4(mx, my, vz, __int_as_float(my_radius));
  return true;
}

// gsr_inputs.raw_params: the kernel was handed _opacity / _scaling / _rotation (scene/gaussian_model.py:95-106) and
// applies sigmoid / exp / normalize itself -- the values the separate activation pass would have written.
__device__ __forceinline__ void apply_activations(float& opacity, float* sc, float4& q) {
  opacity = act_sigmoid(opacity);
#pragma unroll
  for (int k = 0; k < 3; k++) sc[k] = expf(sc[k]);
  q = act_normalize(q);
}

constexpr int PF_THREADS = 128;
#ifndef GSR_PF_MIN_CTAS
#define GSR_PF_MIN_CTAS 7
#endif
#ifndef GSR_PF_EARLY_SH
#define GSR_PF_EARLY_SH 1
#endif

// Loads of the 12-byte rows of means3D / scales of the CTA's Gaussians: one contiguous block each, moved with
// coalesced 128-bit loads into shared memory (STAGE_IN: both arrays 16-byte aligned; stride-3 reads are conflict
// free); else plain per-thread loads.
template <bool STAGE_IN, int THREADS>
__device__ __forceinline__ void load_mean_scale(int P, int base, int tid, const float* __restrict__ means3D,
                                                const float* __restrict__ scales, float* s_mean, float* s_scale,
                                                float* m, float* sc) {
  const int i = base + tid;
  if (STAGE_IN) {
    const int nfl = min(THREADS, P - base) * 3;
    const float4* __restrict__ m4 = reinterpret_cast<const float4*>(means3D + (size_t)base * 3);
    const float4* __restrict__ s4 = reinterpret_cast<const float4*>(scales + (size_t)base * 3);
    if (tid < (nfl >> 2)) {
      reinterpret_cast<float4*>(s_mean)[tid] = __ldg(m4 + tid);
      if (scales != nullptr) reinterpret_cast<float4*>(s_scale)[tid] = __ldg(s4 + tid);
    } else if (tid < (nfl >> 2) + (nfl & 3)) {  // the last CTA's 1..3 trailing floats
      const int j = ((nfl >> 2) << 2) + tid - (nfl >> 2);
      s_mean[j] = __ldg(means3D + (size_t)base * 3 + j);
      if (scales != nullptr) s_scale[j] = __ldg(scales + (size_t)base * 3 + j);
    }
    __syncthreads();
    if (i < P) {
#pragma unroll
      for (int k = 0; k < 3; k++) {
        m[k] = s_mean[3 * tid + k];
        sc[k] = scales != nullptr ? s_scale[3 * tid + k] : 0.f;
      }
    }
  } else if (i < P) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      m[k] = __ldg(means3D + 3 * (size_t)i + k);
      sc[k] = scales != nullptr ? __ldg(scales + 3 * (size_t)i + k) : 0.f;
    }
  }
}

// Generic kernel: any M / alignment, colours precomputed, split SH storage.  One thread per Gaussian; coefficient rows
// are read straight from global memory (VEC_SH: as float4).
template <bool VEC_SH, bool STAGE_IN>
__global__ void __launch_bounds__(256) preprocess_fwd_kernel(
    int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float* __restrict__ opacities, const float* __restrict__ shs,
    const float* __restrict__ sh_rest, const float* __restrict__ colors_precomp,
    const float* __restrict__ cov3D_precomp, Cam cam, int prefiltered,
    int* __restrict__ radii, GeomView geom, uint32_t* __restrict__ ctrl) {
  __shared__ __align__(16) float s_mean[STAGE_IN ? 256 * 3 : 4], s_scale[STAGE_IN ? 256 * 3 : 4];
  const int tid = threadIdx.x, base = blockIdx.x * 256;
  const int i = base + tid;
  float4 q_in = make_float4(0.f, 0.f, 0.f, 0.f);
  float opac_in = 0.f;
  if (i < P) {  // issued before anything waits: one DRAM round trip for all per-Gaussian inputs
    if (rotations != nullptr) q_in = __ldg(reinterpret_cast<const float4*>(rotations) + i);
    opac_in = __ldg(opacities + i);
  }
  float m[3] = {0.f, 0.f, 0.f}, sc[3] = {0.f, 0.f, 0.f};
  load_mean_scale<STAGE_IN, 256>(P, base, tid, means3D, scales, s_mean, s_scale, m, sc);
  if (i >= P) return;
  if (cam.raw_params) apply_activations(opac_in, sc, q_in);
  int out_radius = 0, tiles = 0;
  float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0;
  float conx, cony, conz;
  if (pre_geometry(i, m[0], m[1], m[2], sc[0], sc[1], sc[2], q_in, cov3D_precomp, cam, prefiltered, ctrl, o0, conx,
                   cony, conz, out_radius, tiles)) {
    float r, g, bl;
    uint32_t clamped = 0;
    if (colors_precomp != nullptr) {
      r = __ldg(colors_precomp + 3 * (size_t)i);
      g = __ldg(colors_precomp + 3 * (size_t)i + 1);
      bl = __ldg(colors_precomp + 3 * (size_t)i + 2);
    } else {
      const float dx = m[0] - __ldg(cam.campos), dy = m[1] - __ldg(cam.campos + 1), dz = m[2] - __ldg(cam.campos + 2);
      if (sh_rest != nullptr)
        sh_to_rgb<false>(D, M, shs + (size_t)i * 3, sh_rest + (size_t)i * (M - 1) * 3, dx, dy, dz, r, g, bl, clamped);
      else
        sh_to_rgb<VEC_SH>(D, M, shs + (size_t)i * M * 3, nullptr, dx, dy, dz, r, g, bl, clamped);
    }
    o1 = make_float4(conx, cony, conz, opac_in);
    o2 = make_float4(r, g, bl, __uint_as_float(((uint32_t)tiles << 3) | clamped));
  } else {
    out_radius = 0;
    o0 = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  radii[i] = out_radius;
  geom.keys[0][i] = out_radius > 0 ? __float_as_uint(o0.z) : 0xffffffffu;  // depth-sort key
  geom.g0[i] = o0;
  geom.g1[i] = o1;
  geom.g2[i] = o2;
}

// Main kernel: SH rows of MT coefficients (MT = 16 or 4) in one 16-byte aligned (P, MT, 3) array.  A CTA owns 128
// consecutive Gaussians.  The geometry half runs first; the coefficient rows of the Gaussians that survived it --
// contiguous in global memory -- are then copied into shared memory with COALESCED 16-byte asynchronous copies
// (LDGSTS), 8 lanes per 128-byte line, into rows padded to an odd number of float4 so that every thread reads its own
// row without bank conflicts.  (One thread reading its 192-byte row straight from global memory costs 32 L1
// wavefronts per load instruction: at 1 M Gaussians the L1 data pipe was 59 % busy for the whole kernel and bounded
// it at 3.4 TB/s.)
template <int MT, bool STAGE_IN>
__global__ void __launch_bounds__(PF_THREADS, GSR_PF_MIN_CTAS) preprocess_fwd_staged_kernel(
    int P, int D, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float* __restrict__ opacities, const float* __restrict__ shs,
    const float* __restrict__ cov3D_precomp, Cam cam, int prefiltered, int* __restrict__ radii, GeomView geom,
    uint32_t* __restrict__ ctrl) {
  constexpr int NF = 3 * MT, F4 = NF / 4, RS4 = F4 | 1;
  __shared__ float4 s_tile[PF_THREADS * RS4];
  __shared__ __align__(16) float s_mean[STAGE_IN ? PF_THREADS * 3 : 4], s_scale[STAGE_IN ? PF_THREADS * 3 : 4];
  __shared__ uint32_t s_vis[PF_THREADS / 32];
  const int tid = threadIdx.x, base = blockIdx.x * PF_THREADS;
  const int i = base + tid;
  const int nG = min(PF_THREADS, P - base);
  float4 q_in = make_float4(0.f, 0.f, 0.f, 0.f);
  float opac_in = 0.f;
  if (i < P) {
    if (rotations != nullptr) q_in = __ldg(reinterpret_cast<const float4*>(rotations) + i);
    opac_in = __ldg(opacities + i);
  }
  // chunks (float4) of a row that hold the (D+1)^2 active coefficients
  const int K = (D + 1) * (D + 1), c4 = (3 * K + 3) >> 2;
  auto issue_sh_copies = [&](bool only_visible) {
    const uint32_t tile_addr = (uint32_t)__cvta_generic_to_shared(s_tile);
    const float4* __restrict__ src4 = reinterpret_cast<const float4*>(shs + (size_t)base * NF);
#pragma unroll 4
    for (int j = tid; j < nG * F4; j += PF_THREADS) {
      const int row = j / F4, col = j - row * F4;
      if (col < c4 && (!only_visible || ((s_vis[row >> 5] >> (row & 31)) & 1u)))
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile_addr + (uint32_t)(row * RS4 + col) * 16u),
                     "l"(src4 + j)
                     : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#if GSR_PF_EARLY_SH
  // The rows of ALL the CTA's Gaussians are requested before the geometry runs, so that the copy overlaps it (the
  // price: the rows of the culled ones, 14 % at config 2).
  issue_sh_copies(false);
#endif
  float m[3] = {0.f, 0.f, 0.f}, sc[3] = {0.f, 0.f, 0.f};
  load_mean_scale<STAGE_IN, PF_THREADS>(P, base, tid, means3D, scales, s_mean, s_scale, m, sc);
  if (cam.raw_params) apply_activations(opac_in, sc, q_in);
  int out_radius = 0, tiles = 0;
  float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0;
  float conx = 0.f, cony = 0.f, conz = 0.f;
  bool visible = false;
  if (i < P)
    visible = pre_geometry(i, m[0], m[1], m[2], sc[0], sc[1], sc[2], q_in, cov3D_precomp, cam, prefiltered, ctrl, o0,
                           conx, cony, conz, out_radius, tiles);
  const unsigned vm = __ballot_sync(0xffffffffu, visible);
  if ((tid & 31) == 0) s_vis[tid >> 5] = vm;
  __syncthreads();
#if !GSR_PF_EARLY_SH
  issue_sh_copies(true);
#endif
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();
  if (i >= P) return;
  if (visible) {
    const float dx = m[0] - __ldg(cam.campos), dy = m[1] - __ldg(cam.campos + 1), dz = m[2] - __ldg(cam.campos + 2);
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float basis[16];
    sh_basis(D, dx / len, dy / len, dz / len, basis);
    const float4* row4 = s_tile + tid * RS4;
    float acc[3] = {0.f, 0.f, 0.f};
    // chunk by chunk (float j of the row belongs to coefficient j / 3, channel j % 3; same accumulation order per
    // channel as forward.cu:30-61), so that only one float4 of the row is live at a time
#pragma unroll
    for (int k4 = 0; k4 < F4; k4++) {
      if (k4 < c4) {
        const float4 v = row4[k4];
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int j = 4 * k4 + u;
          if (j / 3 < K) acc[j % 3] = fmaf(basis[j / 3], e[u], acc[j % 3]);
        }
      }
    }
    const float rr = acc[0] + 0.5f, gg = acc[1] + 0.5f, bb = acc[2] + 0.5f;
    const uint32_t clamped = (rr < 0.f ? 1u : 0u) | (gg < 0.f ? 2u : 0u) | (bb < 0.f ? 4u : 0u);
    // glm::max(x, 0) is (x < 0) ? 0 : x, so a NaN colour stays NaN (forward.cu:70)
    o1 = make_float4(conx, cony, conz, opac_in);
    o2 = make_float4(rr < 0.f ? 0.f : rr, gg < 0.f ? 0.f : gg, bb < 0.f ? 0.f : bb,
                     __uint_as_float(((uint32_t)tiles << 3) | clamped));
  } else {
    out_radius = 0;
    o0 = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  radii[i] = out_radius;
  geom.keys[0][i] = out_radius > 0 ? __float_as_uint(o0.z) : 0xffffffffu;  // depth-sort key (onesweep.cuh)
  geom.g0[i] = o0;
  geom.g1[i] = o1;
  geom.g2[i] = o2;
}

// ------------------------------------------------------------------------------------------------------------------
// Persistent variant with bulk-asynchronous staging (TMA 1-D bulk copies, SASS UBLKCP, completion on an mbarrier).
// A CTA walks tiles of 128 consecutive Gaussians; while it computes tile k from one shared-memory stage, the five
// input blocks of tile k + gridDim.x -- SH rows 128 x 192 B, means3D / scales 128 x 12 B, rotations 128 x 16 B,
// opacities 128 x 4 B, each contiguous in global memory -- are already in flight into the other stage, issued by one
// thread as five cp.async.bulk instructions.  So a CTA's bytes are in flight during its WHOLE life, not only while
// its own geometry runs (the closed loop that held the one-tile-per-CTA kernel at ~4 TB/s).
// Price: the SH rows of culled Gaussians are fetched too (14 % at config 2), and a thread reads its 192-byte row from
// the linear tile with 4-way bank conflicts (the shared-memory pipe is otherwise idle here).
// Used for full tiles when M = 16, scales + rotations given, and every array 16-byte aligned; the host launches the
// one-tile kernel for the last partial tile.
// MEASURED SLOWER than the one-tile kernel (85 vs 70 us at config 2, 208 vs 177 us at 3 M) and therefore not the
// default (gsr_set_option("pre_fwd", 2) selects it): two 30 KB stages allow 3 CTAs = 12 warps per SM, and the geometry
// half -- ~600 dependent instructions per Gaussian -- needs ~24 warps per SM to keep the issue slots busy; at 192 B of
// SH per Gaussian shared memory cannot hold a second tile in flight for that many warps.  Kept as the record of the
// experiment and as a bit-exact second implementation (tests/test_gpu_api.py).
constexpr int PFB_STAGE_SH = PF_THREADS * 192, PFB_STAGE_MEAN = PF_THREADS * 12, PFB_STAGE_ROT = PF_THREADS * 16,
              PFB_STAGE_OPAC = PF_THREADS * 4;
constexpr int PFB_OFF_MEAN = PFB_STAGE_SH, PFB_OFF_SCALE = PFB_OFF_MEAN + PFB_STAGE_MEAN,
              PFB_OFF_ROT = PFB_OFF_SCALE + PFB_STAGE_MEAN, PFB_OFF_OPAC = PFB_OFF_ROT + PFB_STAGE_ROT,
              PFB_STAGE_BYTES = PFB_OFF_OPAC + PFB_STAGE_OPAC;  // 30,208
constexpr int PFB_SMEM = 2 * PFB_STAGE_BYTES;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(mbar),
      "r"(parity)
      : "memory");
}

__global__ void __launch_bounds__(PF_THREADS, 3) preprocess_fwd_bulk_kernel(
    int n_tiles, int D, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float* __restrict__ opacities, const float* __restrict__ shs, Cam cam,
    int prefiltered, int* __restrict__ radii, GeomView geom, uint32_t* __restrict__ ctrl) {
  extern __shared__ __align__(128) unsigned char pfb_smem[];
  __shared__ __align__(8) unsigned long long s_mbar[2];
  const int tid = threadIdx.x;
  const uint32_t smem0 = (uint32_t)__cvta_generic_to_shared(pfb_smem);
  const uint32_t mbar0 = (uint32_t)__cvta_generic_to_shared(s_mbar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar0));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar0 + 8u));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int tile, int stage) {  // one thread: arm the stage's barrier, then the five bulk copies
    const uint32_t mb = mbar0 + 8u * stage, dst = smem0 + (uint32_t)stage * PFB_STAGE_BYTES;
    const size_t g = (size_t)tile * PF_THREADS;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"((uint32_t)PFB_STAGE_BYTES)
                 : "memory");
    bulk_g2s(dst, shs + g * 48, PFB_STAGE_SH, mb);
    bulk_g2s(dst + PFB_OFF_MEAN, means3D + g * 3, PFB_STAGE_MEAN, mb);
    bulk_g2s(dst + PFB_OFF_SCALE, scales + g * 3, PFB_STAGE_MEAN, mb);
    bulk_g2s(dst + PFB_OFF_ROT, rotations + g * 4, PFB_STAGE_ROT, mb);
    bulk_g2s(dst + PFB_OFF_OPAC, opacities + g, PFB_STAGE_OPAC, mb);
  };
  if (tid == 0 && (int)blockIdx.x < n_tiles) issue(blockIdx.x, 0);
  const int K = (D + 1) * (D + 1), c4 = (3 * K + 3) >> 2;
  int it = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
    const int stage = it & 1;
    const int next = tile + gridDim.x;
    if (tid == 0 && next < n_tiles) issue(next, stage ^ 1);  // that stage was released by the barrier below
    mbar_wait(mbar0 + 8u * stage, (uint32_t)((it >> 1) & 1));
    const unsigned char* st = pfb_smem + (size_t)stage * PFB_STAGE_BYTES;
    const float* s_mean = reinterpret_cast<const float*>(st + PFB_OFF_MEAN);
    const float* s_scale = reinterpret_cast<const float*>(st + PFB_OFF_SCALE);
    const int i = tile * PF_THREADS + tid;
    float m[3], sc[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { m[k] = s_mean[3 * tid + k]; sc[k] = s_scale[3 * tid + k]; }
    float4 q_in = reinterpret_cast<const float4*>(st + PFB_OFF_ROT)[tid];
    float opac_in = reinterpret_cast<const float*>(st + PFB_OFF_OPAC)[tid];
    if (cam.raw_params) apply_activations(opac_in, sc, q_in);
    int out_radius = 0, tiles = 0;
    float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0;
    float conx = 0.f, cony = 0.f, conz = 0.f;
    if (pre_geometry(i, m[0], m[1], m[2], sc[0], sc[1], sc[2], q_in, nullptr, cam, prefiltered, ctrl, o0, conx, cony,
                     conz, out_radius, tiles)) {
      const float dx = m[0] - __ldg(cam.campos), dy = m[1] - __ldg(cam.campos + 1), dz = m[2] - __ldg(cam.campos + 2);
      const float len = sqrtf(dx * dx + dy * dy + dz * dz);
      float basis[16];
      sh_basis(D, dx / len, dy / len, dz / len, basis);
      const float4* row4 = reinterpret_cast<const float4*>(st) + tid * 12;
      float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int k4 = 0; k4 < 12; k4++) {
        if (k4 < c4) {
          const float4 v = row4[k4];
          const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int j = 4 * k4 + u;
            if (j / 3 < K) acc[j % 3] = fmaf(basis[j / 3], e[u], acc[j % 3]);
          }
        }
      }
      const float rr = acc[0] + 0.5f, gg = acc[1] + 0.5f, bb = acc[2] + 0.5f;
      const uint32_t clamped = (rr < 0.f ? 1u : 0u) | (gg < 0.f ? 2u : 0u) | (bb < 0.f ? 4u : 0u);
      o1 = make_float4(conx, cony, conz, opac_in);
      o2 = make_float4(rr < 0.f ? 0.f : rr, gg < 0.f ? 0.f : gg, bb < 0.f ? 0.f : bb,
                       __uint_as_float(((uint32_t)tiles << 3) | clamped));
    } else {
      out_radius = 0;
      o0 = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    radii[i] = out_radius;
    geom.keys[0][i] = out_radius > 0 ? __float_as_uint(o0.z) : 0xffffffffu;
    geom.g0[i] = o0;
    geom.g1[i] = o1;
    geom.g2[i] = o2;
    __syncthreads();  // every thread is done with this stage before the next iteration's issue refills it
  }
}

// DGR/cuda_rasterizer/rasterizer_impl.cu:54-66 checkFrustum
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float vz = xf_row(view, 2, __ldg(means3D + 3 * (size_t)i), __ldg(means3D + 3 * (size_t)i + 1),
                          __ldg(means3D + 3 * (size_t)i + 2));
  present[i] = !(vz <= 0.2f);
}

}  // namespace gsr
