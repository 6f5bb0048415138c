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

template <bool VEC_SH>
__global__ void __launch_bounds__(256) preprocess_fwd_kernel(
    int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
    const float* __restrict__ rotations, const float* __restrict__ opacities, const float* __restrict__ shs,
    const float* __restrict__ sh_rest, const float* __restrict__ colors_precomp,
    const float* __restrict__ cov3D_precomp, Cam cam, int prefiltered,
    int* __restrict__ radii, GeomView geom, uint32_t* __restrict__ ctrl) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  int out_radius = 0;
  float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0;
  do {
    const float px3 = __ldg(means3D + 3 * (size_t)i), py3 = __ldg(means3D + 3 * (size_t)i + 1),
                pz3 = __ldg(means3D + 3 * (size_t)i + 2);
    const float* __restrict__ view = cam.view;
    const float* __restrict__ proj = cam.proj;
    // in_frustum: cull iff p_view.z <= 0.2 (NaN passes, like the reference)
    const float vz = xf_row(view, 2, px3, py3, pz3);
    if (vz <= 0.2f) {
      if (prefiltered) atomicOr(&ctrl[CTRL_FLAGS], 1u);
      break;
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
      const float s0 = __ldg(scales + 3 * (size_t)i), s1 = __ldg(scales + 3 * (size_t)i + 1),
                  s2 = __ldg(scales + 3 * (size_t)i + 2);
      const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + i);
      cov3d_from_scale_rot(s0, s1, s2, cam.scale_modifier, q.x, q.y, q.z, q.w, c3);
    }
    const float tx = xf_row(view, 0, px3, py3, pz3);
    const float ty = xf_row(view, 1, px3, py3, pz3);
    float a, b, c;
    cov2d(tx, ty, vz, cam, c3, view, a, b, c);
    const float det = __fmaf_rn(a, c, -__fmul_rn(b, b));
    if (det == 0.0f) break;
    const float det_inv = __frcp_rn(det);
    const float conx = __fmul_rn(c, det_inv), cony = __fmul_rn(b, -det_inv), conz = __fmul_rn(a, det_inv);
    const float mid = __fmul_rn(__fadd_rn(a, c), 0.5f);
    const float disc = fmaxf(__fmaf_rn(mid, mid, -det), 0.1f);
    const float sq = __fsqrt_rn(disc);
    const float lmax = fmaxf(__fadd_rn(mid, sq), __fsub_rn(mid, sq));
    const int my_radius = __float2int_rz(ceilf(__fmul_rn(__fsqrt_rn(lmax), 3.0f)));
    // lambda_max >= 0.3 for finite input, so radius >= 2; radius 0 only arises from a NaN covariance.  The
    // reference then still counts the Gaussian's tile but never writes its sort key (duplicateWithKeys skips
    // radii <= 0, rasterizer_impl.cu:84), i.e. it renders an uninitialised key.  Defined behaviour here: culled.
    if (my_radius <= 0) break;
    const float mx = ndc2pix(projx, cam.W), my = ndc2pix(projy, cam.H);
    int minx, miny, maxx, maxy;
    get_rect(mx, my, my_radius, cam.gx, cam.gy, minx, miny, maxx, maxy);
    const int tiles = (maxx - minx) * (maxy - miny);
    if (tiles == 0) break;
    float r, g, bl;
    uint32_t clamped = 0;
    if (colors_precomp != nullptr) {
      r = __ldg(colors_precomp + 3 * (size_t)i);
      g = __ldg(colors_precomp + 3 * (size_t)i + 1);
      bl = __ldg(colors_precomp + 3 * (size_t)i + 2);
    } else {
      const float dx = px3 - __ldg(cam.campos), dy = py3 - __ldg(cam.campos + 1), dz = pz3 - __ldg(cam.campos + 2);
      if (sh_rest != nullptr)
        sh_to_rgb<false>(D, M, shs + (size_t)i * 3, sh_rest + (size_t)i * (M - 1) * 3, dx, dy, dz, r, g, bl, clamped);
      else
        sh_to_rgb<VEC_SH>(D, M, shs + (size_t)i * M * 3, nullptr, dx, dy, dz, r, g, bl, clamped);
    }
    out_radius = my_radius;
    o0 = make_float4(mx, my, vz, __int_as_float(my_radius));
    o1 = make_float4(conx, cony, conz, __ldg(opacities + i));
    o2 = make_float4(r, g, bl, __uint_as_float(((uint32_t)tiles << 3) | clamped));
  } while (0);
  radii[i] = out_radius;
  geom.keys[0][i] = out_radius > 0 ? __float_as_uint(o0.z) : 0xffffffffu;  // depth-sort key (depth_sort.cuh)
  geom.g0[i] = o0;
  geom.g1[i] = o1;
  geom.g2[i] = o2;
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
