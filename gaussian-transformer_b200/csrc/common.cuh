// csrc/common.cuh -- shared constants, scratch layout and small device helpers
// for libgsrast (B200 / sm_100a).  Own code; the reference
// (DGR = /root/reference/submodules/diff-gaussian-rasterization) is cited only
// for the behaviour each piece has to reproduce.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsr {

// DGR/cuda_rasterizer/config.h:15-17 -- tile ids / ranges must be bit-exact, so
// the tile stays 16x16 and there are 3 colour channels.
constexpr int TILE = 16;
constexpr int TILE_PIX = TILE * TILE;
constexpr int NCH = 3;

// DGR/cuda_rasterizer/auxiliary.h:22-39
__device__ constexpr float SH_C0 = 0.28209479177387814f;
__device__ constexpr float SH_C1 = 0.4886025119029199f;
__device__ constexpr float SH_C2_0 = 1.0925484305920792f;
__device__ constexpr float SH_C2_1 = -1.0925484305920792f;
__device__ constexpr float SH_C2_2 = 0.31539156525252005f;
__device__ constexpr float SH_C2_3 = -1.0925484305920792f;
__device__ constexpr float SH_C2_4 = 0.5462742152960396f;
__device__ constexpr float SH_C3_0 = -0.5900435899266435f;
__device__ constexpr float SH_C3_1 = 2.890611442640554f;
__device__ constexpr float SH_C3_2 = -0.4570457994644658f;
__device__ constexpr float SH_C3_3 = 0.3731763325901154f;
__device__ constexpr float SH_C3_4 = -0.4570457994644658f;
__device__ constexpr float SH_C3_5 = 1.445305721320277f;
__device__ constexpr float SH_C3_6 = -0.5900435899266435f;

constexpr size_t ALIGN = 256;
__host__ __device__ inline size_t align_up(size_t v) { return (v + ALIGN - 1) & ~(ALIGN - 1); }

// ---------------------------------------------------------------- scratch layout
// geom  (per Gaussian, written by preprocess_fwd, read by binning / render / bwd)
//   g0 float4[P]  mean2D.x, mean2D.y, depth, radius (int bits)
//   g1 float4[P]  conic.x, conic.y, conic.z, opacity
//   g2 float4[P]  r, g, b, flags (bits 0..2 = clamped r/g/b, bits 3.. = tiles_touched)
//   keys[2] u32[P], vals[2] u32[P]  depth-sort ping-pong (keys = depth bits, 0xffffffff for culled)
//   cta_hist u32[sort_scratch_words(P)]  sort scratch: onesweep.cuh digit histograms + look-back status words
//                 (or depth_sort.cuh's digit-major per-CTA histograms + digit totals)
//   rects uint2[P]  clipped tile rectangle of the k-th Gaussian in depth order (binning.cuh)
// image
//   final_T f32[N], n_contrib u32[N], ranges uint2[T], tile_count u32[T], ctrl u32[64],
//   matrix u32[n_chunks * T] (binning.cuh)
// binning
//   point_list u32[R] (depth-sorted gaussian ids, tile after tile)
// bwd
//   acc0 float4[P]: dL_dcolor.rgb, dL_dopacity;  acc1 float4[P]: sum q dx, sum q dy, sum q dx dx, sum q dx dy;
//   acc2 float[P]: sum q dy dy   (SoA so that preprocess_bwd reads them with coalesced 128-bit loads).
//   acc1 / acc2 are raw moments of q = G * dL/dalpha; finish_2d_grads() turns them into dL_dmean2D / dL_dconic.
struct GeomView {
  float4* g0;
  float4* g1;
  float4* g2;
  uint32_t* keys[2];
  uint32_t* vals[2];
  uint32_t* cta_hist;
  uint2* rects;
};
struct ImageView {
  float* final_T;
  uint32_t* n_contrib;
  uint2* ranges;
  uint32_t* tile_count;
  uint32_t* ctrl;
  uint32_t* matrix;
};
struct BinView {
  uint32_t* point_list;
};
struct BwdView {
  float4* acc0;
  float4* acc1;
  float* acc2;
  unsigned long long* counters;  // 3 test counters in the buffer's tail (render_bwd2_kernel<true> only)
};

// ctrl[] slots
enum { CTRL_R_LO = 4, CTRL_R_HI = 5, CTRL_FLAGS = 6, CTRL_WORDS = 64 };  // words 4..7 go to the host in one copy

#ifndef GSR_SORT_ROUNDS
#define GSR_SORT_ROUNDS 4
#endif
constexpr int SORT_TILE_KEYS = 256 * GSR_SORT_ROUNDS;  // keys per CTA in depth_sort.cuh (RS_TILE)
inline size_t sort_ctas(size_t P) { return (P + SORT_TILE_KEYS - 1) / SORT_TILE_KEYS; }
// onesweep.cuh: 4096 keys per tile; scratch words = [4][256] digit histograms, 64 words of tile counters,
// [4][tiles + groups][256] look-back status words
// onesweep.cuh: tiles of 256 threads x 16 keys (4096), or x 4 keys (1024) for inputs of at most OS_SMALL_N keys
// (small scenes: more, shorter tiles -- the pass is latency bound there).  Scratch words = [4][256] digit histograms,
// 64 words of tile counters, [4][tiles + groups][256] look-back status words
constexpr size_t OS_SMALL_N = 131072;
constexpr int OS_GROUP_TILES = 16;
inline int os_kpt(size_t n) { return n <= OS_SMALL_N ? 4 : 16; }
inline size_t os_tiles(size_t n) { const size_t t = 256 * (size_t)os_kpt(n); return (n + t - 1) / t; }
inline size_t os_groups(size_t n) { return (os_tiles(n) + OS_GROUP_TILES - 1) / OS_GROUP_TILES; }
inline size_t os_scratch_words(size_t n) {
  // both tilings' needs are covered for any n (the small tiling only ever sees n <= OS_SMALL_N)
  const size_t small_n = n < OS_SMALL_N ? n : OS_SMALL_N;
  const size_t a = (n + 4095) / 4096, b = (small_n + 1023) / 1024;
  const size_t t = a > b ? a : b;
  return 4 * 256 + 64 + 4 * (t + (t + OS_GROUP_TILES - 1) / OS_GROUP_TILES + 1) * 256;
}
// words of sort scratch behind keys / vals: whichever of the two sorts needs more
inline size_t sort_scratch_words(size_t P) {
  const size_t a = 256 * (sort_ctas(P) + 1), b = os_scratch_words(P);
  return a > b ? a : b;
}
// Gaussians per binning chunk (binning.cuh): a power of two >= 1024 with at most 1024 chunks; scenes of up to 64 K
// Gaussians get chunks of 128 (more parallel walks: 33.7 K Gaussians at 800x450 forward 0.201 -> 0.171 ms; at 200 K
// Gaussians finer chunks measured 5 % slower, hence the threshold) while the chunks x tiles matrix stays < 32 MB.
constexpr size_t BIN_MATRIX_SOFT_CAP = (size_t)32 << 20;
inline int bin_chunk_shift(size_t P, size_t T) {
  int shift = P <= 65536 ? 7 : 10;
  while (((size_t)1 << shift) * 1024 < P) shift++;
  while (shift < 10 && ((P + ((size_t)1 << shift) - 1) >> shift) * T * 4 > BIN_MATRIX_SOFT_CAP) shift++;
  return shift;
}
inline size_t bin_chunks(size_t P, size_t T) {
  const int shift = bin_chunk_shift(P, T);
  return P ? (P + ((size_t)1 << shift) - 1) >> shift : 0;
}

inline size_t geom_bytes(size_t P) {
  return 3 * align_up(16 * P) + 4 * align_up(4 * P) + align_up(4 * sort_scratch_words(P)) + align_up(8 * P) + ALIGN;
}
inline GeomView geom_view(void* base, size_t P) {
  char* p = (char*)base;
  GeomView v;
  v.g0 = (float4*)p; p += align_up(16 * P);
  v.g1 = (float4*)p; p += align_up(16 * P);
  v.g2 = (float4*)p; p += align_up(16 * P);
  v.keys[0] = (uint32_t*)p; p += align_up(4 * P);
  v.keys[1] = (uint32_t*)p; p += align_up(4 * P);
  v.vals[0] = (uint32_t*)p; p += align_up(4 * P);
  v.vals[1] = (uint32_t*)p; p += align_up(4 * P);
  v.cta_hist = (uint32_t*)p; p += align_up(4 * sort_scratch_words(P));
  v.rects = (uint2*)p;
  return v;
}
inline size_t image_bytes(size_t N, size_t T, size_t P) {
  return 2 * align_up(4 * N) + align_up(8 * T) + align_up(4 * T) + align_up(4 * CTRL_WORDS) +
         align_up(4 * T * bin_chunks(P, T)) + ALIGN;
}
inline ImageView image_view(void* base, size_t N, size_t T) {
  char* p = (char*)base;
  ImageView v;
  v.final_T = (float*)p; p += align_up(4 * N);
  v.n_contrib = (uint32_t*)p; p += align_up(4 * N);
  v.ranges = (uint2*)p; p += align_up(8 * T);
  v.tile_count = (uint32_t*)p; p += align_up(4 * T);
  v.ctrl = (uint32_t*)p; p += align_up(4 * CTRL_WORDS);
  v.matrix = (uint32_t*)p;
  return v;
}
inline size_t bin_bytes(size_t R) { return align_up(4 * R) + ALIGN; }
inline BinView bin_view(void* base, size_t R) {
  (void)R;
  BinView v;
  v.point_list = (uint32_t*)base;
  return v;
}
inline size_t bwd_bytes(size_t P) { return 2 * align_up(16 * P) + align_up(4 * P) + ALIGN; }
inline BwdView bwd_view(void* base, size_t P) {
  char* p = (char*)base;
  BwdView v;
  v.acc0 = (float4*)p; p += align_up(16 * P);
  v.acc1 = (float4*)p; p += align_up(16 * P);
  v.acc2 = (float*)p; p += align_up(4 * P);
  v.counters = (unsigned long long*)p;  // inside the trailing ALIGN bytes of bwd_bytes()
  return v;
}

// Camera / raster constants handed to kernels by value.
struct Cam {
  const float* view;  // 16 floats, column-major
  const float* proj;  // 16 floats
  const float* campos;
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  int W, H, gx, gy;
  int raw_params;  // opacities / scales / rotations are the RAW parameters: the kernels apply the activations
};

// Bounds-checked build (make checked: -DGSR_BOUNDS_CHECK -> lib/libgsrast_checked.so): every index into a
// shared-memory table and every scattered global store of the binning / sort / staging code is checked and the
// kernel traps on a violation.  compute-sanitizer is closed on the GPU pool this was developed on; running the small
// parity cases against the checked library (GSR_LIB_PATH) is the substitute (profiles/r02_bounds_check.txt).
#ifdef GSR_BOUNDS_CHECK
#define GSR_CHECK(cond)        \
  do {                         \
    if (!(cond)) __trap();     \
  } while (0)
#else
#define GSR_CHECK(cond) ((void)0)
#endif

// ---------------------------------------------------------------- device helpers
// m[r]*x + m[4+r]*y + m[8+r]*z + m[12+r] with the rounding points of the
// reference's compiled transformPoint4x3/4x4 (DGR/cuda_rasterizer/auxiliary.h:58-77;
// SASS: FADD(FFMA(z,m8,FFMA(x,m0,FMUL(y,m4))),m12)).  Explicit intrinsics so the
// compiler can neither contract nor re-associate.
__device__ __forceinline__ float xf_row(const float* __restrict__ m, int r, float x, float y, float z) {
  float t = __fmul_rn(y, m[4 + r]);
  t = __fmaf_rn(x, m[r], t);
  t = __fmaf_rn(z, m[8 + r], t);
  return __fadd_rn(t, m[12 + r]);
}

// a0*b0 + a1*b1 + a2*b2 as the reference's mat3 products compile:
// FFMA(a2,b2,FFMA(a0,b0,FMUL(a1,b1))).
__device__ __forceinline__ float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
  return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}

// DGR/cuda_rasterizer/auxiliary.h:41-44 ndc2Pix, evaluated in FP64 like the
// reference (DFMA((v+1), S, -1) * 0.5, then rounded to float).
__device__ __forceinline__ float ndc2pix(float v, int S) {
  double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
  return __double2float_rn(__dmul_rn(t, 0.5));
}

// DGR/cuda_rasterizer/auxiliary.h:46-56 getRect (C truncation = F2I.TRUNC,
// saturating, NaN -> 0).
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy,
                                         int& minx, int& miny, int& maxx, int& maxy) {
  float r = (float)radius;
  minx = min(gx, max(0, __float2int_rz(__fmul_rn(__fsub_rn(px, r), 0.0625f))));
  miny = min(gy, max(0, __float2int_rz(__fmul_rn(__fsub_rn(py, r), 0.0625f))));
  maxx = min(gx, max(0, __float2int_rz(__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(px, r), 16.0f), 1.0f), 0.0625f))));
  maxy = min(gy, max(0, __float2int_rz(__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(py, r), 16.0f), 1.0f), 0.0625f))));
}

// Gaussian exponent at pixel offset (dx,dy) with the reference's rounding
// (DGR/cuda_rasterizer/forward.cu:337-339; SASS:
// FFMA(FFMA(dx, dx*A, dy*(dy*C)), -0.5, -(dy*(dx*B)))).
__device__ __forceinline__ float gauss_power(float dx, float dy, float A, float B, float C) {
  float t1 = __fmul_rn(dy, __fmul_rn(dy, C));
  float s = __fmaf_rn(dx, __fmul_rn(dx, A), t1);
  float v = __fmul_rn(dy, __fmul_rn(dx, B));
  return __fmaf_rn(s, -0.5f, -v);
}

// Same value, same roundings, with Blackwell's packed fp32 pipe (FADD2 / FMUL2: two IEEE fp32 operations per
// issued instruction; the blend kernels are issue-bound).  m = (mean.x, mean.y), npix = (-pix.x, -pix.y),
// AC = (conic A, conic C).  d = (dx, dy) and p = (dx*A, dy*C) are returned for reuse by the backward.
__device__ __forceinline__ float gauss_power2(float2 m, float2 npix, float2 AC, float B, float2& d, float2& p) {
  d = __fadd2_rn(m, npix);
  p = __fmul2_rn(d, AC);
  const float t1 = __fmul_rn(d.y, p.y);
  const float s = __fmaf_rn(d.x, p.x, t1);
  const float v = __fmul_rn(d.y, __fmul_rn(d.x, B));
  return __fmaf_rn(s, -0.5f, -v);
}

// The parameter activations of scene/gaussian_model.py:43-51,95-106 as torch's CUDA kernels evaluate them (bit for
// bit on this GPU, tests/test_gpu_activations.py): sigmoid = 1 / (1 + expf(-x)), exp = expf,
// F.normalize = x / max(||x||_2, 1e-12) with ||q||_2 = sqrt((x^2 + z^2) + (y^2 + w^2)), every product rounded.
__device__ __forceinline__ float quat_norm(float4 q) {
  return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.x, q.x), __fmul_rn(q.z, q.z)),
                              __fadd_rn(__fmul_rn(q.y, q.y), __fmul_rn(q.w, q.w))));
}
__device__ __forceinline__ float act_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }
__device__ __forceinline__ float4 act_normalize(float4 q) {
  const float n = fmaxf(quat_norm(q), 1e-12f);
  return make_float4(__fdiv_rn(q.x, n), __fdiv_rn(q.y, n), __fdiv_rn(q.z, n), __fdiv_rn(q.w, n));
}

__device__ __forceinline__ float warp_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

// render_bwd accumulates the raw moments  sum q dx, sum q dy, sum q dx dx, sum q dx dy, sum q dy dy  with
// q = G * dL/dalpha and (dx, dy) = mean2D - pixel; the reference's 2-D gradients (backward.cu:541-551) are linear in
// them with per-Gaussian coefficients (conic, opacity, -1, -0.5, 0.5*W, 0.5*H), applied here once per Gaussian.
// con_o = (conic A, B, C, opacity) as the forward saw them.
__device__ __forceinline__ void finish_2d_grads(float4 con_o, int W, int H, float4& B, float& Cyy) {
  const float mo = -con_o.w;
  const float sx = B.x, sy = B.y;
  B.x = mo * (0.5f * W) * (con_o.x * sx + con_o.y * sy);
  B.y = mo * (0.5f * H) * (con_o.y * sx + con_o.z * sy);
  B.z *= 0.5f * mo;
  B.w *= 0.5f * mo;
  Cyy *= 0.5f * mo;
}

}  // namespace gsr
