// csrc/loss.cuh -- the training loss that consumes the rasterizer's image and produces its dL/dpixel
// (SURVEY.md section 8f, rank 2):  loss = (1 - lambda) * L1(img, gt) + lambda * (1 - SSIM(img, gt)).
// Behaviour: utils/loss_utils.py:17-18 (l1_loss), :24-63 (ssim: 11x11 Gaussian window, sigma 1.5, zero padding 5,
// per-channel, C1 = 0.01^2, C2 = 0.03^2, mean over all elements), combined as in train.py:91-92.
//
// B200 design: the reference runs five depthwise 11x11 convolutions and ~15 elementwise kernels forward, and
// autograd replays them backward.  Here forward AND backward are two HBM-bound passes:
//   pass 1 (per pixel p): the five window means via separable 11-tap filters on a shared-memory tile (32x32 + 5
//     halo; only x and y are staged, the products are formed on the fly), the SSIM value, block-reduced sums of
//     SSIM and |img - gt|, and three derivative maps dS/dmu1, dS/dE[x^2], dS/dE[xy] (gt needs no gradient);
//   pass 2 (per pixel q): dL/dimg(q) = (1-l)/n * sign(x - y) - l/n * (conv(dS/dmu1) + 2 x conv(dS/dE11) + y conv(dS/dE12))
//     -- the window is symmetric and the padding zero, so the transpose convolution is the same filter.
// Both filter directions are register-blocked sliding windows (8 outputs per thread along the row, 4 along the
// column: 18 / 14 shared loads feed 88 / 44 FMAs per plane) with odd shared-memory row strides, so the kernels sit
// near the FP32 issue rate instead of the shared-memory load rate.
// Floating point, tolerance-checked (1e-5 relative on the loss values, 1e-4 of max on the gradient).
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int LT = 32;              // output tile edge
constexpr int LH = 5;               // window half width
constexpr int LE = LT + 2 * LH;     // staged tile edge (42)
constexpr int LIS = LE + 1;         // s_in row stride (odd: lanes that walk down a column hit distinct banks)
constexpr int LMS = LT + 1;         // s_mid row stride (odd, same reason)
constexpr int LOSS_THREADS = 256;
constexpr int HSEG = 8;             // horizontal pass: outputs per thread
constexpr int VSEG = LT * LT / LOSS_THREADS;  // vertical pass: outputs per thread (4)
constexpr int LOSS_SLOTS = 32;      // partial sums are spread over this many accumulators
static_assert(LT % HSEG == 0 && VSEG * LOSS_THREADS == LT * LT && LOSS_THREADS / 32 * VSEG == LT, "loss tiling");

// utils/loss_utils.py:24-26 evaluated in float32 exactly as torch does (tests/golden/data/loss_window.npy)
#define LOSS_G0 0.001028380123898387f
#define LOSS_G1 0.0075987582094967365f
#define LOSS_G2 0.036000773310661316f
#define LOSS_G3 0.10936068743467331f
#define LOSS_G4 0.21300552785396576f
#define LOSS_G5 0.26601171493530273f
__device__ __forceinline__ constexpr float loss_g(int k) {
  constexpr float g[11] = {LOSS_G0, LOSS_G1, LOSS_G2, LOSS_G3, LOSS_G4, LOSS_G5, LOSS_G4, LOSS_G3, LOSS_G2, LOSS_G1, LOSS_G0};
  return g[k];
}

struct LossAcc {  // device accumulators, zeroed by the host before pass 1
  double sum_abs, sum_ssim;
};

// Horizontal 11-tap filter of the staged tile into s_mid[NP][LE][LMS].  PRODUCTS: s_in holds x, y and the five
// filtered planes are x, y, x*x, y*y, x*y; otherwise NP planes of s_in are filtered as they are.
template <int NP, bool PRODUCTS>
__device__ __forceinline__ void filter_rows(const float* __restrict__ s_in, float* __restrict__ s_mid, int tid) {
  for (int it = tid; it < LE * (LT / HSEG); it += LOSS_THREADS) {
    const int seg = it / LE, r = it - seg * LE;  // lanes walk down rows
    float acc[NP][HSEG];
#pragma unroll
    for (int p = 0; p < NP; p++)
#pragma unroll
      for (int o = 0; o < HSEG; o++) acc[p][o] = 0.f;
    const float* in = s_in + r * LIS + seg * HSEG;
#pragma unroll
    for (int j = 0; j < HSEG + 2 * LH; j++) {
      float v[NP];
      if (PRODUCTS) {
        const float x = in[j], y = in[LE * LIS + j];
        v[0] = x; v[1] = y; v[2] = x * x; v[3] = y * y; v[4] = x * y;
      } else {
#pragma unroll
        for (int p = 0; p < NP; p++) v[p] = in[p * LE * LIS + j];
      }
#pragma unroll
      for (int o = 0; o < HSEG; o++) {
        if (j - o >= 0 && j - o < 11) {
#pragma unroll
          for (int p = 0; p < NP; p++) acc[p][o] = fmaf(loss_g(j - o), v[p], acc[p][o]);
        }
      }
    }
    float* out = s_mid + r * LMS + seg * HSEG;
#pragma unroll
    for (int p = 0; p < NP; p++)
#pragma unroll
      for (int o = 0; o < HSEG; o++) out[p * LE * LMS + o] = acc[p][o];
  }
}

// Vertical 11-tap filter: thread (column tid & 31, rows (tid >> 5) * VSEG ..) -> out[NP][VSEG]
template <int NP>
__device__ __forceinline__ void filter_cols(const float* __restrict__ s_mid, int tid, float (*out)[VSEG]) {
  const int c = tid & 31, r0 = (tid >> 5) * VSEG;
#pragma unroll
  for (int p = 0; p < NP; p++)
#pragma unroll
    for (int o = 0; o < VSEG; o++) out[p][o] = 0.f;
#pragma unroll
  for (int j = 0; j < VSEG + 2 * LH; j++) {
#pragma unroll
    for (int p = 0; p < NP; p++) {
      const float v = s_mid[p * LE * LMS + (r0 + j) * LMS + c];
#pragma unroll
      for (int o = 0; o < VSEG; o++)
        if (j - o >= 0 && j - o < 11) out[p][o] = fmaf(loss_g(j - o), v, out[p][o]);
    }
  }
}

// grid (ceil(W/32), ceil(H/32), C), block 256
__global__ void __launch_bounds__(LOSS_THREADS) loss_pass1_kernel(int H, int W, const float* __restrict__ img,
                                                                  const float* __restrict__ gt,
                                                                  float* __restrict__ maps, LossAcc* __restrict__ acc) {
  __shared__ float s_in[2 * LE * LIS];   // x, y
  __shared__ float s_mid[5 * LE * LMS];  // rows of x, y, x*x, y*y, x*y filtered horizontally
  __shared__ float s_red[2][LOSS_THREADS / 32];
  const int tid = threadIdx.x, ch = blockIdx.z;
  const int x0 = blockIdx.x * LT - LH, y0 = blockIdx.y * LT - LH;
  const size_t plane = (size_t)H * W;
  const float* ip = img + ch * plane;
  const float* gp = gt + ch * plane;
  for (int i = tid; i < LE * LE; i += LOSS_THREADS) {
    const int r = i / LE, c = i - r * LE, gy = y0 + r, gx = x0 + c;
    float x = 0.f, y = 0.f;  // zero padding (F.conv2d padding = 5)
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      x = __ldg(ip + (size_t)gy * W + gx);
      y = __ldg(gp + (size_t)gy * W + gx);
    }
    s_in[r * LIS + c] = x;
    s_in[LE * LIS + r * LIS + c] = y;
  }
  __syncthreads();
  filter_rows<5, true>(s_in, s_mid, tid);
  __syncthreads();
  float m[5][VSEG];
  filter_cols<5>(s_mid, tid, m);
  const int tx = tid & 31, ty0 = (tid >> 5) * VSEG, px = blockIdx.x * LT + tx;
  float v_abs = 0.f, v_ssim = 0.f;
  const size_t cp = (size_t)gridDim.z * plane;
#pragma unroll
  for (int o = 0; o < VSEG; o++) {
    const int py = blockIdx.y * LT + ty0 + o;
    if (px < W && py < H) {
      const float mu1 = m[0][o], mu2 = m[1][o];
      const float s11 = m[2][o] - mu1 * mu1, s22 = m[3][o] - mu2 * mu2, s12 = m[4][o] - mu1 * mu2;
      const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
      const float A1 = 2.f * mu1 * mu2 + C1, A2 = 2.f * s12 + C2, B1 = mu1 * mu1 + mu2 * mu2 + C1, B2 = s11 + s22 + C2;
      const float iB1 = 1.f / B1, iB2 = 1.f / B2, inv = iB1 * iB2;
      const float S = A1 * A2 * inv;
      v_ssim += S;
      const int ci = (ty0 + o + LH) * LIS + tx + LH;
      v_abs += fabsf(s_in[ci] - s_in[LE * LIS + ci]);
      // total derivatives w.r.t. the window means of x, x^2, x*y (the sigma terms depend on mu1 too)
      const size_t q = ch * plane + (size_t)py * W + px;
      maps[q] = 2.f * mu2 * (A2 - A1) * inv + 2.f * mu1 * S * (iB2 - iB1);
      maps[cp + q] = -S * iB2;
      maps[2 * cp + q] = 2.f * A1 * inv;
    }
  }
  v_abs = warp_sum(v_abs);
  v_ssim = warp_sum(v_ssim);
  if ((tid & 31) == 0) { s_red[0][tid >> 5] = v_abs; s_red[1][tid >> 5] = v_ssim; }
  __syncthreads();
  if (tid == 0) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int w = 0; w < LOSS_THREADS / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
    LossAcc* slot = acc + ((blockIdx.x + 7 * blockIdx.y + 13 * blockIdx.z) & (LOSS_SLOTS - 1));
    atomicAdd(&slot->sum_abs, (double)a);
    atomicAdd(&slot->sum_ssim, (double)b);
  }
}

// out[0] = L1, out[1] = SSIM, out[2] = (1 - lambda) * L1 + lambda * (1 - SSIM)   (train.py:91-92)
__global__ void loss_finish_kernel(const LossAcc* __restrict__ acc, double n, float lambda, float* __restrict__ out) {
  double a = 0.0, b = 0.0;
  for (int s = 0; s < LOSS_SLOTS; s++) { a += acc[s].sum_abs; b += acc[s].sum_ssim; }
  const float l1 = (float)(a / n), ss = (float)(b / n);
  out[0] = l1;
  out[1] = ss;
  out[2] = (1.0f - lambda) * l1 + lambda * (1.0f - ss);
}

__global__ void __launch_bounds__(LOSS_THREADS) loss_pass2_kernel(int H, int W, const float* __restrict__ img,
                                                                  const float* __restrict__ gt,
                                                                  const float* __restrict__ maps, float w_l1,
                                                                  float w_ssim, const float* __restrict__ upstream,
                                                                  float* __restrict__ dL_dimg) {
  __shared__ float s_in[3 * LE * LIS];
  __shared__ float s_mid[3 * LE * LMS];
  const int tid = threadIdx.x, ch = blockIdx.z;
  const int x0 = blockIdx.x * LT - LH, y0 = blockIdx.y * LT - LH;
  const size_t plane = (size_t)H * W, cp = (size_t)gridDim.z * plane;
  for (int i = tid; i < LE * LE; i += LOSS_THREADS) {
    const int r = i / LE, c = i - r * LE, gy = y0 + r, gx = x0 + c;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
    const size_t o = ch * plane + (size_t)(in ? gy : 0) * W + (in ? gx : 0);
    s_in[r * LIS + c] = in ? __ldg(maps + o) : 0.f;
    s_in[LE * LIS + r * LIS + c] = in ? __ldg(maps + cp + o) : 0.f;
    s_in[2 * LE * LIS + r * LIS + c] = in ? __ldg(maps + 2 * cp + o) : 0.f;
  }
  __syncthreads();
  filter_rows<3, false>(s_in, s_mid, tid);
  __syncthreads();
  float m[3][VSEG];
  filter_cols<3>(s_mid, tid, m);
  const int tx = tid & 31, ty0 = (tid >> 5) * VSEG, px = blockIdx.x * LT + tx;
  if (upstream != nullptr) {  // d(loss)/d(img) scaled by the upstream gradient of the loss (a device scalar)
    const float u = __ldg(upstream);
    w_l1 *= u;
    w_ssim *= u;
  }
#pragma unroll
  for (int o = 0; o < VSEG; o++) {
    const int py = blockIdx.y * LT + ty0 + o;
    if (px < W && py < H) {
      const size_t q = ch * plane + (size_t)py * W + px;
      const float x = __ldg(img + q), y = __ldg(gt + q);
      const float d = x - y;
      const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);  // torch.abs backward: sign(0) = 0
      dL_dimg[q] = w_l1 * sgn - w_ssim * (m[0][o] + 2.f * x * m[1][o] + y * m[2][o]);
    }
  }
}

}  // namespace gsr
