// csrc/loss.cuh -- the training loss that consumes the rasterizer's image and produces its dL/dpixel
// (SURVEY.md section 8f, rank 2):  loss = (1 - lambda) * L1(img, gt) + lambda * (1 - SSIM(img, gt)).
// Behaviour: utils/loss_utils.py:17-18 (l1_loss), :24-63 (ssim: 11x11 Gaussian window, sigma 1.5, zero padding 5,
// per-channel, C1 = 0.01^2, C2 = 0.03^2, mean over all elements), combined as in train.py:91-92.
//
// B200 design: the reference runs five depthwise 11x11 convolutions and ~15 elementwise kernels forward, and
// autograd replays them backward.  Here forward AND backward are two HBM-bound passes:
//   pass 1 (per pixel p): the five window means via separable 11-tap filters on a shared-memory tile (TILE + 10
//     halo), the SSIM value, block-reduced sums of SSIM and |img - gt|, and three derivative maps
//     dS/dmu1, dS/dE[x^2], dS/dE[xy] (gt needs no gradient);
//   pass 2 (per pixel q): dL/dimg(q) = (1-l)/n * sign(x - y) - l/n * (conv(dS/dmu1) + 2 x conv(dS/dE11) + y conv(dS/dE12))
//     -- the window is symmetric and the padding zero, so the transpose convolution is the same filter.
// Floating point, tolerance-checked (1e-5 relative on the loss values, 1e-4 of max on the gradient).
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int LT = 16;           // output tile edge
constexpr int LH = 5;            // window half width
constexpr int LE = LT + 2 * LH;  // staged tile edge (26)

// utils/loss_utils.py:24-26 evaluated in float32 exactly as torch does (values printed from the reference code)
__device__ __constant__ float LOSS_G[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f,
                                            0.10936068743467331f,  0.21300552785396576f,   0.26601171493530273f,
                                            0.21300552785396576f,  0.10936068743467331f,   0.036000773310661316f,
                                            0.0075987582094967365f, 0.001028380123898387f};

struct LossAcc {  // device accumulators, zeroed by the host before pass 1
  double sum_abs, sum_ssim;
};

// Separable 11-tap filter of NP staged planes: s_in[NP][LE][LE] -> out[NP] at (tx, ty) of the tile.
// s_mid[NP][LE][LT] holds the horizontally filtered rows.
template <int NP>
__device__ __forceinline__ void sep_filter(const float (*s_in)[LE][LE], float (*s_mid)[LE][LT], int tid, float* out) {
  for (int i = tid; i < LE * LT; i += LT * LT) {
    const int r = i / LT, c = i - r * LT;
#pragma unroll
    for (int p = 0; p < NP; p++) {
      float a = 0.f;
#pragma unroll
      for (int k = 0; k < 11; k++) a = fmaf(LOSS_G[k], s_in[p][r][c + k], a);
      s_mid[p][r][c] = a;
    }
  }
  __syncthreads();
  const int tx = tid & (LT - 1), ty = tid / LT;
#pragma unroll
  for (int p = 0; p < NP; p++) {
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) a = fmaf(LOSS_G[k], s_mid[p][ty + k][tx], a);
    out[p] = a;
  }
}

// grid (ceil(W/16), ceil(H/16), C), block 256
__global__ void __launch_bounds__(LT * LT) loss_pass1_kernel(int H, int W, const float* __restrict__ img,
                                                             const float* __restrict__ gt, float* __restrict__ maps,
                                                             LossAcc* __restrict__ acc) {
  __shared__ float s_in[5][LE][LE];   // x, y, x*x, y*y, x*y
  __shared__ float s_mid[5][LE][LT];
  __shared__ float s_red[2][LT * LT / 32];
  const int tid = threadIdx.x, ch = blockIdx.z;
  const int x0 = blockIdx.x * LT - LH, y0 = blockIdx.y * LT - LH;
  const size_t plane = (size_t)H * W;
  const float* ip = img + ch * plane;
  const float* gp = gt + ch * plane;
  for (int i = tid; i < LE * LE; i += LT * LT) {
    const int r = i / LE, c = i - r * LE, gy = y0 + r, gx = x0 + c;
    float x = 0.f, y = 0.f;  // zero padding (F.conv2d padding = 5)
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      x = __ldg(ip + (size_t)gy * W + gx);
      y = __ldg(gp + (size_t)gy * W + gx);
    }
    s_in[0][r][c] = x; s_in[1][r][c] = y; s_in[2][r][c] = x * x; s_in[3][r][c] = y * y; s_in[4][r][c] = x * y;
  }
  __syncthreads();
  float m[5];
  sep_filter<5>(s_in, s_mid, tid, m);
  const int tx = tid & (LT - 1), ty = tid / LT, px = blockIdx.x * LT + tx, py = blockIdx.y * LT + ty;
  float v_abs = 0.f, v_ssim = 0.f;
  if (px < W && py < H) {
    const float mu1 = m[0], mu2 = m[1];
    const float s11 = m[2] - mu1 * mu1, s22 = m[3] - mu2 * mu2, s12 = m[4] - mu1 * mu2;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float A1 = 2.f * mu1 * mu2 + C1, A2 = 2.f * s12 + C2, B1 = mu1 * mu1 + mu2 * mu2 + C1, B2 = s11 + s22 + C2;
    const float inv = 1.f / (B1 * B2);
    const float S = A1 * A2 * inv;
    v_ssim = S;
    v_abs = fabsf(s_in[0][ty + LH][tx + LH] - s_in[1][ty + LH][tx + LH]);
    // total derivatives w.r.t. the window means of x, x^2, x*y (sigma terms depend on mu1 too)
    const float dmu = 2.f * mu2 * (A2 - A1) * inv + 2.f * mu1 * S * (1.f / B2 - 1.f / B1);
    const float d11 = -S / B2;
    const float d12 = 2.f * A1 * inv;
    const size_t o = (size_t)py * W + px, cp = (size_t)gridDim.z * plane;
    maps[ch * plane + o] = dmu;
    maps[cp + ch * plane + o] = d11;
    maps[2 * cp + ch * plane + o] = d12;
  }
  v_abs = warp_sum(v_abs);
  v_ssim = warp_sum(v_ssim);
  if ((tid & 31) == 0) { s_red[0][tid >> 5] = v_abs; s_red[1][tid >> 5] = v_ssim; }
  __syncthreads();
  if (tid == 0) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int w = 0; w < LT * LT / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
    atomicAdd(&acc->sum_abs, (double)a);
    atomicAdd(&acc->sum_ssim, (double)b);
  }
}

// out[0] = L1, out[1] = SSIM, out[2] = (1 - lambda) * L1 + lambda * (1 - SSIM)   (train.py:91-92)
__global__ void loss_finish_kernel(const LossAcc* __restrict__ acc, double n, float lambda, float* __restrict__ out) {
  const float l1 = (float)(acc->sum_abs / n), ss = (float)(acc->sum_ssim / n);
  out[0] = l1;
  out[1] = ss;
  out[2] = (1.0f - lambda) * l1 + lambda * (1.0f - ss);
}

__global__ void __launch_bounds__(LT * LT) loss_pass2_kernel(int H, int W, const float* __restrict__ img,
                                                             const float* __restrict__ gt,
                                                             const float* __restrict__ maps, float w_l1, float w_ssim,
                                                             float* __restrict__ dL_dimg) {
  __shared__ float s_in[3][LE][LE];
  __shared__ float s_mid[3][LE][LT];
  const int tid = threadIdx.x, ch = blockIdx.z;
  const int x0 = blockIdx.x * LT - LH, y0 = blockIdx.y * LT - LH;
  const size_t plane = (size_t)H * W, cp = (size_t)gridDim.z * plane;
  for (int i = tid; i < LE * LE; i += LT * LT) {
    const int r = i / LE, c = i - r * LE, gy = y0 + r, gx = x0 + c;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
    const size_t o = ch * plane + (size_t)(in ? gy : 0) * W + (in ? gx : 0);
    s_in[0][r][c] = in ? __ldg(maps + o) : 0.f;
    s_in[1][r][c] = in ? __ldg(maps + cp + o) : 0.f;
    s_in[2][r][c] = in ? __ldg(maps + 2 * cp + o) : 0.f;
  }
  __syncthreads();
  float m[3];
  sep_filter<3>(s_in, s_mid, tid, m);
  const int tx = tid & (LT - 1), ty = tid / LT, px = blockIdx.x * LT + tx, py = blockIdx.y * LT + ty;
  if (px < W && py < H) {
    const size_t o = ch * plane + (size_t)py * W + px;
    const float x = __ldg(img + o), y = __ldg(gt + o);
    const float d = x - y;
    const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);  // torch.abs backward: sign(0) = 0
    dL_dimg[o] = w_l1 * sgn - w_ssim * (m[0] + 2.f * x * m[1] + y * m[2]);
  }
}

}  // namespace gsr
