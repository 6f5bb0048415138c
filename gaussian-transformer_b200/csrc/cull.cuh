// csrc/cull.cuh -- conservative "can this Gaussian reach alpha >= 1/255 anywhere in
// this pixel rectangle?" test, shared by render_fwd and render_bwd.
//
// The reference (DGR/cuda_rasterizer/forward.cu:325-362) evaluates every Gaussian
// of a tile's list at every pixel and skips it per pixel when power > 0 or
// alpha < 1/255.  The lists come from the bounding square of a 3-sigma circle
// of the LARGEST eigenvalue (forward.cu:226-231), so for anisotropic or
// low-opacity Gaussians most (tile, Gaussian) instances contribute to no pixel.
// Lists, ranges and n_contrib stay bit-exact; we only avoid *evaluating*
// instances that the reference's own per-pixel tests would reject at every
// pixel of the rectangle.  The test must therefore never drop an instance that
// the reference's fp32 arithmetic would accept somewhere: it bounds the
// exponent from above over the continuous rectangle (a superset of the pixel
// centres) and adds a margin that dominates the rounding error of the
// reference's evaluation (<= ~8 ulp of the sum of the magnitudes of its terms).
#pragma once
#include "common.cuh"

namespace gsr {

struct CullPre {   // per (tile, Gaussian), computed once at staging
  float thr;       // cull iff min over rect of (-power) > thr ; +inf = never cull
  float rBC, rBA;  // -B/C, -B/A
};

// Tile-level precomputation.  [x0,x1]x[y0,y1] is the tile's pixel rectangle.
// Returns false when the instance is provably irrelevant for the whole tile.
__device__ __forceinline__ bool cull_prepare(float mx, float my, float A, float B, float C, float o, float x0,
                                             float x1, float y0, float y1, CullPre& pre) {
  pre.thr = __int_as_float(0x7f800000);  // +inf
  pre.rBC = 0.f;
  pre.rBA = 0.f;
  // alpha = min(0.99, o * G) with G <= 1 wherever power <= 0, and fp multiplication is monotone: alpha <= o.
  // (1 - 1e-5 guards an expf(power <= 0) that rounds a hair above 1)
  if (o < (1.0f / 255.0f) * 0.99999f) return false;
  const float det = A * C - B * B;
  // convex (-power) needed for the edge search; anything else (incl. NaN) is never culled
  if (!(A > 0.f && C > 0.f && det > 1e-6f * A * C)) return true;
  const float Dx = fmaxf(fabsf(x0 - mx), fabsf(x1 - mx)), Dy = fmaxf(fabsf(y0 - my), fabsf(y1 - my));
  const float smax = 0.5f * (A * Dx * Dx + C * Dy * Dy) + fabsf(B) * Dx * Dy;
  const float thr = __logf(255.0f * o) + 1e-3f + 1e-5f * smax;
  if (!(thr < 3.0e38f)) return true;  // inf / NaN: keep, never cull
  pre.thr = thr;
  pre.rBC = -B / C;
  pre.rBA = -B / A;
  return true;
}

// min over the rectangle of f(q) = 0.5*(A qx^2 + C qy^2) + B qx qy (q = pixel - mean), f convex.
// Returns true when min f > thr, i.e. no point of the rectangle can reach alpha >= 1/255.
__device__ __forceinline__ bool cull_rect(float mx, float my, float A, float B, float C, const CullPre& pre,
                                          float x0, float x1, float y0, float y1) {
  const float ax0 = x0 - mx, ax1 = x1 - mx, ay0 = y0 - my, ay1 = y1 - my;
  const float cx = fminf(fmaxf(0.f, ax0), ax1), cy = fminf(fmaxf(0.f, ay0), ay1);  // nearest point to the mean
  float fmin = 3.0e38f;
  if (cx != 0.f) {  // mean outside in x: the edge qx = cx is visible
    const float qy = fminf(fmaxf(pre.rBC * cx, ay0), ay1);
    fmin = 0.5f * (A * cx * cx + C * qy * qy) + B * cx * qy;
  }
  if (cy != 0.f) {
    const float qx = fminf(fmaxf(pre.rBA * cy, ax0), ax1);
    fmin = fminf(fmin, 0.5f * (A * qx * qx + C * cy * cy) + B * qx * cy);
  }
  if (cx == 0.f && cy == 0.f) fmin = 0.f;  // mean inside the rectangle
  return fmin > pre.thr;                   // false for NaN
}

}  // namespace gsr
