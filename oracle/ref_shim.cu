// oracle/ref_shim.cu -- TEST INFRASTRUCTURE, not product code.
//
// A thin extern "C" wrapper (our own code) around the UNMODIFIED reference
// rasterizer, which is compiled from its sources where they lie under
// /root/reference (see oracle/Makefile; nothing is copied into this repo).
// The result, oracle/_ref/libref_rasterizer.so, is the real reference CUDA
// implementation of the hot path; it is used
//   * by tests/ (-m gpu) as the parity checker (bit-exact radii / tiles /
//     ranges / point_list, image <= 1e-5, gradients <= 1e-3 rel), and
//   * by bench.py --impl reference as the "reference on the same box" arm.
// The product (libgsrast.so) never links or loads this file.
//
// The wrapper mirrors the torch glue of the reference
// (rasterize_points.cu:35-115 forward, :117-196 backward): it owns the three
// growable byte buffers (geometry / binning / image), zero-fills the outputs
// exactly where the glue does (out_color, radii: :66-67; the nine gradient
// tensors: :151-159) and forwards every argument unchanged to
// CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (cuda_rasterizer/rasterizer.h:20-85).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>

#include "cuda_rasterizer/config.h"
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {

struct GrowBuf {
  char* ptr = nullptr;
  size_t cap = 0;
  size_t size = 0;
  char* resize(size_t n) {
    if (n > cap) {
      if (ptr) cudaFree(ptr);
      size_t want = n + n / 4 + 4096;
      if (cudaMalloc(&ptr, want) != cudaSuccess) {
        ptr = nullptr;
        cap = 0;
        throw std::runtime_error("ref_shim: cudaMalloc failed");
      }
      cap = want;
    }
    size = n;
    return ptr;
  }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    cap = size = 0;
  }
};

struct RefCtx {
  GrowBuf geom, binning, image;
  int P = 0, R = 0, W = 0, H = 0;
  std::string err;
};

thread_local std::string g_err;

}  // namespace

extern "C" {

void* ref_create() { return new RefCtx(); }

void ref_destroy(void* h) {
  RefCtx* c = (RefCtx*)h;
  if (!c) return;
  c->geom.release();
  c->binning.release();
  c->image.release();
  delete c;
}

const char* ref_last_error() { return g_err.c_str(); }

// Forward. Returns num_rendered (>= 0) or -1 on error.
// zero_outputs != 0 replicates torch::full(...,0) on out_color and radii.
int ref_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
                const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities,
                const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix,
                const float* campos, float tanfovx, float tanfovy,
                int prefiltered, float* out_color, int* radii, int debug,
                int zero_outputs) {
  RefCtx* c = (RefCtx*)h;
  try {
    if (zero_outputs) {
      cudaMemsetAsync(out_color, 0, sizeof(float) * NUM_CHANNELS * (size_t)W * H, 0);
      cudaMemsetAsync(radii, 0, sizeof(int) * (size_t)P, 0);
    }
    c->P = P;
    c->W = W;
    c->H = H;
    c->R = 0;
    if (P == 0) return 0;
    std::function<char*(size_t)> gf = [c](size_t n) { return c->geom.resize(n); };
    std::function<char*(size_t)> bf = [c](size_t n) { return c->binning.resize(n); };
    std::function<char*(size_t)> imf = [c](size_t n) { return c->image.resize(n); };
    int R = CudaRasterizer::Rasterizer::forward(
        gf, bf, imf, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
        scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
        projmatrix, campos, tanfovx, tanfovy, prefiltered != 0, out_color,
        radii, debug != 0);
    c->R = R;
    return R;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

// Backward. The nine gradient arrays are zero-filled first when zero_grads != 0
// (rasterize_points.cu:151-159).  dL_dconic is (P,2,2), dL_dmean2D is (P,3).
int ref_backward(void* h, int P, int D, int M, int R, const float* bg, int W,
                 int H, const float* means3D, const float* shs,
                 const float* colors_precomp, const float* scales,
                 float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* campos, float tanfovx,
                 float tanfovy, const int* radii, const float* dL_dpix,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot, int debug,
                 int zero_grads) {
  RefCtx* c = (RefCtx*)h;
  try {
    if (zero_grads) {
      size_t p = (size_t)P;
      cudaMemsetAsync(dL_dmean3D, 0, 4 * 3 * p, 0);
      cudaMemsetAsync(dL_dmean2D, 0, 4 * 3 * p, 0);
      cudaMemsetAsync(dL_dcolor, 0, 4 * NUM_CHANNELS * p, 0);
      cudaMemsetAsync(dL_dconic, 0, 4 * 4 * p, 0);
      cudaMemsetAsync(dL_dopacity, 0, 4 * p, 0);
      cudaMemsetAsync(dL_dcov3D, 0, 4 * 6 * p, 0);
      if (M > 0) cudaMemsetAsync(dL_dsh, 0, 4 * 3 * (size_t)M * p, 0);
      cudaMemsetAsync(dL_dscale, 0, 4 * 3 * p, 0);
      cudaMemsetAsync(dL_drot, 0, 4 * 4 * p, 0);
    }
    if (P == 0) return 0;
    CudaRasterizer::Rasterizer::backward(
        P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales,
        scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
        campos, tanfovx, tanfovy, radii, c->geom.ptr, c->binning.ptr,
        c->image.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
        dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

int ref_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, unsigned char* present) {
  cudaMemsetAsync(present, 0, (size_t)P, 0);
  if (P == 0) return 0;
  CudaRasterizer::Rasterizer::markVisible(P, (float*)means3D, (float*)viewmatrix,
                                          (float*)projmatrix, (bool*)present);
  return 0;
}

// Decoded views into the opaque byte buffers of the last forward call
// (layout: rasterizer_impl.cu:155-194).  Device pointers; valid until the
// next ref_forward on this context.
struct RefViews {
  const float* depths;
  const unsigned char* clamped;
  const int* internal_radii;
  const float* means2D;
  const float* cov3D;
  const float* conic_opacity;
  const float* rgb;
  const uint32_t* tiles_touched;
  const uint32_t* point_offsets;
  const float* accum_alpha;
  const uint32_t* n_contrib;
  const uint32_t* ranges;  // uint2 per tile
  const uint32_t* point_list;
  const uint64_t* point_list_keys;
  const uint32_t* point_list_unsorted;
  const uint64_t* point_list_keys_unsorted;
};

// Device-to-device copy helper so that the tests can pull decoded views into
// torch tensors without binding libcudart themselves.
int ref_copy(void* dst, const void* src, size_t nbytes) {
  return (int)cudaMemcpy(dst, src, nbytes, cudaMemcpyDeviceToDevice);
}

int ref_views(void* h, RefViews* v) {
  RefCtx* c = (RefCtx*)h;
  memset(v, 0, sizeof(*v));
  if (c->P == 0 || !c->geom.ptr) return 0;
  char* g = c->geom.ptr;
  CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(g, c->P);
  v->depths = gs.depths;
  v->clamped = (const unsigned char*)gs.clamped;
  v->internal_radii = gs.internal_radii;
  v->means2D = (const float*)gs.means2D;
  v->cov3D = gs.cov3D;
  v->conic_opacity = (const float*)gs.conic_opacity;
  v->rgb = gs.rgb;
  v->tiles_touched = gs.tiles_touched;
  v->point_offsets = gs.point_offsets;
  char* im = c->image.ptr;
  CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(im, (size_t)c->W * c->H);
  v->accum_alpha = is.accum_alpha;
  v->n_contrib = is.n_contrib;
  v->ranges = (const uint32_t*)is.ranges;
  if (c->binning.ptr) {
    char* b = c->binning.ptr;
    CudaRasterizer::BinningState bs = CudaRasterizer::BinningState::fromChunk(b, c->R);
    v->point_list = bs.point_list;
    v->point_list_keys = bs.point_list_keys;
    v->point_list_unsorted = bs.point_list_unsorted;
    v->point_list_keys_unsorted = bs.point_list_keys_unsorted;
  }
  return 0;
}

}  // extern "C"
